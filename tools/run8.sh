TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
MR_RANKS=8 timeout 600 python -m pytest tests/test_multirank.py -m gpu -x -q > gpurun_out/r02_multirank_pytest_8gpu.log 2>&1; tail -2 gpurun_out/r02_multirank_pytest_8gpu.log
timeout 400 $TR --nproc-per-node 8 --master-port 29721 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02_bench_8gpu_a.json 2> gpurun_out/bench8.err; tail -c 600 gpurun_out/bench8.err
timeout 400 $TR --nproc-per-node 4 --master-port 29722 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r02_bench_4gpu_a.json 2> gpurun_out/bench4.err; tail -c 300 gpurun_out/bench4.err
timeout 300 $TR --nproc-per-node 8 --master-port 29723 bench_configs.py rmatcheck > gpurun_out/r02_rmatcheck_8gpu.json 2> gpurun_out/rmatcheck8.err; tail -c 600 gpurun_out/rmatcheck8.err
timeout 600 $TR --nproc-per-node 8 --master-port 29724 bench_configs.py rmat28 > gpurun_out/r02_rmat28_8gpu.json 2> gpurun_out/rmat28.err; tail -c 1500 gpurun_out/rmat28.err
python - <<'PY'
import json
for f in ("r02_bench_8gpu_a","r02_bench_4gpu_a"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d["ms_per_step"], d["value"]/1e9, d["parity_check"]["ok"], d["e2e"]["ms_per_step"], d["reference_layout_run"]["ms_per_step"])
    except Exception as e: print(f, "ERR", e)
for f in ("r02_rmatcheck_8gpu","r02_rmat28_8gpu"):
    try: print(open(f"gpurun_out/{f}.json").read()[:2500])
    except Exception as e: print(f, "ERR", e)
PY
