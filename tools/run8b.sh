# One 8-GPU gpurun call after the R-MAT changes: multi-rank tests on 8 ranks, config 5 (R-MAT 28) and the scale-10
# fixture check on 8 ranks, bench.py on 8 ranks (outputs under gpurun_out/).
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
MR_RANKS=8 timeout 600 python -m pytest tests/test_multirank.py -m gpu -x -q > gpurun_out/r02_multirank_pytest_8gpu_d.log 2>&1; tail -2 gpurun_out/r02_multirank_pytest_8gpu_d.log
timeout 300 $TR --nproc-per-node 8 --master-port 29731 bench_configs.py rmatcheck > gpurun_out/r02_rmatcheck_8gpu_d.json 2> gpurun_out/rc8.err; tail -c 200 gpurun_out/rc8.err
timeout 400 $TR --nproc-per-node 8 --master-port 29732 bench_configs.py rmat28 > gpurun_out/r02_rmat28_8gpu_d.json 2> gpurun_out/r28.err; tail -c 300 gpurun_out/r28.err
timeout 400 $TR --nproc-per-node 8 --master-port 29733 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02_bench_8gpu_d.json 2> gpurun_out/bench8.err; tail -c 300 gpurun_out/bench8.err
python - <<'PY'
import json
for f in ("r02_rmatcheck_8gpu_d", "r02_rmat28_8gpu_d"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1]); print(f, d["secs"], round(d["pairs_per_s"] / 1e9, 2), d["check"][:60])
    except Exception as e: print(f, "ERR", e)
try:
    d = json.load(open("gpurun_out/r02_bench_8gpu_d.json")); print(8, round(d["ms_per_step"], 3), round(d["value"] / 1e9, 1), d["parity_check"]["ok"], "e2e", round(d["e2e"]["ms_per_step"], 1))
except Exception as e: print(8, "ERR", e)
PY
