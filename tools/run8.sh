# One 8-GPU gpurun call: multi-rank tests on 8 ranks, then bench.py at 8 / 4 / 2 ranks (outputs under gpurun_out/).
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
MR_RANKS=8 timeout 600 python -m pytest tests/test_multirank.py -m gpu -x -q > gpurun_out/r02_multirank_pytest_8gpu.log 2>&1; tail -2 gpurun_out/r02_multirank_pytest_8gpu.log
for n in 8 4 2; do
  timeout 400 $TR --nproc-per-node $n --master-port 2972$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r02_bench_${n}gpu_c.json 2> gpurun_out/bench$n.err; tail -c 300 gpurun_out/bench$n.err
done
python - <<'PY'
import json
for n in (8,4,2):
    try:
        d=json.load(open(f"gpurun_out/r02_bench_{n}gpu_c.json")); print(n, round(d["ms_per_step"],3), round(d["value"]/1e9,1), d["parity_check"]["ok"], "e2e", round(d["e2e"]["ms_per_step"],1), d["e2e"].get("host_placement"), "reflayout", round(d["reference_layout_run"]["ms_per_step"],1))
    except Exception as e: print(n, "ERR", e)
PY
