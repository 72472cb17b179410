# One 8-GPU gpurun call at the end of the round: bench.py on 8 and 4 ranks (outputs under gpurun_out/).
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4; do
  timeout 300 $TR --nproc-per-node $n --master-port 2976$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r02_bench_${n}gpu_e.json 2> gpurun_out/bench${n}e.err; tail -c 200 gpurun_out/bench${n}e.err
done
python - <<'PY'
import json
for n in (8, 4):
    try:
        d = json.load(open(f"gpurun_out/r02_bench_{n}gpu_e.json")); print(n, round(d["ms_per_step"], 3), round(d["value"] / 1e9, 1), d["parity_check"]["ok"], "e2e", round(d["e2e"]["ms_per_step"], 1), "reflayout", round(d["reference_layout_run"]["ms_per_step"], 1), "plain", round(d["shuffle"]["plain"]["bus_gbs"]))
    except Exception as e: print(n, "ERR", e)
PY
