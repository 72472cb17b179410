run() { # name nproc envs...
  name=$1; np=$2; shift 2
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $np --steps 5 --warmup 3 > gpurun_out/$name.json 2> gpurun_out/$name.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["value"]/1e9,2), "Gp/s", round(d["ms_per_step"],2), "ms", [(k["tag"],k["ms"]) for k in d["roofline"]["kernels"][:4]], d.get("shuffle",{}) and d["shuffle"].get("bus_gbs"))
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/$name.err").read()[-1500:])
PY
}
