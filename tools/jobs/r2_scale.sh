#!/bin/bash
# GPU job (N GPUs, N = $1): the driver's scaling runs on this build: bench.py at 2 .. N ranks and the reference arm under torchrun
mkdir -p gpurun_out
N=${1:-2}
for n in 2 4 8; do
  [ $n -le $N ] || continue
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2963$n bench.py --gpus $n --steps 20 --warmup 5 --sustain-s 0 > gpurun_out/r2_scale_n$n.json 2> gpurun_out/r2_scale_n$n.err
  python - $n <<'PY'
import json, sys
n = sys.argv[1]
try:
    j = json.loads(open("gpurun_out/r2_scale_n%s.json" % n).read().strip().splitlines()[-1])
    print("bench", n, "device", round(j["value"]), "ms", round(j["ms_per_step"], 3), "e2e", round(j["e2e"]["value"]), "parity", j["e2e"]["parity"]["max_abs"], "clocks", j["clocks"]["sm_mhz"])
except Exception as e:
    print(n, "ERR", e)
PY
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29639 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2_scale_ref.json 2> gpurun_out/r2_scale_ref.err; tail -c 600 gpurun_out/r2_scale_ref.json
