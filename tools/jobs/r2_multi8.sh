#!/bin/bash
# GPU job (8 GPUs): H2D ceiling at N = 8 (and 4 again on this box), bench at N = 4 and 8 (end-to-end leg against the ceiling)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | wc -l; nproc; nvidia-smi topo -m 2>/dev/null | head -14 > gpurun_out/r2_topo8.txt
for n in 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n tools/h2d_ceiling.py > gpurun_out/r2_h2d8_n$n.json 2>> gpurun_out/r2_h2d8.err
done
for n in 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2962$n bench.py --gpus $n --steps 20 --warmup 5 --sustain-s 0 > gpurun_out/r2_bench_n$n.json 2> gpurun_out/r2_bench_n$n.err
done
python - <<'PY'
import json
for n in (4, 8):
    try:
        j = json.loads(open("gpurun_out/r2_h2d8_n%d.json" % n).read().strip().splitlines()[-1])
        print("ceiling", n, {k: round(j[k]["aggregate_GBps"], 1) for k in ("pinned", "pinned_wc", "chunks8", "two_streams")}, "img/s equiv", round(j["pinned"]["images_per_s_equiv"]))
    except Exception as e:
        print(n, "ERR", e)
    try:
        j = json.loads(open("gpurun_out/r2_bench_n%d.json" % n).read().strip().splitlines()[-1])
        print("bench", n, "device", round(j["value"]), "e2e", round(j["e2e"]["value"]), "GB/s", round(j["e2e"]["value"] * 602112 / 1e9, 1))
    except Exception as e:
        print(n, "ERR", e)
PY
