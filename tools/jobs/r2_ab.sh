#!/bin/bash
# GPU job (1 GPU): A/B of one environment switch: usage  r2_ab.sh VAR "test -k expression"
VAR=$1; KEXPR=$2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -q -x -k "$KEXPR" > gpurun_out/r2_ab_ops.log 2>&1; tail -2 gpurun_out/r2_ab_ops.log
timeout 600 python -m pytest tests/test_gpu_parity258.py -q -k "fp16" > gpurun_out/r2_ab_net.log 2>&1; tail -1 gpurun_out/r2_ab_net.log
python bench.py --no-cpu-baseline --sustain-s 0 --layers-out gpurun_out/r2_layers_ab_on.json > gpurun_out/r2_bench_ab_on.json 2> gpurun_out/r2_bench_ab_on.err
env $VAR=0 python bench.py --no-cpu-baseline --sustain-s 0 --layers-out gpurun_out/r2_layers_ab_off.json > gpurun_out/r2_bench_ab_off.json 2> gpurun_out/r2_bench_ab_off.err
python - <<'PY'
import json
L = {}
for f in ("on", "off"):
    try:
        j = json.loads(open("gpurun_out/r2_bench_ab_%s.json" % f).read().strip().splitlines()[-1])
        L[f] = json.load(open("gpurun_out/r2_layers_ab_%s.json" % f))["layers"]
        print(f, round(j["ms_per_step"], 4), round(j["value"]), "b32", round(j["batch32"]["value"]), "layer sum", round(sum(l["ms"] for l in L[f]), 4))
    except Exception as e:
        print(f, "ERR", e)
if len(L) == 2:
    for a, b in zip(L["on"], L["off"]):
        if abs(a["ms"] - b["ms"]) > 0.002:
            print(a["i"], a["name"][:44], "on %.4f off %.4f" % (a["ms"], b["ms"]))
PY
tail -2 gpurun_out/r2_bench_ab_on.err
