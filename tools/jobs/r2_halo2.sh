#!/bin/bash
# GPU job (1 GPU): dual-M halo kernel -- op tests, network parity, per-layer A/B against B200RN_HALO2=0
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -q -x -k "halo or 3x3" > gpurun_out/r2_halo2_ops.log 2>&1; tail -3 gpurun_out/r2_halo2_ops.log
timeout 600 python -m pytest tests/test_gpu_parity258.py -q -k "fp16 or bf16" > gpurun_out/r2_halo2_net.log 2>&1; tail -2 gpurun_out/r2_halo2_net.log
python bench.py --no-cpu-baseline --sustain-s 0 --layers-out gpurun_out/r2_layers_halo2.json > gpurun_out/r2_bench_halo2.json 2> gpurun_out/r2_bench_halo2.err
B200RN_HALO2=0 python bench.py --no-cpu-baseline --sustain-s 0 --layers-out gpurun_out/r2_layers_nohalo2.json > gpurun_out/r2_bench_nohalo2.json 2> gpurun_out/r2_bench_nohalo2.err
python - <<'PY'
import json
for f in ("halo2", "nohalo2"):
    try:
        j = json.loads(open("gpurun_out/r2_bench_%s.json" % f).read().strip().splitlines()[-1])
        L = json.load(open("gpurun_out/r2_layers_%s.json" % f))["layers"]
        print(f, round(j["ms_per_step"], 4), round(j["value"]), "b32", round(j["batch32"]["value"]), "halo layers", [round(L[i]["ms"], 4) for i in (4, 6, 8, 14, 17, 20)])
    except Exception as e:
        print(f, "ERR", e)
PY
tail -2 gpurun_out/r2_bench_halo2.err
