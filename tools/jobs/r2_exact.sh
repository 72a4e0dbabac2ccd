#!/bin/bash
# GPU job: exact mode on the tensor cores (3 x TF32, two-level accumulation) -- op tests, network parity, concurrency, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -q -k "fp32 or cuda_core" > gpurun_out/r2_tf32_ops.log 2>&1; tail -6 gpurun_out/r2_tf32_ops.log | cut -c1-250
timeout 600 python -m pytest tests/test_gpu_parity258.py tests/test_gpu_network.py -q -s -k "exact or two_threads or semantics" > gpurun_out/r2_tf32_net.log 2>&1
grep -E "parity258|max-abs|passed|failed|Error" gpurun_out/r2_tf32_net.log | cut -c1-200
python bench.py --mode exact --no-cpu-baseline --sustain-s 0 --steps 10 --layers-out gpurun_out/r2_layers_exact_tf32.json > gpurun_out/r2_bench_exact_tf32.json 2> gpurun_out/r2_bench_exact_tf32.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2_bench_exact_tf32.json").read().strip().splitlines()[-1])
    print("exact tf32x3:", j["ms_per_step"], j["value"], j["e2e"]["value"], j["e2e"]["parity"]["max_abs"], j["batch32"]["value"])
except Exception as e:
    print("ERR", e)
PY
tail -3 gpurun_out/r2_bench_exact_tf32.err
