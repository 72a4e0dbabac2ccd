#!/bin/bash
# GPU job (1 GPU): exact-mode epilogue A/B (staging depth), and whether concurrent nvidia-smi pollers slow the e2e leg
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ops.py -q -k "fp32" > gpurun_out/r2_tf32_ops2.log 2>&1; tail -2 gpurun_out/r2_tf32_ops2.log
for stg in 1 2 auto; do
  if [ $stg = auto ]; then unset B200RN_T32_STG; else export B200RN_T32_STG=$stg; fi
  python bench.py --mode exact --no-cpu-baseline --sustain-s 0 --steps 10 --layers-out gpurun_out/r2_layers_exact_stg$stg.json > gpurun_out/r2_bench_exact_stg$stg.json 2> gpurun_out/r2_bench_exact_stg$stg.err
  python -c "
import json
j=json.loads(open('gpurun_out/r2_bench_exact_stg$stg.json').read().strip().splitlines()[-1]); print('exact stg=$stg', j['ms_per_step'], j['value'], j['e2e']['parity']['max_abs'])"
done
unset B200RN_T32_STG
python bench.py --no-cpu-baseline --sustain-s 0 > gpurun_out/r2_bench_smi1.json 2> /dev/null
for i in 1 2 3 4 5 6 7; do nvidia-smi -i 0 --query-gpu=clocks.sm,power.draw --format=csv,noheader -lms 100 > /dev/null 2>&1 & done
sleep 1
python bench.py --no-cpu-baseline --sustain-s 0 > gpurun_out/r2_bench_smi8.json 2> /dev/null
kill $(jobs -p) 2>/dev/null
python - <<'PY'
import json
for f in ("smi1", "smi8"):
    j = json.loads(open("gpurun_out/r2_bench_%s.json" % f).read().strip().splitlines()[-1])
    print(f, "device", round(j["value"]), "e2e", round(j["e2e"]["value"]), "b32 p50", j["batch32"]["e2e_sync_call_ms_p50"])
PY
