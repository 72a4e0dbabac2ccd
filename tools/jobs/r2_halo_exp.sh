#!/bin/bash
# GPU job (1 GPU): what bounds conv3x3_halo_kernel -- per-layer times with B200RN_DBG_SKIP knobs (results wrong on purpose);
# plus the new exact-mode shape tests and 3 workers + gateway (sharing the GPU) under the native generator
mkdir -p gpurun_out
for k in 0 8 4 1 2; do
  B200RN_DBG_SKIP=$k python tools/profile_step.py --batch 256 --steps 2 --warmup 2 --layers-out gpurun_out/r2_halo_skip$k.json > gpurun_out/r2_halo_skip$k.log 2>&1
done
python - <<'PY'
import json
base = json.load(open("gpurun_out/r2_halo_skip0.json"))["layers"]
for i in (4, 6, 8, 14, 17, 20):
    row = ["L%d %.4f" % (i, base[i]["ms"])]
    for k in (8, 4, 1, 2):
        try:
            row.append("skip%d %.4f" % (k, json.load(open("gpurun_out/r2_halo_skip%d.json" % k))["layers"][i]["ms"]))
        except Exception as e:
            row.append("skip%d ERR" % k)
    print("  ".join(row))
PY
timeout 300 python -m pytest tests/test_gpu_ops.py -q -k "exact_mode_shapes or stem_conv" > gpurun_out/r2_t32_shapes.log 2>&1; tail -3 gpurun_out/r2_t32_shapes.log
B200RN_INFLIGHT=2 timeout 600 python tools/serve_bench.py --native --gateway --workers 3 --requests 2000 --threads 96 > gpurun_out/r2_serve_native_gateway_1gpu.json 2> gpurun_out/r2_serve_native_gateway_1gpu.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2_serve_native_gateway_1gpu.json").read().strip().splitlines()[-1])
    for k in ("cache_path_3_floats", "inference_path_150528_floats", "inference_path_octet_stream"):
        r = j[k]; print(k, "%.0f req/s p50 %.1f p99 %.1f failed %d" % (r["req_per_s"], r["p50_ms"], r["p99_ms"], r["failed"]))
except Exception as e:
    print("ERR", e)
PY
