#!/bin/bash
# GPU job (4 GPUs): gateway + 3 workers (one per GPU) under the native load generator, default policy and two batches in flight
mkdir -p gpurun_out
B200RN_INFLIGHT=1 timeout 600 python tools/serve_bench.py --native --gateway --workers 3 --requests 3000 --threads 96 > gpurun_out/r2_serve_native_gateway_3gpu.json 2> gpurun_out/r2_serve_native_gateway_3gpu.err
B200RN_INFLIGHT=2 timeout 600 python tools/serve_bench.py --native --gateway --workers 3 --requests 3000 --threads 96 > gpurun_out/r2_serve_native_gateway_3gpu_inflight2.json 2> gpurun_out/r2_serve_native_gateway_3gpu_inflight2.err
python - <<'PY'
import json
for f in ("r2_serve_native_gateway_3gpu", "r2_serve_native_gateway_3gpu_inflight2"):
    try:
        j = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        for k in ("cache_path_3_floats", "inference_path_150528_floats", "inference_path_octet_stream"):
            r = j[k]; print(f, k, "%.0f req/s p50 %.1f p99 %.1f failed %d" % (r["req_per_s"], r["p50_ms"], r["p99_ms"], r["failed"]))
        print(f, "avg batch", [round(h["batch_processor"]["avg_batch_size"], 1) for h in j["health"]], "requests", [h["total_requests"] for h in j["health"]])
    except Exception as e:
        print(f, "ERR", e)
PY
