#!/bin/bash
# GPU job (1 GPU): one worker under the native load generator: default policy, two batches in flight, fill-or-timeout, both
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python tools/serve_bench.py --native --requests 2000 --threads 64 > gpurun_out/r2_serve_native_$name.json 2> gpurun_out/r2_serve_native_$name.err
}
run default B200RN_INFLIGHT=1
run inflight2 B200RN_INFLIGHT=2
run fillwait B200RN_FILL_WAIT=1
run fillwait_inflight2 B200RN_FILL_WAIT=1 B200RN_INFLIGHT=2
python - <<'PY'
import json
for f in ("default", "inflight2", "fillwait", "fillwait_inflight2"):
    try:
        j = json.loads(open("gpurun_out/r2_serve_native_%s.json" % f).read().strip().splitlines()[-1])
        for k in ("cache_path_3_floats", "inference_path_150528_floats", "inference_path_octet_stream"):
            r = j[k]; print(f, k, "%.0f req/s p50 %.1f p99 %.1f failed %d" % (r["req_per_s"], r["p50_ms"], r["p99_ms"], r["failed"]))
        print(f, "health", j["health"][0]["batch_processor"])
    except Exception as e:
        print(f, "ERR", e)
PY
