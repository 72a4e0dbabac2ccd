#!/bin/bash
# GPU job (1 GPU): worker over HTTP -- GPU worker test, the reference's benchmark.py against gateway + 3 workers sharing
# the one GPU (smoke of the 3-GPU job), serve_bench with and without two batches in flight, offline sweep
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_worker.py -q > gpurun_out/r2_gpu_worker.log 2>&1; tail -2 gpurun_out/r2_gpu_worker.log
timeout 900 python tools/run_reference_benchmark.py --out gpurun_out/r2_refbench_1gpu > gpurun_out/r2_refbench_1gpu.log 2>&1; tail -1 gpurun_out/r2_refbench_1gpu.log
timeout 600 python tools/serve_bench.py --requests 640 --threads 50 > gpurun_out/r2_serve_1w_inflight1.json 2> gpurun_out/r2_serve_1w_inflight1.err
B200RN_INFLIGHT=2 timeout 600 python tools/serve_bench.py --requests 640 --threads 50 > gpurun_out/r2_serve_1w_inflight2.json 2> gpurun_out/r2_serve_1w_inflight2.err
python - <<'PY'
import json
for f in ("inflight1", "inflight2"):
    try:
        j = json.loads(open("gpurun_out/r2_serve_1w_%s.json" % f).read().strip().splitlines()[-1])
        for k in ("cache_path_3_floats", "inference_path_150528_floats", "inference_path_octet_stream"):
            r = j[k]; print(f, k, "%.0f req/s p50 %.1f p99 %.1f" % (r["req_per_s"], r["p50_ms"], r["p99_ms"]))
        print(f, "avg batch", [h["batch_processor"]["avg_batch_size"] for h in j["health"]])
    except Exception as e:
        print(f, "ERR", e)
PY
timeout 900 python tools/sweep.py --out gpurun_out/r2_sweep.json > /dev/null 2> gpurun_out/r2_sweep.err; tail -18 gpurun_out/r2_sweep.err
