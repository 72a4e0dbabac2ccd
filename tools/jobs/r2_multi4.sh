#!/bin/bash
# GPU job (4 GPUs, one worker per GPU on three of them): BASELINE configs[2] -- the reference's unmodified benchmark.py
# against gateway + 3 workers, then the full-payload / octet-stream legs through the gateway; H2D ceiling at N = 1, 2, 4
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 900 python tools/run_reference_benchmark.py --out gpurun_out/r2_refbench_3gpu > gpurun_out/r2_refbench_3gpu.log 2>&1; tail -1 gpurun_out/r2_refbench_3gpu.log
timeout 900 python tools/serve_bench.py --gateway --workers 3 --requests 600 --threads 50 > gpurun_out/r2_serve_gateway_3gpu.json 2> gpurun_out/r2_serve_gateway_3gpu.err
B200RN_INFLIGHT=2 timeout 900 python tools/serve_bench.py --gateway --workers 3 --requests 600 --threads 50 > gpurun_out/r2_serve_gateway_3gpu_inflight2.json 2> gpurun_out/r2_serve_gateway_3gpu_inflight2.err
python - <<'PY'
import json
for f in ("r2_serve_gateway_3gpu", "r2_serve_gateway_3gpu_inflight2"):
    try:
        j = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        for k in ("cache_path_3_floats", "inference_path_150528_floats", "inference_path_octet_stream"):
            r = j[k]; print(f, k, "%.0f req/s p50 %.1f p99 %.1f" % (r["req_per_s"], r["p50_ms"], r["p99_ms"]))
        print(f, "avg batch", [h["batch_processor"]["avg_batch_size"] for h in j["health"]])
    except Exception as e:
        print(f, "ERR", e)
PY
python tools/h2d_ceiling.py > gpurun_out/r2_h2d_n1.json 2>> gpurun_out/r2_h2d.err
for n in 2 4; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n tools/h2d_ceiling.py > gpurun_out/r2_h2d_n$n.json 2>> gpurun_out/r2_h2d.err
done
python - <<'PY'
import json
for n in (1, 2, 4):
    try:
        j = json.loads(open("gpurun_out/r2_h2d_n%d.json" % n).read().strip().splitlines()[-1])
        print(n, {k: round(j[k]["aggregate_GBps"], 1) for k in ("pinned", "pinned_wc", "chunks8", "two_streams")})
    except Exception as e:
        print(n, "ERR", e)
PY
