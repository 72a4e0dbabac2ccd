#!/usr/bin/env python3
"""BASELINE configs[2] and [4] exactly as the reference defines them: the reference's OWN load generator, unmodified,
against this repository's gateway + 3 worker_nodes on the ports it defaults to (gateway :8000, workers :8001-8003,
/root/reference/benchmark.py:222-242; --cache-test flow :180-220; 10,000 requests / 50 threads, README.md:276-300).

benchmark.py is not copied into the repository: `--stage` copies it (here, where /root/reference exists) into
baseline/_ref/, which is git-ignored but travels to the GPU box with the snapshot, and the run executes that file
as it is.  Everything it prints is kept; /health and /stats of the servers are saved beside it.

    python tools/run_reference_benchmark.py --stage                      # in the build container
    python tools/run_reference_benchmark.py --out gpurun_out/r2_refbench # on the GPU box (one worker per GPU, round robin)
    python tools/run_reference_benchmark.py --fake --requests 500        # CPU smoke: stub engine
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import time

import requests

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-inference-engine-cpp_b200")
STAGED = os.path.join(ROOT, "baseline", "_ref", "benchmark.py")
sys.path.insert(0, ROOT)


def wait_up(url, tries=1200):
    for _ in range(tries):
        try:
            if requests.get(url, timeout=0.5).status_code == 200:
                return
        except Exception:
            time.sleep(0.1)
    raise SystemExit("%s did not come up" % url)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stage", action="store_true", help="copy /root/reference/benchmark.py to baseline/_ref/ and exit")
    ap.add_argument("--fake", action="store_true", help="worker_node_fake (stub engine, no GPU)")
    ap.add_argument("--requests", type=int, default=10000)
    ap.add_argument("--threads", type=int, default=50)
    ap.add_argument("--mode", default="fp16")
    ap.add_argument("--inflight", type=int, default=1)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r2_refbench"))
    a = ap.parse_args()
    if a.stage:
        os.makedirs(os.path.dirname(STAGED), exist_ok=True)
        shutil.copyfile("/root/reference/benchmark.py", STAGED)
        print("staged", STAGED)
        return
    if not os.path.exists(STAGED):
        raise SystemExit("baseline/_ref/benchmark.py is missing: run this tool with --stage where /root/reference exists")
    ngpu = 1
    model = "unused.onnx"
    if not a.fake:
        import torch
        import bench
        ngpu = max(1, torch.cuda.device_count())
        model = bench.model_path()
    binary = os.path.join(PKG, "worker_node_fake" if a.fake else "worker_node")
    procs = []
    logs = []
    try:
        for i in range(3):
            env = dict(os.environ, B200RN_MODE=a.mode, B200RN_MAX_BATCH="32", B200RN_DEVICE=str(i % ngpu), B200RN_INFLIGHT=str(a.inflight))
            lf = open("%s_worker_%d.log" % (a.out, i + 1), "w")
            logs.append(lf)
            procs.append(subprocess.Popen([binary, str(8001 + i), "worker_%d" % (i + 1), model], env=env, stdout=lf, stderr=subprocess.STDOUT))
        for i in range(3):
            wait_up("http://localhost:%d/health" % (8001 + i))
        lf = open(a.out + "_gateway.log", "w")
        logs.append(lf)
        procs.append(subprocess.Popen([os.path.join(PKG, "gateway")] + ["localhost:%d" % (8001 + i) for i in range(3)],
                                      env=dict(os.environ, GATEWAY_PORT="8000"), stdout=lf, stderr=subprocess.STDOUT))
        wait_up("http://localhost:8000/stats", 200)
        cmd = [sys.executable, STAGED, "--requests", str(a.requests), "--threads", str(a.threads), "--cache-test"]
        t0 = time.time()
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1800)
        wall = time.time() - t0
        with open(a.out + "_stdout.txt", "w") as f:
            f.write("$ python benchmark.py --requests %d --threads %d --cache-test    (unmodified /root/reference/benchmark.py, exit %d, %.1f s)\n"
                    % (a.requests, a.threads, r.returncode, wall))
            f.write(r.stdout)
        summary = {"command": " ".join(cmd[1:]), "exit": r.returncode, "wall_s": wall, "workers": 3, "gpus": ngpu, "mode": a.mode,
                   "engine": "stub" if a.fake else "b200", "inflight": a.inflight,
                   "health": [requests.get("http://localhost:%d/health" % (8001 + i), timeout=5).json() for i in range(3)],
                   "stats": requests.get("http://localhost:8000/stats", timeout=5).json()}
        for line in r.stdout.splitlines():
            for key, name in (("Requests/sec:", "req_per_s"), ("  p50:", "p50_ms"), ("  p99:", "p99_ms"), ("Success rate:", "success_pct"),
                              ("Cache speedup:", "cache_speedup_x")):
                if key in line:
                    try:
                        summary[name] = float(line.split(":")[1].strip().rstrip("%x"))
                    except ValueError:
                        pass
        with open(a.out + "_summary.json", "w") as f:
            json.dump(summary, f, indent=1)
        print(r.stdout[-3000:])
        print(json.dumps({k: summary.get(k) for k in ("req_per_s", "p50_ms", "p99_ms", "success_pct", "cache_speedup_x")}))
    finally:
        for p in procs:
            p.terminate()
        for p in procs:
            try:
                p.wait(timeout=10)
            except Exception:
                p.kill()
        for lf in logs:
            lf.close()


if __name__ == "__main__":
    main()
