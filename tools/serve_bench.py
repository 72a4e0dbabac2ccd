#!/usr/bin/env python3
"""Closed-loop load on one worker_node over HTTP, in the manner of the reference's benchmark.py
(threads x requests, request_id "req_{i}", percentiles by sorted index int(n*p): benchmark.py:49-52,113-116),
once with its native 3-float payload (cache path) and once with full 3x224x224 payloads, unique per request
(inference path).  Prints one JSON object.   usage: serve_bench.py [--requests N] [--threads T] [--mode fp16]"""
import argparse
import json
import os
import socket
import subprocess
import sys
import threading
import time

import numpy as np
import requests

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200rn_loader  # noqa: E402
import bench  # noqa: E402


def run(url, payloads, threads):
    lat, lock = [], threading.Lock()
    per = len(payloads) // threads

    def work(t):
        s = requests.Session()
        for k in range(per):
            i = t * per + k
            t0 = time.time()
            r = s.post(url + "/infer", data=payloads[i], headers={"Content-Type": "application/json"}, timeout=60)
            dt = (time.time() - t0) * 1000
            if r.status_code == 200:
                with lock:
                    lat.append(dt)

    t0 = time.time()
    ts = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    total = time.time() - t0
    lat.sort()
    n = len(lat)
    return {"requests": per * threads, "ok": n, "req_per_s": n / total, "p50_ms": lat[int(n * 0.50)], "p90_ms": lat[int(n * 0.90)],
            "p99_ms": lat[min(n - 1, int(n * 0.99))], "mean_ms": sum(lat) / n}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--requests", type=int, default=640)
    ap.add_argument("--threads", type=int, default=32)
    ap.add_argument("--mode", default="fp16")
    a = ap.parse_args()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, B200RN_MODE=a.mode, B200RN_MAX_BATCH="32")
    proc = subprocess.Popen([os.path.join(b200rn_loader.PKG_DIR, "worker_node"), str(port), "worker_1", bench.model_path()],
                            env=env, stdout=subprocess.DEVNULL, stderr=subprocess.STDOUT)
    url = "http://127.0.0.1:%d" % port
    try:
        for _ in range(1200):
            try:
                requests.get(url + "/health", timeout=0.5)
                break
            except Exception:
                time.sleep(0.1)
        small = [json.dumps({"request_id": "req_%d" % i, "input_data": [float(i % 10), float(i % 10 + 1), float(i % 10 + 2)]})
                 for i in range(a.requests * 4)]
        out = {"cache_path_3_floats": run(url, small, a.threads)}
        rs = np.random.RandomState(0)
        base = rs.randn(150528).astype(np.float32)
        big = []
        for i in range(a.requests):
            x = base.copy()
            x[0] = float(i)                      # unique first element -> unique cache key, no hits
            big.append(json.dumps({"request_id": "req_%d" % i, "input_data": x.tolist()}))
        out["payload_bytes"] = len(big[0])
        out["inference_path_150528_floats"] = run(url, big, a.threads)
        out["health"] = requests.get(url + "/health", timeout=5).json()
        print(json.dumps(out))
    finally:
        proc.terminate()
        proc.wait(timeout=10)


if __name__ == "__main__":
    main()
