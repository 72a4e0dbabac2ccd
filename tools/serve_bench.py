#!/usr/bin/env python3
"""Closed-loop load on one worker_node over HTTP, in the manner of the reference's benchmark.py
(threads x requests, request_id "req_{i}", percentiles by sorted index int(n*p): benchmark.py:49-52,113-116),
once with its native 3-float payload (cache path) and once with full 3x224x224 payloads, unique per request
(inference path).  Prints one JSON object.
   usage: serve_bench.py [--requests N] [--threads T] [--mode fp16] [--workers W] [--gateway]
--workers W starts W worker_node processes, worker i on GPU i % (number of GPUs); --gateway puts the gateway in
front of them and sends every request there (BASELINE configs[2]: gateway + 3 workers, 50 client threads); without
--gateway only the first worker is loaded."""
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


def run(url, payloads, threads, raw_ids=None):
    lat, lock = [], threading.Lock()
    per = len(payloads) // threads

    def work(t):
        s = requests.Session()
        for k in range(per):
            i = t * per + k
            t0 = time.time()
            hdr = ({"Content-Type": "application/json"} if raw_ids is None else
                   {"Content-Type": "application/octet-stream", "X-Request-Id": raw_ids[i]})
            r = s.post(url + "/infer", data=payloads[i], headers=hdr, timeout=60)
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
    ap.add_argument("--workers", type=int, default=1)
    ap.add_argument("--gateway", action="store_true")
    ap.add_argument("--native", action="store_true",
                    help="drive the load with the C++ generator (distributed-inference-engine-cpp_b200/loadgen) instead of Python threads")
    a = ap.parse_args()
    a_workers = max(1, a.workers)
    import torch
    ngpu = max(1, torch.cuda.device_count())

    def free_port():
        s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
        return p

    procs, addrs = [], []
    for i in range(a_workers):
        port = free_port()
        env = dict(os.environ, B200RN_MODE=a.mode, B200RN_MAX_BATCH="32", B200RN_DEVICE=str(i % ngpu))
        procs.append(subprocess.Popen([os.path.join(b200rn_loader.PKG_DIR, "worker_node"), str(port), "worker_%d" % (i + 1),
                                       bench.model_path()], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.STDOUT))
        addrs.append("127.0.0.1:%d" % port)
    url = "http://" + addrs[0]
    proc = procs[0]
    try:
        for ad in addrs:
            for _ in range(1200):
                try:
                    requests.get("http://" + ad + "/health", timeout=0.5)
                    break
                except Exception:
                    time.sleep(0.1)
        if a.gateway:
            gport = free_port()
            procs.append(subprocess.Popen([os.path.join(b200rn_loader.PKG_DIR, "gateway")] + addrs,
                                          env=dict(os.environ, GATEWAY_PORT=str(gport)), stdout=subprocess.DEVNULL,
                                          stderr=subprocess.STDOUT))
            url = "http://127.0.0.1:%d" % gport
            for _ in range(200):
                try:
                    requests.get(url + "/stats", timeout=0.5)
                    break
                except Exception:
                    time.sleep(0.05)
        # untimed warm-up: the first batches of a worker build its per-batch-size launch plans
        for ad in addrs:
            for k in range(4):
                w = np.full(150528, 1000.0 + k, dtype=np.float32)
                requests.post("http://" + ad + "/infer", data=json.dumps({"request_id": "warm_%d" % k, "input_data": w.tolist()}),
                              timeout=120)
        if a.native:
            host, port = url[len("http://"):].split(":")
            lg = os.path.join(b200rn_loader.PKG_DIR, "loadgen")

            def native(kind, n):
                r = subprocess.run([lg, host, port, str(a.threads), str(n), kind], stdout=subprocess.PIPE, text=True, timeout=900)
                return json.loads(r.stdout.strip().splitlines()[-1])

            out = {"client": "native (host/loadgen_main.cpp), one keep-alive connection per thread",
                   "cache_path_3_floats": native("small", a.requests * 16),
                   "inference_path_150528_floats": native("json", a.requests),
                   "inference_path_octet_stream": native("octet", a.requests * 4)}
            out["config"] = {"workers": a_workers, "gpus": ngpu, "gateway": bool(a.gateway), "threads": a.threads, "mode": a.mode,
                             "inflight": os.environ.get("B200RN_INFLIGHT", "1"), "fill_wait": os.environ.get("B200RN_FILL_WAIT", "0")}
            out["health"] = [requests.get("http://" + ad + "/health", timeout=5).json() for ad in addrs]
            if a.gateway:
                out["stats"] = requests.get(url + "/stats", timeout=5).json()
            print(json.dumps(out))
            return
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
        # the additive raw-float32 form of the same requests (octet-stream body, id in X-Request-Id)
        raws, ids = [], []
        for i in range(a.requests):
            x = base.copy()
            x[0] = float(i) + 0.5                # unique again: no cache hits
            raws.append(x.tobytes())
            ids.append("raw_%d" % i)
        out["inference_path_octet_stream"] = run(url, raws, a.threads, ids)
        out["config"] = {"workers": a_workers, "gpus": ngpu, "gateway": bool(a.gateway), "threads": a.threads, "mode": a.mode}
        out["health"] = [requests.get("http://" + ad + "/health", timeout=5).json() for ad in addrs]
        if a.gateway:
            out["stats"] = requests.get(url + "/stats", timeout=5).json()
        print(json.dumps(out))
    finally:
        for pr in procs:
            pr.terminate()
        for pr in procs:
            pr.wait(timeout=10)


if __name__ == "__main__":
    main()
