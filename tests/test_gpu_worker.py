"""The real worker (worker_node: HTTP shell + WorkerNode + BatchProcessor + LRU + B200 engine) on a GPU:
results over HTTP equal the engine's own, repeated inputs come from the cache, concurrent requests get batched."""
import os
import socket
import subprocess
import threading
import time

import numpy as np
import pytest
import requests

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import b200rn_loader  # noqa: E402
from tools import make_resnet50v2_onnx as gen  # noqa: E402

pkg = b200rn_loader.load_package()
BIN = os.path.join(b200rn_loader.PKG_DIR, "worker_node")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_worker_over_http(tmp_path):
    if not os.path.exists(BIN):
        pytest.skip("worker_node not built")
    model = str(tmp_path / "s.onnx")
    gen.write_model(model, "small64", seed=8)
    port = _free_port()
    env = dict(os.environ, B200RN_MODE="fp16", B200RN_MAX_BATCH="32")
    proc = subprocess.Popen([BIN, str(port), "gpu_worker_0", model], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    url = "http://127.0.0.1:%d" % port
    try:
        for _ in range(600):
            try:
                requests.get(url + "/health", timeout=0.5)
                break
            except Exception:
                assert proc.poll() is None, proc.stdout.read().decode()[-2000:]
                time.sleep(0.1)
        rs = np.random.RandomState(0)
        xs = [rs.randn(3 * 64 * 64).astype(np.float32) for _ in range(24)]
        eng = pkg.Engine(model, mode="fp16", max_batch=32)
        want = eng.batch_predict(xs)
        eng.close()
        got = [None] * len(xs)

        def post(i):
            r = requests.post(url + "/infer", json={"request_id": "req_%d" % i, "input_data": xs[i].tolist()}, timeout=60)
            assert r.status_code == 200, r.text[:300]
            got[i] = r.json()

        ts = [threading.Thread(target=post, args=(i,)) for i in range(len(xs))]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        for i, j in enumerate(got):
            assert j["request_id"] == "req_%d" % i and j["cached"] is False and j["node_id"] == "gpu_worker_0"
            # batch composition differs from the direct call: same tolerance as test_batch_semantics
            assert np.abs(np.asarray(j["output_data"], dtype=np.float32) - want[i]).max() <= 2e-2
        again = requests.post(url + "/infer", json={"request_id": "x", "input_data": xs[3].tolist()}, timeout=60).json()
        assert again["cached"] is True and again["inference_time_us"] == 50
        assert again["output_data"] == got[3]["output_data"]            # the cached row is returned bit for bit
        short = requests.post(url + "/infer", json={"request_id": "s", "input_data": [1.0, 2.0, 3.0]}, timeout=60)
        assert short.status_code == 200 and len(short.json()["output_data"]) == 100     # zero-padded, as the reference
        h = requests.get(url + "/health", timeout=5).json()
        assert h["total_requests"] == 26 and h["cache_hits"] == 1 and h["cache_size"] == 25
        bp = h["batch_processor"]
        assert bp["total_batches"] < 25 and bp["timeout_batches"] == 0 and bp["avg_batch_size"] > 1.0
    finally:
        proc.terminate()
        proc.wait(timeout=10)
