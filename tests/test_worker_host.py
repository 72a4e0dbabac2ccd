"""The worker's request path and wire contract on CPU: worker_node_fake is the real WorkerNode / BatchProcessor /
LRUCache / HTTP shell with a stub engine (row = [sum, len, 0.1f, 1e-8f]).  Checks the JSON bytes against what the
reference emits with nlohmann 3.11.3: alphabetical keys, floats widened to double and printed
shortest-round-trip, cache hit -> cached:true and inference_time_us 50, missing key -> HTTP 500 with the
type_error.302 text, /health fields."""
import json
import os
import socket
import subprocess
import time

import pytest
import requests

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FAKE = os.path.join(ROOT, "distributed-inference-engine-cpp_b200", "worker_node_fake")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.fixture(scope="module")
def worker():
    if not os.path.exists(FAKE):
        pytest.skip("worker_node_fake not built (nlohmann/json.hpp not found)")
    port = _free_port()
    proc = subprocess.Popen([FAKE, str(port), "worker_1", "unused.onnx"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    url = "http://127.0.0.1:%d" % port
    for _ in range(100):
        try:
            requests.get(url + "/health", timeout=0.5)
            break
        except Exception:
            time.sleep(0.05)
    yield url
    proc.terminate()
    proc.wait(timeout=5)


def test_infer_miss_then_hit_bytes(worker):
    body = {"request_id": "req_7", "input_data": [1.0, 2.0, 3.0]}
    r = requests.post(worker + "/infer", json=body, timeout=5)
    assert r.status_code == 200 and r.headers["Content-Type"] == "application/json"
    j = r.json()
    assert list(j.keys()) == ["cached", "inference_time_us", "node_id", "output_data", "request_id"]   # alphabetical
    assert j["cached"] is False and j["node_id"] == "worker_1" and j["request_id"] == "req_7"
    assert '"output_data":[6.0,3.0,0.10000000149011612,9.99999993922529e-09]' in r.text
    r2 = requests.post(worker + "/infer", json={"request_id": "again", "input_data": [1.0, 2.0, 3.0]}, timeout=5)
    assert r2.text == ('{"cached":true,"inference_time_us":50,"node_id":"worker_1",'
                       '"output_data":[6.0,3.0,0.10000000149011612,9.99999993922529e-09],"request_id":"again"}')


def test_errors_are_http_500_with_reference_text(worker):
    r = requests.post(worker + "/infer", json={"input_data": [1.0]}, timeout=5)
    assert r.status_code == 500
    assert r.json() == {"error": "[json.exception.type_error.302] type must be string, but is null"}
    r = requests.post(worker + "/infer", data="{not json", timeout=5)
    assert r.status_code == 500 and "parse_error" in r.json()["error"]
    assert requests.get(worker + "/nope", timeout=5).status_code == 404


def test_health_fields_and_counters(worker):
    for i in range(5):
        requests.post(worker + "/infer", json={"request_id": "h%d" % i, "input_data": [float(i), 9.0]}, timeout=5)
    requests.post(worker + "/infer", json={"request_id": "h_again", "input_data": [0.0, 9.0]}, timeout=5)
    h = requests.get(worker + "/health", timeout=5).json()
    assert list(h.keys()) == ["batch_processor", "cache_hit_rate", "cache_hits", "cache_size", "healthy", "node_id", "total_requests"]
    assert list(h["batch_processor"].keys()) == ["avg_batch_size", "full_batches", "timeout_batches", "total_batches"]
    assert h["healthy"] is True and h["node_id"] == "worker_1"
    assert h["cache_hits"] >= 1 and h["cache_size"] >= 5 and 0 < h["cache_hit_rate"] < 1
    assert h["batch_processor"]["timeout_batches"] == 0 and h["batch_processor"]["full_batches"] == h["batch_processor"]["total_batches"]


def test_large_body_and_keep_alive(worker):
    s = requests.Session()
    x = [float(i % 97) / 97 for i in range(150528)]          # a full 3x224x224 payload as JSON text (~3 MB)
    r = s.post(worker + "/infer", json={"request_id": "big", "input_data": x}, timeout=30)
    assert r.status_code == 200 and r.json()["output_data"][1] == 150528.0
    r = s.post(worker + "/infer", json={"request_id": "big2", "input_data": x}, timeout=30)
    assert r.json()["cached"] is True
