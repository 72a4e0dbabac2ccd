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


# ---- the fast request scanner must be indistinguishable from the nlohmann path ---------------------------------

@pytest.fixture(scope="module")
def worker_pair():
    """Two stub workers: the default one (fast scanner first) and one forced onto the parse-then-convert path."""
    if not os.path.exists(FAKE):
        pytest.skip("worker_node_fake not built (nlohmann/json.hpp not found)")
    procs, urls = [], []
    for fast in ("1", "0"):
        port = _free_port()
        procs.append(subprocess.Popen([FAKE, str(port), "w", "unused.onnx"], env=dict(os.environ, B200RN_FAST_JSON=fast),
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
        urls.append("http://127.0.0.1:%d" % port)
    for u in urls:
        for _ in range(100):
            try:
                requests.get(u + "/health", timeout=0.5)
                break
            except Exception:
                time.sleep(0.05)
    yield urls
    for p in procs:
        p.terminate()
        p.wait(timeout=5)


FAST_JSON_BODIES = [
    '{"request_id":"a","input_data":[1,2,3]}',
    '{ "input_data" : [ 1.5 , -2e3 , 0.1 , 1E-2 , 3.0e+1 ] , "request_id" : "key order and spaces" }',
    '{"request_id":"ints","input_data":[-0,0,7,-7,16777217,123456789012345678]}',          # int64 route, -0 -> +0
    '{"request_id":"big ints","input_data":[1234567890123456789,18446744073709551615,123456789012345678901234567890]}',
    '{"request_id":"denormals and limits","input_data":[1e-45,1.4e-45,3.4028235e38,3.5e38,1e-320,2.2250738585072014e-308]}',
    '{"request_id":"rounding","input_data":[0.1,0.2,0.30000000000000004,16777216.5,16777217.5,1.0000000596046448]}',
    '{"request_id":"empty","input_data":[]}',
    '{"request_id":"esc\\u0041ped","input_data":[1]}',
    '{"request_id":"caf\\u00e9 \xc3\xa9","input_data":[1]}',
    '{"request_id":"x","input_data":[1],"extra":true}',
    '{"request_id":"x","request_id":"dup","input_data":[1]}',
    '{"request_id":"x","input_data":[1,"2"]}',
    '{"request_id":"x","input_data":[1,null]}',
    '{"request_id":"x","input_data":[true]}',
    '{"request_id":"x","input_data":[01]}',
    '{"request_id":"x","input_data":[1.]}',
    '{"request_id":"x","input_data":[.5]}',
    '{"request_id":"x","input_data":[+1]}',
    '{"request_id":"x","input_data":[1e999]}',
    '{"request_id":"x","input_data":[NaN]}',
    '{"request_id":"x","input_data":[1,]}',
    '{"request_id":"x","input_data":[1] }  trailing',
    '{"request_id":"x","input_data":{"a":1}}',
    '{"request_id":7,"input_data":[1]}',
    '{"input_data":[1]}',
    '{"request_id":"x"}',
    '[1,2]',
    '',
    '{"request_id":"x","input_data":[1]',
]


@pytest.mark.parametrize("body", FAST_JSON_BODIES)
def test_fast_scanner_equals_nlohmann_path(worker_pair, body):
    fast, slow = worker_pair
    a = requests.post(fast + "/infer", data=body.encode("utf-8"), timeout=5)
    b = requests.post(slow + "/infer", data=body.encode("utf-8"), timeout=5)
    assert a.status_code == b.status_code

    def norm(t):                      # only the measured time may differ between two processes
        import re
        return re.sub(r'"inference_time_us":\d+', '"inference_time_us":0', t)
    assert norm(a.text) == norm(b.text)


def test_fast_scanner_random_numbers(worker_pair):
    import random
    fast, slow = worker_pair
    rnd = random.Random(5)
    for trial in range(30):
        vals = []
        for _ in range(rnd.randint(1, 40)):
            kind = rnd.random()
            if kind < 0.3:
                vals.append(str(rnd.randint(-10**rnd.randint(1, 17), 10**rnd.randint(1, 17))))
            elif kind < 0.7:
                vals.append(repr(rnd.uniform(-1e3, 1e3)))
            else:
                vals.append("%.*e" % (rnd.randint(0, 17), rnd.uniform(-1, 1) * 10.0 ** rnd.randint(-50, 39)))
        body = '{"request_id":"t%d","input_data":[%s]}' % (trial, ",".join(vals))
        a = requests.post(fast + "/infer", data=body, timeout=5).json()
        b = requests.post(slow + "/infer", data=body, timeout=5).json()
        a.pop("inference_time_us", None), b.pop("inference_time_us", None)
        assert a == b, body


def test_octet_stream_side_channel_equals_json(worker_pair):
    """Additive raw-float32 request form: same response bytes as the JSON form of the same sample (cache hit, so the
    timing field is the constant 50), its own error texts, JSON untouched."""
    import numpy as np
    url = worker_pair[0]
    x = np.array([0.25, -3.5, 1e-8, 16777217.0, 0.1], dtype=np.float32)
    j = requests.post(url + "/infer", json={"request_id": "sc", "input_data": [float(v) for v in x]}, timeout=5)
    assert j.status_code == 200
    hdr = {"Content-Type": "application/octet-stream", "X-Request-Id": "sc"}
    r = requests.post(url + "/infer", data=x.tobytes(), headers=hdr, timeout=5)
    j2 = requests.post(url + "/infer", json={"request_id": "sc", "input_data": [float(v) for v in x]}, timeout=5)
    assert r.status_code == 200 and r.text == j2.text and '"cached":true' in r.text
    r = requests.post(url + "/infer", data=x.tobytes()[:-1], headers=hdr, timeout=5)
    assert r.status_code == 500 and "whole number of float32" in r.json()["error"]
    r = requests.post(url + "/infer", data=x.tobytes(), headers={"Content-Type": "application/octet-stream"}, timeout=5)
    assert r.status_code == 500 and "X-Request-Id" in r.json()["error"]


def test_idle_and_stalled_connections_do_not_hold_request_threads():
    """Two request threads, eight idle keep-alive connections and one peer stalled mid-header: a health probe is
    still answered at once (threads bound requests, not connections), the stalled peer is dropped after the read
    timeout, an idle connection is closed after the idle timeout, and a parked connection still serves its next
    request."""
    if not os.path.exists(FAKE):
        pytest.skip("worker_node_fake not built (nlohmann/json.hpp not found)")
    port = _free_port()
    env = dict(os.environ, B200RN_HTTP_THREADS="2", B200RN_HTTP_TIMEOUTS="1,3")
    proc = subprocess.Popen([FAKE, str(port), "worker_t", "unused.onnx"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, env=env)
    url = "http://127.0.0.1:%d" % port
    try:
        for _ in range(100):
            try:
                requests.get(url + "/health", timeout=0.5)
                break
            except Exception:
                time.sleep(0.05)
        sessions = [requests.Session() for _ in range(8)]
        for i, s in enumerate(sessions):        # each does one request and then sits idle with its connection open
            assert s.post(url + "/infer", json={"request_id": "k%d" % i, "input_data": [float(i)]}, timeout=5).status_code == 200
        stalled = socket.create_connection(("127.0.0.1", port))
        stalled.sendall(b"POST /infer HTTP/1.1\r\nContent-Length: 10\r\n")     # never finishes its header block
        t0 = time.time()
        assert requests.get(url + "/health", timeout=2).status_code == 200
        assert time.time() - t0 < 0.5
        assert sessions[3].post(url + "/infer", json={"request_id": "again", "input_data": [3.0]}, timeout=5).json()["cached"] is True
        stalled.settimeout(4)
        assert stalled.recv(16) == b""                                             # dropped after ~1 s, not kept for ever
        idle = socket.create_connection(("127.0.0.1", port))
        idle.settimeout(8)
        assert idle.recv(16) == b""                                                # closed after the 3 s idle timeout
        assert requests.get(url + "/health", timeout=2).json()["total_requests"] == 9
    finally:
        proc.terminate()
        proc.wait(timeout=5)


def test_native_load_generator_against_the_worker(worker):
    """host/loadgen_main.cpp (the closed-loop generator behind tools/serve_bench.py --native): every payload kind is
    answered, the cache path hits after ten distinct keys, and the summary is well-formed JSON."""
    lg = os.path.join(ROOT, "distributed-inference-engine-cpp_b200", "loadgen")
    if not os.path.exists(lg):
        pytest.skip("loadgen not built")
    host, port = worker[len("http://"):].split(":")
    before = requests.get(worker + "/health", timeout=2).json()
    for kind, n, floats in (("small", 200, None), ("json", 24, 3000), ("octet", 24, 3000)):
        cmd = [lg, host, port, "4", str(n), kind] + ([str(floats)] if floats else [])
        out = subprocess.run(cmd, stdout=subprocess.PIPE, text=True, timeout=60).stdout
        j = json.loads(out.strip().splitlines()[-1])
        assert j["ok"] == n and j["failed"] == 0 and j["req_per_s"] > 0 and j["p50_ms"] <= j["p99_ms"], j
    after = requests.get(worker + "/health", timeout=2).json()
    assert after["total_requests"] - before["total_requests"] == 248
    # 200 small requests over 10 distinct vectors; the four threads ask for the same key at the same moment, so each key
    # can miss up to four times before its first put lands
    assert after["cache_hits"] - before["cache_hits"] >= 160
