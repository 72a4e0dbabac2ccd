"""The gateway on CPU, in front of three worker_node_fake processes: routing equals the reference ring's placement
(checked against the oracle restatement of src/consistent_hash.cpp), bodies pass through byte for byte, /stats has
the reference's shape, fail-over and the circuit-breaker life cycle (CLOSED -> OPEN after 5 failures -> HALF_OPEN
probe after the timeout -> CLOSED after 2 successes), and the HTTP 500 texts."""
import os
import socket
import subprocess
import time

import pytest
import requests

from oracle import routing

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-inference-engine-cpp_b200")
FAKE, GATEWAY = os.path.join(PKG, "worker_node_fake"), os.path.join(PKG, "gateway")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _wait_up(url, path, tries=100):
    for _ in range(tries):
        try:
            requests.get(url + path, timeout=0.5)
            return
        except Exception:
            time.sleep(0.05)
    raise RuntimeError("%s did not come up" % url)


class Cluster:
    def __init__(self, n=3):
        self.ports = [_free_port() for _ in range(n)]
        self.addr = ["127.0.0.1:%d" % p for p in self.ports]             # the ring's node names, as on the command line
        self.node_id = {a: "worker_%d" % i for i, a in enumerate(self.addr)}
        self.procs = {}
        for a in self.addr:
            self.start_worker(a)
        self.gw_port = _free_port()
        env = dict(os.environ, GATEWAY_PORT=str(self.gw_port), GATEWAY_BREAKER_TIMEOUT_MS="600", GATEWAY_IO_TIMEOUT_S="1")
        self.gw = subprocess.Popen([GATEWAY] + self.addr, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        self.url = "http://127.0.0.1:%d" % self.gw_port
        _wait_up(self.url, "/stats")
        self.ring = routing.OracleRing(self.addr, 150)

    def start_worker(self, a):
        port = a.split(":")[1]
        self.procs[a] = subprocess.Popen([FAKE, port, self.node_id[a], "unused.onnx"], stdout=subprocess.PIPE,
                                         stderr=subprocess.STDOUT)
        _wait_up("http://" + a, "/health")

    def stop_worker(self, a):
        self.procs[a].terminate()
        self.procs[a].wait(timeout=5)

    def close(self):
        for p in list(self.procs.values()) + [self.gw]:
            if p.poll() is None:
                p.terminate()
                p.wait(timeout=5)

    def breaker(self, a):
        st = requests.get(self.url + "/stats", timeout=5).json()
        return [b for b in st["circuit_breakers"] if b["node"] == a][0]


@pytest.fixture()
def cluster():
    if not (os.path.exists(FAKE) and os.path.exists(GATEWAY)):
        pytest.skip("worker_node_fake / gateway not built (nlohmann/json.hpp not found)")
    c = Cluster()
    yield c
    c.close()


def test_routing_matches_reference_ring_and_bodies_pass_through(cluster):
    # ids chosen with the oracle ring so that every worker is some id's home (sequential ids cluster on the FNV
    # ring -- the reference's skew, SURVEY 8e -- so a fixed list may miss a worker for some port numbers)
    by_node = {a: [] for a in cluster.addr}
    for i in range(5000):
        rid = "req_%d" % i
        home = cluster.ring.get(rid)
        if len(by_node[home]) < 15:
            by_node[home].append(rid)
    assert all(by_node.values())
    for home, rids in by_node.items():
        for n, rid in enumerate(rids):
            r = requests.post(cluster.url + "/infer", json={"request_id": rid, "input_data": [float(n), 2.0, 3.0]}, timeout=5)
            assert r.status_code == 200 and r.headers["Content-Type"] == "application/json"
            j = r.json()
            assert j["node_id"] == cluster.node_id[home], (rid, home)
            assert j["request_id"] == rid and j["output_data"][:2] == [float(n) + 5.0, 3.0]
    # a cache hit is a constant response: the gateway's bytes are the worker's bytes
    rid = by_node[cluster.addr[0]][0]
    a = cluster.ring.get(rid)
    body = '{"input_data": [0.0, 2.0, 3.0],   "request_id": "%s"}' % rid      # odd spacing and key order survive the hop
    via = requests.post(cluster.url + "/infer", data=body, timeout=5)
    direct = requests.post("http://" + a + "/infer", data=body, timeout=5)
    assert via.text == direct.text and '"cached":true' in via.text and '"inference_time_us":50' in via.text


def test_stats_shape(cluster):
    r = requests.get(cluster.url + "/stats", timeout=5)
    want = '{"circuit_breakers":[' + ",".join(
        '{"failures":0,"node":"%s","state":"CLOSED","successes":0}' % a for a in sorted(cluster.addr)) + '],"total_workers":3}'
    assert r.text == want


def test_error_texts(cluster):
    bad = requests.post(cluster.url + "/infer", data="{nope", timeout=5)
    direct = requests.post("http://" + cluster.addr[0] + "/infer", data="{nope", timeout=5)
    assert bad.status_code == 500 and bad.text == direct.text and "parse_error.101" in bad.text
    r = requests.post(cluster.url + "/infer", json={"input_data": [1.0]}, timeout=5)
    assert r.status_code == 500 and r.json() == {"error": "[json.exception.type_error.302] type must be string, but is null"}
    r = requests.post(cluster.url + "/infer", json={"request_id": 7, "input_data": [1.0]}, timeout=5)
    assert r.json() == {"error": "[json.exception.type_error.302] type must be string, but is number"}
    r = requests.post(cluster.url + "/infer", json=[1, 2], timeout=5)
    assert r.status_code == 500 and "type_error.305" in r.text and "array" in r.text
    # a worker's own HTTP 500 (missing input_data) is a failure for the gateway: every node is tried, then the 500
    r = requests.post(cluster.url + "/infer", json={"request_id": "x"}, timeout=5)
    assert r.status_code == 500 and r.json() == {"error": "All workers failed or circuit breakers open"}
    st = requests.get(cluster.url + "/stats", timeout=5).json()
    assert [b["failures"] for b in st["circuit_breakers"]] == [1, 1, 1]


def test_failover_and_breaker_life_cycle(cluster):
    dead = cluster.addr[1]
    keys = [k for k in ("k%d" % i for i in range(400)) if cluster.ring.get(k) == dead][:8]
    assert len(keys) == 8
    cluster.stop_worker(dead)
    for n, k in enumerate(keys[:5]):
        r = requests.post(cluster.url + "/infer", json={"request_id": k, "input_data": [1.0]}, timeout=10)
        assert r.status_code == 200 and r.json()["node_id"] != cluster.node_id[dead]      # answered by another worker
        b = cluster.breaker(dead)
        assert b["failures"] == n + 1 and b["state"] == ("OPEN" if n + 1 >= 5 else "CLOSED")
    # OPEN: the dead worker is skipped without a connection attempt (failure count stays)
    r = requests.post(cluster.url + "/infer", json={"request_id": keys[5], "input_data": [1.0]}, timeout=10)
    assert r.status_code == 200 and cluster.breaker(dead)["failures"] == 5
    # bring it back; after the breaker's timeout the next request is the probe
    cluster.start_worker(dead)
    time.sleep(0.7)
    r = requests.post(cluster.url + "/infer", json={"request_id": keys[6], "input_data": [1.0]}, timeout=10)
    assert r.json()["node_id"] == cluster.node_id[dead]
    b = cluster.breaker(dead)
    assert b["state"] == "HALF_OPEN" and b["successes"] == 1 and b["failures"] == 5
    r = requests.post(cluster.url + "/infer", json={"request_id": keys[7], "input_data": [1.0]}, timeout=10)
    b = cluster.breaker(dead)
    assert b["state"] == "CLOSED" and b["successes"] == 2 and b["failures"] == 0


def test_large_body_through_the_gateway(cluster):
    data = [float(i % 97) for i in range(150528)]
    r = requests.post(cluster.url + "/infer", json={"request_id": "big", "input_data": data}, timeout=30)
    assert r.status_code == 200 and r.json()["output_data"][1] == 150528.0


def test_octet_stream_through_the_gateway(cluster):
    import numpy as np
    x = np.arange(5, dtype=np.float32)
    rid = "raw_7"
    hdr = {"Content-Type": "application/octet-stream", "X-Request-Id": rid}
    r = requests.post(cluster.url + "/infer", data=x.tobytes(), headers=hdr, timeout=5)
    assert r.status_code == 200
    j = r.json()
    assert j["node_id"] == cluster.node_id[cluster.ring.get(rid)] and j["request_id"] == rid
    assert j["output_data"][:2] == [10.0, 5.0]
    r = requests.post(cluster.url + "/infer", data=x.tobytes(), headers={"Content-Type": "application/octet-stream"}, timeout=5)
    assert r.status_code == 500 and "X-Request-Id" in r.json()["error"]
