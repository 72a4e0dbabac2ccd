"""Pins oracle/routing_oracle.c (and the LRU model) to the reference: golden vectors harvested from
the reference's own sources and, when oracle/_ref is built, the reference objects."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from oracle import routing


@pytest.fixture(scope="module")
def gold(golden_dir):
    with open(os.path.join(golden_dir, "routing_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def lib():
    return routing.oracle_lib()


def test_fnv_golden(lib, gold):
    for k, v in gold["fnv"].items():
        assert lib.ro_fnv1a(k.encode(), len(k.encode())) == v, k
    for hx, v in gold["fnv_bytes_hex"].items():
        b = bytes.fromhex(hx)  # bytes >= 0x80: the sign-extension case
        assert lib.ro_fnv1a(b, len(b)) == v


def test_ring_golden(gold):
    ring = routing.OracleRing(gold["ring3_nodes"])
    for k, v in gold["ring3"].items():
        assert ring.get(k) == v, k
    assert ring.all_nodes() == gold["ring3_all_nodes_order"]
    split = {n: 0 for n in gold["ring3_nodes"]}
    for i in range(10000):
        split[ring.get("req_%d" % i)] += 1
    assert [split[n] for n in gold["ring3_nodes"]] == gold["ring3_split_req_0_9999"]
    nodes8 = ["localhost:%d" % p for p in range(8001, 8009)]
    ring8 = routing.OracleRing(nodes8)
    split = {n: 0 for n in nodes8}
    for i in range(10000):
        split[ring8.get("req_%d" % i)] += 1
    assert [split[n] for n in nodes8] == gold["ring8_split_req_0_9999"]
    assert routing.OracleRing([]).get("x") == ""


def test_hash_float_and_vector_hash_golden(lib, gold):
    for k, v in gold["hash_float"].items():
        assert lib.ro_hash_float(float(k)) == v, k
    for vec, h in gold["vector_hash"]:
        a = np.asarray(vec, dtype=np.float32)
        assert lib.ro_vector_hash(routing.fptr(a), a.size) == h, vec
    i = np.arange(150528, dtype=np.uint64)
    # SURVEY 8(c) writes "i*2654435761 mod 2^32"; the value it recorded is reproduced by the 64-bit
    # product (no 2^32 wrap) -- checked against the reference's VectorHash in oracle/_ref.
    big = ((i * np.uint64(2654435761)) % np.uint64(1000)).astype(np.float32) / np.float32(1000)
    assert lib.ro_vector_hash(routing.fptr(big), big.size) == gold["vector_hash_150528"]
    a, b = [np.asarray(v, dtype=np.float32) for v in gold["vector_hash_collision"]]
    assert lib.ro_vector_hash(routing.fptr(a), a.size) == lib.ro_vector_hash(routing.fptr(b), b.size)


def test_lru_model_golden():
    c = routing.OracleLRU(2)
    A, B, Cc = [1.0, 2.0], [3.0], [4.0, 5.0, 6.0]
    c.put(A, "a"); c.put(B, "b")
    assert c.get(A) == "a"
    c.put(Cc, "c")
    assert c.get(B) is None and c.get(A) == "a" and c.get(Cc) == "c" and c.size() == 2
    # colliding hashes, unequal vectors -> distinct entries
    c2 = routing.OracleLRU(4)
    c2.put([1, 5, 2, 9, 3], "x"); c2.put([1, 7, 2, 8, 3], "y")
    assert c2.get([1, 5, 2, 9, 3]) == "x" and c2.get([1, 7, 2, 8, 3]) == "y"
    # -0.0 == +0.0 under operator==
    c2.put([0.0, 1.0], "z")
    assert c2.get([-0.0, 1.0]) == "z"


# ---- against the reference's own objects (only where oracle/_ref could be built) ----

ref = routing.ref_lib()
needs_ref = pytest.mark.skipif(ref is None, reason="oracle/_ref not built (no /root/reference on this machine)")


@needs_ref
def test_oracle_vs_reference_hashes(lib):
    rs = np.random.RandomState(7)
    r = ref.ref_ring_new(150)
    try:
        for _ in range(2000):
            n = rs.randint(0, 40)
            key = bytes(rs.randint(1, 256, size=n).astype(np.uint8))  # includes bytes >= 0x80, no NUL
            assert lib.ro_fnv1a(key, len(key)) == ref.ref_ring_hash(r, key, len(key))
    finally:
        ref.ref_ring_free(r)
    vals = np.concatenate([rs.randn(2000).astype(np.float32),
                           np.asarray([0.0, -0.0, np.inf, -np.inf, 1e-45, 3.4e38], dtype=np.float32)])
    for v in vals:
        assert lib.ro_hash_float(float(v)) == ref.ref_hash_float(float(v))
    for n in [0, 1, 2, 3, 4, 7, 100, 150528]:
        a = rs.randn(n).astype(np.float32)
        assert lib.ro_vector_hash(routing.fptr(a), n) == ref.ref_vector_hash(routing.fptr(a), n)


@needs_ref
def test_oracle_vs_reference_ring():
    for nodes in (["localhost:8001", "localhost:8002", "localhost:8003"],
                  ["gpu%d" % i for i in range(8)], ["only"], ["a", "a"]):
        r = ref.ref_ring_new(150)
        for n in nodes:
            ref.ref_ring_add(r, n.encode())
        ring = routing.OracleRing(nodes)
        buf = C.create_string_buffer(256)
        for i in range(3000):
            k = ("req_%d" % i).encode()
            ref.ref_ring_get(r, k, len(k), buf, 256)
            assert buf.value.decode() == ring.get(k)
        big = C.create_string_buffer(4096)
        ref.ref_ring_all(r, big, 4096)
        assert big.value.decode().split("\n")[:-1] == ring.all_nodes()
        ref.ref_ring_free(r)


@needs_ref
def test_lru_model_vs_reference_sequence():
    rs = np.random.RandomState(3)
    keys = [rs.randint(0, 4, size=rs.randint(1, 5)).astype(np.float32) for _ in range(40)]
    c = ref.ref_lru_new(8)
    m = routing.OracleLRU(8)
    out = np.zeros(16, dtype=np.float32)
    for step in range(3000):
        k = keys[rs.randint(len(keys))]
        if rs.rand() < 0.5:
            v = rs.randn(3).astype(np.float32)
            ref.ref_lru_put(c, routing.fptr(k), k.size, routing.fptr(v), 3)
            m.put(k, v)
        else:
            n = ref.ref_lru_get(c, routing.fptr(k), k.size, routing.fptr(out), 16)
            got = m.get(k)
            assert (n < 0) == (got is None)
            if got is not None:
                assert np.array_equal(out[:n], got)
        assert ref.ref_lru_size(c) == m.size()
    assert ref.ref_lru_hits(c) == m.hits and ref.ref_lru_misses(c) == m.misses
    ref.ref_lru_free(c)
