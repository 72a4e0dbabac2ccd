#!/usr/bin/env python3
"""Regenerates (default) or verifies (--check) tests/golden/routing_golden.json from the reference's OWN routing and
cache-key code: /root/reference/src/consistent_hash.cpp and include/lru_cache.h compiled where they lie into
oracle/_ref/libref_routing.so by oracle/Makefile (`make -C oracle ref`; needs /root/reference, so this runs in the
build container, not on the GPU box).  The keys are the ones SURVEY.md 8(c) lists.

    python tests/golden/make_routing_golden.py           # rewrite the file
    python tests/golden/make_routing_golden.py --check   # exit 1 if the committed file differs"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import routing  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "routing_golden.json")
FNV_KEYS = ["", "a", "req_0", "req_1", "req_9999", "cache_miss_0", "cache_hit_99", "diag_test", "test1",
            "localhost:8001#0", "localhost:8001#149"]
RING3_KEYS = ["req_%d" % i for i in range(10)] + ["req_9999", "cache_miss_0", "cache_hit_99", "diag_test", "test1", ""]
FLOATS = ["0.0", "-0.0", "1.0", "2.0", "3.0", "0.5", "-1.0"]
VECTORS = [[], [1], [0, 1, 2], [1, 2, 3], [9, 10, 11], [1, 2, 3, 4], [0, 0, 0]]


def main():
    ref = routing.ref_lib()
    if ref is None:
        sys.exit("oracle/_ref/libref_routing.so is not built (make -C oracle ref; needs /root/reference)")

    def ring(nodes):
        r = ref.ref_ring_new(150)
        for n in nodes:
            ref.ref_ring_add(r, n.encode())
        return r

    def get(r, key):
        buf = C.create_string_buffer(256)
        k = key if isinstance(key, bytes) else key.encode()
        ref.ref_ring_get(r, k, len(k), buf, 256)
        return buf.value.decode()

    def vhash(v):
        a = np.asarray(v, dtype=np.float32)
        return int(ref.ref_vector_hash(a.ctypes.data_as(C.POINTER(C.c_float)), a.size))

    nodes3 = ["localhost:800%d" % i for i in (1, 2, 3)]
    nodes8 = ["localhost:800%d" % i for i in range(1, 9)]
    r3, r8 = ring(nodes3), ring(nodes8)
    buf = C.create_string_buffer(4096)
    ref.ref_ring_all(r3, buf, 4096)
    split = lambda r, nodes: [sum(1 for i in range(10000) if get(r, "req_%d" % i) == n) for n in nodes]
    # SURVEY 8(c) writes "i*2654435761 mod 2^32"; the value it recorded comes from the 64-bit product (no 2^32 wrap)
    idx = np.arange(150528, dtype=np.uint64)
    big = ((idx * np.uint64(2654435761)) % np.uint64(1000)).astype(np.float32) / np.float32(1000)
    g = {
        "_source": "Produced by tests/golden/make_routing_golden.py from the reference's own sources compiled where they lie "
                   "(src/consistent_hash.cpp, include/lru_cache.h via oracle/_ref/libref_routing.so, g++ -std=c++17 -O2); "
                   "the keys are those of SURVEY.md section 8(c).",
        "fnv": {k: int(ref.ref_ring_hash(r3, k.encode(), len(k.encode()))) for k in FNV_KEYS},
        "fnv_bytes_hex": {"c3a9": int(ref.ref_ring_hash(r3, b"\xc3\xa9", 2))},
        "ring3_nodes": nodes3,
        "ring3": {k: get(r3, k) for k in RING3_KEYS},
        "ring3_all_nodes_order": [s for s in buf.value.decode().split("\n") if s],
        "ring3_split_req_0_9999": split(r3, nodes3),
        "ring8_split_req_0_9999": split(r8, nodes8),
        "hash_float": {f: int(ref.ref_hash_float(float(f))) for f in FLOATS},
        "vector_hash": [[v, vhash(v)] for v in VECTORS],
        "vector_hash_150528": vhash(big),
        "vector_hash_collision": [[1, 5, 2, 9, 3], [1, 7, 2, 8, 3]],
    }
    assert vhash(g["vector_hash_collision"][0]) == vhash(g["vector_hash_collision"][1])
    ref.ref_ring_free(r3)
    ref.ref_ring_free(r8)
    if "--check" in sys.argv:
        old = json.load(open(OUT))
        bad = [k for k in g if k != "_source" and old.get(k) != json.loads(json.dumps(g[k]))]
        if bad:
            sys.exit("routing_golden.json differs from the reference in: %s" % bad)
        print("routing_golden.json matches the reference's own code (%d sections)" % (len(g) - 1))
        return
    with open(OUT, "w") as f:
        json.dump(g, f, indent=1)
        f.write("\n")
    print("wrote", OUT)


if __name__ == "__main__":
    main()
