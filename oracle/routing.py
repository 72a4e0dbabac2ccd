"""TEST INFRASTRUCTURE ONLY (oracle): ctypes access to oracle/liboracle_routing.so (the plain-C
restatement, routing_oracle.c) and, if built, oracle/_ref/libref_routing.so (the reference's own
consistent_hash.cpp / lru_cache.h / batch_processor.h behind ref_shim.cpp).  Also a small
pure-Python LRU model following include/lru_cache.h:18-48 for operation-sequence tests.
"""
import ctypes as C
import os
from collections import OrderedDict

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _load(path):
    return C.CDLL(path) if os.path.exists(path) else None


def oracle_lib():
    lib = _load(os.path.join(_HERE, "liboracle_routing.so"))
    if lib is None:
        raise RuntimeError("oracle/liboracle_routing.so missing: run `make -C oracle` or __graft_entry__.build()")
    lib.ro_fnv1a.restype = C.c_uint32
    lib.ro_fnv1a.argtypes = [C.c_char_p, C.c_size_t]
    lib.ro_ring_build.restype = C.c_size_t
    lib.ro_ring_build.argtypes = [C.POINTER(C.c_char_p), C.c_size_t, C.c_int, C.POINTER(C.c_uint32),
                                  C.POINTER(C.c_int32)]
    lib.ro_ring_lookup.restype = C.c_int32
    lib.ro_ring_lookup.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_int32), C.c_size_t, C.c_char_p, C.c_size_t]
    lib.ro_hash_float.restype = C.c_uint64
    lib.ro_hash_float.argtypes = [C.c_float]
    lib.ro_vector_hash.restype = C.c_uint64
    lib.ro_vector_hash.argtypes = [C.POINTER(C.c_float), C.c_size_t]
    return lib


def ref_lib():
    """The reference's own code, or None when oracle/_ref was not built (no /root/reference)."""
    lib = _load(os.path.join(_HERE, "_ref", "libref_routing.so"))
    if lib is None:
        return None
    lib.ref_ring_new.restype = C.c_void_p
    lib.ref_ring_new.argtypes = [C.c_int]
    lib.ref_ring_free.argtypes = [C.c_void_p]
    lib.ref_ring_add.argtypes = [C.c_void_p, C.c_char_p]
    lib.ref_ring_remove.argtypes = [C.c_void_p, C.c_char_p]
    lib.ref_ring_hash.restype = C.c_uint32
    lib.ref_ring_hash.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
    lib.ref_ring_get.restype = C.c_int
    lib.ref_ring_get.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    lib.ref_ring_all.restype = C.c_int
    lib.ref_ring_all.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
    lib.ref_hash_float.restype = C.c_uint64
    lib.ref_hash_float.argtypes = [C.c_float]
    lib.ref_vector_hash.restype = C.c_uint64
    lib.ref_vector_hash.argtypes = [C.POINTER(C.c_float), C.c_size_t]
    lib.ref_lru_new.restype = C.c_void_p
    lib.ref_lru_new.argtypes = [C.c_size_t]
    lib.ref_lru_free.argtypes = [C.c_void_p]
    lib.ref_lru_get.restype = C.c_long
    lib.ref_lru_get.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_size_t, C.POINTER(C.c_float), C.c_size_t]
    lib.ref_lru_put.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_size_t, C.POINTER(C.c_float), C.c_size_t]
    for f in ("ref_lru_size", "ref_lru_hits", "ref_lru_misses"):
        getattr(lib, f).restype = C.c_size_t
        getattr(lib, f).argtypes = [C.c_void_p]
    lib.ref_batch_run.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_longlong), C.POINTER(C.c_double)]
    return lib


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class OracleRing:
    """ConsistentHash restated (src/consistent_hash.cpp:16-45) on top of routing_oracle.c."""

    def __init__(self, nodes, vnodes=150):
        self.lib = oracle_lib()
        self.nodes = list(nodes)
        arr = (C.c_char_p * len(nodes))(*[n.encode() for n in nodes])
        cap = max(1, len(nodes) * vnodes)
        self.h = (C.c_uint32 * cap)()
        self.owner = (C.c_int32 * cap)()
        self.m = self.lib.ro_ring_build(arr, len(nodes), vnodes, self.h, self.owner)

    def get(self, key):
        if isinstance(key, str):
            key = key.encode()
        i = self.lib.ro_ring_lookup(self.h, self.owner, self.m, key, len(key))
        return "" if i < 0 else self.nodes[i]

    def all_nodes(self):
        seen, out = set(), []
        for k in range(self.m):
            n = self.nodes[self.owner[k]]
            if n not in seen:
                seen.add(n)
                out.append(n)
        return out


class OracleLRU:
    """include/lru_cache.h:18-48: get moves to front and counts hit/miss; put on an existing key
    updates + moves to front; put on a new key at capacity evicts the least recently used first.
    Key equality is vector<float>::operator== (so -0.0 == +0.0 and NaN never matches)."""

    def __init__(self, capacity):
        self.cap = capacity
        self.d = OrderedDict()  # most recent last
        self.hits = self.misses = 0

    @staticmethod
    def _k(key):
        a = np.asarray(key, dtype=np.float32)
        if np.isnan(a).any():
            return None  # never equal to anything, itself included
        a = a + np.float32(0.0)  # -0.0 -> +0.0, the same equivalence class operator== uses
        return (a.size, a.tobytes())

    def get(self, key):
        k = self._k(key)
        if k is None or k not in self.d:
            self.misses += 1
            return None
        self.d.move_to_end(k)
        self.hits += 1
        return self.d[k]

    def put(self, key, value):
        k = self._k(key)
        if k is not None and k in self.d:
            self.d[k] = value
            self.d.move_to_end(k)
            return
        if len(self.d) >= self.cap:
            self.d.popitem(last=False)
        if k is None:
            k = ("nan", id(value), len(self.d), self.hits + self.misses)  # a distinct entry every time
        self.d[k] = value

    def size(self):
        return len(self.d)
