"""CPU-side tests of the product's host code (no GPU): the C-ABI library loads and exports every
symbol include/b200rn.h declares, the ONNX parser + planner agree with the oracle's independent
reader, and the routing / cache-key / LRU / batching restatements match the oracle, the golden
vectors and (where built) the reference's own objects."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

import b200rn_loader
from oracle import onnx_mini, routing
from tools import make_resnet50v2_onnx as gen

pkg = b200rn_loader.load_package()
lib = pkg.lib
ROOT = b200rn_loader.ROOT


@pytest.fixture(scope="module")
def gold(golden_dir):
    with open(os.path.join(golden_dir, "routing_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def tiny_model(tmp_path_factory):
    p = str(tmp_path_factory.mktemp("m") / "tiny.onnx")
    gen.write_model(p, "tiny", seed=0)
    return p


@pytest.fixture(scope="module")
def r50_model(tmp_path_factory):
    p = str(tmp_path_factory.mktemp("m") / "r50.onnx")
    gen.write_model(p, "resnet50", seed=0)
    return p


def test_every_declared_symbol_is_exported():
    hdr = open(os.path.join(ROOT, "include", "b200rn.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(b200rn_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 35
    for n in sorted(names):
        assert hasattr(lib, n), "libb200rn.so does not export %s" % n


def test_no_gpu_means_error_not_fallback(tiny_model):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.B200RNError, match="no CPU path"):
        pkg.Engine(tiny_model, mode="exact")


def test_missing_or_malformed_model_is_an_error(tmp_path):
    assert lib.b200rn_describe_plan(str(tmp_path / "nope.onnx").encode(), 0, None, 0) == -1
    assert b"cannot open" in lib.b200rn_global_error()
    bad = tmp_path / "bad.onnx"
    bad.write_bytes(b"\x3a\xff\xff\xff\x0fgarbage")
    assert lib.b200rn_describe_plan(str(bad).encode(), 0, None, 0) == -1


@pytest.mark.parametrize("prologue", [False, True])
def test_plan_resnet50(r50_model, prologue, monkeypatch):
    monkeypatch.setenv("B200RN_PROLOGUE", "1" if prologue else "0")
    monkeypatch.setenv("B200RN_TAIL", "0")
    monkeypatch.setenv("B200RN_DUAL_INPUT", "0")
    monkeypatch.setenv("B200RN_STEM_POOL", "0")
    text = pkg.describe_plan(r50_model, "fp16")
    lines = text.strip().splitlines()
    head, layers = lines[0], lines[1:]
    # 4,087,136,256 conv MACs + 2,048,000 FC MACs (8.178368512 GFLOP / image)
    assert "macs/sample=4089184256" in head
    kinds = [l.split()[1] for l in layers]
    assert kinds.count("Conv") == 53 and kinds.count("MaxPool") == 1 and kinds.count("Gemm") == 1
    assert kinds.count("GlobalAvgPool") == 1 and kinds[0] == "InputXform"
    # every BN / ReLU / Add of the graph was fused away: 57 launches for 174 nodes
    assert len(layers) == 57 and "Affine" not in kinds and "Add" not in kinds
    convs = [l for l in layers if l.split()[1] == "Conv"]
    assert sum("in2=t" in l for l in convs) == 16           # residual adds in the epilogue
    if prologue:
        # the BN->ReLU after each residual sum: applied in the reading 1x1 convs' operand staging where the raw
        # sum is written anyway (12 blocks), else written as the only output of the producer (last block of a stage)
        assert sum(" pre=1" in l for l in convs) == 12
        assert sum("out_act=t" in l for l in convs) == 4 and sum(" out=- " in l for l in convs) == 4
    else:
        assert sum(" pre=1" in l for l in convs) == 0
        assert sum("out_act=t" in l for l in convs) == 16      # raw sum + BN-ReLU'd copy from the producer
        assert sum(" out=- " in l for l in convs) == 4          # raw copy not written when unused
    shapes = set(re.search(r"k=(\dx\d) s=(\d) p=(\d) cin=(\d+) cout=(\d+)", l).groups() +
                 (re.search(r"in=t\d+\[\d+x(\d+)x", l).group(1),) for l in convs)
    assert len(shapes) == 23                                # 23 unique conv shapes


def test_plan_resnet50_tail_fusion(r50_model, monkeypatch):
    """16-bit plan with tail fusion only: at 56x56 the wide 1x1 conv (+residual) that ends a block and the narrow 1x1
    conv that starts the next run as one launch; the exact (fp32 CUDA-core) plan is never fused that way."""
    monkeypatch.delenv("B200RN_TAIL", raising=False)
    monkeypatch.setenv("B200RN_DUAL_INPUT", "0")
    monkeypatch.setenv("B200RN_STEM_POOL", "0")
    monkeypatch.setenv("B200RN_PROLOGUE", "0")
    lines = pkg.describe_plan(r50_model, "fp16").strip().splitlines()
    head, layers = lines[0], lines[1:]
    assert "macs/sample=4089184256" in head and len(layers) == 55
    fused = [l for l in layers if "tail_out=t" in l]
    assert len(fused) == 2
    for l in fused:
        assert "cin=64 cout=256" in l and "tail_cout=64" in l and "out_act=-" in l and "in2=t" in l
        assert re.search(r"in=t\d+\[64x56x56\]", l) and re.search(r"tail_out=t\d+\[64x56x56\]", l)
        assert "stage1_conv" in l and l.count("_conv") == 2       # two ONNX Conv nodes in one launch
    exact = pkg.describe_plan(r50_model, "exact").strip().splitlines()[1:]
    assert len(exact) == 57 and not any("tail_out" in l or "in_b=t" in l for l in exact)


def test_plan_resnet50_default_16bit(r50_model, monkeypatch):
    """Default 16-bit plan: the four projection blocks run 'main 1x1 conv + shortcut 1x1 conv + Add' as one two-input
    conv (the shortcut tensor is never materialised), on top of the tail fusion: 51 launches for 53 convs."""
    for v in ("B200RN_TAIL", "B200RN_DUAL_INPUT", "B200RN_PROLOGUE", "B200RN_STEM_POOL", "B200RN_STEM_V2"):
        monkeypatch.delenv(v, raising=False)
    lines = pkg.describe_plan(r50_model, "fp16").strip().splitlines()
    head, layers = lines[0], lines[1:]
    assert "macs/sample=4089184256" in head and len(layers) == 51
    two = [l for l in layers if "in_b=t" in l]
    assert len(two) == 4
    want = [("64", "64", "1", "56"), ("128", "256", "2", "28"), ("256", "512", "2", "14"), ("512", "1024", "2", "7")]
    got = [(re.search(r" cin=(\d+) ", l).group(1), re.search(r"cin_b=(\d+)", l).group(1), re.search(r"stride_b=(\d)", l).group(1),
            re.search(r"out=t\d+\[\d+x(\d+)x", l).group(1)) for l in two]
    assert got == want
    for l in two:
        assert "in2=-" in l and "k=1x1 s=1 p=0" in l and "out_act=t" not in l.split("tail_out")[0] or "tail_out=t" in l or "out_act=t" in l
        assert l.count("_conv") >= 2 and "_plus0" in l
    assert sum("in2=t" in l for l in layers) == 12                  # the other residual adds stay epilogue adds
    tails = [l for l in layers if "tail_out=t" in l]
    assert len(tails) == 2 and sum("in_b=t" in l for l in tails) == 1   # stage 1's projection block also carries the tail


def test_plan_resnet50_stem_pool_opt_in(r50_model, monkeypatch):
    """B200RN_STEM_POOL=1: the stem conv carries the max-pool (its outputs are the pooled 56x56 tensors, no MaxPool
    launch remains).  Opt-in: measured no faster than the two separate launches."""
    for v in ("B200RN_TAIL", "B200RN_DUAL_INPUT", "B200RN_PROLOGUE", "B200RN_STEM_V2"):
        monkeypatch.delenv(v, raising=False)
    monkeypatch.setenv("B200RN_STEM_POOL", "1")
    lines = pkg.describe_plan(r50_model, "fp16").strip().splitlines()
    head, layers = lines[0], lines[1:]
    assert "macs/sample=4089184256" in head and len(layers) == 50
    stem = [l for l in layers if "fused_pool=1" in l]
    assert len(stem) == 1 and "k=7x7 s=2 p=3" in stem[0] and re.search(r"out_act=t\d+\[64x56x56\]", stem[0]) and "pool0" in stem[0]
    assert not any(l.split()[1] == "MaxPool" for l in layers)
    monkeypatch.setenv("B200RN_STEM_V2", "0")          # without the dedicated stem kernel there is nothing to fuse into
    assert len(pkg.describe_plan(r50_model, "fp16").strip().splitlines()) == 52


def test_cpp_parser_matches_python_reader(tiny_model):
    m = onnx_mini.load(tiny_model)
    text = pkg.describe_plan(tiny_model, "exact")
    assert m["inputs"][0][0] == "data" and m["inputs"][0][1] == [-1, 3, 32, 32]
    assert "input data [3,32,32]" in text and "[10]" in text.splitlines()[0]
    n_conv = sum(n["op"] == "Conv" for n in m["nodes"])
    assert sum(l.split()[1] == "Conv" for l in text.splitlines()[1:]) == n_conv
    macs = 0
    shapes = {"data": (3, 32, 32)}
    # independent MAC count from the Python-parsed graph
    for n in m["nodes"]:
        i = n["input"][0]
        if n["op"] == "Conv":
            w = m["initializers"][n["input"][1]]
            c, h, wd = shapes[i]
            s, p = n["attr"]["strides"][0], n["attr"]["pads"][0]
            ho, wo = (h + 2 * p - w.shape[2]) // s + 1, (wd + 2 * p - w.shape[3]) // s + 1
            macs += w.size * ho * wo
            shapes[n["output"][0]] = (w.shape[0], ho, wo)
        elif n["op"] == "MaxPool":
            c, h, wd = shapes[i]
            shapes[n["output"][0]] = (c, (h + 2 - 3) // 2 + 1, (wd + 2 - 3) // 2 + 1)
        elif n["op"] == "Gemm":
            macs += m["initializers"][n["input"][1]].size
        elif n["op"] in ("GlobalAveragePool", "Flatten"):
            shapes[n["output"][0]] = (shapes[i][0], 1, 1)
        else:
            shapes[n["output"][0]] = shapes[i]
    assert "macs/sample=%d" % macs in text


def test_unsupported_operator_fails_loudly(tmp_path):
    # a Softmax-terminated graph is supported; a graph with an unknown op is an error, not a fallback
    p = str(tmp_path / "sm.onnx")
    gen.write_model(p, "tiny", seed=1, softmax=True)
    assert "Softmax" in pkg.describe_plan(p, "exact")
    data = open(p, "rb").read().replace(b"Softmax", b"Sxftmax")
    q = str(tmp_path / "unk.onnx")
    open(q, "wb").write(data)
    with pytest.raises(pkg.B200RNError, match="not supported"):
        pkg.describe_plan(q, "exact")


# ---- routing / cache keys: product vs golden vs oracle vs reference -----------------------------

def _ring(nodes, vnodes=150):
    r = lib.b200rn_ring_create(vnodes)
    for n in nodes:
        lib.b200rn_ring_add(r, n.encode())
    return r


def _get(r, key):
    if isinstance(key, str):
        key = key.encode()
    buf = C.create_string_buffer(256)
    lib.b200rn_ring_get(r, key, len(key), buf, 256)
    return buf.value.decode()


def test_product_fnv_and_ring_golden(gold):
    for k, v in gold["fnv"].items():
        assert lib.b200rn_fnv1a(k.encode(), len(k.encode())) == v
    for hx, v in gold["fnv_bytes_hex"].items():
        b = bytes.fromhex(hx)
        assert lib.b200rn_fnv1a(b, len(b)) == v
    r = _ring(gold["ring3_nodes"])
    for k, v in gold["ring3"].items():
        assert _get(r, k) == v
    buf = C.create_string_buffer(1024)
    lib.b200rn_ring_nodes(r, buf, 1024)
    assert buf.value.decode().split("\n")[:-1] == gold["ring3_all_nodes_order"]
    split = {n: 0 for n in gold["ring3_nodes"]}
    for i in range(10000):
        split[_get(r, "req_%d" % i)] += 1
    assert [split[n] for n in gold["ring3_nodes"]] == gold["ring3_split_req_0_9999"]
    lib.b200rn_ring_destroy(r)
    nodes8 = ["localhost:%d" % p for p in range(8001, 8009)]
    r8 = _ring(nodes8)
    split = {n: 0 for n in nodes8}
    for i in range(10000):
        split[_get(r8, "req_%d" % i)] += 1
    assert [split[n] for n in nodes8] == gold["ring8_split_req_0_9999"]
    lib.b200rn_ring_destroy(r8)
    e = lib.b200rn_ring_create(150)
    assert _get(e, "anything") == ""
    lib.b200rn_ring_destroy(e)


def test_product_ring_vs_oracle_with_removal():
    rs = np.random.RandomState(5)
    nodes = ["gpu-worker-%d:80%02d" % (i, i) for i in range(8)]
    r = _ring(nodes)
    for victim in (None, nodes[3], nodes[0]):
        if victim:
            lib.b200rn_ring_remove(r, victim.encode())
            nodes = [n for n in nodes if n != victim]
        o = routing.OracleRing(nodes)
        for _ in range(3000):
            key = bytes(rs.randint(1, 256, size=rs.randint(0, 24)).astype(np.uint8))
            assert _get(r, key) == o.get(key)
    lib.b200rn_ring_destroy(r)


def test_product_cache_keys_golden_and_oracle(gold):
    o = routing.oracle_lib()
    for k, v in gold["hash_float"].items():
        assert lib.b200rn_hash_float(float(k)) == v
    for vec, h in gold["vector_hash"]:
        a = np.asarray(vec, dtype=np.float32)
        assert lib.b200rn_vector_hash(pkg.fptr(a), a.size) == h
    rs = np.random.RandomState(9)
    for n in [0, 1, 2, 3, 5, 64, 150528]:
        a = (rs.randn(n) * 10 ** rs.uniform(-3, 3)).astype(np.float32)
        assert lib.b200rn_vector_hash(pkg.fptr(a), n) == o.ro_vector_hash(routing.fptr(a), n)
    specials = np.asarray([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 3.4028235e38], dtype=np.float32)
    for v in specials:
        assert lib.b200rn_hash_float(C.c_float(v)) == o.ro_hash_float(C.c_float(v))


def test_product_lru_vs_oracle_and_reference():
    ref = routing.ref_lib()
    rs = np.random.RandomState(3)
    keys = [rs.randint(0, 4, size=rs.randint(1, 5)).astype(np.float32) for _ in range(40)]
    keys.append(np.asarray([0.0, 1.0], dtype=np.float32))
    keys.append(np.asarray([-0.0, 1.0], dtype=np.float32))   # equal to the previous under operator==
    c = lib.b200rn_lru_create(8)
    rc = ref.ref_lru_new(8) if ref else None
    m = routing.OracleLRU(8)
    out = np.zeros(16, dtype=np.float32)
    out2 = np.zeros(16, dtype=np.float32)
    for _ in range(4000):
        k = keys[rs.randint(len(keys))]
        if rs.rand() < 0.5:
            v = rs.randn(3).astype(np.float32)
            lib.b200rn_lru_put(c, pkg.fptr(k), k.size, pkg.fptr(v), 3)
            m.put(k, v)
            if rc:
                ref.ref_lru_put(rc, routing.fptr(k), k.size, routing.fptr(v), 3)
        else:
            n = lib.b200rn_lru_get(c, pkg.fptr(k), k.size, pkg.fptr(out), 16)
            got = m.get(k)
            assert (n < 0) == (got is None)
            if got is not None:
                assert np.array_equal(out[:n], got)
            if rc:
                n2 = ref.ref_lru_get(rc, routing.fptr(k), k.size, routing.fptr(out2), 16)
                assert n2 == n and (n < 0 or np.array_equal(out[:n], out2[:n]))
        assert lib.b200rn_lru_size(c) == m.size()
    assert lib.b200rn_lru_hits(c) == m.hits and lib.b200rn_lru_misses(c) == m.misses
    if rc:
        assert ref.ref_lru_hits(rc) == m.hits and ref.ref_lru_size(rc) == m.size()
        ref.ref_lru_free(rc)
    lib.b200rn_lru_destroy(c)


def test_lru_golden_sequence():
    c = lib.b200rn_lru_create(2)
    A = np.asarray([1, 2], np.float32); B = np.asarray([3], np.float32); Cc = np.asarray([4, 5, 6], np.float32)
    v = np.asarray([7.0], np.float32)
    out = np.zeros(4, np.float32)
    lib.b200rn_lru_put(c, pkg.fptr(A), 2, pkg.fptr(v), 1)
    lib.b200rn_lru_put(c, pkg.fptr(B), 1, pkg.fptr(v), 1)
    assert lib.b200rn_lru_get(c, pkg.fptr(A), 2, pkg.fptr(out), 4) == 1
    lib.b200rn_lru_put(c, pkg.fptr(Cc), 3, pkg.fptr(v), 1)
    assert lib.b200rn_lru_get(c, pkg.fptr(B), 1, pkg.fptr(out), 4) == -1     # B evicted
    assert lib.b200rn_lru_get(c, pkg.fptr(A), 2, pkg.fptr(out), 4) == 1
    assert lib.b200rn_lru_get(c, pkg.fptr(Cc), 3, pkg.fptr(out), 4) == 1
    assert lib.b200rn_lru_size(c) == 2
    lib.b200rn_lru_destroy(c)


def test_batch_processor_behaviour_matches_reference():
    out = (C.c_longlong * 5)()
    avg = C.c_double()
    lib.b200rn_batch_selftest(32, 20, 30, 64, 1, out, C.byref(avg))
    total, batches, timeout_b, full_b, wrong = list(out)
    # opportunistic drain -> no timeout batches, every batch "full", batches < requests
    assert total == 64 and wrong == 0 and timeout_b == 0 and full_b == batches and 2 <= batches < 64
    assert abs(avg.value * batches - 64) < 1e-6
    ref = routing.ref_lib()
    if ref:
        out2 = (C.c_longlong * 5)()
        avg2 = C.c_double()
        ref.ref_batch_run(32, 20, 30, 64, 1, out2, C.byref(avg2))
        assert out2[0] == 64 and out2[4] == 0 and out2[2] == 0 and out2[3] == out2[1] and 2 <= out2[1] < 64


def test_batch_processor_fill_wait_and_two_in_flight():
    """The two optional batching policies of SURVEY 8(f) rank 3 (neither is in the reference, both are off by default):
    fill-or-timeout counts deadline batches in timeout_batches and full ones in full_batches; two consumer threads
    run two callbacks at once and still answer every request correctly."""
    out = (C.c_longlong * 6)()
    avg = C.c_double()
    # three requests, batch cap 8, 60 ms to fill: one batch of 3 closed by the deadline
    lib.b200rn_batch_selftest2(8, 200, 1, 3, 1, 1, 1, out, C.byref(avg))
    total, batches, timeout_b, full_b, wrong = list(out)[:5]
    # (a loaded host may start a submitter thread late: then a second deadline batch, never a "full" one)
    assert total == 3 and wrong == 0 and full_b == 0 and 1 <= timeout_b == batches <= 2 and abs(avg.value * batches - 3.0) < 1e-9
    # eight requests fill the batch before the deadline: one full batch, no timeout batch
    lib.b200rn_batch_selftest2(8, 2000, 1, 8, 1, 1, 1, out, C.byref(avg))
    assert list(out)[:5] == [8, 1, 0, 1, 0]
    # default policy, one consumer: never more than one callback at a time
    lib.b200rn_batch_selftest2(4, 20, 20, 16, 2, 0, 1, out, C.byref(avg))
    assert out[0] == 32 and out[4] == 0 and out[5] == 1
    # two in flight: two callbacks overlap, all 32 answers right, batch accounting still adds up
    lib.b200rn_batch_selftest2(4, 20, 20, 16, 2, 0, 2, out, C.byref(avg))
    assert out[0] == 32 and out[4] == 0 and out[5] == 2
    assert abs(avg.value * out[1] - 32) < 1e-6 and out[2] == 0 and out[3] == out[1]
