"""Seeded random convolution shapes through b200rn_conv2d against the float64 CPU conv (tests/gpu_helpers.py): the
planner picks among nine tensor-core kernels by geometry (generic, dual-M, halo, dual-M halo, pair, wide, pair-wide,
flattened or patch tiling, stationary or streamed weights), and the hand-picked cases of test_gpu_ops.py sit on the
ResNet-50 shapes; these sit beside them -- odd spatial sizes, batches that leave ragged last tiles, every epilogue
combination -- so that a route taken for a shape nobody listed is still checked.  Sizes are kept small enough for the
float64 reference to take well under a second each."""
import os
import random
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def _cases(seed, n):
    rnd = random.Random(seed)
    out = []
    while len(out) < n:
        k = rnd.choice([1, 1, 1, 3])
        cin = rnd.choice([64, 128, 256, 512, 1024]) if k == 1 else rnd.choice([64, 128, 256])
        cout = rnd.choice([64, 128, 256, 512, 1024])
        stride = rnd.choice([1, 1, 1, 2])
        hw = rnd.choice([5, 7, 9, 13, 14, 20, 28])
        h, w = hw, rnd.choice([hw, hw + 1, max(3, hw - 2)])
        b = rnd.choice([1, 2, 3, 5, 9, 17, 33, 40, 65])
        pad = 1 if k == 3 else 0
        ho, wo = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
        if b * ho * wo * cout * cin * k * k > 1.5e10:           # keep the float64 reference cheap
            continue
        residual = rnd.random() < 0.5
        want_act = rnd.random() < 0.5
        want_out = True if not want_act else rnd.random() < 0.7
        relu = 0 if residual else rnd.choice([0, 1])
        out.append(dict(B=b, H=h, W=w, Cin=cin, Cout=cout, k=k, stride=stride, pad=pad, relu=relu, residual=residual,
                        want_out=want_out, want_act=want_act, act_relu=rnd.choice([0, 1]), seed=len(out)))
    return out


CASES = _cases(20261107, 72)


def _id(kw):
    return "%dx%d_s%d_%d-%d_@%dx%d_B%d%s%s%s" % (kw["k"], kw["k"], kw["stride"], kw["Cin"], kw["Cout"], kw["H"], kw["W"], kw["B"],
                                                 "_res" if kw["residual"] else "", "_out" if kw["want_out"] else "",
                                                 "_act" if kw["want_act"] else "")


@pytest.mark.parametrize("dtype", [0, 1, 2], ids=["exact", "bf16", "fp16"])
@pytest.mark.parametrize("kw", CASES, ids=[_id(c) for c in CASES])
def test_random_conv_shapes(dtype, kw):
    import gpu_helpers as gh
    r = gh.conv_case(dtype, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (kw, key, r[key])
