"""GPU parity of the conv kernels, called through the C ABI (b200rn_conv2d / b200rn_stem_conv), against a
float64 CPU convolution of the same (rounded) operands.  Tolerances are in gpu_helpers.TOL: fp32 path
2e-5 abs+rel; 16-bit paths are bounded by the rounding of the stored output (bf16 2^-8, fp16 2^-11)."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import gpu_helpers as gh  # noqa: E402


@pytest.mark.parametrize("dtype", [0, 1, 2], ids=["fp32", "bf16", "fp16"])
@pytest.mark.parametrize("name,kw", gh.CONV_CASES, ids=[c[0] for c in gh.CONV_CASES])
def test_conv2d(dtype, name, kw, monkeypatch):
    if name.startswith("wide"):
        monkeypatch.setenv("B200RN_WIDE", "2")      # also the residual-free shapes go through conv_wide_kernel
    r = gh.conv_case(dtype, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


@pytest.mark.parametrize("dtype", [1, 2], ids=["bf16", "fp16"])
@pytest.mark.parametrize("shape", [(2, 224, 224), (3, 64, 64), (1, 32, 32)])
def test_stem_conv(dtype, shape):
    r = gh.stem_case(dtype, *shape)
    assert r["rc"] == 0, r["msg"]
    assert r["out"]["viol"] == 0, r["out"]


TAIL_CASES = [
    # (B, H, W, tail_cout, relu): 56x56 is the ResNet stage the planner fuses; the small and ragged ones exercise
    # partial patches, a single tile, and more tiles than SMs
    (2, 56, 56, 64, 0), (1, 8, 8, 64, 0), (3, 13, 9, 64, 1), (1, 4, 4, 128, 0), (8, 56, 56, 128, 0), (40, 28, 28, 64, 0),
]


@pytest.mark.parametrize("dtype", [1, 2], ids=["bf16", "fp16"])
@pytest.mark.parametrize("case", TAIL_CASES, ids=["%dx%dx%d_t%d_r%d" % c for c in TAIL_CASES])
def test_conv1x1_tail(dtype, case):
    B, H, W, tc, relu = case
    r = gh.tail_case(dtype, B, H, W, tail_cout=tc, relu=relu)
    assert r["rc"] == 0, r["msg"]
    assert r["out"]["viol"] == 0, r["out"]
    assert r["tail"]["viol"] == 0, r["tail"]


def test_conv1x1_tail_rejects_unsupported_shapes():
    x = torch.zeros(1, 8, 8, 64, dtype=torch.float16, device="cuda")
    import numpy as np
    z = lambda *s: np.zeros(s, dtype=np.float32)
    o = torch.zeros(1, 8, 8, 256, dtype=torch.float16, device="cuda")
    t = torch.zeros(1, 8, 8, 192, dtype=torch.float16, device="cuda")
    rc = gh.lib.b200rn_conv1x1_tail(2, x.data_ptr(), 1, 8, 8, 64, gh.pkg.fptr(z(256, 64)), None, 0, o.data_ptr(), o.data_ptr(),
                                    gh.pkg.fptr(z(256)), gh.pkg.fptr(z(256)), 1, gh.pkg.fptr(z(192, 256)), None, 192, 1, t.data_ptr())
    assert rc != 0 and b"tail" in gh.lib.b200rn_global_error()


def test_conv2d_rejects_bad_shapes():
    # the tensor-core path needs channel multiples of 64; it must say so rather than fall back
    x = torch.zeros(1, 8, 8, 32, dtype=torch.float16, device="cuda")
    o = torch.zeros(1, 8, 8, 64, dtype=torch.float16, device="cuda")
    import numpy as np
    w = np.zeros((64, 1, 1, 32), dtype=np.float32)
    rc = gh.lib.b200rn_conv2d(2, x.data_ptr(), 1, 8, 8, 32, gh.pkg.fptr(w), None, 64, 1, 1, 1, 0, 0, None,
                              o.data_ptr(), None, None, None, 0, None, None, 0)
    assert rc != 0 and b"% 64" in gh.lib.b200rn_global_error()
