"""GPU parity of the conv kernels, called through the C ABI (b200rn_conv2d / b200rn_stem_conv), against a
float64 CPU convolution of the same (rounded) operands.  Tolerances are in gpu_helpers.TOL: fp32 path
2e-5 abs+rel; 16-bit paths are bounded by the rounding of the stored output (bf16 2^-8, fp16 2^-11)."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import gpu_helpers as gh  # noqa: E402


@pytest.mark.parametrize("dtype", [0, 1, 2], ids=["fp32", "bf16", "fp16"])
@pytest.mark.parametrize("name,kw", gh.CONV_CASES, ids=[c[0] for c in gh.CONV_CASES])
def test_conv2d(dtype, name, kw):
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


def test_conv2d_rejects_bad_shapes():
    # the tensor-core path needs channel multiples of 64; it must say so rather than fall back
    x = torch.zeros(1, 8, 8, 32, dtype=torch.float16, device="cuda")
    o = torch.zeros(1, 8, 8, 64, dtype=torch.float16, device="cuda")
    import numpy as np
    w = np.zeros((64, 1, 1, 32), dtype=np.float32)
    rc = gh.lib.b200rn_conv2d(2, x.data_ptr(), 1, 8, 8, 32, gh.pkg.fptr(w), None, 64, 1, 1, 1, 0, 0, None,
                              o.data_ptr(), None, None, None, 0, None, None, 0)
    assert rc != 0 and b"% 64" in gh.lib.b200rn_global_error()
