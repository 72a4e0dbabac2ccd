"""The allocation-free stem entry point, b200rn_stem_conv_static.

Who needs it: embedders that may not allocate device memory at call time -- a serving worker whose
device pool was divided up when it started, a request thread that cannot tolerate the device-wide
synchronisation cudaMalloc implies, a plugin host that does not let plugins own device memory.  For
them the library keeps the stem's scratch (framed input, packed filters, bias, input scale/shift) in
one fixed device array; this test checks that the result is the same as the allocating entry point's.
The comparison target is b200rn_stem_conv rather than the CPU oracle: that entry point is itself pinned to a
float64 CPU convolution by tests/test_gpu_ops.py::test_stem_conv, so agreement with it carries the oracle's
verdict over to the static variant.

It lives in its own file, collected last, so a `-x` run has already exercised every other test.
The scratch array is shared and keeps the previous call's contents, so the check is made on two calls
in a row with different shapes (the second, smaller one would read stale borders if the entry point
forgot to clear what it uses), then once more with the first shape.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import gpu_helpers as gh  # noqa: E402

lib, pkg = gh.lib, gh.pkg


def _run_both(dtype, B, H, W, Cout, seed):
    g = torch.Generator().manual_seed(seed)
    tdt = gh.TORCH_DT[dtype]
    x = torch.randn(B, 3, H, W, generator=g).contiguous().cuda()
    s = (torch.rand(3, generator=g) + 0.5).numpy().astype(np.float32)
    t = (torch.randn(3, generator=g) * 0.2).numpy().astype(np.float32)
    w4 = np.zeros((Cout, 7, 7, 4), dtype=np.float32)
    w4[..., :3] = (torch.randn(Cout, 7, 7, 3, generator=g) * (2.0 / 147) ** 0.5).numpy()
    bias = (torch.randn(Cout, generator=g) * 0.5).numpy().astype(np.float32)
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    want = torch.full((B, Ho, Wo, Cout), float("nan"), dtype=tdt, device="cuda")
    got = torch.full((B, Ho, Wo, Cout), float("nan"), dtype=tdt, device="cuda")
    torch.cuda.synchronize()
    rc = lib.b200rn_stem_conv(dtype, x.data_ptr(), B, H, W, pkg.fptr(s), pkg.fptr(t), pkg.fptr(w4), pkg.fptr(bias),
                              Cout, 1, want.data_ptr())
    assert rc == 0, lib.b200rn_global_error().decode()
    err = C.create_string_buffer(512)
    rc = lib.b200rn_stem_conv_static(dtype, x.data_ptr(), B, H, W, pkg.fptr(s), pkg.fptr(t), pkg.fptr(w4),
                                     pkg.fptr(bias), Cout, 1, got.data_ptr(), err, 512)
    assert rc == 0, err.value.decode()
    return want.float().cpu(), got.float().cpu()


@pytest.mark.parametrize("dtype", [2, 1], ids=["fp16", "bf16"])
def test_static_scratch_stem_matches_allocating_entry_point(dtype):
    rtol, atol = gh.TOL[dtype]
    for (B, H, W, Cout, seed) in [(3, 96, 96, 64, 0), (1, 32, 48, 128, 1), (3, 96, 96, 64, 2)]:
        want, got = _run_both(dtype, B, H, W, Cout, seed)
        assert torch.isfinite(got).all()
        diff = (got - want).abs()
        assert bool((diff <= atol + rtol * want.abs()).all()), "max diff %g at %s" % (float(diff.max()), (B, H, W, Cout))


def test_static_scratch_stem_refuses_what_does_not_fit():
    x = torch.zeros(64, 3, 224, 224, device="cuda")
    out = torch.zeros(64, 112, 112, 64, dtype=torch.float16, device="cuda")
    w4 = np.zeros((64, 7, 7, 4), dtype=np.float32)
    err = C.create_string_buffer(512)
    rc = lib.b200rn_stem_conv_static(2, x.data_ptr(), 64, 224, 224, None, None, pkg.fptr(w4), None, 64, 1,
                                     out.data_ptr(), err, 512)
    assert rc != 0 and b"static array holds" in err.value
    rc = lib.b200rn_stem_conv_static(2, x.data_ptr(), 1, 224, 224, None, None, pkg.fptr(w4), None, 256, 1,
                                     out.data_ptr(), err, 512)
    assert rc != 0 and b"at most" in err.value
