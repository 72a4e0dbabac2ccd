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
    if name.startswith("pair"):
        monkeypatch.setenv("B200RN_PAIR", "2")      # also Cout = 128 layers go through conv_pair_kernel (N = 128 tile)
    if "stationary" in name:
        monkeypatch.setenv("B200RN_PAIRWIDE_BSTAT", "1")   # opt-in form of conv_pairwide_kernel: half slabs kept in each CTA
    r = gh.conv_case(dtype, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


ONE_BY_ONE_CASES = [c for c in gh.CONV_CASES if c[1]["k"] == 1 and c[1]["stride"] == 1 and c[1]["pad"] == 0]


@pytest.mark.parametrize("dtype", [1, 2], ids=["bf16", "fp16"])
@pytest.mark.parametrize("name,kw", ONE_BY_ONE_CASES, ids=[c[0] for c in ONE_BY_ONE_CASES])
def test_conv1x1_with_spatial_patches(dtype, name, kw, monkeypatch):
    """1x1 / stride-1 convs are planned as one row of B*H*W pixels (an M tile = 128 consecutive rows); B200RN_FLATTEN=0
    keeps the spatial patch tiling every other conv uses.  Both must give the same answers."""
    monkeypatch.setenv("B200RN_FLATTEN", "0")
    if name.startswith("wide"):
        monkeypatch.setenv("B200RN_WIDE", "2")
    if name.startswith("pair"):
        monkeypatch.setenv("B200RN_PAIR", "2")
    r = gh.conv_case(dtype, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


def test_exact_mode_raw_tile_is_a_hi(monkeypatch):
    """kind::tf32 ignores the low 13 mantissa bits of its operands, so the 3 x TF32 kernel leaves the raw fp32 activation
    tile in place as A_hi and writes only A_lo.  With B200RN_TF32_RAW_HI=0 the transform warps write the masked hi as
    the first version did: both must give the same bits."""
    import numpy as np
    import torch
    g = torch.Generator().manual_seed(3)
    B, H, W, Cin, Cout = 5, 14, 14, 128, 128
    x = torch.randn(B, H, W, Cin, generator=g).cuda()
    res = torch.randn(B, H, W, Cout, generator=g).cuda()
    w = np.ascontiguousarray((torch.randn(Cout, 3, 3, Cin, generator=g) * 0.03).numpy())
    bias = torch.randn(Cout, generator=g).numpy().astype(np.float32)
    one, zero = np.ones(Cout, np.float32), np.zeros(Cout, np.float32)
    outs = []
    for raw in ("1", "0"):
        monkeypatch.setenv("B200RN_TF32_RAW_HI", raw)
        out = torch.full((B, H, W, Cout), float("nan"), device="cuda")
        rc = gh.lib.b200rn_conv2d(0, x.data_ptr(), B, H, W, Cin, gh.pkg.fptr(w), gh.pkg.fptr(bias), Cout, 3, 3, 1, 1, 0,
                                  res.data_ptr(), out.data_ptr(), None, gh.pkg.fptr(one), gh.pkg.fptr(zero), 1, None, None, 1)
        assert rc == 0, gh.lib.b200rn_global_error()
        torch.cuda.synchronize()
        outs.append(out.cpu())
    assert torch.isfinite(outs[0]).all()
    assert torch.equal(outs[0], outs[1])


LONG_K_CASES = [c for c in gh.CONV_CASES + gh.FP32_CASES
                if c[1]["k"] * c[1]["k"] * (c[1]["Cin"] // 32) >= 8 and c[1]["Cin"] % 32 == 0 and c[1]["Cout"] % 64 == 0]


@pytest.mark.parametrize("fused", ["0", "-8"], ids=["three-instructions", "fused-any-filter"])
@pytest.mark.parametrize("name,kw", LONG_K_CASES, ids=[c[0] for c in LONG_K_CASES])
def test_conv2d_exact_mode_hi_product_forms(name, kw, fused, monkeypatch):
    """3 x TF32 kernel: A_hi * W_hi and A_hi * W_lo as one MMA of width 2N (default for 3x3 convs with >= 16 K blocks) or
    as two.  B200RN_TF32_FUSED=0 forces the three-instruction form everywhere, -8 the fused one for every layer with
    at least 8 K blocks, 1x1 included."""
    monkeypatch.setenv("B200RN_TF32_FUSED", fused)
    r = gh.conv_case(0, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


RESIDUAL_CASES = [c for c in gh.CONV_CASES + gh.FP32_CASES if c[1].get("residual")]


@pytest.mark.parametrize("name,kw", RESIDUAL_CASES, ids=[c[0] for c in RESIDUAL_CASES])
def test_conv2d_exact_mode_residual_from_global(name, kw, monkeypatch):
    """Exact mode (3 x TF32) with B200RN_TF32_RES_TMA=0: the residual read by per-thread global loads instead of the TMA
    landing buffers that are the default."""
    monkeypatch.setenv("B200RN_TF32_RES_TMA", "0")
    r = gh.conv_case(0, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


PAIRWIDE_CASES = [c for c in gh.CONV_CASES if c[0].startswith(("pairwide", "wide 512->2048", "wide 256->1024"))]


@pytest.mark.parametrize("dtype", [1, 2], ids=["bf16", "fp16"])
@pytest.mark.parametrize("name,kw", PAIRWIDE_CASES, ids=[c[0] for c in PAIRWIDE_CASES])
def test_conv2d_large_k_residual_on_one_cta(dtype, name, kw, monkeypatch):
    """The same large-K residual shapes with the CTA-pair route off (B200RN_PAIRWIDE=0): conv_wide_kernel with two
    48 KB operand stages, the path these layers took before conv_pairwide_kernel."""
    monkeypatch.setenv("B200RN_PAIRWIDE", "0")
    r = gh.conv_case(dtype, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


@pytest.mark.parametrize("name,kw", gh.FP32_CASES, ids=[c[0] for c in gh.FP32_CASES])
def test_conv2d_exact_mode_shapes(name, kw):
    r = gh.conv_case(0, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


SIMT_CASES = [c for c in gh.CONV_CASES if c[0] in ("1x1 s1 64->64 @56 B2", "3x3 s1 128->128 @28 B5", "1x1 s2 256->512 @56 B2",
                                                   "1x1 64->256 @14 res+dual", "3x3 s2 512->512 @14 B33")]


@pytest.mark.parametrize("name,kw", SIMT_CASES, ids=[c[0] for c in SIMT_CASES])
def test_conv2d_fp32_cuda_core_kernel(name, kw, monkeypatch):
    """B200RN_EXACT_SIMT=1: the fp32 CUDA-core conv (round 1's exact path; the stem and the FC still run on it)."""
    monkeypatch.setenv("B200RN_EXACT_SIMT", "1")
    r = gh.conv_case(0, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


HALO_CASES = [c for c in gh.CONV_CASES if "halo" in c[0] or c[0] in ("3x3 s1 64->64 @56 B2", "3x3 s1 128->128 @28 B5")]


@pytest.mark.parametrize("dtype", [1, 2], ids=["bf16", "fp16"])
@pytest.mark.parametrize("name,kw", HALO_CASES, ids=[c[0] for c in HALO_CASES])
def test_conv3x3_single_m_halo_kernel(dtype, name, kw, monkeypatch):
    """B200RN_HALO2=0: the one-M-tile halo kernel also where the weight slab is streamed (the dual-M form's predecessor)."""
    monkeypatch.setenv("B200RN_HALO2", "0")
    r = gh.conv_case(dtype, **kw)
    assert r["rc"] == 0, r["msg"]
    assert r["out"]["viol"] == 0, (name, r["out"])


SINGLE_SM_CASES = [c for c in gh.CONV_CASES if "dual-M" in c[0] or c[0].startswith("pair")]


@pytest.mark.parametrize("dtype", [1, 2], ids=["bf16", "fp16"])
@pytest.mark.parametrize("name,kw", SINGLE_SM_CASES, ids=[c[0] for c in SINGLE_SM_CASES])
def test_conv2d_large_k_without_cta_pairs(dtype, name, kw, monkeypatch):
    """B200RN_PAIR=0: the same large-K shapes on the single-SM kernels (dual-M / deep / generic) the pair kernel replaced."""
    monkeypatch.setenv("B200RN_PAIR", "0")
    r = gh.conv_case(dtype, **kw)
    assert r["rc"] == 0, r["msg"]
    for key in ("out", "act"):
        if key in r:
            assert r[key]["viol"] == 0, (name, key, r[key])


@pytest.mark.parametrize("dtype", [0, 1, 2], ids=["fp32", "bf16", "fp16"])
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


TWO_INPUT_CASES = [
    # (B, H, W, Cin, Cin_b, stride_b, Cout): the four projection blocks of ResNet-50 at small batch, then ragged / odd
    # shapes: a patch that does not fill the tile, an even second input (last row / column unused), many N tiles
    (4, 56, 56, 64, 64, 1, 256), (6, 28, 28, 128, 256, 2, 512), (9, 14, 14, 256, 512, 2, 1024), (20, 7, 7, 512, 1024, 2, 2048),
    (3, 13, 9, 64, 128, 1, 256), (2, 5, 7, 128, 64, 2, 256), (1, 3, 3, 64, 64, 3, 512),
]


@pytest.mark.parametrize("dtype", [1, 2], ids=["bf16", "fp16"])
@pytest.mark.parametrize("case", TWO_INPUT_CASES, ids=["%dx%dx%d_%d+%ds%d_%d" % c for c in TWO_INPUT_CASES])
@pytest.mark.parametrize("pairwide", ["1", "0"], ids=["pair", "one-cta"])
def test_conv1x1_two_input(dtype, case, pairwide, monkeypatch):
    monkeypatch.setenv("B200RN_PAIRWIDE", pairwide)      # K >= 512 shapes: conv_pairwide_kernel or conv_wide_kernel
    B, H, W, Cin, Cin_b, sb, Cout = case
    r = gh.two_input_case(dtype, B, H, W, Cin, Cin_b, sb, Cout=Cout)
    assert r["rc"] == 0, r["msg"]
    assert r["out"]["viol"] == 0, r["out"]
    assert r["act"]["viol"] == 0, r["act"]


def test_conv1x1_two_input_even_second_input_and_bad_geometry():
    # Hb = 2*H: the shortcut's last row and column are never read (as 56 -> 28 in the network)
    r = gh.two_input_case(2, 2, 6, 6, 64, 64, 2, Hb=12, Wb=12, want_act=False)
    assert r["rc"] == 0 and r["out"]["viol"] == 0, r
    import numpy as np
    x = torch.zeros(1, 6, 6, 64, dtype=torch.float16, device="cuda")
    xb = torch.zeros(1, 10, 10, 64, dtype=torch.float16, device="cuda")     # 10 at stride 2 gives 5, not the 6x6 output grid
    o = torch.zeros(1, 6, 6, 256, dtype=torch.float16, device="cuda")
    w = np.zeros((256, 128), dtype=np.float32)
    rc = gh.lib.b200rn_conv1x1_two_input(2, x.data_ptr(), 1, 6, 6, 64, xb.data_ptr(), 10, 10, 64, 2, gh.pkg.fptr(w), None, 256,
                                         o.data_ptr(), None, None, None, 0)
    assert rc != 0 and b"second input" in gh.lib.b200rn_global_error()


def test_conv2d_rejects_bad_shapes():
    # the tensor-core path needs channel multiples of 64; it must say so rather than fall back
    x = torch.zeros(1, 8, 8, 32, dtype=torch.float16, device="cuda")
    o = torch.zeros(1, 8, 8, 64, dtype=torch.float16, device="cuda")
    import numpy as np
    w = np.zeros((64, 1, 1, 32), dtype=np.float32)
    rc = gh.lib.b200rn_conv2d(2, x.data_ptr(), 1, 8, 8, 32, gh.pkg.fptr(w), None, 64, 1, 1, 1, 0, 0, None,
                              o.data_ptr(), None, None, None, 0, None, None, 0)
    assert rc != 0 and b"% 64" in gh.lib.b200rn_global_error()
