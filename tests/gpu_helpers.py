"""Shared helpers for the GPU parity tests: torch provides device memory; the CPU oracle (torch-CPU
float64 conv on the same rounded operands) provides the expected values."""
import ctypes as C

import numpy as np
import torch
import torch.nn.functional as F

import b200rn_loader

pkg = b200rn_loader.load_package()
lib = pkg.lib

TORCH_DT = {0: torch.float32, 1: torch.bfloat16, 2: torch.float16}
# relative / absolute tolerance on conv outputs of O(1): storage rounding of the output dominates
TOL = {0: (2e-5, 2e-5), 1: (1.2e-2, 1.2e-2), 2: (1.5e-3, 1.5e-3)}


def conv_case(dtype, B, H, W, Cin, Cout, k, stride, pad, relu=0, residual=False, want_out=True, want_act=False,
              act_relu=1, seed=0, prologue=False):
    """Runs b200rn_conv2d on random data; returns dict(err_out, err_act, ...) vs a float64 CPU conv."""
    g = torch.Generator().manual_seed(seed)
    tdt = TORCH_DT[dtype]
    x = torch.randn(B, Cin, H, W, generator=g).to(tdt)
    w = (torch.randn(Cout, Cin, k, k, generator=g) * (2.0 / (Cin * k * k)) ** 0.5).to(tdt)
    bias = torch.randn(Cout, generator=g) * 0.5
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    res = torch.randn(B, Cout, Ho, Wo, generator=g).to(tdt) if residual else None
    s2 = torch.rand(Cout, generator=g) + 0.5
    t2 = torch.randn(Cout, generator=g) * 0.3
    ps = torch.rand(Cin, generator=g) + 0.5
    pt = torch.randn(Cin, generator=g) * 0.3
    xin = x.double()
    if prologue:
        # what the kernel feeds the MMA: relu(ps*x+pt) rounded back to the storage type
        xin = (x.float() * ps.view(1, -1, 1, 1) + pt.view(1, -1, 1, 1)).clamp_min(0).to(tdt).double()
    # expected (float64 on the rounded operands)
    y = F.conv2d(xin, w.double(), bias.double(), stride=stride, padding=pad)
    if relu:
        y = y.clamp_min(0)
    if residual:
        y = y + res.double()
    ya = y * s2.double().view(1, -1, 1, 1) + t2.double().view(1, -1, 1, 1)
    if act_relu:
        ya = ya.clamp_min(0)
    # device operands, NHWC
    d_x = x.permute(0, 2, 3, 1).contiguous().cuda()
    d_res = res.permute(0, 2, 3, 1).contiguous().cuda() if residual else None
    d_out = torch.full((B, Ho, Wo, Cout), float("nan"), dtype=tdt, device="cuda") if want_out else None
    d_act = torch.full((B, Ho, Wo, Cout), float("nan"), dtype=tdt, device="cuda") if want_act else None
    w_khwc = np.ascontiguousarray(w.float().permute(0, 2, 3, 1).numpy())   # [Cout][kh][kw][Cin]
    b_np = bias.numpy().astype(np.float32)
    s_np, t_np = s2.numpy().astype(np.float32), t2.numpy().astype(np.float32)
    ps_np, pt_np = ps.numpy().astype(np.float32), pt.numpy().astype(np.float32)
    torch.cuda.synchronize()
    rc = lib.b200rn_conv2d(dtype, d_x.data_ptr(), B, H, W, Cin, pkg.fptr(w_khwc), pkg.fptr(b_np), Cout, k, k, stride,
                           pad, relu, d_res.data_ptr() if residual else None,
                           d_out.data_ptr() if want_out else None, d_act.data_ptr() if want_act else None,
                           pkg.fptr(s_np), pkg.fptr(t_np), act_relu,
                           pkg.fptr(ps_np) if prologue else None, pkg.fptr(pt_np) if prologue else None, 1)
    out = {"rc": rc, "msg": lib.b200rn_global_error().decode() if rc else ""}
    if rc:
        return out
    rtol, atol = TOL[dtype]

    def cmp(dev, ref):
        got = dev.float().cpu().permute(0, 3, 1, 2).double()
        bad = ~torch.isfinite(got)
        err = (got - ref).abs()
        lim = atol + rtol * ref.abs()
        return {"max_err": float(err[~bad].max()) if (~bad).any() else float("nan"), "nonfinite": int(bad.sum()),
                "viol": int((err > lim).sum() + bad.sum()), "n": got.numel(),
                "worst_ratio": float((err / lim)[~bad].max()) if (~bad).any() else float("inf")}
    if want_out:
        out["out"] = cmp(d_out, y)
    if want_act:
        out["act"] = cmp(d_act, ya)
    return out


def tail_case(dtype, B, H, W, tail_cout=64, relu=0, act_relu=1, tail_relu=1, seed=0, Cin=64):
    """Runs b200rn_conv1x1_tail (1x1 conv to 256 ch + residual, then the next 1x1 conv on relu(bn(.))) on random
    data; compares both outputs with a float64 CPU evaluation on the same rounded operands."""
    g = torch.Generator().manual_seed(seed)
    tdt = TORCH_DT[dtype]
    Cout = 256
    x = torch.randn(B, Cin, H, W, generator=g).to(tdt)
    w = (torch.randn(Cout, Cin, 1, 1, generator=g) * (2.0 / Cin) ** 0.5).to(tdt)
    bias = torch.randn(Cout, generator=g) * 0.5
    res = torch.randn(B, Cout, H, W, generator=g).to(tdt)
    s2 = torch.rand(Cout, generator=g) + 0.5
    t2 = torch.randn(Cout, generator=g) * 0.3
    w1 = (torch.randn(tail_cout, Cout, 1, 1, generator=g) * (2.0 / Cout) ** 0.5).to(tdt)
    bias1 = torch.randn(tail_cout, generator=g) * 0.5
    y = F.conv2d(x.double(), w.double(), bias.double())
    if relu:
        y = y.clamp_min(0)
    y = y + res.double()
    a = y * s2.double().view(1, -1, 1, 1) + t2.double().view(1, -1, 1, 1)
    if act_relu:
        a = a.clamp_min(0)
    a = a.to(tdt).double()                     # the activated tile is held in the 16-bit storage type
    u = F.conv2d(a, w1.double(), bias1.double())
    if tail_relu:
        u = u.clamp_min(0)
    d_x = x.permute(0, 2, 3, 1).contiguous().cuda()
    d_res = res.permute(0, 2, 3, 1).contiguous().cuda()
    d_out = torch.full((B, H, W, Cout), float("nan"), dtype=tdt, device="cuda")
    d_tail = torch.full((B, H, W, tail_cout), float("nan"), dtype=tdt, device="cuda")
    f32 = lambda t: np.ascontiguousarray(t.float().numpy().astype(np.float32))
    w_np, w1_np = f32(w.view(Cout, Cin)), f32(w1.view(tail_cout, Cout))
    b_np, b1_np, s_np, t_np = f32(bias), f32(bias1), f32(s2), f32(t2)
    torch.cuda.synchronize()
    rc = lib.b200rn_conv1x1_tail(dtype, d_x.data_ptr(), B, H, W, Cin, pkg.fptr(w_np), pkg.fptr(b_np), relu, d_res.data_ptr(),
                                 d_out.data_ptr(), pkg.fptr(s_np), pkg.fptr(t_np), act_relu, pkg.fptr(w1_np), pkg.fptr(b1_np),
                                 tail_cout, tail_relu, d_tail.data_ptr())
    out = {"rc": rc, "msg": lib.b200rn_global_error().decode() if rc else ""}
    if rc:
        return out
    rtol, atol = TOL[dtype]

    def cmp(dev, ref):
        got = dev.float().cpu().permute(0, 3, 1, 2).double()
        bad = ~torch.isfinite(got)
        err = (got - ref).abs()
        lim = atol + rtol * ref.abs()
        return {"max_err": float(err[~bad].max()) if (~bad).any() else float("nan"), "nonfinite": int(bad.sum()),
                "viol": int((err > lim).sum() + bad.sum()), "n": got.numel()}
    out["out"] = cmp(d_out, y)
    out["tail"] = cmp(d_tail, u)
    return out


def two_input_case(dtype, B, H, W, Cin, Cin_b, stride_b, Cout=256, Hb=None, Wb=None, want_out=True, want_act=True, seed=0):
    """Runs b200rn_conv1x1_two_input (main 1x1 conv + strided shortcut 1x1 conv summed in one GEMM) on random data;
    compares with a float64 CPU evaluation of the two convs on the same rounded operands."""
    g = torch.Generator().manual_seed(seed)
    tdt = TORCH_DT[dtype]
    Hb = Hb if Hb is not None else (H - 1) * stride_b + 1
    Wb = Wb if Wb is not None else (W - 1) * stride_b + 1
    x = torch.randn(B, Cin, H, W, generator=g).to(tdt)
    xb = torch.randn(B, Cin_b, Hb, Wb, generator=g).to(tdt)
    w1 = (torch.randn(Cout, Cin, 1, 1, generator=g) * (1.0 / Cin) ** 0.5).to(tdt)
    w2 = (torch.randn(Cout, Cin_b, 1, 1, generator=g) * (1.0 / Cin_b) ** 0.5).to(tdt)
    bias = torch.randn(Cout, generator=g) * 0.5
    s2 = torch.rand(Cout, generator=g) + 0.5
    t2 = torch.randn(Cout, generator=g) * 0.3
    y = F.conv2d(x.double(), w1.double()) + F.conv2d(xb.double(), w2.double(), stride=stride_b) + bias.double().view(1, -1, 1, 1)
    ya = (y * s2.double().view(1, -1, 1, 1) + t2.double().view(1, -1, 1, 1)).clamp_min(0)
    d_x = x.permute(0, 2, 3, 1).contiguous().cuda()
    d_xb = xb.permute(0, 2, 3, 1).contiguous().cuda()
    d_out = torch.full((B, H, W, Cout), float("nan"), dtype=tdt, device="cuda") if want_out else None
    d_act = torch.full((B, H, W, Cout), float("nan"), dtype=tdt, device="cuda") if want_act else None
    f32 = lambda t: np.ascontiguousarray(t.float().numpy().astype(np.float32))
    wcat = np.ascontiguousarray(np.concatenate([f32(w1.view(Cout, Cin)), f32(w2.view(Cout, Cin_b))], axis=1))
    torch.cuda.synchronize()
    rc = lib.b200rn_conv1x1_two_input(dtype, d_x.data_ptr(), B, H, W, Cin, d_xb.data_ptr(), Hb, Wb, Cin_b, stride_b, pkg.fptr(wcat),
                                      pkg.fptr(f32(bias)), Cout, d_out.data_ptr() if want_out else None,
                                      d_act.data_ptr() if want_act else None, pkg.fptr(f32(s2)), pkg.fptr(f32(t2)), 1)
    out = {"rc": rc, "msg": lib.b200rn_global_error().decode() if rc else ""}
    if rc:
        return out
    rtol, atol = TOL[dtype]

    def cmp(dev, ref):
        got = dev.float().cpu().permute(0, 3, 1, 2).double()
        bad = ~torch.isfinite(got)
        err = (got - ref).abs()
        lim = atol + rtol * ref.abs()
        return {"max_err": float(err[~bad].max()) if (~bad).any() else float("nan"), "nonfinite": int(bad.sum()),
                "viol": int((err > lim).sum() + bad.sum()), "n": got.numel()}
    if want_out:
        out["out"] = cmp(d_out, y)
    if want_act:
        out["act"] = cmp(d_act, ya)
    return out


def stem_case(dtype, B, H, W, Cout=64, seed=0):
    g = torch.Generator().manual_seed(seed)
    tdt = TORCH_DT[dtype]
    x = torch.randn(B, 3, H, W, generator=g)
    s = torch.rand(3, generator=g) + 0.5
    t = torch.randn(3, generator=g) * 0.2
    w = (torch.randn(Cout, 3, 7, 7, generator=g) * (2.0 / 147) ** 0.5).to(tdt)
    bias = torch.randn(Cout, generator=g) * 0.5
    xa = (x * s.view(1, 3, 1, 1) + t.view(1, 3, 1, 1)).to(tdt)      # what the re-layout kernel stores
    y = F.conv2d(xa.double(), w.double(), bias.double(), stride=2, padding=3).clamp_min(0)
    Ho, Wo = y.shape[2], y.shape[3]
    d_x = x.contiguous().cuda()
    d_out = torch.full((B, Ho, Wo, Cout), float("nan"), dtype=tdt, device="cuda")
    w4 = np.zeros((Cout, 7, 7, 4), dtype=np.float32)
    w4[..., :3] = w.float().permute(0, 2, 3, 1).numpy()
    b_np = bias.numpy().astype(np.float32)
    s_np, t_np = s.numpy().astype(np.float32), t.numpy().astype(np.float32)
    torch.cuda.synchronize()
    rc = lib.b200rn_stem_conv(dtype, d_x.data_ptr(), B, H, W, pkg.fptr(s_np), pkg.fptr(t_np), pkg.fptr(w4),
                              pkg.fptr(b_np), Cout, 1, d_out.data_ptr())
    out = {"rc": rc, "msg": lib.b200rn_global_error().decode() if rc else ""}
    if rc:
        return out
    rtol, atol = TOL[dtype]
    got = d_out.float().cpu().permute(0, 3, 1, 2).double()
    bad = ~torch.isfinite(got)
    err = (got - y).abs()
    lim = 2 * atol + 2 * rtol * y.abs()     # the fused input affine is rounded once more than the reference
    out["out"] = {"max_err": float(err[~bad].max()) if (~bad).any() else float("nan"), "nonfinite": int(bad.sum()),
                  "viol": int((err > lim).sum() + bad.sum()), "n": got.numel(),
                  "worst_ratio": float((err / lim)[~bad].max()) if (~bad).any() else float("inf")}
    return out


CONV_CASES = [
    # name, kwargs
    ("1x1 s1 64->64 @56 B2", dict(B=2, H=56, W=56, Cin=64, Cout=64, k=1, stride=1, pad=0)),
    ("1x1 s1 256->128 @14 B3", dict(B=3, H=14, W=14, Cin=256, Cout=128, k=1, stride=1, pad=0, relu=1)),
    ("3x3 s1 64->64 @56 B2", dict(B=2, H=56, W=56, Cin=64, Cout=64, k=3, stride=1, pad=1, relu=1)),
    ("3x3 s1 128->128 @28 B5", dict(B=5, H=28, W=28, Cin=128, Cout=128, k=3, stride=1, pad=1, relu=1)),
    ("1x1 s2 256->512 @56 B2", dict(B=2, H=56, W=56, Cin=256, Cout=512, k=1, stride=2, pad=0)),
    ("3x3 s2 128->128 @56 B2", dict(B=2, H=56, W=56, Cin=128, Cout=128, k=3, stride=2, pad=1, relu=1)),
    ("3x3 s1 512->512 @7 B4", dict(B=4, H=7, W=7, Cin=512, Cout=512, k=3, stride=1, pad=1, relu=1)),
    ("1x1 64->256 @14 res+dual", dict(B=3, H=14, W=14, Cin=64, Cout=256, k=1, stride=1, pad=0, residual=True,
                                      want_out=True, want_act=True)),
    ("1x1 512->2048 @7 res act-only", dict(B=2, H=7, W=7, Cin=512, Cout=2048, k=1, stride=1, pad=0, residual=True,
                                           want_out=False, want_act=True)),
    ("3x3 s2 512->512 @14 B33", dict(B=33, H=14, W=14, Cin=512, Cout=512, k=3, stride=2, pad=1, relu=1)),
    ("1x1 256->64 @56 prologue", dict(B=2, H=56, W=56, Cin=256, Cout=64, k=1, stride=1, pad=0, relu=1, prologue=True)),
    ("1x1 s2 512->1024 @28 prologue", dict(B=3, H=28, W=28, Cin=512, Cout=1024, k=1, stride=2, pad=0, prologue=True)),
    ("1x1 1024->256 @14 prologue B40", dict(B=40, H=14, W=14, Cin=1024, Cout=256, k=1, stride=1, pad=0, relu=1, prologue=True)),
    ("3x3 s1 64->128 @30x17 B3 (halo, ragged)", dict(B=3, H=30, W=17, Cin=64, Cout=128, k=3, stride=1, pad=1)),
    ("3x3 s1 128->64 @14 B7 (halo)", dict(B=7, H=14, W=14, Cin=128, Cout=64, k=3, stride=1, pad=1, relu=1)),
    ("3x3 s1 64->64 @56 B9 (halo, many tiles)", dict(B=9, H=56, W=56, Cin=64, Cout=64, k=3, stride=1, pad=1, relu=1)),
    ("1x1 512->256 @7 B1024 (dual-M)", dict(B=1024, H=7, W=7, Cin=512, Cout=256, k=1, stride=1, pad=0, relu=1)),
    ("3x3 64->256 @7 B1000 (dual-M, odd tile count)", dict(B=1000, H=7, W=7, Cin=64, Cout=256, k=3, stride=1, pad=1, relu=1)),
    ("1x1 s2 1024->512 @14 B300 (dual-M, no relu)", dict(B=300, H=14, W=14, Cin=1024, Cout=512, k=1, stride=2, pad=0)),
    ("3x3 s1 256->256 @14 B64 (N=256 tiles)", dict(B=64, H=14, W=14, Cin=256, Cout=256, k=3, stride=1, pad=1, relu=1)),
    # wide-epilogue kernel (16-bit modes): every combination of residual / raw / activated outputs, one and many N
    # tiles, fewer M tiles than CTAs per N tile, ragged patches, more tiles than a CTA's rings hold
    ("wide 64->256 @56 B8 res raw+act", dict(B=8, H=56, W=56, Cin=64, Cout=256, k=1, stride=1, pad=0, residual=True,
                                             want_out=True, want_act=True)),
    ("wide 64->256 @56 B8 shortcut", dict(B=8, H=56, W=56, Cin=64, Cout=256, k=1, stride=1, pad=0)),
    ("wide 256->1024 @14 B40 res raw+act", dict(B=40, H=14, W=14, Cin=256, Cout=1024, k=1, stride=1, pad=0, residual=True,
                                                want_out=True, want_act=True)),
    ("wide 256->1024 @14 B40 res act-only no-relu", dict(B=40, H=14, W=14, Cin=256, Cout=1024, k=1, stride=1, pad=0,
                                                         residual=True, want_out=False, want_act=True, act_relu=0)),
    ("wide 128->512 @28 B20 act-only no res", dict(B=20, H=28, W=28, Cin=128, Cout=512, k=1, stride=1, pad=0, relu=1,
                                                   want_out=False, want_act=True)),
    ("wide 512->2048 @7 B70 res raw+act (8 N tiles)", dict(B=70, H=7, W=7, Cin=512, Cout=2048, k=1, stride=1, pad=0,
                                                           residual=True, want_out=True, want_act=True)),
    ("wide s2 512->1024 @28 B9 shortcut", dict(B=9, H=28, W=28, Cin=512, Cout=1024, k=1, stride=2, pad=0)),
    ("wide 64->256 @13x9 B3 ragged res raw", dict(B=3, H=13, W=9, Cin=64, Cout=256, k=1, stride=1, pad=0, residual=True)),
    # dual-M halo kernel (weight slab streamed: two M tiles per weight tile): an odd number of M tiles (the item's second
    # tile lies outside the batch), more items than SMs x 2 (both accumulator pairs reused)
    ("halo2 3x3 128->128 @24x14 B1 (3 M tiles)", dict(B=1, H=24, W=14, Cin=128, Cout=128, k=3, stride=1, pad=1, relu=1)),
    ("halo2 3x3 128->128 @28 B80 (640 M tiles)", dict(B=80, H=28, W=28, Cin=128, Cout=128, k=3, stride=1, pad=1, relu=1)),
    # CTA-pair kernel (tcgen05.mma.cta_group::2, one 256 x N tile per 2-SM cluster): N = 128 halves, an odd number of
    # M tiles (the pair's second tile lies wholly outside the batch), one pair only, more items than clusters x 2
    # (both accumulators of the double buffer reused), strided 3x3 taps through the pair's TMA loads
    ("pair 1x1 512->128 @28 B20 (N=128)", dict(B=20, H=28, W=28, Cin=512, Cout=128, k=1, stride=1, pad=0, relu=1)),
    ("pair 3x3 256->256 @14 B9 (odd M tiles)", dict(B=9, H=14, W=14, Cin=256, Cout=256, k=3, stride=1, pad=1, relu=1)),
    ("pair 1x1 1024->256 @14 B1 (one pair)", dict(B=1, H=14, W=14, Cin=1024, Cout=256, k=1, stride=1, pad=0)),
    ("pair 1x1 1024->256 @14 B256 (196 items)", dict(B=256, H=14, W=14, Cin=1024, Cout=256, k=1, stride=1, pad=0, relu=1)),
    ("pair 3x3 s2 256->256 @28 B70", dict(B=70, H=28, W=28, Cin=256, Cout=256, k=3, stride=2, pad=1, relu=1)),
    ("pair 1x1 2048->512 @7 B200 (two N tiles)", dict(B=200, H=7, W=7, Cin=2048, Cout=512, k=1, stride=1, pad=0, relu=1)),
    # CTA-pair kernel with the wide epilogue (conv_pairwide_kernel: K >= 512, streamed weights, residual): many tiles per
    # cluster (rings and both accumulators reused), an odd number of M tiles, act-only / raw-only outputs, K = 1024
    ("pairwide 512->2048 @7 B256 res raw+act (392 items)", dict(B=256, H=7, W=7, Cin=512, Cout=2048, k=1, stride=1, pad=0,
                                                                residual=True, want_out=True, want_act=True)),
    ("pairwide 512->1024 @14 B3 res raw (odd M tiles)", dict(B=3, H=14, W=14, Cin=512, Cout=1024, k=1, stride=1, pad=0,
                                                             residual=True)),
    ("pairwide 1024->512 @14 B40 res act-only no-relu", dict(B=40, H=14, W=14, Cin=1024, Cout=512, k=1, stride=1, pad=0,
                                                             residual=True, want_out=False, want_act=True, act_relu=0)),
    ("pairwide 512->256 @13x9 B5 ragged res raw+act relu", dict(B=5, H=13, W=9, Cin=512, Cout=256, k=1, stride=1, pad=0, relu=1,
                                                                residual=True, want_out=True, want_act=True)),
    ("pairwide 256->1024 @14 B256 res raw+act (stationary half slabs, 784 items)",
     dict(B=256, H=14, W=14, Cin=256, Cout=1024, k=1, stride=1, pad=0, residual=True, want_out=True, want_act=True)),
    ("pairwide 256->256 @9x5 B3 res raw (stationary, one pair, ragged)", dict(B=3, H=9, W=5, Cin=256, Cout=256, k=1, stride=1, pad=0,
                                                                              residual=True)),
    # large-K plain 1x1 conv with at least two waves of tiles: the wide kernel's four-stage configuration
    ("deep 512->256 @28 B50 (306 tiles)", dict(B=50, H=28, W=28, Cin=512, Cout=256, k=1, stride=1, pad=0, relu=1)),
]


# exact mode only (fp32 storage, 3 x TF32 on the tensor cores): channel counts the 16-bit kernels do not take (32-multiples),
# the 64-wide N tile with one chunk per epilogue group, K that is not a whole number of 4-block accumulation chunks,
# residual + both outputs on a ragged patch, a strided 3x3, and a shape the tensor-core form refuses (falls back to CUDA cores)
FP32_CASES = [
    ("t32 1x1 96->64 @28 B3 (3 K blocks)", dict(B=3, H=28, W=28, Cin=96, Cout=64, k=1, stride=1, pad=0, relu=1)),
    ("t32 3x3 32->192 @14 B5 (9 K blocks, N=64 x3)", dict(B=5, H=14, W=14, Cin=32, Cout=192, k=3, stride=1, pad=1, relu=1)),
    ("t32 1x1 160->128 @13x9 B2 res raw+act", dict(B=2, H=13, W=9, Cin=160, Cout=128, k=1, stride=1, pad=0, residual=True,
                                                   want_out=True, want_act=True)),
    ("t32 3x3 s2 96->128 @30x17 B3 act-only", dict(B=3, H=30, W=17, Cin=96, Cout=128, k=3, stride=2, pad=1, relu=1,
                                                   want_out=False, want_act=True)),
    ("t32 1x1 2048->64 @7 B40 (64 K blocks, 16 chunks)", dict(B=40, H=7, W=7, Cin=2048, Cout=64, k=1, stride=1, pad=0)),
    ("fallback 1x1 48->40 @14 B2 (CUDA cores)", dict(B=2, H=14, W=14, Cin=48, Cout=40, k=1, stride=1, pad=0, relu=1)),
]
