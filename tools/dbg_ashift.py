#!/usr/bin/env python3
"""Descriptor experiment: does a K-major SWIZZLE_128B A operand whose start address is shifted by whole
128-byte rows (not 1024-aligned) read the rows one would expect?  One 128-pixel tile (B=1, H=8, W=16),
1x1 conv 64->64 with identity weights, so out[r] must equal in[r+shift] bit for bit.
Measured on B200: yes for every shift, with the descriptor's base_offset field left at 0 (setting it to
(addr >> 7) & 7 scrambles rows) -- the swizzle is applied to absolute shared-memory addresses.  The 3x3 halo
kernel (csrc/conv_umma.cu) is built on this."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200rn_loader  # noqa: E402

pkg = b200rn_loader.load_package()
lib = pkg.lib
for shift in (0, 1, 2, 3, 9, 17):
    for baseoff in (0,):
        os.environ["B200RN_DBG_ASHIFT"] = str(shift)
        g = torch.Generator().manual_seed(shift)
        x = torch.randn(1, 8, 16, 64, generator=g)
        x[..., 0] = torch.arange(128, dtype=torch.float32).view(1, 8, 16)                     # channel 0 = pixel index
        d_x = x.to(torch.float16).cuda().contiguous()
        w = np.zeros((64, 1, 1, 64), dtype=np.float32)
        for o in range(64):
            w[o, 0, 0, o] = 1.0
        out = torch.full((1, 8, 16, 64), float("nan"), dtype=torch.float16, device="cuda")
        rc = lib.b200rn_conv2d(2, d_x.data_ptr(), 1, 8, 16, 64, pkg.fptr(w), None, 64, 1, 1, 1, 0, 0, None, out.data_ptr(),
                               None, None, None, 0, None, None, 0)
        o = out.float().cpu().view(128, 64)
        rows = o[:, 0].round().int().tolist()
        xin = d_x.float().cpu().view(128, 64)
        exact = bool(torch.equal(o[:128 - shift], xin[shift:]))
        expect = [r + shift for r in range(128 - shift)]
        print("shift %d rc %d: rows as expected: %s; all 64 channels bit-exact: %s; first rows %s" %
              (shift, rc, rows[:128 - shift] == expect, exact, rows[:8]))
