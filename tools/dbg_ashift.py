#!/usr/bin/env python3
"""Descriptor experiment: does a K-major SWIZZLE_128B A operand whose start address is shifted by whole
128-byte rows (not 1024-aligned) read the rows one would expect, with or without the base_offset field?
One 128-pixel tile (B=1, H=8, W=16), 1x1 conv 64->64, identity-like weights so that out[r] = in[r+shift]."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200rn_loader  # noqa: E402

pkg = b200rn_loader.load_package()
lib = pkg.lib
for shift in (0, 1, 2, 3, 9):
    for baseoff in (0, 1):
        os.environ["B200RN_DBG_ASHIFT"] = str(shift)
        os.environ["B200RN_DBG_BASEOFF"] = str(baseoff)
        x = torch.arange(128, dtype=torch.float32).view(1, 8, 16, 1).repeat(1, 1, 1, 64)      # pixel r holds value r
        x = x + torch.arange(64, dtype=torch.float32).view(1, 1, 1, 64) / 64.0                # + channel/64
        d_x = x.to(torch.float16).cuda().contiguous()
        w = np.zeros((64, 1, 1, 64), dtype=np.float32)
        for o in range(64):
            w[o, 0, 0, o] = 1.0
        out = torch.full((1, 8, 16, 64), float("nan"), dtype=torch.float16, device="cuda")
        rc = lib.b200rn_conv2d(2, d_x.data_ptr(), 1, 8, 16, 64, pkg.fptr(w), None, 64, 1, 1, 1, 0, 0, None, out.data_ptr(),
                               None, None, None, 0, None, None, 0)
        o = out.float().cpu().view(128, 64)
        rows = o[:, 0].round().int().tolist()
        frac_ok = bool(torch.allclose(o - o[:, :1], (torch.arange(64) / 64.0).view(1, 64).expand(128, 64), atol=2e-2))
        expect = [r + shift for r in range(128 - shift)]
        print("shift %d baseoff %d rc %d: rows match expected shift: %s; channel order intact: %s; first rows %s" %
              (shift, baseoff, rc, rows[:128 - shift] == expect, frac_ok, rows[:12]))
