#!/usr/bin/env python3
"""Prints CTA 0's per-tile timeline of the 3x3 halo kernel (64->64 @56x56, batch 256): B200RN_DBG_TRACE=1."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200RN_DBG_TRACE"] = "1"
import b200rn_loader
pkg = b200rn_loader.load_package(); lib = pkg.lib
B, H, W, C = 256, 56, 56, 64
x = torch.randn(B, H, W, C, device="cuda").half()
out = torch.empty(B, H, W, C, device="cuda", dtype=torch.float16)
w = (np.random.RandomState(0).randn(C, 3, 3, C) * 0.05).astype(np.float32)
for _ in range(2):
    rc = lib.b200rn_conv2d(2, x.data_ptr(), B, H, W, C, pkg.fptr(w), None, C, 3, 3, 1, 1, 1, None, out.data_ptr(), None, None, None, 0, None, None, 0)
print("rc", rc)
