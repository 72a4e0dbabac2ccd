#!/usr/bin/env python3
"""Prints CTA 0's per-tile timeline of the generic conv kernel on the stem (batch 256): B200RN_DBG_TRACE=1."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200RN_DBG_TRACE"] = "1"
import b200rn_loader
pkg = b200rn_loader.load_package(); lib = pkg.lib
B = 256
x = torch.randn(B, 3, 224, 224, device="cuda")
out = torch.empty(B, 112, 112, 64, device="cuda", dtype=torch.float16)
w = (np.random.RandomState(0).randn(64, 7, 7, 4) * 0.05).astype(np.float32)
for _ in range(2):
    rc = lib.b200rn_stem_conv(2, x.data_ptr(), B, 224, 224, None, None, pkg.fptr(w), None, 64, 1, out.data_ptr())
print("rc", rc)
