#!/usr/bin/env python3
"""Prints CTA 0's per-tile timeline of one conv through b200rn_conv2d (B200RN_DBG_TRACE=1).
   usage: dbg_conv_trace.py B H W Cin Cout k stride pad"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200RN_DBG_TRACE"] = "1"
import b200rn_loader
pkg = b200rn_loader.load_package(); lib = pkg.lib
B, H, W, Cin, Cout, k, stride, pad = [int(a) for a in sys.argv[1:9]]
Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
x = torch.randn(B, H, W, Cin, device="cuda").half()
out = torch.empty(B, Ho, Wo, Cout, device="cuda", dtype=torch.float16)
w = (np.random.RandomState(0).randn(Cout, k, k, Cin) * 0.05).astype(np.float32)
for _ in range(2):
    rc = lib.b200rn_conv2d(2, x.data_ptr(), B, H, W, Cin, pkg.fptr(w), None, Cout, k, k, stride, pad, 1, None, out.data_ptr(), None, None, None, 0, None, None, 0)
print("rc", rc)
