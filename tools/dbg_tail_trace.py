#!/usr/bin/env python3
"""Prints CTA 0's per-chunk epilogue timeline of the fused 1x1+tail kernel (B200RN_DBG_TRACE=1).
   usage: dbg_tail_trace.py [B] [H] [W] [tail_cout]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200RN_DBG_TRACE"] = "1"
import gpu_helpers as gh
a = [int(v) for v in sys.argv[1:]] + [256, 56, 56, 64][len(sys.argv) - 1:]
for _ in range(2):
    r = gh.tail_case(2, a[0], a[1], a[2], tail_cout=a[3])
print({k: v for k, v in r.items() if k != "msg"})
