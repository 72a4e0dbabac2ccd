#!/usr/bin/env python3
"""Runs every conv / stem case of tests/gpu_helpers.py in all three dtypes and prints one line per case.
Unlike pytest -x it never stops early: one GPU call shows everything that is wrong."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_helpers as gh  # noqa: E402

names = {0: "fp32", 1: "bf16", 2: "fp16"}
only = [int(a) for a in sys.argv[1:]] or [0, 1, 2]
for dt in only:
    for name, kw in gh.CONV_CASES:
        t = time.time()
        r = gh.conv_case(dt, **kw)
        print("%-5s %-32s rc=%d %s %s %s  (%.2fs)" % (names[dt], name, r["rc"], r.get("msg", ""), r.get("out", ""), r.get("act", ""),
                                                      time.time() - t), flush=True)
    if dt:
        for (B, H, W) in [(2, 224, 224), (3, 64, 64), (1, 32, 32)]:
            r = gh.stem_case(dt, B, H, W)
            print("%-5s stem B%d %dx%d rc=%d %s %s" % (names[dt], B, H, W, r["rc"], r.get("msg", ""), r.get("out", "")), flush=True)
