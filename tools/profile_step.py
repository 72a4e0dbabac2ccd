#!/usr/bin/env python3
"""Minimal driver for ncu: W warm-up + K timed forwards of one batch, inputs resident in HBM.
    python tools/profile_step.py --batch 256 --steps 2 --warmup 1 [--mode fp16]
Prints the CUDA-event time per step (a number printed under ncu is not a bench value)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import b200rn_loader  # noqa: E402
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--warmup", type=int, default=1)
ap.add_argument("--mode", default="fp16")
ap.add_argument("--layers-out", default="")
a = ap.parse_args()
pkg = b200rn_loader.load_package()
eng = pkg.Engine(bench.model_path(), mode=a.mode, max_batch=a.batch)
x = [torch.randn(a.batch, 3, 224, 224, device="cuda") for _ in range(2)]
ptrs = [t.data_ptr() for t in x]
if a.warmup:
    eng.time_forward(ptrs, a.batch, a.warmup)
ms = eng.time_forward(ptrs, a.batch, a.steps)
print("batch %d mode %s: %.3f ms/step, %d launches/step" % (a.batch, a.mode, ms / a.steps, eng.launch_count()))
if a.layers_out:
    import json
    layers = eng.profile_layers(ptrs[0], a.batch, iters=5)
    for l in layers:
        l["tflops"] = l["flops"] / (l["ms"] * 1e-3) / 1e12 if l["ms"] > 0 else 0.0
        l["gbs"] = l["bytes"] / (l["ms"] * 1e-3) / 1e9 if l["ms"] > 0 else 0.0
    json.dump({"batch": a.batch, "mode": a.mode, "layers": layers}, open(a.layers_out, "w"), indent=1)
eng.close()
