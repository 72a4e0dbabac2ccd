#!/usr/bin/env python3
"""Generates tests/golden/resnet50_parity258_seed0.npz: what the CPU oracle says about the 258-sample network
parity set (tests/parity_set.py) on the deterministic synthetic ResNet-50 v2 fixture
(tools/make_resnet50v2_onnx.py, seed 0, calibrated classifier -- the model bench.py times).

Stored (small on purpose; the GPU tests recompute the fp32 oracle on the box and check it against this file):
  top1_fp32 / top1_fp64     [258]  arg-max class of the fp32 and the fp64 oracle run
  margin_fp64               [258]  top-1 minus top-2 logit of the fp64 run
  fp32_vs_fp64_maxabs       [258]  per-sample max |fp32 - fp64| logit difference
  logits_fp32 / logits_fp64 [18, 1000] full rows for samples 0..15, 256 (zeros) and 257 ([1,2,3] padded)
  rows                      [18]   which samples those are

Run from the repo root (a few minutes of CPU):  python tests/golden/make_parity258_golden.py"""
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity_set  # noqa: E402
from oracle import onnx_mini, resnet_oracle  # noqa: E402
from tools import make_resnet50v2_onnx as gen  # noqa: E402

with tempfile.TemporaryDirectory() as td:
    p = os.path.join(td, "r50.onnx")
    gen.write_model(p, "resnet50", seed=0)
    m = onnx_mini.load(p)
    x = torch.from_numpy(parity_set.padded())
    y32, y64 = [], []
    with torch.no_grad():
        for i in range(0, len(x), 32):
            y32.append(resnet_oracle.run_graph(m, x[i:i + 32], torch.float32).numpy())
            y64.append(resnet_oracle.run_graph(m, x[i:i + 32], torch.float64).numpy())
            print("oracle: %d / %d" % (min(i + 32, len(x)), len(x)), flush=True)
y32 = np.concatenate(y32).astype(np.float32)
y64 = np.concatenate(y64).astype(np.float64)
srt = np.sort(y64, axis=1)
rows = np.asarray(list(range(16)) + [256, 257])
out = os.path.join(ROOT, "tests", "golden", "resnet50_parity258_seed0.npz")
np.savez_compressed(out, top1_fp32=y32.argmax(1), top1_fp64=y64.argmax(1), margin_fp64=srt[:, -1] - srt[:, -2],
                    fp32_vs_fp64_maxabs=np.abs(y32 - y64).max(1), rows=rows, logits_fp32=y32[rows], logits_fp64=y64[rows])
print("wrote", out)
print("fp32 vs fp64: max-abs %.3e, top-1 agreement %d/%d" % (np.abs(y32 - y64).max(), (y32.argmax(1) == y64.argmax(1)).sum(), len(y32)))
mg = srt[:, -1] - srt[:, -2]
print("fp64 top-1/top-2 margin: min %.4g, 5%% %.4g, median %.4g; distinct top-1 classes %d; logit std %.3f" %
      (mg.min(), np.percentile(mg, 5), np.median(mg), len(set(y64.argmax(1))), y64.std()))
print("margins below 0.2:", np.sort(mg[mg < 0.2]))
