#!/usr/bin/env python3
"""Generates tests/golden/resnet50_logits_seed0.npz: the CPU oracle's fp32 and fp64 logits for 8 parity
inputs on the deterministic synthetic ResNet-50 v2 fixture (tools/make_resnet50v2_onnx.py, seed 0).
Run from the repo root:  python tests/golden/make_logits_golden.py"""
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import onnx_mini, resnet_oracle  # noqa: E402
from tools import make_resnet50v2_onnx as gen  # noqa: E402


def parity_inputs(n, shape, seed):
    import importlib.util
    spec = importlib.util.spec_from_file_location("tgn", os.path.join(ROOT, "tests", "test_gpu_network.py"))
    src = open(spec.origin).read()
    ns = {"np": np}
    start = src.index("def parity_inputs")
    end = src.index("@pytest.fixture", start)
    exec(src[start:end], ns)
    return ns["parity_inputs"](n, shape, seed)


with tempfile.TemporaryDirectory() as td:
    p = os.path.join(td, "r50.onnx")
    gen.write_model(p, "resnet50", seed=0)
    m = onnx_mini.load(p)
    x = parity_inputs(8, (3, 224, 224), 0)
    with torch.no_grad():
        y32 = resnet_oracle.run_graph(m, torch.from_numpy(x), torch.float32).numpy()
        y64 = resnet_oracle.run_graph(m, torch.from_numpy(x), torch.float64).numpy()
out = os.path.join(ROOT, "tests", "golden", "resnet50_logits_seed0.npz")
np.savez_compressed(out, n=8, seed=0, logits_fp32=y32.astype(np.float32), logits_fp64=y64.astype(np.float64))
print("wrote", out, "fp32-vs-fp64 max-abs", np.abs(y32 - y64).max(), "top-1", y64.argmax(1))
