"""CPU: the oracle (oracle/resnet_oracle.py) still says what the committed golden files say.  The goldens were produced
by the oracle itself (tests/golden/make_*_golden.py) -- the reference has no golden outputs for this path and ONNX
Runtime is not on the image (PARITY UNPINNED for the arithmetic) -- so this guards the checker
against drifting, nothing more."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")

import parity_set  # noqa: E402
from oracle import resnet_oracle  # noqa: E402
from tools import make_resnet50v2_onnx as gen  # noqa: E402


def test_oracle_reproduces_parity258_golden_rows(tmp_path, golden_dir):
    g = np.load(os.path.join(golden_dir, "resnet50_parity258_seed0.npz"))
    assert g["margin_fp64"].shape == (258,) and (g["top1_fp32"] == g["top1_fp64"]).all()
    assert g["fp32_vs_fp64_maxabs"].max() < 1e-3          # the restatement's own rounding is far below the 1e-3 bar
    p = str(tmp_path / "r50.onnx")
    gen.write_model(p, "resnet50", seed=0)
    xs = parity_set.samples()
    rows = [0, 1, 256, 257]                               # two random images, all-zeros, [1,2,3] zero-padded
    out = np.stack(resnet_oracle.OracleEngine(p).batch_predict([xs[i] for i in rows]))
    want = g["logits_fp32"][[list(g["rows"]).index(i) for i in rows]]
    assert np.abs(out - want).max() <= 2e-3               # fp32 reduction order differs between hosts
    assert (out.argmax(1) == g["top1_fp64"][rows]).all()
    assert len(xs) == 258 and xs[257].size == 3 and not xs[256].any()
