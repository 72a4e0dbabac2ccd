"""A model file the repository's own ONNX writer did not produce (tools/make_torch_export_onnx.py: torch.nn modules
serialised by PyTorch's TorchScript exporter) through both ONNX readers, the planner, the oracle and -- on a GPU -- the
engine.  PyTorch's eager forward of the module is the expected value: it shares no graph-walking code with the oracle or
the engine, and the exporter has already folded some BatchNorms into biased convolutions, which the repository's
generator never does."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

import b200rn_loader  # noqa: E402
from oracle import onnx_mini, resnet_oracle  # noqa: E402
from tools import make_torch_export_onnx as te  # noqa: E402

pkg = b200rn_loader.load_package()


@pytest.fixture(scope="module")
def exported(tmp_path_factory):
    path = str(tmp_path_factory.mktemp("te") / "torch_export_preact.onnx")
    net = te.export(path, seed=0)
    x = np.random.RandomState(4).randn(5, 3, 64, 64).astype(np.float32)
    with torch.no_grad():
        want = net(torch.from_numpy(x)).numpy()
    return path, net, x, want


def test_readers_planner_and_oracle_on_a_torch_exported_file(exported):
    path, net, x, want = exported
    m = onnx_mini.load(path)
    ops = sorted(set(n["op"] for n in m["nodes"]))
    assert ops == ["Add", "BatchNormalization", "Conv", "Flatten", "Gemm", "GlobalAveragePool", "MaxPool", "Relu"]
    assert any(len(n["input"]) == 3 for n in m["nodes"] if n["op"] == "Conv")       # exporter-folded BN -> conv bias
    assert m["inputs"][0][1] == [-1, 3, 64, 64] and m["outputs"][0][1] == [-1, 100]
    got = np.stack(resnet_oracle.OracleEngine(path).batch_predict(list(x.reshape(len(x), -1))))
    assert np.abs(got - want).max() <= 1e-5                                         # the oracle against PyTorch eager
    # the C++ reader + planner: same I/O, every Conv / Gemm MAC accounted for, everything fused into 13 launches
    plan = pkg.describe_plan(path, "fp16")
    head = plan.splitlines()[0]
    macs = 0
    hw = 64
    # MACs from a shape walk of the module
    def conv_macs(c, h):
        ho = (h + 2 * c.padding[0] - c.kernel_size[0]) // c.stride[0] + 1
        return c.out_channels * c.in_channels * c.kernel_size[0] * c.kernel_size[1] * ho * ho, ho
    k, h = conv_macs(net[1], hw); macs += k
    h = (h + 2 - 3) // 2 + 1
    for blk in (net[5], net[6], net[7]):
        k1, _ = conv_macs(blk.conv1, h); k2, h2 = conv_macs(blk.conv2, h); k3, _ = conv_macs(blk.conv3, h2)
        macs += k1 + k2 + k3
        if blk.shortcut is not None:
            macs += conv_macs(blk.shortcut, h)[0]
        h = h2
    macs += net[12].in_features * net[12].out_features
    assert "input data [3,64,64] -> output logits [100]" in head and "macs/sample=%d" % macs in head
    assert "layers=13" in head


@pytest.mark.gpu
@pytest.mark.parametrize("mode,tol", [("exact", 1e-3), ("exact-simt", 1e-3), ("fp16", 0.05), ("bf16", 0.4)])
def test_engine_on_a_torch_exported_file(exported, mode, tol, monkeypatch):
    path, net, x, want = exported
    if mode == "exact-simt":
        monkeypatch.setenv("B200RN_EXACT_SIMT", "1")
        mode = "exact"
    eng = pkg.Engine(path, mode=mode, max_batch=4)                                  # 5 samples: two slices
    got = np.stack(eng.batch_predict(list(x.reshape(len(x), -1))))
    eng.close()
    err = float(np.abs(got - want).max())
    print("\n[torch-export/%s] max-abs vs PyTorch eager %.3e (logit std %.2f)" % (mode, err, want.std()))
    assert err <= tol
    if mode == "exact":
        assert (got.argmax(1) == want.argmax(1)).all()
