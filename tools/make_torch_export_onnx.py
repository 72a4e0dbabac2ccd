#!/usr/bin/env python3
"""An ONNX file this repository's own writer (tools/make_resnet50v2_onnx.py) did NOT produce: a small pre-activation
bottleneck network built from torch.nn modules and serialised by PyTorch's TorchScript ONNX exporter, i.e. by
libtorch's C++ protobuf writer with its node naming, attribute set (ceil_mode, dilations, group, kernel_shape ...),
initializer encoding and constant folding.  Both ONNX readers here (csrc/onnx_reader.cpp and oracle/onnx_mini.py) had
only ever seen files emitted by the repository's generator; the real resnet50-v2-7.onnx cannot be had on this image.

The `onnx` Python package is absent, and the exporter imports it for one step only -- splicing onnx-script custom
functions into the model, of which this network has none -- so that step is replaced by the identity for the
duration of the export.  PyTorch's own eager forward of the same module is a second, independent oracle.

    from tools import make_torch_export_onnx as te
    module = te.export(path, seed=0)        # writes the file, returns the eval-mode module
"""
import contextlib

import torch
import torch.nn as nn


class Bottleneck(nn.Module):
    """Pre-activation (v2) bottleneck with the stride on the 3x3 and the projection taken from the activated input."""

    def __init__(self, cin, mid, cout, stride, project):
        super().__init__()
        self.bn1 = nn.BatchNorm2d(cin)
        self.conv1 = nn.Conv2d(cin, mid, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(mid)
        self.conv2 = nn.Conv2d(mid, mid, 3, stride, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(mid)
        self.conv3 = nn.Conv2d(mid, cout, 1, bias=False)
        self.shortcut = nn.Conv2d(cin, cout, 1, stride, bias=False) if project else None

    def forward(self, x):
        a = torch.relu(self.bn1(x))
        sc = self.shortcut(a) if self.shortcut is not None else x
        y = self.conv1(a)
        y = self.conv2(torch.relu(self.bn2(y)))
        y = self.conv3(torch.relu(self.bn3(y)))
        return y + sc


def build(seed=0, classes=100):
    torch.manual_seed(seed)
    net = nn.Sequential(
        nn.BatchNorm2d(3), nn.Conv2d(3, 64, 7, 2, 3, bias=False), nn.BatchNorm2d(64), nn.ReLU(), nn.MaxPool2d(3, 2, 1),
        Bottleneck(64, 64, 256, 1, True), Bottleneck(256, 64, 256, 1, False), Bottleneck(256, 128, 512, 2, True),
        nn.BatchNorm2d(512), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(512, classes))
    g = torch.Generator().manual_seed(seed + 1)
    for m in net.modules():                       # running statistics and affine terms away from their defaults
        if isinstance(m, nn.BatchNorm2d):
            m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.2)
            m.running_var.copy_(torch.rand(m.num_features, generator=g) * 0.8 + 0.6)
            m.weight.data.copy_(torch.rand(m.num_features, generator=g) * 0.4 + 0.8)
            m.bias.data.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    return net.eval()


@contextlib.contextmanager
def _without_onnx_package():
    from torch.onnx._internal.torchscript_exporter import onnx_proto_utils as opu
    orig = opu._add_onnxscript_fn
    opu._add_onnxscript_fn = lambda model_bytes, custom_opsets: model_bytes      # no onnx-script functions in this model
    try:
        yield
    finally:
        opu._add_onnxscript_fn = orig


def export(path, seed=0, hw=64, opset=11):
    import warnings
    net = build(seed)
    x = torch.randn(2, 3, hw, hw, generator=torch.Generator().manual_seed(seed + 2))
    with _without_onnx_package(), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        torch.onnx.export(net, (x,), path, dynamo=False, opset_version=opset, input_names=["data"], output_names=["logits"],
                          dynamic_axes={"data": {0: "N"}, "logits": {0: "N"}})
    return net


if __name__ == "__main__":
    import sys
    export(sys.argv[1] if len(sys.argv) > 1 else "/tmp/torch_export_preact.onnx")
