#!/usr/bin/env python3
"""Write a synthetic ResNet-v2 ONNX file with the topology and operator set of the
model-zoo `resnet50-v2-7.onnx` the reference loads (reference: src/worker_node.cpp:156-171,
README "models/resnet50-v2-7.onnx"; the file itself is absent: .MISSING_LARGE_BLOBS:1).

There is no `onnx` package and no network on this image, so this tool encodes the
protobuf wire format by hand (ModelProto / GraphProto / NodeProto / TensorProto field
numbers from onnx.proto3).  It is a *data generator*, not part of the product path and
not the oracle: the C++ parser (csrc/onnx_reader.cpp) and the oracle's Python reader
(oracle/onnx_mini.py) both parse what this writes, independently of each other.

Topology (Gluon `resnet50_v2` as exported by MXNet, opset 7): data -> BatchNorm (gamma
fixed to 1) -> Conv 7x7/2 -> BN -> ReLU -> MaxPool 3x3/2 -> [3,4,6,3] pre-activation
bottlenecks (stride on the 3x3, 1x1 projection shortcut taken from the post-BN-ReLU
tensor on the first unit of each stage) -> BN -> ReLU -> GlobalAveragePool -> Flatten ->
Gemm(transB=1) with bias.  All convs are bias-free.  Initializers are also listed as graph
inputs (IR version 3 convention), after the real input `data`.

Weights are seeded (numpy RandomState, bit-reproducible across machines): He-normal
convs; BN gamma in [0.8,1.2], beta ~ 0.1*N(0,1), running mean ~ 0.1*N(0,1), running var
chosen from an analytic variance-propagation estimate (x a factor in [0.8,1.25]) so that
activations stay O(1..10) through all 50 layers.
"""
import argparse
import struct
import sys

import numpy as np

# ---------------------------------------------------------------- protobuf writer


def _varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def f_varint(field: int, v: int) -> bytes:
    return _key(field, 0) + _varint(v)


def f_bytes(field: int, b: bytes) -> bytes:
    return _key(field, 2) + _varint(len(b)) + b


def f_str(field: int, s: str) -> bytes:
    return f_bytes(field, s.encode("utf-8"))


def f_float(field: int, v: float) -> bytes:
    return _key(field, 5) + struct.pack("<f", v)


# AttributeProto.type enum
A_FLOAT, A_INT, A_INTS = 1, 2, 7


def attr_i(name, v):
    return f_str(1, name) + f_varint(3, v) + f_varint(20, A_INT)


def attr_f(name, v):
    return f_str(1, name) + f_float(2, v) + f_varint(20, A_FLOAT)


def attr_ints(name, vs):
    # unpacked repeated int64 (what the proto2-era MXNet exporter emitted)
    return f_str(1, name) + b"".join(f_varint(8, v) for v in vs) + f_varint(20, A_INTS)


def node(op, inputs, outputs, name, attrs=()):
    b = b"".join(f_str(1, i) for i in inputs)
    b += b"".join(f_str(2, o) for o in outputs)
    b += f_str(3, name) + f_str(4, op)
    b += b"".join(f_bytes(5, a) for a in attrs)
    return b


def tensor(name, arr, use_raw=True):
    arr = np.ascontiguousarray(arr, dtype=np.float32)
    b = b"".join(f_varint(1, d) for d in arr.shape)
    b += f_varint(2, 1)  # FLOAT
    if use_raw:
        b += f_str(8, name) + f_bytes(9, arr.tobytes())
    else:
        b += f_bytes(4, arr.tobytes()) + f_str(8, name)  # packed float_data
    return b


def value_info(name, dims):
    """dims: list of int or str (str -> dim_param, i.e. dynamic)."""
    shape = b""
    for d in dims:
        if isinstance(d, str):
            shape += f_bytes(1, f_str(2, d))
        else:
            shape += f_bytes(1, f_varint(1, d))
    ttype = f_varint(1, 1) + f_bytes(2, shape)  # elem_type FLOAT, shape
    return f_str(1, name) + f_bytes(2, f_bytes(1, ttype))


# ---------------------------------------------------------------- model builder


class Builder:
    def __init__(self, seed):
        self.rs = np.random.RandomState(seed)
        self.nodes = []
        self.inits = []  # (name, array)
        self.ops = []    # (kind, inputs, output, params) for the calibration forward

    def init(self, name, arr):
        self.inits.append((name, np.asarray(arr, dtype=np.float32)))
        return name

    def bn(self, x, c, name, var_est, fix_gamma=False, eps=1e-5):
        rs = self.rs
        gamma = np.ones(c) if fix_gamma else rs.uniform(0.8, 1.2, c)
        beta = 0.1 * rs.randn(c)
        mean = 0.1 * np.sqrt(var_est) * rs.randn(c)
        var = var_est * rs.uniform(0.8, 1.25, c)
        ins = [x, self.init(name + "_gamma", gamma), self.init(name + "_beta", beta),
               self.init(name + "_running_mean", mean), self.init(name + "_running_var", var)]
        out = name + "_fwd"
        self.nodes.append(node("BatchNormalization", ins, [out], name + "_fwd",
                               [attr_f("epsilon", eps), attr_f("momentum", 0.9), attr_i("spatial", 1)]))
        self.ops.append(("bn", [x], out, dict(g=gamma, b=beta, m=mean, v=var, eps=eps)))
        return out

    def conv(self, x, cin, cout, k, stride, pad, name):
        fan_in = cin * k * k
        w = self.rs.randn(cout, cin, k, k) * np.sqrt(2.0 / fan_in)
        ins = [x, self.init(name + "_weight", w)]
        out = name + "_fwd"
        self.nodes.append(node("Conv", ins, [out], name + "_fwd",
                               [attr_ints("dilations", [1, 1]), attr_i("group", 1),
                                attr_ints("kernel_shape", [k, k]),
                                attr_ints("pads", [pad, pad, pad, pad]),
                                attr_ints("strides", [stride, stride])]))
        self.ops.append(("conv", [x], out, dict(w=w, stride=stride, pad=pad)))
        return out

    def relu(self, x, name):
        out = name + "_fwd"
        self.nodes.append(node("Relu", [x], [out], name + "_fwd"))
        self.ops.append(("relu", [x], out, {}))
        return out

    def simple(self, op, ins, name, attrs=()):
        out = name + "_fwd"
        self.nodes.append(node(op, ins, [out], name + "_fwd", attrs))
        self.ops.append((op.lower(), list(ins), out, {}))
        return out

    def pooled_features(self, x):
        """Calibration forward (torch CPU, fp32) of everything recorded so far; returns the last value."""
        import torch
        import torch.nn.functional as F
        env = {"data": torch.from_numpy(x)}
        f32 = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32))
        last = None
        with torch.no_grad():
            for kind, ins, out, p in self.ops:
                a = env[ins[0]]
                if kind == "bn":
                    sh = (1, -1, 1, 1)
                    y = (a - f32(p["m"]).view(sh)) / torch.sqrt(f32(p["v"]).view(sh) + p["eps"]) * f32(p["g"]).view(sh) + f32(p["b"]).view(sh)
                elif kind == "conv":
                    y = F.conv2d(a, f32(p["w"]), None, stride=p["stride"], padding=p["pad"])
                elif kind == "relu":
                    y = torch.relu(a)
                elif kind == "maxpool":
                    y = F.max_pool2d(a, 3, stride=2, padding=1)
                elif kind == "add":
                    y = a + env[ins[1]]
                elif kind == "globalaveragepool":
                    y = a.mean(dim=(2, 3), keepdim=True)
                else:
                    raise ValueError(kind)
                env[out] = y
                last = y
        return last.reshape(last.shape[0], -1).numpy()


def calibration_inputs(n, shape, seed=777):
    """Mixed noise / smooth / scaled images used only to centre the synthetic classifier."""
    rs = np.random.RandomState(seed)
    c, h, w = shape
    yy, xx = np.meshgrid(np.linspace(-1, 1, h), np.linspace(-1, 1, w), indexing="ij")
    xs = []
    for i in range(n):
        k = i % 4
        if k == 0:
            x = rs.randn(c, h, w)
        elif k == 1:
            a = rs.randn(c, 3)
            x = a[:, 0, None, None] * yy + a[:, 1, None, None] * xx + a[:, 2, None, None] + 0.3 * rs.randn(c, h, w)
        elif k == 2:
            x = rs.randn(c, h, w) * rs.uniform(0.2, 3.0)
        else:
            f = rs.uniform(1, 6, size=2)
            x = np.sin(f[0] * np.pi * yy + rs.uniform(0, 6))[None] * np.cos(f[1] * np.pi * xx)[None] * rs.randn(c, 1, 1) * 2
        xs.append(x.astype(np.float32))
    return np.stack(xs)


def build(seed=0, image=224, stem=64, widths=(256, 512, 1024, 2048), blocks=(3, 4, 6, 3),
          classes=1000, batch_dim="N", float_data_for_small=True, softmax=False, calib=None, calibrate_out=None):
    b = Builder(seed)
    p = "resnetv24_"
    x = b.bn("data", 3, p + "batchnorm0", var_est=1.0, fix_gamma=True)
    x = b.conv(x, 3, stem, 7, 2, 3, p + "conv0")
    # conv of a ~N(0,1) tensor with He weights: var ~ 2
    x = b.bn(x, stem, p + "batchnorm1", var_est=2.0)
    x = b.relu(x, p + "relu0")
    x = b.simple("MaxPool", [x], p + "pool0",
                 [attr_ints("kernel_shape", [3, 3]), attr_ints("pads", [1, 1, 1, 1]),
                  attr_ints("strides", [2, 2])])
    # running variance estimate of the un-normalised residual trunk
    trunk_var = 1.2  # max-pooled relu(N(0,1))
    cin = stem
    for s, (cout, nb) in enumerate(zip(widths, blocks)):
        mid = cout // 4
        sp = p + "stage%d_" % (s + 1)
        nconv = nbn = nact = 0  # Gluon-style running name counters within a stage
        for u in range(nb):
            stride = 2 if (u == 0 and s > 0) else 1
            act = b.bn(x, cin, sp + "batchnorm%d" % nbn, var_est=trunk_var)
            act = b.relu(act, sp + "activation%d" % nact)
            y = b.conv(act, cin, mid, 1, 1, 0, sp + "conv%d" % nconv)
            y = b.bn(y, mid, sp + "batchnorm%d" % (nbn + 1), var_est=1.0)
            y = b.relu(y, sp + "activation%d" % (nact + 1))
            y = b.conv(y, mid, mid, 3, stride, 1, sp + "conv%d" % (nconv + 1))
            y = b.bn(y, mid, sp + "batchnorm%d" % (nbn + 2), var_est=1.0)
            y = b.relu(y, sp + "activation%d" % (nact + 2))
            y = b.conv(y, mid, cout, 1, 1, 0, sp + "conv%d" % (nconv + 2))
            nconv += 3
            nbn += 3
            nact += 3
            if u == 0:
                shortcut = b.conv(act, cin, cout, 1, stride, 0, sp + "conv%d" % nconv)
                nconv += 1
                sc_var = 1.0  # He conv of relu(N(0,1))
            else:
                shortcut = x
                sc_var = trunk_var
            x = b.simple("Add", [y, shortcut], sp + "_plus%d" % u)
            trunk_var = sc_var + 1.0
            cin = cout
    x = b.bn(x, cin, p + "batchnorm2", var_est=trunk_var)
    x = b.relu(x, p + "relu1")
    x = b.simple("GlobalAveragePool", [x], p + "pool1")
    if calibrate_out is not None:
        feats = b.pooled_features(calibration_inputs(48, (3, image, image)))
        np.savez_compressed(calibrate_out, mean=feats.mean(0).astype(np.float32), std=feats.std(0).astype(np.float32))
        calib = dict(mean=feats.mean(0), std=feats.std(0))
    b.ops = []
    x = b.simple("Flatten", [x], p + "flatten0", [attr_i("axis", 1)])
    wfc = b.rs.randn(classes, cin) * np.sqrt(1.0 / cin) * 4.0
    bfc = 0.5 * b.rs.randn(classes)
    if calib is not None:
        # Make the classifier read standardised deviations from the mean pooled feature, so that the
        # winning class depends on the input (a random deep net's pooled features barely move, and an
        # uncalibrated FC would crown the same class for every image).
        wfc = wfc / np.maximum(np.asarray(calib["std"], dtype=np.float64), 1e-3)[None, :]
        bfc = bfc - wfc @ np.asarray(calib["mean"], dtype=np.float64)
    out = p + "dense0_fwd"
    b.nodes.append(node("Gemm", [x, b.init(p + "dense0_weight", wfc), b.init(p + "dense0_bias", bfc)],
                        [out], p + "dense0_fwd",
                        [attr_f("alpha", 1.0), attr_f("beta", 1.0), attr_i("transA", 0), attr_i("transB", 1)]))
    if softmax:
        sm = p + "softmax0"
        b.nodes.append(node("Softmax", [out], [sm], sm, [attr_i("axis", 1)]))
        out = sm

    g = b"".join(f_bytes(1, n) for n in b.nodes)
    g += f_str(2, "mxnet_converted_model")
    for name, arr in b.inits:
        # mix the two TensorProto payload encodings so both parser paths are exercised
        use_raw = not (float_data_for_small and arr.ndim == 1)
        g += f_bytes(5, tensor(name, arr, use_raw=use_raw))
    g += f_bytes(11, value_info("data", [batch_dim, 3, image, image]))
    for name, arr in b.inits:
        g += f_bytes(11, value_info(name, list(arr.shape)))
    g += f_bytes(12, value_info(out, [batch_dim, classes]))

    m = f_varint(1, 3)  # ir_version 3
    m += f_str(2, "b200rn-synthetic-fixture") + f_str(3, "1")
    m += f_bytes(7, g)
    m += f_bytes(8, f_str(1, "") + f_varint(2, 7))  # opset 7
    return m


PRESETS = {
    # the BASELINE model: ResNet-50 v2, 3x224x224 -> 1000
    "resnet50": dict(image=224, stem=64, widths=(256, 512, 1024, 2048), blocks=(3, 4, 6, 3), classes=1000),
    # same op set / fusion patterns, tiny, for CPU-side parser/planner tests
    "tiny": dict(image=32, stem=8, widths=(16, 32, 64, 128), blocks=(2, 1, 1, 1), classes=10),
    # channel counts that satisfy the tensor-core path (multiples of 64) at small cost
    "small64": dict(image=64, stem=64, widths=(256, 512), blocks=(2, 2), classes=100),
}


def calib_path(preset, seed):
    import os
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "calib_%s_seed%d.npz" % (preset, seed))


def write_model(path, preset="resnet50", seed=0, calibrate=False, **kw):
    """Writes the model.  If tools/calib_<preset>_seed<seed>.npz exists (committed for resnet50/seed 0)
    the classifier is centred with it, which keeps the weights bit-reproducible across machines;
    calibrate=True recomputes that file here with a torch-CPU forward first."""
    import os
    cfg = dict(PRESETS[preset])
    cfg.update(kw)
    cp = calib_path(preset, seed)
    if calibrate:
        build(seed=seed, calibrate_out=cp, **cfg)
    if os.path.exists(cp) and "calib" not in cfg:
        z = np.load(cp)
        cfg["calib"] = dict(mean=z["mean"], std=z["std"])
    data = build(seed=seed, **cfg)
    with open(path, "wb") as f:
        f.write(data)
    return len(data)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("out")
    ap.add_argument("--preset", default="resnet50", choices=sorted(PRESETS))
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--softmax", action="store_true", help="append a Softmax node (not in the v2-7 graph)")
    ap.add_argument("--static-batch", type=int, default=0, help="write a fixed batch dim instead of 'N'")
    ap.add_argument("--calibrate", action="store_true", help="(re)compute tools/calib_<preset>_seed<seed>.npz first")
    a = ap.parse_args(argv)
    kw = {}
    if a.softmax:
        kw["softmax"] = True
    if a.static_batch:
        kw["batch_dim"] = a.static_batch
    n = write_model(a.out, a.preset, a.seed, calibrate=a.calibrate, **kw)
    print("wrote %s (%d bytes)" % (a.out, n))


if __name__ == "__main__":
    sys.exit(main())
