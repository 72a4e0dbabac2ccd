"""TEST INFRASTRUCTURE ONLY (oracle) -- CPU restatement of the reference's hot path.

PARITY UNPINNED: the reference's arithmetic is ONNX Runtime's (un-vendored, version not
pinned: /root/reference/CMakeLists.txt:11-13,52) on models/resnet50-v2-7.onnx (absent:
.MISSING_LARGE_BLOBS:1).  Neither exists on this image, and the reference has no tests or
golden outputs for this path, so this oracle cannot be checked against the reference's own
numbers.  It executes the ONNX operator definitions literally, unfused, in graph order, in
fp32 (the comparison target) or fp64 (to bound the restatement's own rounding).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this.  The product path never does.

What each function follows:
  batch_predict  <- InferenceEngine::batchPredict, src/inference_engine.cpp:134-209
                    (empty -> {} :137-139; short sample zero-padded :157-159; batch dim set
                    :163-164; one Session::Run :176-183; rows split by product of output dims
                    after dim0 :188-206)
  predict        <- InferenceEngine::predict, src/inference_engine.cpp:89-132 (pad AND truncate
                    via resize :100-103)
  io shapes      <- initializeSession, src/inference_engine.cpp:39-68 (first input / output only,
                    -1 -> 1)
  run_graph      <- what Session::Run executes: ONNX opset-7 definitions of
                    BatchNormalization, Conv, Relu, MaxPool, Add, GlobalAveragePool, Flatten,
                    Reshape, Gemm, Softmax.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import onnx_mini


def _bn(x, scale, bias, mean, var, eps):
    # ONNX BatchNormalization, inference form: scale * (x - mean) / sqrt(var + eps) + bias
    shp = (1, -1) + (1,) * (x.dim() - 2)
    return (x - mean.view(shp)) / torch.sqrt(var.view(shp) + eps) * scale.view(shp) + bias.view(shp)


def run_graph(model, x, dtype=torch.float32, return_all=False):
    """Execute the parsed graph on `x` (torch tensor, NCHW).  Unfused, node by node."""
    env = {}
    for k, v in model["initializers"].items():
        env[k] = torch.from_numpy(v).to(dtype) if v.dtype != np.int64 else torch.from_numpy(v)
    env[model["inputs"][0][0]] = x.to(dtype)
    for n in model["nodes"]:
        a = n["attr"]
        i = [env[k] for k in n["input"]]
        op = n["op"]
        if op == "BatchNormalization":
            y = _bn(i[0], i[1], i[2], i[3], i[4], a.get("epsilon", 1e-5))
        elif op == "Conv":
            pads = a.get("pads", [0, 0, 0, 0])
            assert pads[0] == pads[2] and pads[1] == pads[3], "asymmetric pads"
            y = F.conv2d(i[0], i[1], i[2] if len(i) > 2 else None,
                         stride=tuple(a.get("strides", [1, 1])), padding=(pads[0], pads[1]),
                         dilation=tuple(a.get("dilations", [1, 1])), groups=a.get("group", 1))
        elif op == "Relu":
            y = torch.relu(i[0])
        elif op == "MaxPool":
            pads = a.get("pads", [0, 0, 0, 0])
            # ONNX MaxPool: padding does not take part in the max (torch pads with -inf)
            y = F.max_pool2d(i[0], tuple(a["kernel_shape"]), stride=tuple(a.get("strides", [1, 1])),
                             padding=(pads[0], pads[1]))
        elif op == "Add":
            y = i[0] + i[1]
        elif op == "GlobalAveragePool":
            y = i[0].mean(dim=(2, 3), keepdim=True)
        elif op == "Flatten":
            ax = a.get("axis", 1)
            y = i[0].reshape(int(np.prod(i[0].shape[:ax])), -1)
        elif op == "Reshape":
            shape = [int(s) for s in (i[1].tolist() if len(i) > 1 else a["shape"])]
            shape = [i[0].shape[k] if s == 0 else s for k, s in enumerate(shape)]
            y = i[0].reshape(shape)
        elif op == "Gemm":
            A = i[0].t() if a.get("transA", 0) else i[0]
            B = i[1].t() if a.get("transB", 0) else i[1]
            y = a.get("alpha", 1.0) * (A @ B)
            if len(i) > 2:
                y = y + a.get("beta", 1.0) * i[2]
        elif op == "Softmax":
            ax = a.get("axis", 1)
            flat = i[0].reshape(int(np.prod(i[0].shape[:ax])), -1)
            y = torch.softmax(flat, dim=1).reshape(i[0].shape)
        else:
            raise NotImplementedError("oracle: op %s" % op)
        env[n["output"][0]] = y
    if return_all:
        return env
    return env[model["outputs"][0][0]]


class OracleEngine:
    """Mirror of the reference's InferenceEngine surface (include/inference_engine.h:10-21)."""

    def __init__(self, model_path, shard_id=0, dtype=torch.float32, threads=None):
        self.model_path = model_path
        self.shard_id = shard_id
        self.dtype = dtype
        if threads:
            torch.set_num_threads(threads)
        self.model = onnx_mini.load(model_path)
        # src/inference_engine.cpp:47-51,64-68: dynamic dims (-1) reported as 1
        self.input_shape = [1 if d == -1 else d for d in self.model["inputs"][0][1]]
        self.output_shape = [1 if d == -1 else d for d in self.model["outputs"][0][1]]

    def get_input_shape(self):
        return list(self.input_shape)

    def get_output_shape(self):
        return list(self.output_shape)

    def batch_predict(self, inputs):
        if len(inputs) == 0:
            return []
        per = int(np.prod(self.input_shape[1:]))
        rows = []
        for v in inputs:
            v = np.asarray(v, dtype=np.float32).ravel()
            if v.size > per:
                # The reference appends an over-long sample whole (:155), which mis-sizes the
                # tensor handed to ORT; what ORT then does is unverified.  The replacement
                # rejects it (documented divergence), and so does the oracle.
                raise ValueError("sample longer than the model input (%d > %d)" % (v.size, per))
            r = np.zeros(per, dtype=np.float32)
            r[:v.size] = v
            rows.append(r)
        x = torch.from_numpy(np.stack(rows)).reshape([len(rows)] + self.input_shape[1:])
        with torch.no_grad():
            y = run_graph(self.model, x, self.dtype)
        y = y.reshape(len(rows), -1)
        return [y[k].to(torch.float64 if self.dtype == torch.float64 else torch.float32).numpy().copy()
                for k in range(len(rows))]

    def predict(self, v):
        per = int(np.prod(self.input_shape))
        v = np.asarray(v, dtype=np.float32).ravel()
        r = np.zeros(per, dtype=np.float32)
        r[:min(per, v.size)] = v[:per]
        x = torch.from_numpy(r).reshape(self.input_shape)
        with torch.no_grad():
            y = run_graph(self.model, x, self.dtype)
        return y.reshape(-1).numpy().copy()


def synthetic_inputs(batch, seed, shape=(3, 224, 224)):
    """i.i.d. N(0,1) fp32 NCHW, seed = 1234 + batch_index for bench, or as given."""
    rs = np.random.RandomState(seed)
    return rs.randn(batch, *shape).astype(np.float32)
