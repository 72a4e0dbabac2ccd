"""TEST INFRASTRUCTURE ONLY (oracle).  Minimal ONNX protobuf *reader* in pure Python.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product path (csrc/onnx_reader.cpp) has its own C++ parser;
this one is written independently so the two can be checked against each other.

What it follows: the reference hands `model_path` to ONNX Runtime
(/root/reference/src/inference_engine.cpp:31) and uses only input 0 / output 0
(:39-68).  ONNX Runtime is not on this image, so this reader decodes the wire format
directly (field numbers from onnx.proto3: ModelProto.graph=7, GraphProto.node=1 /
initializer=5 / input=11 / output=12, NodeProto input=1 output=2 name=3 op_type=4
attribute=5, AttributeProto name=1 f=2 i=3 s=4 t=5 floats=7 ints=8 type=20,
TensorProto dims=1 data_type=2 float_data=4 int64_data=7 name=8 raw_data=9).
"""
import struct

import numpy as np


def _read_varint(buf, pos):
    result = 0
    shift = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not (b & 0x80):
            break
        shift += 7
    return result, pos


def _fields(buf):
    """Yield (field_number, wire_type, value) for one message; value is int or memoryview."""
    pos = 0
    n = len(buf)
    while pos < n:
        key, pos = _read_varint(buf, pos)
        field, wire = key >> 3, key & 7
        if wire == 0:
            v, pos = _read_varint(buf, pos)
        elif wire == 1:
            v = bytes(buf[pos:pos + 8])
            pos += 8
        elif wire == 2:
            ln, pos = _read_varint(buf, pos)
            v = buf[pos:pos + ln]
            pos += ln
        elif wire == 5:
            v = bytes(buf[pos:pos + 4])
            pos += 4
        else:
            raise ValueError("unsupported wire type %d" % wire)
        yield field, wire, v


def _s64(v):
    return v - (1 << 64) if v >= (1 << 63) else v


def _packed_or_single_ints(wire, v, out):
    if wire == 0:
        out.append(_s64(v))
    else:
        pos = 0
        while pos < len(v):
            x, pos = _read_varint(v, pos)
            out.append(_s64(x))


def parse_tensor(buf):
    dims, dtype, name, raw = [], 1, "", None
    floats, int64s = [], []
    for f, w, v in _fields(buf):
        if f == 1:
            _packed_or_single_ints(w, v, dims)
        elif f == 2:
            dtype = v
        elif f == 4:
            if w == 5:
                floats.append(struct.unpack("<f", v)[0])
            else:
                floats.extend(np.frombuffer(bytes(v), dtype="<f4").tolist())
        elif f == 7:
            _packed_or_single_ints(w, v, int64s)
        elif f == 8:
            name = bytes(v).decode()
        elif f == 9:
            raw = bytes(v)
    if dtype == 1:
        arr = np.frombuffer(raw, dtype="<f4").copy() if raw is not None else np.asarray(floats, dtype=np.float32)
    elif dtype == 7:
        arr = np.frombuffer(raw, dtype="<i8").copy() if raw is not None else np.asarray(int64s, dtype=np.int64)
    else:
        raise ValueError("tensor %s: unsupported data_type %d" % (name, dtype))
    return name, arr.reshape(dims)


def parse_attr(buf):
    name, atype = "", 0
    f_, i_, s_, t_ = None, None, None, None
    floats, ints = [], []
    for f, w, v in _fields(buf):
        if f == 1:
            name = bytes(v).decode()
        elif f == 2:
            f_ = struct.unpack("<f", v)[0]
        elif f == 3:
            i_ = _s64(v)
        elif f == 4:
            s_ = bytes(v)
        elif f == 5:
            t_ = parse_tensor(v)[1]
        elif f == 7:
            if w == 5:
                floats.append(struct.unpack("<f", v)[0])
            else:
                floats.extend(np.frombuffer(bytes(v), dtype="<f4").tolist())
        elif f == 8:
            _packed_or_single_ints(w, v, ints)
        elif f == 20:
            atype = v
    if atype == 1 or (atype == 0 and f_ is not None):
        return name, f_
    if atype == 2 or (atype == 0 and i_ is not None):
        return name, i_
    if atype == 3:
        return name, s_
    if atype == 4:
        return name, t_
    if atype == 6:
        return name, floats
    if atype == 7 or ints:
        return name, ints
    return name, None


def parse_node(buf):
    n = {"input": [], "output": [], "name": "", "op": "", "attr": {}}
    for f, w, v in _fields(buf):
        if f == 1:
            n["input"].append(bytes(v).decode())
        elif f == 2:
            n["output"].append(bytes(v).decode())
        elif f == 3:
            n["name"] = bytes(v).decode()
        elif f == 4:
            n["op"] = bytes(v).decode()
        elif f == 5:
            k, val = parse_attr(v)
            n["attr"][k] = val
    return n


def parse_value_info(buf):
    name, dims = "", []
    for f, w, v in _fields(buf):
        if f == 1:
            name = bytes(v).decode()
        elif f == 2:  # TypeProto
            for f2, _, v2 in _fields(v):
                if f2 == 1:  # tensor_type
                    for f3, _, v3 in _fields(v2):
                        if f3 == 2:  # shape
                            for f4, _, v4 in _fields(v3):
                                if f4 == 1:  # dim
                                    d = -1  # dim_param / unknown -> -1, as ORT reports it
                                    for f5, _, v5 in _fields(v4):
                                        if f5 == 1:
                                            d = _s64(v5)
                                    dims.append(d)
    return name, dims


def load(path):
    """-> dict(nodes=[...], initializers={name: ndarray}, inputs=[(name, dims)], outputs=[(name, dims)], opset=int)"""
    with open(path, "rb") as fh:
        data = memoryview(fh.read())
    graph = None
    opset = 0
    for f, w, v in _fields(data):
        if f == 7:
            graph = v
        elif f == 8:
            for f2, _, v2 in _fields(v):
                if f2 == 2:
                    opset = max(opset, v2)
    if graph is None:
        raise ValueError("no graph in %s" % path)
    nodes, inits, inputs, outputs = [], {}, [], []
    for f, w, v in _fields(graph):
        if f == 1:
            nodes.append(parse_node(v))
        elif f == 5:
            k, arr = parse_tensor(v)
            inits[k] = arr
        elif f == 11:
            inputs.append(parse_value_info(v))
        elif f == 12:
            outputs.append(parse_value_info(v))
    # ORT's session inputs exclude initializers (IR<4 lists them among graph inputs)
    real_inputs = [(n, d) for n, d in inputs if n not in inits]
    return {"nodes": nodes, "initializers": inits, "inputs": real_inputs, "outputs": outputs, "opset": opset}
