"""ctypes binding of libb200rn.so (C ABI: include/b200rn.h) -- host-side plumbing for tests and
bench.py.  The directory name carries hyphens (it mirrors the upstream repository name), so load it
through `b200rn_loader.load_package()` at the repo root rather than a plain `import`.

There is no CPU fallback: if the shared library is missing this module raises at import time, and
`Engine(...)` raises when no CUDA device is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200RN_LIB") or os.path.join(_HERE, "libb200rn.so")   # B200RN_LIB: A/B another build

MODE_EXACT, MODE_BF16, MODE_FP16 = 0, 1, 2
MODES = {"exact": MODE_EXACT, "fp32": MODE_EXACT, "bf16": MODE_BF16, "fp16": MODE_FP16}
LAYER_KINDS = ["InputXform", "Conv", "MaxPool", "Affine", "Add", "GlobalAvgPool", "Gemm", "Softmax"]

if not os.path.exists(LIB_PATH):
    raise ImportError("%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback)" % LIB_PATH)

lib = C.CDLL(LIB_PATH)

_fp = C.POINTER(C.c_float)
_vp = C.c_void_p


def _sig(name, restype, argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = argtypes
    return f


_sig("b200rn_create", C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)])
_sig("b200rn_destroy", None, [_vp])
_sig("b200rn_input_shape", C.c_int, [_vp, C.POINTER(C.c_int64), C.c_int])
_sig("b200rn_output_shape", C.c_int, [_vp, C.POINTER(C.c_int64), C.c_int])
_sig("b200rn_input_elems", C.c_int64, [_vp])
_sig("b200rn_output_elems", C.c_int64, [_vp])
_sig("b200rn_infer", C.c_int, [_vp, C.POINTER(_fp), C.POINTER(C.c_size_t), C.c_size_t, _fp])
_sig("b200rn_infer_packed", C.c_int, [_vp, _vp, C.c_size_t, _vp])
_sig("b200rn_infer_device", C.c_int, [_vp, _vp, C.c_size_t, _vp])
_sig("b200rn_submit_packed", C.c_int, [_vp, C.c_int, _vp, C.c_size_t, _vp])
_sig("b200rn_wait", C.c_int, [_vp, C.c_int])
_sig("b200rn_predict", C.c_int, [_vp, _fp, C.c_size_t, _fp])
_sig("b200rn_host_alloc", _vp, [C.c_size_t])
_sig("b200rn_host_free", None, [_vp])
_sig("b200rn_last_error", C.c_char_p, [_vp])
_sig("b200rn_global_error", C.c_char_p, [])
_sig("b200rn_describe_plan", C.c_int64, [C.c_char_p, C.c_int, C.c_char_p, C.c_size_t])
_sig("b200rn_launch_count", C.c_int, [_vp])
_sig("b200rn_layer_count", C.c_int, [_vp])
_sig("b200rn_layer_name", C.c_int, [_vp, C.c_int, C.c_char_p, C.c_size_t])
_sig("b200rn_profile_layers", C.c_int, [_vp, _vp, C.c_size_t, C.c_int, _fp, C.POINTER(C.c_int),
                                        C.POINTER(C.c_double), C.POINTER(C.c_double)])
_sig("b200rn_time_forward", C.c_int, [_vp, C.POINTER(_vp), C.c_int, C.c_size_t, C.c_int, _fp])
_sig("b200rn_conv2d", C.c_int, [C.c_int, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _fp, _fp, C.c_int, C.c_int,
                                C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _fp, _fp, C.c_int, _fp, _fp, C.c_int])
_sig("b200rn_conv1x1_tail", C.c_int, [C.c_int, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _fp, _fp, C.c_int, _vp, _vp, _fp, _fp,
                                      C.c_int, _fp, _fp, C.c_int, C.c_int, _vp])
_sig("b200rn_conv1x1_two_input", C.c_int, [C.c_int, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, C.c_int, C.c_int, C.c_int, C.c_int,
                                           _fp, _fp, C.c_int, _vp, _vp, _fp, _fp, C.c_int])
_sig("b200rn_stem_conv", C.c_int, [C.c_int, _vp, C.c_int, C.c_int, C.c_int, _fp, _fp, _fp, _fp, C.c_int, C.c_int, _vp])
_sig("b200rn_stem_conv_static", C.c_int, [C.c_int, _vp, C.c_int, C.c_int, C.c_int, _fp, _fp, _fp, _fp, C.c_int, C.c_int, _vp,
                                          C.c_char_p, C.c_size_t])
_sig("b200rn_fnv1a", C.c_uint32, [C.c_char_p, C.c_size_t])
_sig("b200rn_ring_create", _vp, [C.c_int])
_sig("b200rn_ring_destroy", None, [_vp])
_sig("b200rn_ring_add", None, [_vp, C.c_char_p])
_sig("b200rn_ring_remove", None, [_vp, C.c_char_p])
_sig("b200rn_ring_get", C.c_int, [_vp, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t])
_sig("b200rn_ring_nodes", C.c_int, [_vp, C.c_char_p, C.c_size_t])
_sig("b200rn_hash_float", C.c_uint64, [C.c_float])
_sig("b200rn_vector_hash", C.c_uint64, [_fp, C.c_size_t])
_sig("b200rn_lru_create", _vp, [C.c_size_t])
_sig("b200rn_lru_destroy", None, [_vp])
_sig("b200rn_lru_get", C.c_long, [_vp, _fp, C.c_size_t, _fp, C.c_size_t])
_sig("b200rn_lru_put", None, [_vp, _fp, C.c_size_t, _fp, C.c_size_t])
_sig("b200rn_lru_size", C.c_size_t, [_vp])
_sig("b200rn_lru_hits", C.c_size_t, [_vp])
_sig("b200rn_lru_misses", C.c_size_t, [_vp])
_sig("b200rn_batch_selftest", None, [C.c_int] * 5 + [C.POINTER(C.c_longlong), C.POINTER(C.c_double)])
_sig("b200rn_batch_selftest2", None, [C.c_int] * 7 + [C.POINTER(C.c_longlong), C.POINTER(C.c_double)])


class B200RNError(RuntimeError):
    pass


def fptr(a):
    return a.ctypes.data_as(_fp)


def describe_plan(onnx_path, mode="exact"):
    cap = 1 << 20
    buf = C.create_string_buffer(cap)
    n = lib.b200rn_describe_plan(onnx_path.encode(), MODES[mode], buf, cap)
    if n < 0:
        raise B200RNError(lib.b200rn_global_error().decode())
    return buf.value.decode()


class Engine:
    """Python mirror of the reference's InferenceEngine surface (include/inference_engine.h:10-21)
    over the C ABI: predict / batch_predict / get_input_shape / get_output_shape, plus the packed,
    device-resident and double-buffered forms the bench uses."""

    def __init__(self, model_path, shard_id=0, mode="fp16", device=0, max_batch=256):
        self.model_path = model_path
        self.shard_id = shard_id
        self.mode = mode
        self._h = _vp()
        rc = lib.b200rn_create(model_path.encode(), device, MODES[mode], max_batch, C.byref(self._h))
        if rc != 0:
            self._h = None
            raise B200RNError("b200rn_create(%s): %s" % (model_path, lib.b200rn_global_error().decode()))
        self.in_elems = lib.b200rn_input_elems(self._h)
        self.out_elems = lib.b200rn_output_elems(self._h)

    def close(self):
        if getattr(self, "_h", None):
            lib.b200rn_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc, what):
        if rc != 0:
            raise B200RNError("%s failed (%d): %s" % (what, rc, lib.b200rn_last_error(self._h).decode()))

    def _shape(self, fn):
        d = (C.c_int64 * 8)()
        r = fn(self._h, d, 8)
        return [int(d[i]) for i in range(r)]

    def get_input_shape(self):
        return self._shape(lib.b200rn_input_shape)

    def get_output_shape(self):
        return self._shape(lib.b200rn_output_shape)

    def batch_predict(self, inputs):
        """inputs: list of 1-D float sequences (ragged allowed) -> list of np.float32 rows."""
        n = len(inputs)
        if n == 0:
            return []
        arrs = [np.ascontiguousarray(np.asarray(v, dtype=np.float32).ravel()) for v in inputs]
        ptrs = (_fp * n)(*[fptr(a) for a in arrs])
        lens = (C.c_size_t * n)(*[a.size for a in arrs])
        out = np.empty((n, self.out_elems), dtype=np.float32)
        self._check(lib.b200rn_infer(self._h, ptrs, lens, n, fptr(out)), "b200rn_infer")
        return [out[i] for i in range(n)]

    def predict(self, v):
        a = np.ascontiguousarray(np.asarray(v, dtype=np.float32).ravel())
        out = np.empty(self.out_elems, dtype=np.float32)
        self._check(lib.b200rn_predict(self._h, fptr(a), a.size, fptr(out)), "b200rn_predict")
        return out

    def infer_packed(self, x):
        """x: float32 array [B, C, H, W] (host) -> [B, out_elems]."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        b = x.shape[0]
        out = np.empty((b, self.out_elems), dtype=np.float32)
        self._check(lib.b200rn_infer_packed(self._h, x.ctypes.data, b, out.ctypes.data), "b200rn_infer_packed")
        return out

    # raw-pointer forms (pinned host or device memory owned by the caller)
    def infer_packed_ptr(self, in_ptr, batch, out_ptr):
        self._check(lib.b200rn_infer_packed(self._h, in_ptr, batch, out_ptr), "b200rn_infer_packed")

    def infer_device_ptr(self, d_in, batch, d_out):
        self._check(lib.b200rn_infer_device(self._h, d_in, batch, d_out), "b200rn_infer_device")

    def submit_ptr(self, slot, in_ptr, batch, out_ptr):
        self._check(lib.b200rn_submit_packed(self._h, slot, in_ptr, batch, out_ptr), "b200rn_submit_packed")

    def wait(self, slot):
        self._check(lib.b200rn_wait(self._h, slot), "b200rn_wait")

    def time_forward(self, d_inputs, batch, iters):
        """d_inputs: list of device pointers (int) to [batch, C, H, W] fp32 batches, used round-robin."""
        arr = (_vp * len(d_inputs))(*d_inputs)
        ms = C.c_float()
        self._check(lib.b200rn_time_forward(self._h, arr, len(d_inputs), batch, iters, C.byref(ms)), "b200rn_time_forward")
        return ms.value

    def layer_count(self):
        return lib.b200rn_layer_count(self._h)

    def launch_count(self):
        return lib.b200rn_launch_count(self._h)

    def profile_layers(self, d_in, batch, iters=5):
        n = self.layer_count()
        ms = np.zeros(n, dtype=np.float32)
        kinds = (C.c_int * n)()
        flops = (C.c_double * n)()
        nbytes = (C.c_double * n)()
        self._check(lib.b200rn_profile_layers(self._h, d_in, batch, iters, fptr(ms), kinds, flops, nbytes),
                    "b200rn_profile_layers")
        out = []
        buf = C.create_string_buffer(2048)
        for i in range(n):
            lib.b200rn_layer_name(self._h, i, buf, 2048)
            out.append({"i": i, "kind": LAYER_KINDS[kinds[i]], "ms": float(ms[i]), "flops": flops[i],
                        "bytes": nbytes[i], "name": buf.value.decode()})
        return out
