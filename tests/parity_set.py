"""The network parity set of SURVEY.md 8(d): 256 images N(0,1), image i drawn with seed i, plus an all-zeros
image and the 3-float payload [1, 2, 3] that the reference's load generator really sends
(/root/reference/benchmark.py:21-24), which batchPredict zero-pads (src/inference_engine.cpp:157-159).

258 samples, returned ragged exactly as a caller of batchPredict would hand them over."""
import numpy as np

N_RANDOM = 256
SHAPE = (3, 224, 224)


def samples(n_random=N_RANDOM, shape=SHAPE):
    per = int(np.prod(shape))
    out = [np.random.RandomState(seed).randn(per).astype(np.float32) for seed in range(n_random)]
    out.append(np.zeros(per, dtype=np.float32))
    out.append(np.asarray([1.0, 2.0, 3.0], dtype=np.float32))
    return out


def padded(n_random=N_RANDOM, shape=SHAPE):
    """The same set as one [n, C, H, W] array (short samples zero-padded)."""
    per = int(np.prod(shape))
    rows = np.zeros((n_random + 2, per), dtype=np.float32)
    for i, v in enumerate(samples(n_random, shape)):
        rows[i, :v.size] = v
    return rows.reshape((n_random + 2,) + tuple(shape))
