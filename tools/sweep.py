#!/usr/bin/env python3
"""BASELINE configs[3]: offline throughput sweep over batch 1..256 on one B200, per arithmetic mode.
For every B in {1,2,4,8,16,32,64,128,256}:
  back_to_back   K forwards queued on the engine's compute stream, CUDA events around all of them (b200rn_time_forward),
                 four rotating device inputs; at small B the working set (weights 51 MB + activations) stays in the
                 126 MB L2, as it would under sustained serving load -- so a second figure is taken:
  l2_flushed     single forwards, each timed by its own event pair after a 256 MB device write has evicted L2
                 (median of 9)
  e2e            the same batches from pinned host memory through b200rn_submit_packed / b200rn_wait, wall clock
One JSON object on stdout (and --out).   usage: sweep.py [--modes fp16,exact] [--out profiles/r2_sweep.json]"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import b200rn_loader  # noqa: E402

BATCHES = [1, 2, 4, 8, 16, 32, 64, 128, 256]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--modes", default="fp16,exact")
    ap.add_argument("--out", default="")
    ap.add_argument("--seconds", type=float, default=0.25, help="length of each back-to-back leg")
    a = ap.parse_args()
    import numpy as np
    import torch
    pkg = b200rn_loader.load_package()
    torch.cuda.set_device(0)
    path = bench.model_path()
    pk = bench.peaks()
    g = torch.Generator(device="cuda").manual_seed(77)
    d_in = [torch.randn(256, *bench.IMG, generator=g, device="cuda", dtype=torch.float32) for _ in range(4)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {"gpu": torch.cuda.get_device_name(0), "peaks": pk, "flop_per_image": bench.FLOP_PER_IMAGE, "modes": {}}
    for mode in a.modes.split(","):
        eng = pkg.Engine(path, mode=mode, max_batch=256)
        in_elems, out_elems = eng.in_elems, eng.out_elems
        h_in = [pkg.lib.b200rn_host_alloc(256 * in_elems * 4) for _ in range(2)]
        h_out = [pkg.lib.b200rn_host_alloc(256 * out_elems * 4) for _ in range(2)]
        for p in h_in:
            arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(256 * in_elems,))
            arr[:] = np.random.RandomState(3).randn(256 * in_elems).astype(np.float32)
        rows = []
        for B in BATCHES:
            ptrs = [t.data_ptr() for t in d_in]                  # the first B images of each rotating batch
            eng.time_forward(ptrs, B, 8)                          # eager run, graph capture, replays
            est = eng.time_forward(ptrs, B, 8) / 8
            K = max(16, int(a.seconds * 1e3 / est))
            ms = eng.time_forward(ptrs, B, K) / K
            fl = []
            for i in range(9):
                flush.fill_(i)
                torch.cuda.synchronize()
                fl.append(eng.time_forward([ptrs[i % 4]], B, 1))
            fl.sort()

            def e2e(n):
                eng.submit_ptr(0, h_in[0], B, h_out[0])
                for i in range(1, n):
                    eng.submit_ptr(i & 1, h_in[i & 1], B, h_out[i & 1])
                    eng.wait((i - 1) & 1)
                eng.wait((n - 1) & 1)

            e2e(4)
            n = max(8, int(a.seconds * 1e3 / est))
            t0 = time.perf_counter()
            e2e(n)
            dt = time.perf_counter() - t0
            ips = B / (ms * 1e-3)
            tf = ips * bench.FLOP_PER_IMAGE / 1e12
            rows.append({"batch": B, "ms_per_step": ms, "images_per_s": ips, "tflops": tf, "frac_of_sustained_bf16_peak": tf / pk["tflops"],
                         "frac_of_burst_bf16_peak": tf / pk["tflops_burst"], "l2_flushed_ms_median": fl[len(fl) // 2],
                         "l2_flushed_images_per_s": B / (fl[len(fl) // 2] * 1e-3), "e2e_images_per_s": n * B / dt, "steps": K})
            sys.stderr.write("%s B=%d: %.3f ms  %.0f img/s  (flushed %.3f ms, e2e %.0f img/s)\n" % (mode, B, ms, ips, fl[len(fl) // 2], n * B / dt))
        out["modes"][mode] = rows
        for p in h_in + h_out:
            pkg.lib.b200rn_host_free(p)
        eng.close()
    text = json.dumps(out, indent=1)
    if a.out:
        with open(a.out, "w") as f:
            f.write(text + "\n")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
