#!/usr/bin/env python3
"""bench.py -- ResNet-50 v2 images/s of the B200 backend (and, with --impl reference, of the CPU
restatement of the reference's path), one JSON line on stdout.

A "step" is one forward pass of the hot path over one batch of synthetic 3x224x224 images
(BASELINE.json metric: images/s at batch 32/256).  Launch:
    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
One replica per GPU, no data-path collective (the path shards by request: SURVEY 8e) -> "weak".

  value     whole-job images/s with the input batches already resident in HBM (kernel path only),
            timed with CUDA events on the engine's compute stream, max over ranks.
  e2e       the same metric through the public C-ABI call with HOST buffers: pinned host input ->
            cudaMemcpyAsync -> forward -> cudaMemcpyAsync -> pinned host output every step, using the
            double-buffered submit/wait entry points; wall clock, max over ranks.
  roofline  the conv kernels (all 47 tcgen05 conv launches of a step): algorithmic conv FLOPs per step over
            the sum of their CUDA-event durations, against the measured sustained bf16 peak.
  cpu_baseline  the CPU oracle (torch-CPU fp32, unfused; "port" -- ONNX Runtime is not on this image)
            on a bounded sample, all host threads.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "resnet50_v2_images_per_s"
UNIT = "images/s"
IMG = (3, 224, 224)
FLOP_PER_IMAGE = 8.178368512e9          # 2 x (4,087,136,256 conv + 2,048,000 FC MACs)
PARITY_BOUND = {"exact": 1e-3, "fp16": 0.2, "bf16": 1.5}   # max-abs logit error, same constants as tests/test_gpu_parity258.py


def model_path():
    """The synthetic ResNet-50 v2 ONNX fixture (seed 0), generated once per machine under /tmp."""
    from tools import make_resnet50v2_onnx as gen
    p = os.path.join(tempfile.gettempdir(), "b200rn_resnet50v2_seed0_%d.onnx" % os.getuid())
    if not os.path.exists(p):
        tmp = p + ".%d.tmp" % os.getpid()
        gen.write_model(tmp, "resnet50", seed=0)
        os.replace(tmp, p)
    return p


def kernel_source_hash():
    """sha256 over the CUDA sources of libb200rn.so: identifies the kernel set a committed ncu profile was taken on."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "distributed-inference-engine-cpp_b200", "csrc")
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".cpp", ".h")):
            with open(os.path.join(d, name), "rb") as f:
                h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        return {"tflops": m.get("bf16_tflops_sustained", m["bf16_tflops"]), "tflops_burst": m["bf16_tflops"],
                "hbm_gbs": m["hbm_gbs"], "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    except Exception:
        return {"tflops": 1400.0, "tflops_burst": 1590.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def host_threads():
    """Threads this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1 to every rank, which would run
    the CPU arm on one core; the arm runs on rank 0 alone, so it takes every core the process is allowed on."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def cpu_baseline_run(images, iters, threads=None):
    """The CPU oracle on `images` synthetic inputs, `iters` timed passes after one warm-up."""
    import numpy as np
    import torch
    from oracle import onnx_mini, resnet_oracle
    torch.set_num_threads(threads or host_threads())
    cores = torch.get_num_threads()
    model = onnx_mini.load(model_path())
    x = torch.from_numpy(resnet_oracle.synthetic_inputs(images, 1234, IMG))
    with torch.no_grad():
        resnet_oracle.run_graph(model, x[:1])
        t0 = time.perf_counter()
        for _ in range(iters):
            resnet_oracle.run_graph(model, x)
        dt = time.perf_counter() - t0
    return images * iters / dt, cores, dt


def reference_arm(args):
    """--impl reference: the reference's path on the host cores.  ONNX Runtime and the model file are
    absent from this image, so this is the oracle port (torch-CPU fp32 restatement), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    from oracle import onnx_mini, resnet_oracle
    sample = 8                                    # images per step: a bounded sample of the 256-image batch
    model = onnx_mini.load(model_path())
    x = torch.from_numpy(resnet_oracle.synthetic_inputs(sample, 1234, IMG))
    torch.set_num_threads(host_threads())         # torchrun sets OMP_NUM_THREADS=1; this arm owns the host
    cores = torch.get_num_threads()
    with torch.no_grad():
        for _ in range(max(1, min(args.warmup, 2))):
            resnet_oracle.run_graph(model, x)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            resnet_oracle.run_graph(model, x)
        dt = time.perf_counter() - t0
    v = sample * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "ResNet-50 v2 forward, synthetic 3x224x224, batch 256 per step on the GPU arm; "
                                   "CPU arm runs a bounded sample of %d images per step" % sample,
                       "batch": args.batch, "sample_images_per_step": sample},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "%d images x %d steps, torch-CPU fp32 unfused restatement (oracle/resnet_oracle.py); "
                                       "ONNX Runtime CPU is not runnable on this image" % (sample, args.steps)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def whole_job_value(world, steps, batch, max_ms):
    """images/s of the whole job: every rank ran `steps` batches; time is the max over ranks."""
    return world * steps * batch / (max_ms * 1e-3)


def dry_run_dist(args):
    """No GPU: each rank pretends its K steps took (10 + rank) ms and the ranks agree on the aggregate."""
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        dist.init_process_group("gloo")
        dist.barrier()
    t = torch.tensor([10.0 + rank], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps({"metric": METRIC, "unit": UNIT, "dry_run": True, "n_gpus": world, "steps": args.steps,
                          "max_ms": float(t.item()), "value": whole_job_value(world, args.steps, args.batch, float(t.item())),
                          "scaling": "weak"}), flush=True)
    return 0


def bind_to_gpu_numa_node(device):
    """One replica per GPU: run this rank on the CPUs of the GPU's own NUMA node before any pinned host buffer is
    allocated, so that the end-to-end leg's host->device copies read node-local memory (with eight ranks the copies
    add up to ~320 GB/s; un-placed buffers capped the 8-GPU e2e number at 178 GB/s).  Returns the node or None."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(device)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--mode", default="fp16", choices=["fp16", "bf16", "exact"])
    ap.add_argument("--exact-simt", action="store_true", help="exact mode on the fp32 CUDA-core conv kernel (B200RN_EXACT_SIMT=1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sustain-s", type=float, default=5.0, help="length of the sustained leg in seconds (0 = skip)")
    ap.add_argument("--layers-out", default="", help="write the per-layer roofline table (JSON) here")
    ap.add_argument("--dry-run-dist", action="store_true",
                    help="CPU-only check of the multi-rank plumbing (gloo): barrier, max-over-ranks, one line from rank 0")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return reference_arm(args)
    if args.dry_run_dist:
        return dry_run_dist(args)

    if args.exact_simt:
        os.environ["B200RN_EXACT_SIMT"] = "1"
    import numpy as np
    import torch
    import b200rn_loader
    pkg = b200rn_loader.load_package()       # raises if libb200rn.so is missing: no fallback

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 backend has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    saved_stdout = None
    if world > 1:
        import torch.distributed as dist
        # rank 0 prints ONE line on stdout.  NCCL writes its version banner there on the first collective (at
        # NCCL_DEBUG=WARN too), so until the JSON line is ready stdout is pointed at stderr.
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if rank == 0:
        model_path()
    barrier()
    path = model_path()
    B, K, W = args.batch, args.steps, args.warmup
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    eng = pkg.Engine(path, mode=args.mode, device=local, max_batch=max(B, 32))
    in_elems = eng.in_elems

    # ---- device-resident inputs: 4 distinct batches (4 x 154 MB at B=256 > 126 MB L2) ------------
    n_rot = 4
    g = torch.Generator(device="cuda").manual_seed(1234 + rank)
    d_in = [torch.randn(B, *IMG, generator=g, device="cuda", dtype=torch.float32) for _ in range(n_rot)]
    ptrs = [t.data_ptr() for t in d_in]
    eng.time_forward(ptrs, B, W)                          # warm-up (also builds TMA descriptors)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ms = eng.time_forward(ptrs, B, K)                     # EXACTLY K steps, CUDA events on the compute stream
    barrier()
    ms = max_over_ranks(ms)
    value = whole_job_value(world, K, B, ms)

    # ---- end to end through the C ABI with pinned host buffers ------------------------------------
    in_bytes, out_bytes = B * in_elems * 4, B * eng.out_elems * 4
    h_in = [pkg.lib.b200rn_host_alloc(in_bytes) for _ in range(2)]
    h_out = [pkg.lib.b200rn_host_alloc(out_bytes) for _ in range(2)]
    if not all(h_in) or not all(h_out):
        raise SystemExit("bench.py: pinned host allocation failed")
    rs = np.random.RandomState(99 + rank)
    for p in h_in:
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(B * in_elems,))
        a[:] = torch.randn(B * in_elems, generator=torch.Generator().manual_seed(int(rs.randint(1 << 30)))).numpy()
    # Rows 0..7 of host batch 0 are parity samples (six N(0,1) images with seeds 0..5, all-zeros, [1,2,3] zero-padded:
    # SURVEY 8d), so that the logits the timed e2e leg itself produces can be checked against the oracle below.
    n_par = min(8, B)
    par_x = np.zeros((n_par, in_elems), dtype=np.float32)
    for i in range(min(6, n_par)):
        par_x[i] = np.random.RandomState(i).randn(in_elems).astype(np.float32)
    if n_par == 8:
        par_x[7, :3] = (1.0, 2.0, 3.0)
    np.ctypeslib.as_array(C.cast(h_in[0], C.POINTER(C.c_float)), shape=(B * in_elems,))[:n_par * in_elems] = par_x.ravel()

    def e2e_steps(n):
        eng.submit_ptr(0, h_in[0], B, h_out[0])
        for i in range(1, n):
            eng.submit_ptr(i & 1, h_in[i & 1], B, h_out[i & 1])
            eng.wait((i - 1) & 1)
        eng.wait((n - 1) & 1)

    e2e_steps(W)
    barrier()
    t0 = time.perf_counter()
    e2e_steps(K)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    clocks = sampler.stop()
    e2e_value = world * K * B / e2e_s
    logits = np.ctypeslib.as_array(C.cast(h_out[0], C.POINTER(C.c_float)), shape=(B, eng.out_elems)).copy()
    finite = bool(np.isfinite(logits).all())
    # parity of the timed configuration: the e2e leg's own output rows against the CPU oracle (checker only)
    parity, parity_bad = None, 0.0
    if rank == 0:
        from oracle import resnet_oracle
        torch.set_num_threads(host_threads())
        ref = np.stack(resnet_oracle.OracleEngine(path).batch_predict(list(par_x)))
        perr = float(np.abs(logits[:n_par] - ref).max())
        bound = PARITY_BOUND[args.mode]
        parity = {"samples": n_par, "max_abs": perr, "bound": bound, "top1_agree": int((logits[:n_par].argmax(1) == ref.argmax(1)).sum()),
                  "checked": "rows 0..%d of the e2e leg's last batch-0 result vs oracle/resnet_oracle.py (fp32)" % (n_par - 1)}
        if not finite or not (perr <= bound):
            parity_bad = 1.0
            sys.stderr.write("bench.py: the timed configuration's logits are wrong (finite=%s, max-abs %.3e > %.1e): no throughput reported\n"
                             % (finite, perr, bound))
    if max_over_ranks(parity_bad) > 0:     # every rank leaves together
        raise SystemExit(3)

    # ---- batch 32 (configs[1]: the served batch cap), same two measurements ------------------------
    b32 = {}
    if B != 32:
        p32 = [t.data_ptr() for t in d_in]                # first 32 images of each rotating batch
        eng.time_forward(p32, 32, W)
        barrier()
        ms32 = max_over_ranks(eng.time_forward(p32, 32, K))
        lat = []
        x32 = np.ctypeslib.as_array(C.cast(h_in[0], C.POINTER(C.c_float)), shape=(B * in_elems,))[:32 * in_elems]
        for i in range(W + K):
            t1 = time.perf_counter()
            eng.infer_packed_ptr(x32.ctypes.data, 32, h_out[1])
            if i >= W:
                lat.append(time.perf_counter() - t1)
        lat.sort()
        b32 = {"value": world * K * 32 / (ms32 * 1e-3), "ms_per_step": ms32 / K,
               "e2e_sync_call_ms_p50": 1e3 * lat[int(len(lat) * 0.50)], "e2e_sync_call_ms_p99": 1e3 * lat[min(len(lat) - 1, int(len(lat) * 0.99))]}

    # ---- sustained leg: the same device-resident forward back to back for >= 5 s with its own clock record, so that
    # the sustained bf16 peak (measured over seconds at the power cap) has something comparable to divide -------------
    sustained = None
    if args.sustain_s > 0:
        per_step_ms = ms / K
        n_sus = max(K, int(args.sustain_s * 1e3 / per_step_ms) + 1)
        s2 = ClockSampler(local)
        barrier()
        s2.start()
        ms_sus = max_over_ranks(eng.time_forward(ptrs, B, n_sus))
        barrier()
        c2 = s2.stop()
        sustained = {"seconds": ms_sus * 1e-3, "steps": n_sus, "images_per_s": world * n_sus * B / (ms_sus * 1e-3),
                     "ms_per_step": ms_sus / n_sus, "sm_mhz_median": c2.get("sm_mhz"), "power_w_max": c2.get("power_w_max"),
                     "reasons": c2.get("reasons")}

    # ---- roofline of the dominant kernel: per-launch CUDA events, rank 0 ---------------------------
    roof, layers = None, []
    pk = peaks()
    if rank == 0:
        layers = eng.profile_layers(ptrs[0], B, iters=5)
        conv = [l for l in layers if l["kind"] == "Conv"]
        conv_ms = sum(l["ms"] for l in conv)
        conv_flops = sum(l["flops"] for l in conv)
        total_ms = sum(l["ms"] for l in layers)
        achieved = conv_flops / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else 0.0
        # DRAM bytes of the same conv launches from the committed ncu pass (profiles/, tools/ncu_summary.py).  The
        # profile is stamped with the hash of the kernel sources and the launch count it was taken on; a stale stamp
        # or another batch / mode gives null instead of a number that no longer describes these kernels.
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(ROOT, "profiles", "r2_conv_traffic.json")) as f:
                tj = json.load(f)
            if (B == tj.get("batch", 256) and args.mode == tj.get("mode", "fp16") and tj.get("kernel_source_sha256") == kernel_source_hash()
                    and tj.get("conv_launches_per_step") == len(conv) + sum(1 for l in layers if l["kind"] == "Gemm")):   # the profile counts the FC's tensor-core launch too
                traffic = tj["conv_dram_bytes_per_step"]
                traffic_src = ("profiles/r2_conv_traffic.json (ncu dram__bytes_read+write, sum over the %d conv launches of one step; "
                               "kernel sources %s)" % (len(conv), tj["kernel_source_sha256"][:12]))
            else:
                traffic_src = "profiles/r2_conv_traffic.json is stale for this build / configuration: not reported"
        except Exception:
            pass
        # frac divides by the SUSTAINED bf16 peak (seconds-long cuBLAS loop at the power cap, ~1.34 GHz); the timed
        # region here is a sub-second burst at up to 1.965 GHz, for which the burst peak is the like-for-like
        # denominator: frac_burst.  frac_sustained_leg divides the >= 5 s leg's whole-step rate by the sustained peak.
        if args.mode == "exact":
            # exact mode multiplies in TF32 (half the bf16 rate; no separate measured peak on this pool) and issues three
            # TF32 products per algorithmic product, so frac <= 1/3 by construction
            pk = dict(pk, tflops=pk["tflops"] / 2, tflops_burst=pk["tflops_burst"] / 2,
                      source=pk["source"] + " / 2 for kind::tf32; 3 TF32 MMAs per algorithmic product")
        roof = {"bound": "tensor", "kernel": "tcgen05 conv kernels (%d launches/step: conv_pair, conv_pairwide, conv_wide, conv_tail, conv_dualm, conv_umma, conv3x3_halo, conv3x3_halo2, stem7x7)" % len(conv),
                "achieved": achieved, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": achieved / pk["tflops"],
                "peak_kind": "bf16_tflops_sustained", "peak_burst": pk["tflops_burst"], "frac_burst": achieved / pk["tflops_burst"],
                "frac_sustained_leg": (sustained["images_per_s"] / world * FLOP_PER_IMAGE / 1e12 / pk["tflops"]) if sustained else None,
                "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_bytes": sum(l["bytes"] for l in conv),
                "peak_source": pk["source"], "conv_share_of_step": conv_ms / total_ms if total_ms else None,
                "conv_ms_per_step": conv_ms, "all_layers_ms_per_step": total_ms}
        # the memory-bound side, against the measured copy bandwidth
        mem = [l for l in layers if l["kind"] not in ("Conv", "Gemm")]
        for l in layers:
            l["tflops"] = l["flops"] / (l["ms"] * 1e-3) / 1e12 if l["ms"] > 0 else 0.0
            l["gbs"] = l["bytes"] / (l["ms"] * 1e-3) / 1e9 if l["ms"] > 0 else 0.0
        roof["memory_bound_layers"] = [{"kind": l["kind"], "ms": l["ms"], "GB/s": l["gbs"], "frac_of_hbm": l["gbs"] / pk["hbm_gbs"]} for l in mem]
        if args.layers_out:
            with open(args.layers_out, "w") as f:
                json.dump({"batch": B, "mode": args.mode, "peaks": pk, "layers": layers}, f, indent=1)

    # ---- CPU baseline: bounded sample on rank 0 at N=1 ---------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores, dt = cpu_baseline_run(images=32, iters=30)
        # second row: the reference's own thread setting (SetIntraOpNumThreads(4), src/inference_engine.cpp:19), and
        # BASELINE configs[0] (batch 1) at both thread counts -- all bounded
        v4, c4, dt4 = cpu_baseline_run(images=32, iters=4, threads=4)
        v1, c1, dt1 = cpu_baseline_run(images=1, iters=12)
        v14, c14, dt14 = cpu_baseline_run(images=1, iters=8, threads=4)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "32 images x 30 passes (%.1f s) of the same synthetic workload; torch-CPU fp32 unfused restatement "
                         "(oracle/resnet_oracle.py) -- ONNX Runtime CPU is not runnable on this image" % dt,
               "rows": [{"batch": 32, "threads": cores, "images_per_s": v, "seconds": dt},
                        {"batch": 32, "threads": c4, "images_per_s": v4, "seconds": dt4,
                         "note": "the reference's SetIntraOpNumThreads(4), src/inference_engine.cpp:19"},
                        {"batch": 1, "threads": c1, "images_per_s": v1, "seconds": dt1, "note": "BASELINE configs[0]"},
                        {"batch": 1, "threads": c14, "images_per_s": v14, "seconds": dt14, "note": "BASELINE configs[0] at 4 threads"}]}

    for p in h_in + h_out:
        pkg.lib.b200rn_host_free(p)
    launches = eng.launch_count()
    eng.close()
    if dist is not None:
        dist.destroy_process_group()
    if rank != 0:
        return 0
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp16": "fp16", "bf16": "bf16", "exact": "f32"}[args.mode], "data": "synthetic",
            "config": {"workload": "offline batched ResNet-50 v2 forward (BASELINE configs[3] at batch %d; batch 32 = configs[1] "
                                   "cap reported under batch32), synthetic N(0,1) 3x224x224 fp32 NCHW inputs, random-init "
                                   "weights of the resnet50-v2-7 topology" % B,
                       "batch": B, "mode": args.mode + ("-simt" if args.exact_simt else ""), "replicas": world, "parallelism": "replica per GPU, no collective", "numa_node_rank0": numa,
                       "l2": "inputs larger than L2: %d rotating device batches of %.0f MB" % (n_rot, in_bytes / 1e6),
                       "flop_per_image": FLOP_PER_IMAGE},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": out_bytes,
                    "api": "b200rn_submit_packed/b200rn_wait (double-buffered), pinned host buffers", "outputs_finite": finite,
                    "parity": parity},
            "gpu_launches": launches * K, "launches_per_step": launches,
            "tflops_whole_step": value / world * FLOP_PER_IMAGE / 1e12,
            "roofline": roof, "sustained": sustained, "cpu_baseline": cpu, "clocks": clocks, "batch32": b32}
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
