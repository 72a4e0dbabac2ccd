#!/usr/bin/env python3
"""Bare host->device copy ceiling of the box: N ranks (one per GPU, launched like bench.py under torch.distributed.run)
each loop cudaMemcpyAsync of one 256-image fp32 batch (154 MB) from pinned host memory to device memory, nothing else
running -- what the end-to-end leg of bench.py could reach at most if the forward were free.  One cudaMemcpyAsync per
copy (or per chunk); rank 0 prints one JSON object.

    python tools/h2d_ceiling.py                                   # N = 1
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/h2d_ceiling.py

Variants per run (each its own 1.5 s leg, same barrier discipline):
  pinned        cudaHostAlloc default
  pinned_wc     cudaHostAllocWriteCombined (host writes bypass the cache; DMA reads need no snoop)
  chunks8       the batch as eight 19 MB copies back to back
  two_streams   two copies in flight on two streams (two transfer slots)
Host buffers are allocated after the rank is bound to the GPU's NUMA node (bench.bind_to_gpu_numa_node)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    from cuda.bindings import runtime as rt
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    numa = bench.bind_to_gpu_numa_node(local) if world > 1 else None
    nbytes = 256 * 3 * 224 * 224 * 4
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    dev2 = torch.empty(nbytes, dtype=torch.uint8, device="cuda")

    def ck(r):
        if isinstance(r, tuple):
            err, *rest = r
        else:
            err, rest = r, []
        if int(err) != 0:
            raise SystemExit("CUDA error %s" % err)
        return rest[0] if len(rest) == 1 else rest

    h_def = ck(rt.cudaHostAlloc(nbytes, rt.cudaHostAllocDefault))
    h_def2 = ck(rt.cudaHostAlloc(nbytes, rt.cudaHostAllocDefault))
    h_wc = ck(rt.cudaHostAlloc(nbytes, rt.cudaHostAllocWriteCombined))
    for h in (h_def, h_def2, h_wc):
        ck(rt.cudaMemset(dev.data_ptr(), 1, 1))
        ck(rt.cudaMemcpy(h, dev.data_ptr(), nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost))   # touch every page
    s1 = ck(rt.cudaStreamCreateWithFlags(rt.cudaStreamNonBlocking))
    s2 = ck(rt.cudaStreamCreateWithFlags(rt.cudaStreamNonBlocking))
    H2D = rt.cudaMemcpyKind.cudaMemcpyHostToDevice

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def leg(issue, seconds=1.5):
        """issue(i) queues one batch's worth of copies; returns aggregate and this rank's GB/s."""
        for i in range(3):
            issue(i)
        barrier()
        n = 0
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < seconds:
            for _ in range(4):
                issue(n)
                n += 1
            ck(rt.cudaStreamSynchronize(s1))
            ck(rt.cudaStreamSynchronize(s2))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        mine = n * nbytes / dt / 1e9
        t = torch.tensor([mine], dtype=torch.float64, device="cuda")
        lo = t.clone()
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        barrier()
        return {"aggregate_GBps": float(t.item()), "min_rank_GBps": float(lo.item()), "rank0_GBps": mine,
                "images_per_s_equiv": float(t.item()) * 1e9 / (3 * 224 * 224 * 4)}

    out = {"n_gpus": world, "bytes_per_copy": nbytes, "numa_node_rank0": numa, "host_cpus": len(os.sched_getaffinity(0))}
    out["pinned"] = leg(lambda i: ck(rt.cudaMemcpyAsync(dev.data_ptr(), h_def, nbytes, H2D, s1)))
    out["pinned_wc"] = leg(lambda i: ck(rt.cudaMemcpyAsync(dev.data_ptr(), h_wc, nbytes, H2D, s1)))
    c8 = nbytes // 8

    def chunks(i):
        for k in range(8):
            ck(rt.cudaMemcpyAsync(dev.data_ptr() + k * c8, h_def + k * c8, c8, H2D, s1))
    out["chunks8"] = leg(chunks)

    def two(i):
        if i & 1:
            ck(rt.cudaMemcpyAsync(dev2.data_ptr(), h_def2, nbytes, H2D, s2))
        else:
            ck(rt.cudaMemcpyAsync(dev.data_ptr(), h_def, nbytes, H2D, s1))
    out["two_streams"] = leg(two)
    if dist is not None:
        dist.destroy_process_group()
        sys.stdout.flush()
        os.dup2(saved, 1)
    if rank == 0:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
