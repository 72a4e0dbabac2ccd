"""The N>1 plumbing of bench.py on CPU (gloo, world_size 2): the path shards by request, one replica per
rank and no data-path collective, so what there is to test is the launch contract -- every rank joins the
barrier, the time is the max over ranks, and rank 0 alone prints one JSON line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(extra, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2"] + extra
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    return [l for l in r.stdout.splitlines() if l.startswith("{")]


def test_two_ranks_agree_and_rank0_prints_once():
    lines = _torchrun(["--steps", "4", "--warmup", "3", "--batch", "256", "--dry-run-dist"], 29541)
    assert len(lines) == 1
    j = json.loads(lines[0])
    # rank 1 "took" 11 ms, rank 0 10 ms: the job time is the slower rank's, the work is both ranks'
    assert j["n_gpus"] == 2 and j["max_ms"] == 11.0 and j["scaling"] == "weak"
    assert abs(j["value"] - 2 * 4 * 256 / 11e-3) < 1e-6


def test_reference_arm_under_torchrun_runs_on_rank0_only():
    lines = _torchrun(["--impl", "reference", "--steps", "1", "--warmup", "1"], 29542)
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["cpu_baseline"]["kind"] == "port" and j["value"] > 0
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["unit"] == "images/s"
    # torchrun exports OMP_NUM_THREADS=1 (2 here) to its ranks; the arm must still use every core it is allowed on
    assert j["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
