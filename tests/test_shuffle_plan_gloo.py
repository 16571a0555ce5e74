"""N>1 host logic on CPU: two gloo ranks run the product's shuffle plan (cyl_shuffle_plan, no GPU needed) exactly as
shuffle.cpp wires it and must reproduce the oracle's per-rank shuffle result. Also checks the rank plumbing of
`bench.py --impl reference` under torchrun (rank 0 prints the line, the other rank exits 0 without work)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(script_args, nproc, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port)] + script_args
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)


def test_shuffle_plan_world2_gloo():
    p = _torchrun([os.path.join(ROOT, "tests", "gloo_worker.py")], 2, 29731)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert "GLOO_OK world=2" in p.stdout


def test_shuffle_plan_world3_gloo():
    p = _torchrun([os.path.join(ROOT, "tests", "gloo_worker.py")], 3, 29732)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert "GLOO_OK world=3" in p.stdout


def test_reference_arm_rank_plumbing():
    p = _torchrun([os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--ref-rows", "20000"],
                  2, 29733)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1  # only rank 0 prints
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "rows/s" and d["value"] > 0 and d["cpu_baseline"]["kind"] in ("reference", "port")
