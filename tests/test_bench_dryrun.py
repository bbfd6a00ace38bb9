"""bench.py is the driver's contract; `--dry-run-cpu` runs the whole script (configs, engine calls, e2e loop, JSON
line) on the CPU path with tiny shapes so that a broken flag or key shows up here and not on the GPU box.  The line it
prints is marked `dry_run` - it is not a measurement."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.join(os.path.dirname(__file__), "..")
KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
        "dtype", "data", "config", "clocks", "e2e", "gpu_launches"}


def _run(args, nproc=1):
    cmd = [sys.executable]
    if nproc > 1:
        cmd += ["-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
                "--master-port", "29571"]
    cmd += [os.path.join(ROOT, "bench.py"), "--dry-run-cpu", "--steps", "3", "--warmup", "1", "--gpus", str(nproc)] + args
    p = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, p.stdout[-2000:]            # ONE json line, from rank 0
    if nproc == 1:       # and nothing else on stdout: whatever libraries print goes to stderr (torchrun adds its own banner)
        assert p.stdout.strip() == lines[0], p.stdout[-2000:]
    return json.loads(lines[0])


@pytest.mark.parametrize("args", [
    ["--config", "orr", "--batch", "64"],
    ["--config", "orr", "--batch", "64", "--wire", "f16", "--hyper", "solver=cg"],
    ["--config", "pa", "--batch", "64"],
    ["--config", "fgm", "--batch", "64"],
    ["--config", "kmeans", "--batch", "600"],
    ["--config", "mcpa", "--batch", "32"],
    ["--config", "mcpa", "--batch", "32", "--sparse-direct"],
])
def test_bench_line_contract(args):
    out = _run(args)
    assert KEYS <= set(out), sorted(KEYS - set(out))
    assert out["dry_run"] and out["metric"] == "records/sec" and out["n_gpus"] == 1 and out["steps"] == 3
    assert out["value"] > 0 and out["higher_is_better"] is True and out["scaling"] == "weak"
    e = out["e2e"]
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(e), e
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    if "--wire" in args:
        assert e["wire"] == "f16" and e["h2d_bytes_per_step"] == 64 * (1000 + 2) * 2
    if "--sparse-direct" in args:
        assert out["config"]["sparse_direct"] is True


def test_bench_reference_arm_and_two_ranks():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1"],
                       capture_output=True, text=True, cwd=ROOT, timeout=120)
    ref = json.loads(p.stdout.strip().splitlines()[-1])
    assert p.returncode == 0 and ref["impl"] == "reference" and "unavailable" in ref
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29573", os.path.join(ROOT, "bench.py"), "--impl", "reference",
                        "--gpus", "2"], capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert p.returncode == 0 and len([l for l in p.stdout.splitlines() if l.startswith("{")]) == 1      # rank 0 only
    out = _run(["--config", "orr", "--batch", "64", "--no-e2e"], nproc=2)       # gloo, world 2: one line, aggregate value
    assert out["n_gpus"] == 2 and out["config"]["global_batch"] == 128 and out["e2e"] is None
    out = _run(["--config", "orr", "--batch", "64", "--no-e2e", "--tc", "syncFusion=scatter"], nproc=2)
    assert out["config"]["sync_fusion"] == "scatter"


def test_graft_entry_build_and_smoke_script():
    """`build()` as the driver runs it on this CPU box, and the steps of `smoke()` on the CPU path."""
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    g.build()
    g.smoke("cpu")
