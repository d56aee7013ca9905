"""bench.py's driver-facing contract, as far as it can be checked without a GPU: the reference arm (the oracle port on the
host cores) prints ONE JSON line with the agreed keys, only rank 0 of a multi-rank launch runs it, and the product arm
refuses to run without a CUDA device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None):
    e = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        e.pop(k, None)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, env=e, capture_output=True, text=True, timeout=600)


def test_reference_arm_line():
    p = _run(["--impl", "reference", "--workload", "c1", "--steps", "2", "--warmup", "1", "--rounds", "20", "--cpu-budget-s", "5"])
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "cv_fits_per_sec" and d["unit"] == "fits/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["unit"] == "fits/s" and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "fits/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_runs_on_rank_zero_only():
    p = _run(["--impl", "reference", "--gpus", "2", "--workload", "c1", "--steps", "1", "--warmup", "0", "--rounds", "20"],
             env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert p.returncode == 0 and p.stdout.strip() == ""


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the behaviour on a box without a GPU")
def test_product_arm_refuses_to_run_without_a_gpu():
    p = _run(["--workload", "c1", "--steps", "1", "--warmup", "0", "--rounds", "5"])
    assert p.returncode != 0
    assert "CUDA" in (p.stderr + p.stdout)
    assert not any(l.startswith("{") for l in p.stdout.splitlines())      # no bench line from a CPU fallback
