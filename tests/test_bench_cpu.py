"""bench.py's contract, as far as it can be checked without a GPU: the reference arm (the CPU restatement of the
reference timed on the host cores) prints one JSON line with the keys the driver reads, and the product arm refuses
to run without a CUDA device instead of falling back to the CPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = ["--size", "512", "--stars", "2000", "--steps", "1", "--warmup", "1"]


def _run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, env=e, capture_output=True,
                          text=True, timeout=600)


def test_reference_arm_prints_the_contract_line():
    p = _run(["--impl", "reference", "--cpu-sample", "400"] + SMALL)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["unit"] == "stars/s"
    assert d["metric"].startswith("stars/sec") and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert d["sample_stars_per_step"] == 400 and "workload" in d["config"] and d["config"]["stars_total"] == 2000
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_without_work():
    p = _run(["--impl", "reference", "--gpus", "2"] + SMALL, env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert p.returncode == 0, p.stderr[-2000:]
    assert not [l for l in p.stdout.splitlines() if l.startswith("{")]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU refusal")
def test_product_arm_refuses_to_run_without_a_gpu():
    p = _run(SMALL)
    assert p.returncode != 0
    assert not [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert "CUDA" in (p.stdout + p.stderr)


def test_reference_arm_ignores_torchrun_omp_setting():
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the reference arm must still use every host core it may run on
    (SCALE_r01: it did not, and timed out at N = 2, 4, 8)."""
    ncpu = len(os.sched_getaffinity(0))
    if ncpu < 2:
        pytest.skip("single-core host")
    p = _run(["--impl", "reference", "--gpus", "2", "--cpu-sample", "400"] + SMALL,
             env={"OMP_NUM_THREADS": "1", "RANK": "0", "LOCAL_RANK": "0", "WORLD_SIZE": "2"})
    assert p.returncode == 0, p.stderr[-2000:]
    d = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][0])
    assert d["cpu_baseline"]["cores"] == ncpu > 1 and d["n_gpus"] == 2
    assert d["config"]["stars_total"] == 4000  # same keys and values as the product arm prints at N = 2
