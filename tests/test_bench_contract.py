"""bench.py's contract, checked on the CPU: the reference arm (`--impl reference`) runs the reference's own
per-particle code on the host cores and prints one JSON line with the keys the driver reads; non-zero ranks of a
torchrun launch exit 0 without work. (The GPU arm needs a B200 and is exercised by the driver itself.)
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(args, env=None):
    e = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        e.pop(k, None)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True,
                          env=e, timeout=600)


def test_reference_arm_prints_the_contract_line():
    r = run_bench(["--impl", "reference", "--cells", "16", "--cpu-cells", "16", "--steps", "1", "--warmup", "1"])
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "IMC particle-segments/sec" and d["unit"] == "segments/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["steps"] == 1 and d["warmup"] == 1
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "segments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0 and "workload" in d["config"]


def test_reference_arm_labels_are_true_under_torchrun_environment():
    """`torch.distributed.run` with nproc > 1 exports OMP_NUM_THREADS=1: the arm must set its thread count itself,
    report what OpenMP then says, name the twin it actually ran in `config`, and not map the CUDA library."""
    r = run_bench(["--impl", "reference", "--gpus", "2", "--cells", "32", "--cpu-cells", "16", "--steps", "1", "--warmup", "1"],
                  env={"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0", "OMP_NUM_THREADS": "1"})
    assert r.returncode == 0, r.stderr
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    want = len(os.sched_getaffinity(0))
    assert d["cpu_baseline"]["cores"] == want, (d["cpu_baseline"]["cores"], want)
    cfg = d["config"]
    assert cfg["cells"] == 16 and cfg["twin_of"]["cells"] == 32 and abs(cfg["photons_per_step"] - cfg["twin_of"]["photons_per_step"] / 8) <= 1
    assert cfg["shrunk_to_fit_time_budget"] is False and cfg["omp_num_threads_env"] == "1"
    assert "16^3 cells" in cfg["workload"] and cfg["cuda_library_mapped"] is False
    assert f"{want} threads" in d["cpu_baseline"]["sample"]


def test_reference_arm_other_ranks_exit_quietly():
    r = run_bench(["--impl", "reference", "--gpus", "2", "--cells", "16"], env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_our_arm_refuses_to_run_without_a_gpu():
    try:
        import torch
        if torch.cuda.is_available():
            import pytest
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    r = run_bench(["--cells", "16", "--steps", "1", "--warmup", "1"])
    assert r.returncode != 0 and "needs a CUDA device" in (r.stderr + r.stdout)
