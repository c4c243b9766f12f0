"""bench.py's reference arm (the reference's CPU path timed on the host: the unmodified reference function when
oracle/_ref/ has been made by oracle/make_ref.py, else the oracle port) on a small grid: one JSON line with the
contract's keys from rank 0, nothing from the other ranks."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(env_extra):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2",
                           "--warmup", "1", "--grid", "24", "--gpus", "2"], capture_output=True, text=True, env=env,
                          timeout=300)


def test_reference_arm_line():
    r = run({"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"})
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "fom_cell_updates_per_s" and d["unit"] == "cell-updates/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 2
    assert d["value"] > 0 and d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0,
                                           "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    made = all(os.path.exists(os.path.join(ROOT, "oracle", "_ref", f)) for f in ("hypernet2D.py", "rbf_utils.py"))
    assert cb["kind"] == ("reference" if made else "port")
    assert cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert cb["cores"] == len(os.sched_getaffinity(0))          # one process per host core
    assert d["gpu_launches"] == 0 and d["config"]["grid"] == 24


def test_reference_arm_other_ranks_are_silent():
    r = run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and r.stdout.strip() == ""
