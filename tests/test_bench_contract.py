"""bench.py's reference arm (the one arm that runs without a GPU): ONE JSON line with the contract's keys, the same
`config` keys as the product's arm prints, every host core in use even when the launcher exported OMP_NUM_THREADS=1
(torchrun does for nproc > 1: round 1's N >= 2 reference lines ran on one thread)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_line():
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="2")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                        "--res", "64", "--spp", "4", "--bounces", "2"], capture_output=True, text=True, env=env, timeout=300, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == "Msamples/s" and d["unit"] == "Msamples/s" and d["n_gpus"] == 2 and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == d["value"] and cb["sample"]
    assert cb["cores"] == cb["host_cores"] >= 1 and cb["all_cores"] is True          # OMP_NUM_THREADS=1 was overridden
    sys.path.insert(0, ROOT)
    import bench
    want = bench.Spec(64, 64, 4, 2).config(2)
    assert d["config"] == want                                                      # the keys and values the product's arm prints


def test_reference_arm_other_ranks_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, env=env, timeout=120, cwd=ROOT)
    assert p.returncode == 0 and p.stdout.strip() == ""
