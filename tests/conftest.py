import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
ASSETS = os.path.join(ROOT, "pathtracerpython_b200", "scenes")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


# ---- measured divergence of the FP32 product path from the FP64 oracle (SURVEY.md 8c check 2: target < 100 ppm).
# Every FP32 parity test reports what it measured, not only that it stayed under its limit; the summary is printed at
# the end of the run (the driver's pytest tail) and written to gpurun_out/divergence.json when that directory exists.
MEASURED = []


def record_divergence(name, divergent, total, limit):
    """divergent path-bounces (discrete history differs from the oracle's) out of `total`; returns the fraction."""
    frac = float(divergent) / max(1, int(total))
    MEASURED.append({"test": name, "divergent": int(divergent), "path_bounces": int(total), "ppm": 1e6 * frac, "limit_ppm": 1e6 * limit})
    return frac


def pytest_terminal_summary(terminalreporter):
    if not MEASURED:
        return
    terminalreporter.write_line("FP32 path-bounces whose discrete history differs from the FP64 oracle's (measured / limit):")
    for m in MEASURED:
        terminalreporter.write_line(f"  {m['test']}: {m['divergent']} of {m['path_bounces']} = {m['ppm']:.1f} ppm (limit {m['limit_ppm']:.0f} ppm)")
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        import json
        with open(os.path.join(out, "divergence.json"), "w") as f:
            json.dump(MEASURED, f, indent=1)
