"""TEST INFRASTRUCTURE (oracle/): recipe for oracle/_ref/ - a verbatim, git-ignored copy of the reference tree.

The reference is 5 Python files and 9 scene files with no build step.  /root/reference exists only in the build
container; the GPU box gets whatever lies under the repo (oracle/_ref/ is git-ignored but travels with gpurun), so
this copies the UNMODIFIED files there - nothing is edited, nothing enters the history - for one purpose: timing the
reference's own `main.py` (its multiprocessing.Pool and all) on the GPU box's host cores next to the GPU numbers
(SURVEY.md 8(d) CPU baseline (i)).  `run_main()` launches it as a subprocess exactly as its README /
header comment says (main.py:3-4), with oracle/stubs on PYTHONPATH for the three imports that are not arithmetic
(ipdb, colorama, pyqtgraph: debug prompt, coloured text, OpenGL viewer).

    python oracle/make_ref.py            # copy (no-op when /root/reference is absent)
"""
import os
import shutil
import subprocess
import sys
import time

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
REF_DST = os.path.join(_HERE, "_ref", "pathtracerpython")
STUBS = os.path.join(_HERE, "stubs")
FILES = ["main.py", "utils.py", "scene_reader.py", "vector.py", "plot.py"]


def make(force=False):
    """Copy the reference's sources and scene files into oracle/_ref/pathtracerpython/.  Returns the path, or None when
    neither the source tree nor an earlier copy exists."""
    if not os.path.isdir(REF_SRC):
        return REF_DST if available() else None
    os.makedirs(os.path.join(REF_DST, "objs"), exist_ok=True)
    for f in FILES:
        dst = os.path.join(REF_DST, f)
        if force or not os.path.exists(dst):
            shutil.copyfile(os.path.join(REF_SRC, f), dst)
    for f in sorted(os.listdir(os.path.join(REF_SRC, "objs"))):
        dst = os.path.join(REF_DST, "objs", f)
        if force or not os.path.exists(dst):
            shutil.copyfile(os.path.join(REF_SRC, "objs", f), dst)
    return REF_DST


def available():
    return all(os.path.exists(os.path.join(REF_DST, f)) for f in FILES) and os.path.exists(os.path.join(REF_DST, "objs", "cornellroom.sdl"))


def run_main(n_rays, n_bounces, out_png=None, timeout=900):
    """`main.py objs/cornellroom.sdl --out F -r N -b N` (main.py:125-139, 165-293), unmodified, all host cores through its
    own Pool(cpu_count()).  -> dict(seconds, samples, samples_per_s, cores)."""
    if not available():
        raise RuntimeError("oracle/_ref is empty: run oracle/make_ref.py where /root/reference is mounted")
    out_png = out_png or os.path.join(REF_DST, "_bench_out.png")
    env = dict(os.environ)
    env["PYTHONPATH"] = STUBS + os.pathsep + env.get("PYTHONPATH", "")
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):     # torchrun sets OMP_NUM_THREADS=1
        env.pop(k, None)
    t = time.perf_counter()
    p = subprocess.run([sys.executable, "main.py", "objs/cornellroom.sdl", "--out", out_png, "-r", str(n_rays), "-b", str(n_bounces)],
                       cwd=REF_DST, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, timeout=timeout)
    dt = time.perf_counter() - t
    if p.returncode != 0:
        raise RuntimeError("reference main.py failed: " + p.stderr[-800:])
    samples = 40 * 40 * n_rays           # objs/cornellroom.sdl: size 40 40
    return {"seconds": dt, "samples": samples, "samples_per_s": samples / dt, "cores": os.cpu_count(),
            "command": f"main.py objs/cornellroom.sdl --out F -r {n_rays} -b {n_bounces}"}


if __name__ == "__main__":
    print(make(force="--force" in sys.argv))
