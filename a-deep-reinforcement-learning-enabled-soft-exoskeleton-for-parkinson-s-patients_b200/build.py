"""Builds libexo_env.so (hand-written sm_100a CUDA + the C-ABI) in-tree with nvcc.

    python <package>/build.py [--force]

One object per translation unit, compiled in parallel; the result lands next to this file so it travels
with the repository snapshot. There is no alternative backend: if nvcc is missing the build fails.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.environ.get("EXO_LIB_OUT", os.path.join(HERE, "libexo_env.so"))
UNITS = ["exo_api.cu", "exo_step_f32.cu", "exo_step_f64.cu", "exo_motor_f32.cu", "exo_motor_f64.cu", "exo_reset_f32.cu", "exo_reset_f64.cu", "exo_lap.cu", "exo_policy.cu", "exo_policy_fp32.cu", "exo_td7.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"] + \
    os.environ.get("EXO_EXTRA_NVCC_FLAGS", "").split()      # tuning experiments only (e.g. -DEXO_STEP_THREADS=32)


def _sources():
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    out.append(os.path.join(HERE, "..", "include", "exo_env.h"))
    out.append(os.path.join(HERE, "..", "include", "exo_lap.h"))
    out.append(os.path.join(HERE, "..", "include", "exo_policy.h"))
    out.append(os.path.join(HERE, "..", "include", "exo_td7.h"))
    return out


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)

    def cc(unit):
        obj = os.path.join(OBJ, unit.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, unit), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (unit, r.stderr))
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=len(UNITS)) as ex:
        res = list(ex.map(cc, UNITS))
    if verbose:
        for _, log in res:
            sys.stderr.write(log)
    cmd = [NVCC, "-shared", "-o", LIB] + [o for o, _ in res] + ["-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
