"""Build libcombatseq.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

One object per translation unit under csrc/ (compiled in parallel), linked into one shared library.  An object is
rebuilt when its source, any header under csrc/ or the public header is newer.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
OUT = os.path.join(HERE, "libcombatseq.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-diag-suppress", "177"]
UNITS = ["combatseq.cu", "tu_apl.cu", "tu_glm.cu", "tu_mq.cu", "tu_mc.cu", "tu_screen.cu", "tu_sim.cu"]


def headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not f.endswith(".cu")]
    return hs + [os.path.join(HERE, "..", "include", "combatseq.h")]


def stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    hs = headers()
    jobs = []
    for u in UNITS:
        src, obj = os.path.join(CSRC, u), os.path.join(OBJ, u[:-3] + ".o")
        if force or stale(obj, [src] + hs):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            results = list(ex.map(compile_one, jobs))
        for src, r in results:
            if verbose or r.returncode != 0:
                sys.stderr.write("== %s\n%s%s" % (os.path.basename(src), r.stdout, r.stderr))
        bad = [src for src, r in results if r.returncode != 0]
        if bad:
            raise RuntimeError("nvcc failed for " + ", ".join(os.path.basename(b) for b in bad))
    objs = [os.path.join(OBJ, u[:-3] + ".o") for u in UNITS]
    if force or jobs or stale(OUT, objs):
        subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT] + objs + ["-ldl"], check=True)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(OUT)
