"""Builds stereo_match_b200/libb2stereo.so in-tree with nvcc for sm_100a (no other target)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libb2stereo.so")
SOURCES = ["api.cu", "sgbm_cost.cu", "sgbm_agg.cu", "sgbm_wave.cu", "sgbm_wave8.cu", "sgbm_wave_g8.cu", "sgbm_wave_g8s.cu", "sgbm_wave_g16.cu", "sgbm_wave_g32.cu", "sgbm_post.cu", "speckle.cu", "bm.cu", "geom.cu", "cloud.cu", "wls.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "--fmad=false",            # reprojection must not contract fp64 mul+add (bitwise parity)
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-Xptxas", "-v"]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    global OBJ, LIB
    extra = os.environ.get("B2S_NVCC_EXTRA", "").split()      # A/B builds: e.g. -DB2S_WAVE_PFD=10
    if extra:
        tag = "".join(c if c.isalnum() else "_" for c in "".join(extra))
        OBJ, LIB = os.path.join(HERE, "_obj" + tag), os.path.join(HERE, "libb2stereo%s.so" % tag)
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "sgbm_wave.cuh"), os.path.join(CSRC, "sgbm_wave_impl.cuh"), os.path.join(HERE, "..", "include", "b2stereo.h")]
    jobs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        r = subprocess.run([_nvcc()] + NVCC_FLAGS + extra + ["-c", src, "-o", obj], capture_output=True, text=True)
        log = obj.replace(".o", ".ptxas.log")
        with open(log, "w") as f:
            f.write(r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, r.stderr[-4000:]))
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4) or 1) as ex:
        list(ex.map(compile_one, jobs))
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if force or jobs or _stale(LIB, objs):
        subprocess.check_call([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB] + objs + ["-lcudart"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
