"""In-tree nvcc build of librtb200.so for sm_100a (cross-compiles without a GPU).

    python unity-raytracer_b200/build.py [--force] [--verbose]

-fmad=false : the parity contract needs strict binary32 (the reference's C# does not contract a*b+c); the only FMAs are
              the explicit ones in the conservative BVH slab test.  -prec-div/-prec-sqrt are nvcc defaults, spelled out.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.environ.get("RTB200_OUT", os.path.join(HERE, "librtb200.so"))
SOURCES = ["rt_build.cu", "rt_render.cu", "rt_abi.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
EXTRA = os.environ.get("RTB200_EXTRA_NVCC_FLAGS", "").split()      # development builds only (e.g. -DRT_TILE_TIMING)
FLAGS = EXTRA + ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false", "-prec-div=true",
         "-prec-sqrt=true", "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "-Xptxas", "-v", "-I", os.path.join(HERE, "..", "include")]


def _newest_dep():
    t = 0.0
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            t = max(t, os.path.getmtime(os.path.join(root, f)))
    return max(t, os.path.getmtime(os.path.abspath(__file__)))


def build(force=False, verbose=False):
    build_microbench(force)
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= _newest_dep():
        return OUT
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, obj, r

    with ThreadPoolExecutor(len(SOURCES)) as ex:
        results = list(ex.map(compile_one, SOURCES))
    log = []
    for src, obj, r in results:
        log.append(f"==== {src}\n{r.stdout}{r.stderr}")
        if r.returncode != 0:
            sys.stderr.write("\n".join(log))
            raise RuntimeError(f"nvcc failed on {src}")
    with open(os.path.join(objdir, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    cmd = [NVCC, "-shared", "-o", OUT] + [o for _, o, _ in results] + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    return OUT


MICROBENCH_OUT = os.path.join(HERE, "librtb200_microbench.so")


def build_microbench(force=False):
    """Measurement infrastructure (L2 read bandwidth, FP32 FMA rate) for bench.py's rooflines; not part of librtb200.so."""
    src = os.path.join(CSRC, "rt_microbench.cu")
    if not force and os.path.exists(MICROBENCH_OUT) and os.path.getmtime(MICROBENCH_OUT) >= os.path.getmtime(src):
        return MICROBENCH_OUT
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-shared",
           src, "-o", MICROBENCH_OUT]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed on rt_microbench.cu")
    return MICROBENCH_OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
