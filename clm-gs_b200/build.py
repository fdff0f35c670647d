"""Build libclmgs.so (sm_100a only) in-tree with nvcc.  No torch involved.

    python clm-gs_b200/build.py [--force] [--verbose]

Per-file flags: proj.cu is the decision-critical translation unit and is
compiled with -fmad=false (see csrc/gs_math.cuh); raster.cu is built with
-use_fast_math like the reference's extension (its parity bar is a tolerance,
not bit equality); everything else uses the default contraction rules and
explicit fast intrinsics where wanted.
"""
import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libclmgs.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "--expt-relaxed-constexpr", "-Xptxas", "-v"]
PER_FILE = {
    "proj.cu": ["-fmad=false"],
    "densify.cu": ["-fmad=false"],
    # the rasteriser follows the reference's own build (--use_fast_math, submodules/gsplat/setup.py:81): flush-to-zero
    # exp, approximate reciprocal -- its parity bar is 1e-4 on the image / 1e-3 on gradients, not bit equality
    "raster.cu": ["-use_fast_math"],
}
# host-only translation units go straight to g++ (no device code; nvcc's front end is not needed for them).
# -ffp-contract=off: the AVX2 and AVX-512 Adam rows must round identically (see cpu_adam.cpp).
HOST_CXX = ["-O3", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-mavx2", "-ffp-contract=off", "-pthread", "-g1"]


def nvcc():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found")
    return p


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def headers_mtime():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "clmgs.h"))
    hs.append(os.path.abspath(__file__))
    return max(os.path.getmtime(h) for h in hs)


def compile_one(src, force, verbose):
    obj = os.path.join(OBJ, src.rsplit(".", 1)[0] + ".o")
    sp = os.path.join(CSRC, src)
    if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(sp), headers_mtime()):
        return obj, ""
    if src.endswith(".cpp"):
        cmd = [shutil.which("g++") or "g++"] + HOST_CXX + ["-c", sp, "-o", obj]
    else:
        cmd = [nvcc()] + ARCH + COMMON + PER_FILE.get(src, []) + ["-x", "cu", "-c", sp, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, " ".join(cmd), r.stderr[-6000:]))
    return obj, r.stderr


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    srcs = sources()
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: compile_one(s, force, verbose), srcs))
    objs = [o for o, _ in results]
    log = "\n".join(l for _, l in results if l)
    if log:
        with open(os.path.join(OBJ, "ptxas.log"), "a") as f:
            f.write(log)
        if verbose:
            print(log)
    if force or not os.path.exists(LIB) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        cmd = [nvcc()] + ARCH + ["-shared", "-o", LIB] + objs + ["-Xlinker", "--no-undefined", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s" % r.stderr[-4000:])
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
