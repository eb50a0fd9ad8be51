"""Builds libsumob200.so (C ABI + engine host code in engine.cu, one object of kernels.cu per kernel set) in-tree with nvcc
for sm_100a.  No torch, no JIT cache."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsumob200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    # bit-exact fp64: no FMA contraction, IEEE division and square root (DESIGN.md "numerics")
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true",
    "-Xcompiler", "-fPIC,-ffp-contract=off", "-shared",
]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", f) for f in os.listdir(os.path.join(HERE, "..", "include"))]
    return any(os.path.getmtime(s) > t for s in srcs)


# kernel sets: kernels.cu compiled once per feature selection (see its header); the engine picks one per scenario
KERNEL_SETS = {"full": [], "krauss": ["-DSB2_NO_IDM"], "plain": ["-DSB2_NO_IDM", "-DSB2_NO_OBSERVERS"]}


def build(force=False, verbose=False, extra_flags=(), out=None):
    out = out or LIB
    if not force and out == LIB and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + list(extra_flags) + (["-Xptxas", "-v"] if verbose else [])
    objdir = os.path.join(HERE, "build", os.path.basename(out))
    os.makedirs(objdir, exist_ok=True)
    jobs = [([nvcc] + flags + ["-c", "-o", os.path.join(objdir, "engine.o"), os.path.join(CSRC, "engine.cu")])]
    for name, defs in KERNEL_SETS.items():
        jobs.append([nvcc] + flags + ["-DSB2_KSET=" + name] + defs + ["-c", "-o", os.path.join(objdir, "kernels_%s.o" % name), os.path.join(CSRC, "kernels.cu")])
    procs = [subprocess.Popen(j) for j in jobs]
    for p, j in zip(procs, jobs):
        if p.wait() != 0:
            raise subprocess.CalledProcessError(p.returncode, j)
    objs = [j[j.index("-o") + 1] for j in jobs]
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out] + objs)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
