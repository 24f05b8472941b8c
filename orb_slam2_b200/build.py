"""Builds orb_slam2_b200/liborbx.so (the C-ABI library) in-tree with nvcc for sm_100a.

Flags that are part of the bit-exactness contract (DESIGN.md hazards H2/H3): --fmad=false,
-prec-div=true, -prec-sqrt=true, never --use_fast_math; host code -ffp-contract=off.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "liborbx.so")
SOURCES = ["orbx_api.cu", "k_pyramid.cu", "k_fast.cu", "k_octree.cu", "k_describe.cu", "k_blur.cu", "k_stereo.cu", "k_match.cu", "k_frame.cu", "k_bow.cu", "k_project.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--fmad=false", "-prec-div=true", "-prec-sqrt=true",
              "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-Wall", "-shared", "-cudart", "static"]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "orbx.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("ORBX_NVCC_EXTRA", "").split()            # experiments only, e.g. -DORBX_PHASE_CLOCKS (phase timings printed by the latency kernels)
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building liborbx.so")
    if verbose:
        sys.stderr.write(res.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force=True, verbose="-v" in sys.argv))
