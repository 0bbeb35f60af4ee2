"""Builds libb200km.so (sm_100a only) in-tree with nvcc.  `python -m spark_rapids_ml_nai_b200.build`."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libb200km.so")
SOURCES = ["api.cu", "kernels_generic.cu", "kernels_tc.cu", "kernels_screen.cu", "kernels_fixup.cu", "kernels_gemm.cu", "kernels_update.cu", "kernels_init.cu", "kernels_small.cu", "noalloc.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-O2,-Wall", "--expt-relaxed-constexpr",
]
if os.environ.get("B200KM_ENABLE_TRACE"):  # development builds: per-tile event clocks (tools/trace_view.py)
    NVCC_FLAGS.append("-DB200KM_ENABLE_TRACE")
if os.environ.get("B200KM_UPD_BATCH"):  # development builds: update-warp batch size experiments
    NVCC_FLAGS.append("-DB200KM_UPD_BATCH=" + os.environ["B200KM_UPD_BATCH"])


def _stale(out, deps):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(p) > t for p in deps)


def build(verbose=False, force=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(LIBDIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(HERE, "..", "include", "b200km.h"))
    # objects built with other flags (a development build with B200KM_ENABLE_TRACE ...) are not reused
    stamp = os.path.join(LIBDIR, ".flags")
    flags = " ".join(NVCC_FLAGS)
    if not os.path.exists(stamp) or open(stamp).read() != flags:
        force = True
    objs = []
    for src in SOURCES:
        path = os.path.join(CSRC, src)
        obj = os.path.join(LIBDIR, src.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [path] + headers):
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
            if verbose:
                print(" ".join(cmd))
            subprocess.check_call(cmd)
    if force or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-ldl"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    with open(stamp, "w") as f:
        f.write(flags)
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
