"""Builds libfluid_b200.so in-tree with nvcc for sm_100a (no torch, no JIT cache)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libfluid_b200.so")
SOURCES = ["fluid_sim.cu", "jfa.cu", "selftest.cu", "render.cu", "dormant.cu"]
HEADERS = ["multi.inl", "field.inl", "kernels_band.cuh", "band_common.cuh", "common.cuh", "radix_sort.cuh", "kernels_hash.cuh", "kernels_cells.cuh", "kernels_count.cuh", "kernels_sph.cuh", "sim_internal.h",
           os.path.join("..", "..", "include", "fluid_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "--fmad=true",  # exact paths use __f*_rn intrinsics explicitly
    "-shared", "-ldl",
]


def _find_nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libfluid_b200 has no prebuilt fallback and no CPU path")


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps += [os.path.normpath(os.path.join(CSRC, h)) for h in HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every CUDA source of the package into libfluid_b200.so (sm_100a only)."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = _find_nvcc()
    ccbin = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else (shutil.which("g++") or "g++")
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    extra = os.environ.get("FLUID_NVCC_EXTRA", "").split()  # tuning sweeps only (scripts/sweep_force.sh)
    cmd = [nvcc, "-ccbin", ccbin] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + srcs
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
