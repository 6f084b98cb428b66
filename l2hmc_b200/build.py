"""Build the C-ABI CUDA library (csrc/*.cu -> csrc/libl2hmc_b200.so) for sm_100a with nvcc.

The library has no torch / Python dependency: plain `extern "C"` entry points declared in
include/l2hmc_b200.h.  It is built IN-TREE so the .so travels to the GPU box with the repo
snapshot.  `python -m l2hmc_b200.build [--force] [--verbose]`.
"""
import glob
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libl2hmc_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    # the update/stencil kernels keep the reference's per-op fp32 rounding; FMAs are explicit
    "-fmad=false",
    "-Xcompiler", "-fPIC", "-shared",
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h")) + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def _includes(path, seen=None):
    """Transitive closure of the quoted #include files of `path` (for per-object staleness)."""
    seen = set() if seen is None else seen
    try:
        text = open(path).read()
    except OSError:
        return seen
    for inc in re.findall(r'^\s*#\s*include\s+"([^"]+)"', text, flags=re.M):
        full = os.path.normpath(os.path.join(os.path.dirname(path), inc))
        if full not in seen and os.path.exists(full):
            seen.add(full)
            _includes(full, seen)
    return seen


def _obj_stale(src, obj):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [src, os.path.abspath(__file__)] + sorted(_includes(src)))


def build(force=False, verbose=False):
    """Compile every CUDA source into one shared library.  Returns the library path."""
    if not force and not _stale():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libl2hmc_b200.so")
    objs = []
    procs = []
    for src in sources():  # one nvcc per translation unit, in parallel
        obj = src[:-3] + ".o"
        objs.append(obj)
        if not force and not verbose and not _obj_stale(src, obj):
            continue
        cmd = [nvcc] + [f for f in NVCC_FLAGS if f != "-shared"] + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", src, "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, pr in procs:
        out, _ = pr.communicate()
        if pr.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {os.path.basename(src)}\n{out}\n")
        failed |= pr.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed building libl2hmc_b200.so")
    tmp = LIB + ".tmp"
    subprocess.check_call([nvcc, "-shared", "-o", tmp] + objs + ["-lcudart"])
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
