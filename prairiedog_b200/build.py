"""Builds prairiedog_b200/libpdgpu.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

``python -m prairiedog_b200.build`` or ``__graft_entry__.build()``.  nvcc cross-compiles without
a GPU; the built ``.so`` is git-ignored but travels to the GPU box with the snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpdgpu.so")
SOURCES = ["api.cu", "ingest.cu", "ingest_batch.cu", "gindex.cu", "gbuild.cu", "ksort.cu", "msd3.cu", "msdlong.cu", "xbuild.cu", "build.cu", "export.cu", "stats.cu",
           "radix.cu", "scan.cu", "query.cu", "comm.cu", "filter.cu", "noalloc.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--cudart", "static"]


def _deps():
    d = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    d.append(os.path.join(HERE, "..", "include", "pdgpu.h"))
    return d


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in _deps())


def build_library(force=False, verbose=True, extra_flags=(), out=None, tag="", csrc=None):
    """extra_flags / out / tag / csrc: tuning variants (tools/tune_msd3.py) built next to the product library;
    csrc = another source directory (for example an earlier revision of the kernels, for A/B timing)."""
    csrc = csrc or CSRC
    if out is None:
        out = LIB
        if not force and not needs_build():
            return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for src in SOURCES:
        obj = os.path.join(HERE, "build", src.replace(".cu", tag + ".o"))
        cmd = [nvcc] + NVCC_FLAGS + list(extra_flags) + ["-c", os.path.join(csrc, src), "-o", obj]
        if verbose:
            print(" ".join(cmd))
        procs.append((src, subprocess.Popen(cmd)))
        objs.append(obj)
    for src, p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed on %s" % src)
    cmd = [nvcc, "-shared", "--cudart", "static", "-o", out] + objs + ["-ldl"]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    build_library(force="--force" in sys.argv)
