#!/usr/bin/env python
"""Recipe that compiles the REAL reference engine into oracle/_ref/ (test infrastructure only).

TEST INFRASTRUCTURE -- never imported by the product package ``prairiedog_b200``.

What is built (sources are compiled where they lie under /root/reference; nothing is copied):

* ``oracle/_ref/liblemongraph.so`` -- the reference's native graph engine:
  lemongraph/lib/lemongraph.c, lemongraph/lib/db.c and the vendored LMDB
  (lemongraph/deps/lmdb/libraries/liblmdb/{mdb.c,midl.c}), linked with zlib, exactly the
  object list of lemongraph/Makefile:11-14.  It is the CPU baseline of kind "reference"
  (driven by oracle/lg_driver.c) and it travels to the GPU box with the snapshot.
* ``oracle/_ref/_lemongraph_cffi*.so`` -- the cffi API-mode module that the reference's Python
  bindings import (lemongraph/LemonGraph/cffi_stubs.py:238-255).  Only usable in THIS container
  (the Python half of the reference stays under /root/reference); oracle/live_reference.py
  loads it to run the unmodified reference pipeline and to generate tests/golden/.
* ``oracle/_ref/lg_driver`` -- oracle/lg_driver.c linked against liblemongraph.so.

If /root/reference is absent (GPU box) this script does nothing and exits 0: the prebuilt
files that travelled with the snapshot are used as they are.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PDG_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
LG = os.path.join(REF, "lemongraph")
SOURCES = [
    "deps/lmdb/libraries/liblmdb/mdb.c",
    "deps/lmdb/libraries/liblmdb/midl.c",
    "lib/lemongraph.c",
    "lib/db.c",
]
INCLUDES = ["lib", "deps/lmdb/libraries/liblmdb"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_native(verbose=True):
    os.makedirs(OUT, exist_ok=True)
    lib = os.path.join(OUT, "liblemongraph.so")
    srcs = [os.path.join(LG, s) for s in SOURCES]
    if _newer(lib, srcs):
        cmd = ["gcc", "-O3", "-fPIC", "-shared", "-pthread", "-w"]
        cmd += ["-I" + os.path.join(LG, i) for i in INCLUDES]
        cmd += srcs + ["-lz", "-o", lib]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    drv_src = os.path.join(HERE, "lg_driver.c")
    drv = os.path.join(OUT, "lg_driver")
    if os.path.exists(drv_src) and _newer(drv, [drv_src, lib]):
        cmd = ["gcc", "-O2", "-w", "-I" + os.path.join(LG, "lib"),
               "-I" + os.path.join(LG, "deps/lmdb/libraries/liblmdb"),
               drv_src, "-o", drv, "-L" + OUT, "-llemongraph",
               "-Wl,-rpath,$ORIGIN", "-lz", "-pthread"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return lib


def build_cffi(verbose=True):
    """Build LemonGraph._lemongraph_cffi from the reference's own cdef (imported, not copied)."""
    import glob
    if glob.glob(os.path.join(OUT, "_lemongraph_cffi*.so")):
        return
    import cffi
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "_pdg_ref_cffi_stubs", os.path.join(LG, "LemonGraph", "cffi_stubs.py"))
    stubs = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(stubs)   # defines CDEF / C_HEADER_SRC / C_KEYWORDS (cffi_stubs.py:3-255)
    ffi = cffi.FFI()
    ffi.set_source(
        "_lemongraph_cffi", stubs.C_HEADER_SRC,
        sources=[os.path.join(LG, s) for s in SOURCES],
        include_dirs=[os.path.join(LG, i) for i in INCLUDES],
        libraries=["z"], extra_compile_args=["-O3", "-w"])
    ffi.cdef(stubs.CDEF)
    tmp = os.path.join(OUT, "_cffi_build")
    os.makedirs(tmp, exist_ok=True)
    built = ffi.compile(tmpdir=tmp, verbose=False)
    dst = os.path.join(OUT, os.path.basename(built))
    os.replace(built, dst)
    if verbose:
        print("built", dst)


def main():
    if not os.path.isdir(LG):
        print("reference not present at %s: using prebuilt oracle/_ref as is" % REF)
        return 0
    build_native()
    build_cffi()
    return 0


if __name__ == "__main__":
    sys.exit(main())
