"""ctypes wrapper of oracle/pdg_oracle.c (the plain-C CPU restatement of the hot path).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs import this.  The product package never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "pdg_oracle.c")
LIB = os.path.join(HERE, "_build", "libpdgoracle.so")
_lib = None


def build(force=False):
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", SRC, "-o", LIB])
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(LIB)
        vp, u64, i64, u32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int64, ctypes.c_uint32
        L.orc_new.restype = vp
        L.orc_new.argtypes = [ctypes.c_int]
        L.orc_free.argtypes = [vp]
        L.orc_add_fasta.restype = i64
        L.orc_add_fasta.argtypes = [vp, ctypes.c_char_p, u32]
        L.orc_add_fasta_mem.restype = i64
        L.orc_add_fasta_mem.argtypes = [vp, vp, u64, u32]
        L.orc_build.argtypes = [vp]
        L.orc_counts.argtypes = [vp, vp, vp, vp]
        L.orc_index.argtypes = [vp, vp, vp, vp]
        L.orc_edges.restype = u64
        L.orc_edges.argtypes = [vp, vp, vp, vp, vp]
        L.orc_checksum.argtypes = [vp, vp, vp]
        L.orc_query.restype = u64
        L.orc_query.argtypes = [vp, u64, u64, i64, ctypes.c_int, vp, u64]
        _lib = L
    return _lib


class COracle(object):
    def __init__(self, k=11):
        self.L = lib()
        self.k = k
        self.h = self.L.orc_new(k)
        self.n_types = 0

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_free(self.h)
            self.h = None

    def add_fasta(self, path):
        r = self.L.orc_add_fasta(self.h, path.encode(), self.n_types)
        if r < 0:
            raise ValueError("pdg_oracle: %s is not well-formed (%d)" % (path, r))
        base = self.n_types
        self.n_types += r
        return base, r

    def add_fasta_bytes(self, buf):
        arr = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else np.ascontiguousarray(buf)
        r = self.L.orc_add_fasta_mem(self.h, arr.ctypes.data, arr.size, self.n_types)
        if r < 0:
            raise ValueError("pdg_oracle: input is not well-formed (%d)" % r)
        base = self.n_types
        self.n_types += r
        return base, r

    def build(self):
        self.L.orc_build(self.h)

    def counts(self):
        a = (ctypes.c_uint64 * 3)()
        base = ctypes.addressof(a)
        self.L.orc_counts(self.h, base, base + 8, base + 16)
        return a[0], a[1], a[2]

    def index(self):
        n_nodes, _, n_kmers = self.counts()
        codes = np.zeros(n_nodes, dtype=np.uint64)
        rowptr = np.zeros(n_nodes + 1, dtype=np.uint32)
        pos = np.zeros(n_kmers, dtype=np.uint32)
        self.L.orc_index(self.h, codes.ctypes.data, rowptr.ctypes.data, pos.ctypes.data)
        return codes, rowptr, pos

    def edges(self):
        n = self.counts()[1]
        src, tgt = np.zeros(n, dtype=np.uint64), np.zeros(n, dtype=np.uint64)
        ty, val = np.zeros(n, dtype=np.uint32), np.zeros(n, dtype=np.int32)
        m = self.L.orc_edges(self.h, src.ctypes.data, tgt.ctypes.data, ty.ctypes.data, val.ctypes.data)
        assert m == n
        return src, tgt, ty, val

    def checksum(self):
        a = (ctypes.c_uint64 * 2)()
        base = ctypes.addressof(a)
        self.L.orc_checksum(self.h, base, base + 8)
        return a[0], a[1]

    def query(self, a, b, max_len=1000, connected=False):
        cap = 1024
        while True:
            hits = np.zeros((cap, 3), dtype=np.int64)
            n = self.L.orc_query(self.h, int(a), int(b), int(max_len), 1 if connected else 0, hits.ctypes.data, cap)
            if n <= cap:
                return hits[:n]
            cap = int(n)
