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
        L.orc_stream_sums.argtypes = [vp, vp, ctypes.c_int, vp, vp, vp]
        L.orc_mark_codes.restype = ctypes.c_int
        L.orc_mark_codes.argtypes = [vp, vp]
        L.orc_bitmap_sum.argtypes = [vp, u64, vp, vp]
        L.orc_emit_codes.restype = u64
        L.orc_emit_codes.argtypes = [vp, vp]
        L.orc_sort_unique.restype = u64
        L.orc_sort_unique.argtypes = [vp, vp, u64, ctypes.c_int]
        L.orc_merge_unique.restype = u64
        L.orc_merge_unique.argtypes = [vp, u64, vp, u64, vp]
        L.orc_codes_sum.restype = u64
        L.orc_codes_sum.argtypes = [vp, u64]
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

    # ---- streaming forms (no occurrence list; inputs too large to sort in one piece) ----
    def stream_sums(self, type_offsets=(0,)):
        """(n_kmers, n_edges, [edge checksum with type ids shifted by every offset]) of the loaded contigs."""
        off = np.ascontiguousarray(type_offsets, dtype=np.uint32)
        sums = np.zeros(off.size, dtype=np.uint64)
        nk, ne = ctypes.c_uint64(), ctypes.c_uint64()
        self.L.orc_stream_sums(self.h, off.ctypes.data, off.size, ctypes.addressof(nk), ctypes.addressof(ne),
                               sums.ctypes.data)
        return nk.value, ne.value, [int(x) for x in sums]

    def mark_codes(self, bitmap):
        """ORs the presence bit of every emitted code into ``bitmap`` (uint64 words, 4^k bits; k <= 15)."""
        if self.L.orc_mark_codes(self.h, bitmap.ctypes.data) != 0:
            raise ValueError("orc_mark_codes: k > 15")

    def distinct_codes(self):
        """Sorted distinct emitted codes of the loaded contigs."""
        n = self.L.orc_emit_codes(self.h, None)
        a, t = np.empty(max(n, 1), dtype=np.uint64), np.empty(max(n, 1), dtype=np.uint64)
        self.L.orc_emit_codes(self.h, a.ctypes.data)
        m = self.L.orc_sort_unique(a.ctypes.data, t.ctypes.data, n, 2 * self.k)
        return a[:m].copy()


def bitmap_sum(bitmap):
    n, s = ctypes.c_uint64(), ctypes.c_uint64()
    lib().orc_bitmap_sum(bitmap.ctypes.data, bitmap.size, ctypes.addressof(n), ctypes.addressof(s))
    return n.value, s.value


def sort_unique(codes, key_bits):
    """Sorted distinct values of a uint64 array (destroyed)."""
    t = np.empty(max(codes.size, 1), dtype=np.uint64)
    m = lib().orc_sort_unique(codes.ctypes.data, t.ctypes.data, codes.size, int(key_bits))
    return codes[:m]


def merge_unique(a, b):
    out = np.empty(a.size + b.size, dtype=np.uint64)
    m = lib().orc_merge_unique(a.ctypes.data, a.size, b.ctypes.data, b.size, out.ctypes.data)
    return out[:m]


def codes_sum(codes):
    codes = np.ascontiguousarray(codes, dtype=np.uint64)
    return int(lib().orc_codes_sum(codes.ctypes.data, codes.size))
