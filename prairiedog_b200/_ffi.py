"""ctypes binding of libpdgpu.so (include/pdgpu.h) -- the only way the Python host reaches CUDA.

There is no CPU implementation behind this module: if the shared library is missing or no CUDA
device can be opened, every entry point raises ``PdgError``.  torch is used by the callers only to
own pinned host / device buffers whose raw pointers are handed to the C ABI.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PDG_LIBPDGPU", os.path.join(HERE, "libpdgpu.so"))   # override: tuning variants only

PDG_OK = 0
PDG_E_INVALID = -1
PDG_E_CUDA = -2
PDG_E_CAPACITY = -3
PDG_E_NOMEM = -4
PDG_E_COMM = -5

PDG_F_FORCE_SORT_NODES = 1
PDG_F_FORCE_GENERAL = 2
PDG_IN_HOST = 1

u8p = ctypes.POINTER(ctypes.c_uint8)
u32p = ctypes.POINTER(ctypes.c_uint32)
i32p = ctypes.POINTER(ctypes.c_int32)
u64p = ctypes.POINTER(ctypes.c_uint64)
c_u64 = ctypes.c_uint64
c_i64 = ctypes.c_int64
c_int = ctypes.c_int
vp = ctypes.c_void_p


class PdgError(RuntimeError):
    def __init__(self, code, msg):
        super(PdgError, self).__init__("libpdgpu error %d: %s" % (code, msg))
        self.code = code


class IngestInfo(ctypes.Structure):
    _fields_ = [("n_bytes", c_u64), ("n_headers", c_u64), ("n_residues", c_u64),
                ("n_dirty", c_u64), ("residue_base", c_u64), ("appended", c_u64)]


INGEST_INFO_DTYPE = [("n_bytes", "<u8"), ("n_headers", "<u8"), ("n_residues", "<u8"),
                     ("n_dirty", "<u8"), ("residue_base", "<u8"), ("appended", "<u8")]


class Hit(ctypes.Structure):
    _fields_ = [("type_id", ctypes.c_uint32), ("i", ctypes.c_int32), ("j", ctypes.c_int32),
                ("aux", ctypes.c_uint32)]


class BuildStats(ctypes.Structure):
    _fields_ = [("n_residues", c_u64), ("n_contigs", c_u64), ("n_kmers", c_u64),
                ("n_edge_records", c_u64), ("n_nodes", c_u64), ("n_edges", c_u64),
                ("node_path", c_u64), ("n_demoted", c_u64),
                ("sort_passes", c_u64), ("sort_records", c_u64),
                ("ms_upload", ctypes.c_float), ("ms_nodes", ctypes.c_float),
                ("ms_sort", ctypes.c_float), ("ms_rows", ctypes.c_float),
                ("ms_explicit", ctypes.c_float), ("ms_total", ctypes.c_float),
                ("ms_sort_kernel_avg", ctypes.c_float), ("pad", ctypes.c_float),
                ("sort_kernel_launches", c_u64), ("kernel_launches", c_u64),
                ("ms_hist", ctypes.c_float), ("ms_scans", ctypes.c_float), ("ms_pass", ctypes.c_float * 3),
                ("pad2", ctypes.c_float),
                ("ms_ingest_a", ctypes.c_float), ("ms_ingest_b", ctypes.c_float), ("ms_ingest_c", ctypes.c_float),
                ("pad3", ctypes.c_float), ("sort_groups", c_u64), ("pipelined", c_u64), ("ingest_bytes", c_u64), ("ingest_launches", c_u64)]


ABI_VERSION = 6      # PDG_ABI_VERSION of include/pdgpu.h

# name -> (restype, argtypes); every symbol of include/pdgpu.h (tests/test_abi.py checks the list
# against the header)
SIGNATURES = {
    "pdg_last_error": (ctypes.c_char_p, []),
    "pdg_abi_version": (c_int, []),
    "pdg_struct_sizes": (c_int, [u64p, u64p, u64p]),
    "pdg_create": (vp, [c_int, c_int]),
    "pdg_destroy": (None, [vp]),
    "pdg_set_stream": (c_int, [vp, vp]),
    "pdg_clear": (c_int, [vp]),
    "pdg_ingest_fasta": (c_int, [vp, vp, c_u64, vp, c_u64, ctypes.POINTER(IngestInfo)]),
    "pdg_ingest_fasta_batch": (c_int, [vp, vp, vp, ctypes.c_uint32, c_int, vp, c_u64, vp]),
    "pdg_add_contigs": (c_int, [vp, vp, vp, vp, c_u64]),
    "pdg_kmer_stats": (c_int, [vp, vp, vp, c_u64, u64p, u64p]),
    "pdg_emit_codes": (c_int, [vp, c_u64, c_u64, vp]),
    "pdg_fetch_residues": (c_int, [vp, c_u64, c_u64, vp]),
    "pdg_snapshot_sizes": (c_int, [vp, vp]),
    "pdg_snapshot_export": (c_int, [vp, vp, c_u64, vp, vp, vp, vp, vp, vp, vp, vp]),
    "pdg_store_import": (c_int, [vp, vp, c_u64]),
    "pdg_add_edges_host": (c_int, [vp, vp, vp, vp, vp, c_u64]),
    "pdg_add_nodes_host": (c_int, [vp, vp, c_u64]),
    "pdg_finalize": (c_int, [vp, c_int]),
    "pdg_counts": (c_int, [vp, u64p, u64p, u64p]),
    "pdg_get_build_stats": (c_int, [vp, ctypes.POINTER(BuildStats)]),
    "pdg_export_nodes": (c_int, [vp, vp]),
    "pdg_export_edges": (c_int, [vp, vp, vp, vp, vp]),
    "pdg_node_ids": (c_int, [vp, vp, c_u64, vp]),
    "pdg_index_size": (c_int, [vp, u64p, u64p]),
    "pdg_export_index": (c_int, [vp, vp, vp, vp]),
    "pdg_checksum": (c_int, [vp, u64p, u64p]),
    "pdg_query_connected": (c_int, [vp, vp, vp, c_u64, u64p]),
    "pdg_query_paths": (c_int, [vp, vp, vp, c_u64, c_i64, u64p]),
    "pdg_query_fetch": (c_int, [vp, vp, vp]),
    "pdg_query_walk": (c_int, [vp, c_u64, ctypes.POINTER(Hit), vp, i32p]),
    "pdg_filter_degrees": (c_int, [vp, ctypes.POINTER(ctypes.c_uint32)]),
    "pdg_filter_select": (c_int, [vp, ctypes.c_uint32, u64p, u64p]),
    "pdg_filter_fetch": (c_int, [vp, vp, vp, vp, vp]),
    "pdg_query_strings_size": (c_int, [vp, u64p]),
    "pdg_query_strings_fetch": (c_int, [vp, vp, vp]),
    "pdg_comm_unique_id": (c_int, [vp]),
    "pdg_comm_init": (c_int, [vp, vp, c_int, c_int]),
    "pdg_codes_checksum_static": (c_int, [c_int, vp, c_u64, vp, vp]),
}
STATIC_MAX_CODES = 1 << 27

_lib = None


def load(build_if_missing=False):
    """Returns the ctypes handle of libpdgpu.so with argtypes set; raises PdgError if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH) and build_if_missing:
        from . import build
        build.build_library()
    if not os.path.exists(LIB_PATH):
        raise PdgError(PDG_E_INVALID, "%s not built: run `python -m prairiedog_b200.build` "
                       "(there is no CPU fallback)" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.pdg_abi_version() != ABI_VERSION:
        raise PdgError(PDG_E_INVALID, "libpdgpu ABI %d, expected %d: rebuild" % (lib.pdg_abi_version(), ABI_VERSION))
    _lib = lib
    return lib


def last_error():
    return load().pdg_last_error().decode("utf-8", "replace")


def check(rc):
    if rc != PDG_OK:
        raise PdgError(rc, last_error())
    return rc


def np_ptr(a):
    """Raw pointer of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data
