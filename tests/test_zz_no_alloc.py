"""`-m gpu`: pdg_codes_checksum_static (include/pdgpu.h K6, csrc/noalloc.cu) -- the node checksum computed with scratch
in a static __device__ array instead of a per-call allocation -- against pdg_checksum, the existing entry point, on
the node tables of two different graphs in two calls in a row (the second table is the smaller one, so partial sums
the first call left in the shared array would show).  Integer arithmetic modulo 2^64: the results must be equal.

Collected last on purpose; needs nothing from the tests before it."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def node_table(k, n_genomes, n_bases):
    """(node codes as exported by pdg_export_nodes, nodes_sum of pdg_checksum) of a small synthetic pangenome"""
    import torch
    from prairiedog_b200 import B200Graph, synth
    g = B200Graph(k=k)
    g.add_fasta_batch([(name, buf, torch.from_numpy(buf).cuda())
                       for name, buf, _ in synth.genomes(n_genomes, n_bases=n_bases, n_contigs=7)])
    g.save()
    codes = g.engine.export_nodes()
    nodes_sum, _ = g.engine.checksum()
    assert codes.size == g.engine.counts()[0] > 0
    return codes, nodes_sum


def static_sum(lib, d_codes, n, d_out, stream=None):
    import torch
    from prairiedog_b200 import _ffi
    _ffi.check(lib.pdg_codes_checksum_static(0, d_codes.data_ptr() if n else None, n, d_out.data_ptr(),
                                             stream.cuda_stream if stream is not None else None))
    (stream or torch.cuda.current_stream()).synchronize()
    return int(d_out.cpu().numpy().view(np.uint64)[0])


def test_static_checksum_equals_pdg_checksum_twice_in_a_row(lib_built):
    import torch
    from prairiedog_b200 import _ffi
    lib = _ffi.load()
    big, big_sum = node_table(31, 3, 60_000)      # ~1.8e5 distinct 31-mers: ~180 partial sums
    small, small_sum = node_table(11, 2, 3_000)   # a few thousand: a handful
    assert big.size > 20 * small.size and big_sum != small_sum
    # everything the calls need exists before the first one: they allocate nothing themselves
    d_big = torch.from_numpy(big.view(np.int64)).cuda()
    d_small = torch.from_numpy(small.view(np.int64)).cuda()
    d_out = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    st = torch.cuda.Stream()
    torch.cuda.synchronize()
    got_big = static_sum(lib, d_big, big.size, d_out)
    got_small = static_sum(lib, d_small, small.size, d_out)
    assert got_big == big_sum
    assert got_small == small_sum
    # and once more the other way round, on a stream of the caller's.  The kernels are loaded by now (the first launch
    # of a lazily loaded module may take device memory), so the device's free memory must not move at all.
    free_before = torch.cuda.mem_get_info()[0]
    assert static_sum(lib, d_small, small.size, d_out, st) == small_sum
    assert static_sum(lib, d_big, big.size, d_out, st) == big_sum
    assert torch.cuda.mem_get_info()[0] == free_before, "the entry point changed the device's free memory"
    # an empty table sums to 0 (and overwrites what the last call left in the output)
    assert static_sum(lib, d_small, 0, d_out) == 0


def test_static_checksum_refuses_what_the_scratch_cannot_hold(lib_built):
    import torch
    from prairiedog_b200 import _ffi
    lib = _ffi.load()
    d_out = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = lib.pdg_codes_checksum_static(0, d_out.data_ptr(), _ffi.STATIC_MAX_CODES + 1, d_out.data_ptr(), None)
    assert rc == _ffi.PDG_E_CAPACITY
    assert "exceed the static scratch" in _ffi.last_error() and str(_ffi.STATIC_MAX_CODES) in _ffi.last_error()
    assert lib.pdg_codes_checksum_static(0, d_out.data_ptr(), 1, None, None) == _ffi.PDG_E_INVALID
    assert lib.pdg_codes_checksum_static(-1, d_out.data_ptr(), 1, d_out.data_ptr(), None) == _ffi.PDG_E_INVALID
    # the limit itself is the header's
    hdr = open(__import__("os").path.join(__import__("conftest").ROOT, "include", "pdgpu.h")).read()
    assert "#define PDG_STATIC_MAX_CODES (1ull << 27)" in hdr and _ffi.STATIC_MAX_CODES == 1 << 27
