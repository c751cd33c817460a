"""`-m gpu`: parity at E. coli scale.  Full-size genomes (5 Mbp, 200 contigs) against the C oracle bit for
bit (index arrays, checksums), and size-independent properties at larger sizes: the order-independent
edge checksum is additive over disjoint genome sets, rows are sorted, every indexed position spells
the k-mer of its row, batch order of independent files does not change the canonical graph."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_gpu(lib_built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"


def _genomes(n, first=0, n_bases=5_000_000, n_contigs=200):
    from prairiedog_b200 import synth
    return [(name, buf, None) for name, buf, _ in synth.genomes(n, n_bases=n_bases, n_contigs=n_contigs, first=first)]


def _build(k, files, first_type=0):
    from prairiedog_b200 import B200Graph
    g = B200Graph(k=k)
    g._next_type = first_type         # edge labels are hashed by type id: keep ids as in the whole graph
    counts = g.add_fasta_batch(files)
    g.save()
    return g, counts


@pytest.mark.parametrize("k", [11, 13, 31])
def test_full_size_genomes_match_the_c_oracle(k):
    from oracle.c_oracle import COracle
    files = _genomes(2)
    g, _ = _build(k, files)
    o = COracle(k)
    for _, buf, _ in files:
        o.add_fasta_bytes(buf)
    o.build()
    assert g.engine.counts() == o.counts()
    assert g.engine.checksum() == o.checksum()
    for got, want in zip(g.engine.export_index(), o.index()):
        assert np.array_equal(got, want)


@pytest.mark.parametrize("k", [11, 31])
def test_edge_checksum_is_additive_and_rows_are_sorted(k):
    files = _genomes(12)
    whole, counts = _build(k, files)
    a, _ = _build(k, files[:5])
    b, _ = _build(k, files[5:], first_type=a._next_type)
    na, ea, ka = a.engine.counts()
    nb, eb, kb = b.engine.counts()
    nw, ew, kw = whole.engine.counts()
    assert (ew, kw) == (ea + eb, ka + kb) and kw == sum(counts)
    assert max(na, nb) <= nw <= na + nb
    mask = (1 << 64) - 1
    assert whole.engine.checksum()[1] == (a.engine.checksum()[1] + b.engine.checksum()[1]) & mask
    codes, rowptr, pos = whole.engine.export_index()
    assert codes.size == nw and int(rowptr[-1]) == kw == pos.size
    assert (np.diff(codes.astype(np.int64)) > 0).all() if k <= 31 else True
    assert (np.diff(rowptr.astype(np.int64)) > 0).all()              # no empty row
    d = np.diff(pos.astype(np.int64))
    d[rowptr[1:-1].astype(np.int64) - 1] = 1                          # row boundaries may go down
    assert (d > 0).all()                                              # positions ascend inside every row
    # sampled rows: the k-mer at every position of the row is the row's code
    rng = np.random.default_rng(7)
    for r in rng.integers(0, nw, size=200).tolist():
        ps = pos[int(rowptr[r]):int(rowptr[r + 1])][:4]
        for p in ps.tolist():
            assert int(whole.engine.emit_codes(p, k)[0]) == int(codes[r])


def test_batch_order_does_not_change_the_graph():
    """type ids follow load order, labels do not: the canonical (string) forms must agree"""
    files = _genomes(3, n_bases=150_000, n_contigs=10)
    g1, _ = _build(11, files)
    g2, _ = _build(11, files[::-1])
    assert g1.engine.checksum()[0] == g2.engine.checksum()[0]          # node set checksum
    assert np.array_equal(g1.engine.export_nodes(), g2.engine.export_nodes())
    assert g1.canonical() == g2.canonical()


@pytest.mark.parametrize("k", [17, 23, 31])
def test_long_k_prefix_path_equals_lsd_path_and_oracle(k):
    """k > 16 (msdlong.cu: MSD on the 11-residue prefix + window sort) against the C oracle and against
    the LSD sort of 128-bit records it replaces (PDG_F_FORCE_SORT_NODES)"""
    from oracle.c_oracle import COracle
    from prairiedog_b200 import B200Graph, _ffi
    files = _genomes(3, n_bases=1_000_000, n_contigs=40)
    g, _ = _build(k, files)
    assert g.engine.stats()["sort_passes"] == 4                       # the prefix path ran
    o = COracle(k)
    for _, buf, _ in files:
        o.add_fasta_bytes(buf)
    o.build()
    assert g.engine.counts() == o.counts()
    assert g.engine.checksum() == o.checksum()
    for got, want in zip(g.engine.export_index(), o.index()):
        assert np.array_equal(got, want)
    h = B200Graph(k=k)
    h.add_fasta_batch(files)
    h._flush()
    h.engine.finalize(_ffi.PDG_F_FORCE_SORT_NODES)
    assert h.engine.stats()["sort_passes"] != 4
    for got, want in zip(g.engine.export_index(), h.engine.export_index()):
        assert np.array_equal(got, want)


@pytest.mark.parametrize("run,prefix_path", [(3_000, True), (6_000, True), (20_000, False)])
def test_long_k_low_complexity(run, prefix_path):
    """a poly-A run puts `run` positions into one prefix bucket.  3 000: the eight-warp tier of the CTA sort (up to
    4096 records); 6 000: the shared-memory network for huge buckets; 20 000: larger than that kernel's shared memory,
    the build takes the LSD path.  The answer stays exact either way."""
    from oracle.c_oracle import COracle
    rng = np.random.default_rng(11)
    seq = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=30_000)) + b"A" * run + \
        bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=10_000))
    lines = [seq[i:i + 80] for i in range(0, len(seq), 80)]
    img = np.frombuffer(b">repeat\n" + b"\n".join(lines) + b"\n", dtype=np.uint8)
    g, _ = _build(21, [("rep.fasta", img, None)])
    assert (g.engine.stats()["sort_passes"] == 4) == prefix_path
    o = COracle(21)
    o.add_fasta_bytes(img)
    o.build()
    assert g.engine.counts() == o.counts()
    for got, want in zip(g.engine.export_index(), o.index()):
        assert np.array_equal(got, want)


def _replicated_image(copies_per_segment, seg_len=3000, sub_rate=0.01, seed=5):
    """One FASTA image: for every entry R of copies_per_segment a random segment of seg_len bases, written R times as
    R contigs with a 1 % substitution rate per copy.  At seg_len << 4^11 nearly every 11-residue prefix of a segment is
    unique to it, so its prefix bucket holds ~R records: a mix of equal codes (stability: positions must stay in
    order) and codes that differ in the suffix only (the bucket sort proper)."""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    parts = []
    for si, copies in enumerate(copies_per_segment):
        base = rng.integers(0, 4, size=seg_len, dtype=np.uint8)
        for c in range(copies):
            d = base.copy()
            hit = rng.random(seg_len) < sub_rate
            d[hit] = (d[hit] + rng.integers(1, 4, size=int(hit.sum()), dtype=np.uint8)) & 3
            parts.append(np.frombuffer((">seg%d_copy%d\n" % (si, c)).encode(), dtype=np.uint8))
            parts.append(acgt[d])
            parts.append(np.frombuffer(b"\n", dtype=np.uint8))
    return np.concatenate(parts)


@pytest.mark.parametrize("k", [17, 31])
def test_long_k_every_bucket_size_class(k):
    """msdlong.cu sorts a prefix bucket by its size class: <= 32, 64, 128, 256 records (power-of-two networks, 1-8 key
    registers per lane), 257-512 (the padding-skipping network over 10, 12, 14 or 16 registers), 513-2048 (one CTA, shared
    memory); the 3-genome test above only ever sees buckets of a few records.  Here segments are repeated 20 ... 1500
    times, one bucket size per class and register count, and the whole index is compared with the C oracle."""
    from oracle.c_oracle import COracle
    img = _replicated_image([20, 50, 100, 200, 270, 300, 330, 400, 440, 500, 700, 1500])
    g, _ = _build(k, [("replicated.fasta", img, None)])
    assert g.engine.stats()["sort_passes"] == 4                       # the prefix path ran
    o = COracle(k)
    o.add_fasta_bytes(img)
    o.build()
    assert g.engine.counts() == o.counts()
    assert g.engine.checksum() == o.checksum()
    for got, want in zip(g.engine.export_index(), o.index()):
        assert np.array_equal(got, want)
