"""`-m gpu`: pdg_ingest_fasta_batch (word-parallel parser, ingest_batch.cu) against pdg_ingest_fasta
(byte automaton, ingest.cu) file by file, and against the reference loader's own results
(tests/golden, prairiedog/kmers.py:51-77).  Bar: identical infos, header rows, store contents."""
import glob
import os

import numpy as np
import pytest

from conftest import golden, data_path, DATA, ROOT

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_gpu(lib_built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"


def _images(paths):
    return [np.fromfile(p, dtype=np.uint8) for p in paths]


def _per_file(k, images):
    from prairiedog_b200.engine import Engine
    e = Engine(k)
    infos, tables, keep = [], [], []
    for img in images:
        d = e.to_device(img)
        keep.append(d)
        info, table = e.ingest(d)
        infos.append(tuple(int(getattr(info, f)) for f in
                           ("n_bytes", "n_headers", "n_residues", "n_dirty", "residue_base", "appended")))
        tables.append(table.copy())
    n_res = sum(i[2] for i in infos if i[5])
    return infos, tables, e.fetch_residues(0, n_res) if n_res else ""


def _batch(k, images, host):
    import torch
    from prairiedog_b200.engine import Engine
    e = Engine(k)
    if host:
        pinned = [torch.from_numpy(img.copy()).pin_memory() if img.size else torch.empty(0, dtype=torch.uint8)
                  for img in images]
        ptrs = [p.data_ptr() if p.numel() else 0 for p in pinned]
    else:
        pinned = [e.to_device(img) for img in images]
        ptrs = [p.data_ptr() if p.numel() else 0 for p in pinned]
    infos, table, row_base = e.ingest_batch(ptrs, [img.size for img in images], host=host)
    out_infos = [tuple(int(infos[f][i]) for f in ("n_bytes", "n_headers", "n_residues", "n_dirty",
                                                    "residue_base", "appended")) for i in range(len(images))]
    tables = [table[int(row_base[i]):int(row_base[i + 1])] for i in range(len(images))]
    n_res = sum(i[2] for i in out_infos if i[5])
    return out_infos, tables, e.fetch_residues(0, n_res) if n_res else ""


def _synthetic(n, n_bases, n_contigs, width=None):
    from prairiedog_b200 import synth
    out = []
    for g, (name, buf, lens) in enumerate(synth.genomes(n, n_bases=n_bases, n_contigs=n_contigs)):
        out.append(buf)
    return out


PURE = sorted(glob.glob(os.path.join(DATA, "*.fa*")))
CORNER = sorted(glob.glob(os.path.join(DATA, "corner", "*.fasta")))


def _check(images, host, k=11):
    want = _per_file(k, images)
    got = _batch(k, images, host)
    assert got[0] == want[0]
    for tb, tw in zip(got[1], want[1]):
        assert np.array_equal(tb[:, :3], tw)
    assert got[2] == want[2]
    return got


@pytest.mark.parametrize("host", [False, True], ids=["device", "host"])
def test_batch_equals_per_file_on_fixtures(host):
    _check(_images(PURE), host)


@pytest.mark.parametrize("host", [False, True], ids=["device", "host"])
def test_batch_with_impure_files_takes_the_exact_path(host):
    """CRLF, trailing blanks, lower case / N, empty records ...: one impure file sends the batch
    through the byte automaton; results equal the per-file calls either way."""
    _check(_images(CORNER + PURE), host, k=4)
    for p in CORNER:                       # and every corner case alone
        _check(_images([p]), host, k=4)


@pytest.mark.parametrize("host", [False, True], ids=["device", "host"])
def test_batch_synthetic_multi_tile(host):
    """files spanning many 16 KiB tiles, odd sizes, an empty file in the middle, headers that
    straddle chunk / tile borders"""
    imgs = _synthetic(5, 300_000, 7)
    imgs.insert(2, np.zeros(0, dtype=np.uint8))
    rng = np.random.default_rng(5)
    # a file whose header lines are long (several chunks) and whose sequence lines are ragged
    parts = []
    for c in range(40):
        parts.append((">" + "h%d " % c + "x" * int(rng.integers(1, 400)) + "\n").encode())
        for _ in range(int(rng.integers(1, 30))):
            parts.append(bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=int(rng.integers(1, 200)))) + b"\n")
    imgs.append(np.frombuffer(b"".join(parts), dtype=np.uint8))
    # no trailing newline, header ended by end of file, text before the first header
    imgs.append(np.frombuffer(b">a\nACGTACGTACGTACGT\n>tail", dtype=np.uint8))
    imgs.append(np.frombuffer(b"ACGTAC\n>b\nACGTACGTACGTACGTTTT", dtype=np.uint8))
    imgs.append(np.frombuffer(b"\n\n>c\n\nACGTACGTACGTA\n\n\nCGT\n", dtype=np.uint8))
    got = _check(imgs, host)
    assert all(i[3] == 0 for i in got[0])


def test_batch_unaligned_device_images():
    """device images at odd addresses use the byte loads of fb_load_chunk"""
    import torch
    from prairiedog_b200.engine import Engine
    imgs = _synthetic(2, 100_000, 3)
    want = _per_file(11, imgs)
    e = Engine(11)
    big = torch.zeros(sum(i.size for i in imgs) + 64, dtype=torch.uint8, device="cuda")
    ptrs, off = [], 3
    for img in imgs:
        big[off:off + img.size] = torch.from_numpy(img).cuda()
        ptrs.append(big.data_ptr() + off)
        off += img.size + 5
    infos, table, row_base = e.ingest_batch(ptrs, [i.size for i in imgs])
    assert [int(x) for x in infos["n_residues"]] == [w[2] for w in want[0]]
    assert np.array_equal(table[:, :3], np.concatenate(want[1]))
    assert e.fetch_residues(0, int(infos["n_residues"].sum())) == want[2]


def test_header_hash_flags_duplicate_labels():
    from prairiedog_b200.engine import Engine
    img = np.frombuffer(b">x 1\nACGTACGTACGTACG\n>y\nACGTTTTTACGTACG\n>x 1  \nCCCCCCCCCCCCCCCCCC\n", dtype=np.uint8)
    e = Engine(11)
    d = e.to_device(img)
    infos, table, _ = e.ingest_batch([d.data_ptr()], [img.size])
    h = table[:, 3]
    assert h[0] == h[2] and h[0] != h[1] and (h != 0).all()


@pytest.mark.parametrize("host", [False, True], ids=["device", "host"])
@pytest.mark.parametrize("gold,name", [("fixtures_k11.json", "all4"), ("synthetic.json", "syn_k11"),
                                       ("synthetic.json", "syn_k31"), ("synthetic.json", "syn_k5_repeats")])
def test_graph_from_batch_matches_reference(gold, name, host):
    """the whole pipeline through add_fasta_batch against the golden vectors of the reference"""
    from prairiedog_b200 import B200Graph
    case = golden(gold)[name]
    g = B200Graph(k=case["k"])
    files = []
    for f in case["files"]:
        img = np.fromfile(data_path(f), dtype=np.uint8)
        files.append((f, img, None if host else g.engine.to_device(img)))
    returned = g.add_fasta_batch(files)
    g.save()
    assert returned == [s["returned"] for s in case["stats"]]
    nodes, edges = g.canonical()
    assert nodes == case["nodes"]
    assert [list(e) for e in edges] == case["edges"]


def test_batch_reports_files_it_cannot_hold():
    from prairiedog_b200 import B200Graph
    g = B200Graph(k=4)
    good = np.fromfile(data_path("15.fa"), dtype=np.uint8)
    bad = np.fromfile(data_path(os.path.join("corner", "lower_and_n.fasta")), dtype=np.uint8)
    with pytest.raises(ValueError) as ei:
        g.add_fasta_batch([("15.fa", good, None), ("lower_and_n.fasta", bad, None)])
    assert ei.value.failed == [1] and ei.value.counts[0] == 12


def test_duplicate_headers_in_batch_keep_set_semantics():
    """two contigs with the same header in one file share an edge label: the batch path must fall
    back to resolved labels exactly like add_fasta (lemongraph.c:674-685 set semantics)"""
    from prairiedog_b200 import B200Graph
    p = data_path(os.path.join("corner", "dup_header.fasta"))
    img = np.fromfile(p, dtype=np.uint8)
    g1 = B200Graph(k=4)
    g1.add_fasta(p)
    g1.save()
    g2 = B200Graph(k=4)
    g2.add_fasta_batch([("dup_header.fasta", img, None), ("15.fa", np.fromfile(data_path("15.fa"), dtype=np.uint8), None)])
    g2.save()
    g3 = B200Graph(k=4)
    g3.add_fasta(p)
    g3.add_fasta(data_path("15.fa"))
    g3.save()
    assert g2.canonical() == g3.canonical()
    assert g1.canonical()[0] != []


def _pipelined_worker(group_bytes, k, n_genomes, n_bases, n_contigs, out):
    """runs in a fresh process: the group size of the pipelined build is read once from the environment"""
    import os
    import sys
    os.environ["PDG_FB_GROUP_BYTES"] = str(group_bytes)
    sys.path.insert(0, ROOT)
    import numpy as np
    from prairiedog_b200 import B200Graph, synth
    from oracle.c_oracle import COracle
    files = [(name, buf, None) for name, buf, _ in synth.genomes(n_genomes, n_bases=n_bases, n_contigs=n_contigs)]
    g = B200Graph(k=k)
    counts = g.add_fasta_batch(files)
    g.save()
    st = g.engine.stats()
    o = COracle(k)
    for _, buf, _ in files:
        o.add_fasta_bytes(buf)
    o.build()
    ok = g.engine.counts() == o.counts() and g.engine.checksum() == o.checksum()
    for got, want in zip(g.engine.export_index(), o.index()):
        ok = ok and bool(np.array_equal(got, want))
    # a second save without changes rebuilds in one shot and must give the same graph
    g._unsaved = True
    g.save()
    ok2 = g.engine.checksum() == o.checksum()
    with open(out, "w") as f:
        f.write("%d %d %d %d %d" % (ok, ok2, st["pipelined"], st["sort_groups"], sum(counts)))


@pytest.mark.parametrize("group_bytes,k", [(60_000, 11), (250_000, 11), (60_000, 7), (1 << 25, 11), (100_000, 4),
                                           (60_000, 21), (250_000, 31), (1 << 25, 17)])
def test_pipelined_build_equals_oracle(tmp_path, group_bytes, k):
    """host images, k <= 11 or k >= 17, empty graph: passes 1-2 run per copy group inside the ingest call and
    pass 3 is launched at its end; small groups force many groups (and groups smaller than a tile).  For
    k >= 17 the staged sort is the one on the 11-residue prefix and pdg_finalize finishes the buckets."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "r.txt")
    ctx = mp.get_context("spawn")
    p = ctx.Process(target=_pipelined_worker, args=(group_bytes, k, 6, 120_000, 9, out))
    p.start()
    p.join(600)
    assert p.exitcode == 0
    ok, ok2, pipelined, groups, n = map(int, open(out).read().split())
    assert ok == 1 and ok2 == 1 and pipelined == 1
    assert groups == (6 if group_bytes < 130_000 else 1 if group_bytes > 1_000_000 else groups) and 1 <= groups <= 6
    assert group_bytes > 1_000_000 or groups >= 3


def test_pipelined_build_falls_back_when_contigs_differ():
    """the ingest assumed one contig per header; the caller declares something else -> one-shot build"""
    from prairiedog_b200.engine import Engine
    from oracle.c_oracle import COracle
    import torch
    imgs = _synthetic(3, 100_000, 5)
    e = Engine(11)
    pinned = [torch.from_numpy(i.copy()).pin_memory() for i in imgs]
    infos, table, row_base = e.ingest_batch([p.data_ptr() for p in pinned], [i.size for i in imgs], host=True)
    # declare only the contigs of the first two files
    nh = int(row_base[2])
    res = table[:nh, 2].astype(np.int64)
    fo = np.repeat(np.arange(2), np.diff(row_base[:3]))
    begin = res + infos["residue_base"].astype(np.int64)[fo]
    end = np.empty_like(res)
    end[:-1] = res[1:]
    end[row_base[1:3] - 1] = infos["n_residues"].astype(np.int64)[:2]
    e.add_contigs(begin, end - res, np.arange(nh, dtype=np.uint32))
    e.finalize()
    assert e.stats()["pipelined"] == 0
    o = COracle(11)
    for i in imgs[:2]:
        o.add_fasta_bytes(i)
    o.build()
    assert e.counts() == o.counts() and e.checksum() == o.checksum()


@pytest.mark.parametrize("host", [False, True], ids=["device", "host"])
@pytest.mark.parametrize("k", [4, 11])
def test_batch_with_contigs_of_length_k_and_shorter(host, k):
    """a middle contig of exactly k residues has one window but yields nothing in the reference
    (kmers.py:110-117: len <= k is skipped); the speculative build of the batch path must agree with
    the per-file path, which only sees the declared contigs"""
    from prairiedog_b200 import B200Graph
    rng = np.random.default_rng(k)

    def seq(n):
        return bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=n))
    a = b">a1\n" + seq(60) + b"\n>a_len_k\n" + seq(k) + b"\n>a3\n" + seq(45) + b"\n"
    b = b">b1\n" + seq(70) + b"\n>b_short\n" + seq(k - 1) + b"\n>b3\n" + seq(k + 1) + b"\n"
    imgs = [("a.fasta", np.frombuffer(a, dtype=np.uint8)), ("b.fasta", np.frombuffer(b, dtype=np.uint8))]
    g1 = B200Graph(k=k)
    for name, img in imgs:
        g1.add_fasta_bytes(name, img)
    g1.save()
    g2 = B200Graph(k=k)
    counts = g2.add_fasta_batch([(name, img, None if host else g2.engine.to_device(img)) for name, img in imgs])
    g2.save()
    assert counts == [60 - k + 1 + 45 - k + 1, 70 - k + 1 + 2]
    assert g2.engine.counts() == g1.engine.counts()
    assert g2.engine.counts()[2] == sum(counts)
    assert g2.canonical() == g1.canonical()
    assert g2.engine.checksum() == g1.engine.checksum()
