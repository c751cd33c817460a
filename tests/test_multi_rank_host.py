"""`not gpu`: the N > 1 host logic (prairiedog_b200/distributed.py) on CPU with world_size 2 over gloo.

Each rank builds the graph of ITS genomes with the CPU oracle (test infrastructure) and the ranks then
run, on the host, the same exchange csrc/comm.cu runs on the device: distinct node keys hash
partitioned by ``owner``, owner-side deduplication, all-gather of the owned sets.  The result must be
the single-process graph: same global node table, edge sets disjoint with the same union, counts
that add up."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT, data_path

FILES = ["syn101.fasta", "syn102.fasta", "syn201.fasta", "syn202.fasta", "syn301.fasta"]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_graph(paths, k):
    from oracle import ref_spec
    g = ref_spec.SpecGraph()
    n = 0
    for p in paths:
        n += g.add_genome(p, k)
    nodes, edges = g.canonical()
    return nodes, edges, n


def _worker(rank, world, port, k, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    from prairiedog_b200 import codec, distributed as D
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        paths = [data_path(f) for f in FILES]
        mine = D.shard(paths, rank, world)
        nodes, edges, n_kmers = _oracle_graph(mine, k)
        # 128-byte id from rank 0 reaches every rank unchanged (the NCCL id path of attach())
        uid = np.arange(128, dtype=np.uint8) if rank == 0 else np.zeros(128, dtype=np.uint8)
        assert D.broadcast_bytes(uid, 0).tolist() == list(range(128))
        # node exchange, host mirror of comm.cu: bucket by owner -> all-to-all -> unique -> all-gather
        codes = np.array([codec.encode(v, k) for v in nodes], dtype=np.uint64)
        own = D.owner(codes, world, k)
        send = [codes[own == r] for r in range(world)]
        recv = [None] * world
        dist.all_gather_object(recv, send)                      # recv[src][dst]
        owned = np.unique(np.concatenate([recv[src][rank] for src in range(world)]))
        assert (D.owner(owned, world, k) == rank).all()
        gathered = [None] * world
        dist.all_gather_object(gathered, owned)
        table = np.concatenate(gathered)                        # ranges ascend with the rank: already sorted
        assert (np.diff(table.astype(np.int64)) > 0).all()
        assert np.unique(table).size == table.size              # owned sets are disjoint
        tot_edges, tot_kmers = D.reduce_counts([len(edges), n_kmers])
        all_edges = [None] * world
        dist.all_gather_object(all_edges, edges)
        if rank == 0:
            g_nodes, g_edges, g_kmers = _oracle_graph(paths, k)
            assert [codec.decode(int(c), k) for c in table] == sorted(g_nodes, key=lambda v: codec.encode(v, k))
            flat = [e for part in all_edges for e in part]
            assert len(flat) == len(set(flat)) == len(g_edges)  # edges never need to move
            assert sorted(flat) == g_edges
            assert (int(tot_edges), int(tot_kmers)) == (len(g_edges), g_kmers)
            open(os.path.join(out_dir, "ok"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def _query_worker(rank, world, port, k, out_dir):
    """The path query over an N-rank graph, host side: every rank answers the same pair batch from the C oracle of
    ITS shard (standing in for its GPU), distributed.exchange_hits merges over gloo; rank 0 compares with the oracle
    of all files in one graph."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    from oracle.c_oracle import COracle
    from prairiedog_b200 import distributed as D
    from prairiedog_b200.engine import HIT_DTYPE
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        paths = [data_path(f) for f in FILES]
        whole = COracle(k)
        for p in paths:
            whole.add_fasta(p)
        whole.build()
        codes, rowptr, pos = whole.index()
        rng = np.random.default_rng(5)                      # same pairs on every rank
        # code of the k-mer starting at every residue position (-1: no k-mer starts there, i.e. a contig tail)
        code_at = np.full(int(pos.max()) + 2, -1, dtype=np.int64)
        code_at[pos] = np.repeat(codes, np.diff(rowptr.astype(np.int64))).astype(np.int64)
        qa, qb = [], []
        while len(qa) < 30:                                 # positives: b starts d residues after a, same contig
            p0, d = int(rng.integers(0, code_at.size - 60)), int(rng.integers(1, 50))
            if (code_at[p0:p0 + d + 1] >= 0).all():
                qa.append(code_at[p0])
                qb.append(code_at[p0 + d])
        for _ in range(10):                                 # and unrelated pairs
            qa.append(code_at[pos[int(rng.integers(0, pos.size))]])
            qb.append(code_at[pos[int(rng.integers(0, pos.size))]])
        qa, qb = np.array(qa, dtype=np.uint64), np.array(qb, dtype=np.uint64)
        mine = COracle(k)
        for p in D.shard(paths, rank, world):
            mine.add_fasta(p)
        mine.build()
        bases = D.type_bases(mine.n_types)
        assert bases[-1] == whole.n_types and bases[0] == 0
        off, rows = [0], []
        for a, b in zip(qa.tolist(), qb.tolist()):
            h = mine.query(a, b, 200)
            rows += [(t, i, j, 0) for t, i, j in h.tolist()]
            off.append(len(rows))
        local = np.array(rows, dtype=HIT_DTYPE) if rows else np.zeros(0, dtype=HIT_DTYPE)
        goff, ghits, owner = D.exchange_hits(np.array(off, dtype=np.uint64), local, int(bases[rank]))
        # the all-on-device form of the same exchange (torch tensors in; CPU tensors over gloo here)
        import torch
        g2 = D.exchange_hits_device(torch.tensor(off, dtype=torch.int64),
                                    torch.from_numpy(local.view(np.int32).reshape(-1, 4).copy()), int(bases[rank]))
        assert g2[0].tolist() == goff.tolist() and np.array_equal(g2[1], ghits) and g2[2].tolist() == owner.tolist()
        # every rank receives the same global answer
        want_off, want = [0], []
        for a, b in zip(qa.tolist(), qb.tolist()):
            want += [tuple(x) for x in whole.query(a, b, 200).tolist()]
            want_off.append(len(want))
        assert goff.tolist() == want_off
        got = [(int(h["type_id"]), int(h["i"]), int(h["j"])) for h in ghits]
        for q in range(40):                                 # per pair: the same hits in the same order
            assert got[want_off[q]:want_off[q + 1]] == want[want_off[q]:want_off[q + 1]], q
        assert len(want) > 0 and set(owner.tolist()) == set(range(world))
        # the owner is the rank whose type range holds the hit
        assert (np.searchsorted(bases, ghits["type_id"].astype(np.int64), side="right") - 1 == owner).all()
        assert D.gather_objects(rank) == list(range(world))
        open(os.path.join(out_dir, "ok%d" % rank), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world_size_2_query_exchange_gloo(tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_query_worker, args=(2, _free_port(), 11, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_merge_hits_orders_by_pair_then_rank():
    from prairiedog_b200 import distributed as D
    from prairiedog_b200.engine import HIT_DTYPE
    r0 = np.array([(0, 1, 2, 0), (0, 3, 4, 0), (1, 5, 6, 0)], dtype=HIT_DTYPE)      # pair 0: 2 hits, pair 2: 1 hit
    r1 = np.array([(7, 0, 0, 0), (8, 1, 1, 0), (9, 2, 2, 0)], dtype=HIT_DTYPE)      # pair 0: 1, pair 1: 2
    off, hits, owner = D.merge_hits([[2, 0, 1], [1, 2, 0]], [r0, r1])
    assert off.tolist() == [0, 3, 5, 6]
    assert hits["type_id"].tolist() == [0, 0, 7, 8, 9, 1]
    assert owner.tolist() == [0, 0, 1, 1, 1, 0]
    off, hits, owner = D.merge_hits([[0, 0]], [np.zeros(0, dtype=HIT_DTYPE)])
    assert off.tolist() == [0, 0, 0] and hits.size == 0
    # the all-on-device form (torch tensors in, here on the CPU) gives the same answer as the host form
    import torch
    want = D.exchange_hits(np.array([0, 2, 3], dtype=np.uint64), r0, 10)
    got = D.exchange_hits_device(torch.tensor([0, 2, 3], dtype=torch.int64),
                                 torch.from_numpy(r0.view(np.int32).reshape(-1, 4).copy()), 10)
    assert got[0].tolist() == want[0].tolist() and np.array_equal(got[1], want[1]) and got[2].tolist() == want[2].tolist()
    # single process: exchange_hits only applies the type base
    off, hits, owner = D.exchange_hits(np.array([0, 2, 3], dtype=np.uint64), r0, 10)
    assert hits["type_id"].tolist() == [10, 10, 11] and off.tolist() == [0, 2, 3] and r0["type_id"].tolist() == [0, 0, 1]


def test_merge_hits_torch_equals_numpy_merge():
    """the device merge of exchange_hits against the plain numpy statement of the same step"""
    import torch
    from prairiedog_b200 import distributed as D
    from prairiedog_b200.engine import HIT_DTYPE
    rng = np.random.default_rng(11)
    for world, n_q in ((1, 5), (2, 40), (3, 17), (8, 300)):
        counts = rng.integers(0, 6, size=(world, n_q))
        counts[rng.integers(0, world), :] = 0 if world > 1 else counts[0]          # an idle rank
        parts, recv = [], []
        cap = max(int(counts.sum(axis=1).max()), 1)
        for r in range(world):
            n = int(counts[r].sum())
            h = np.zeros(n, dtype=HIT_DTYPE)
            h["type_id"], h["i"], h["j"], h["aux"] = 1000 * r + rng.integers(0, 99, n), rng.integers(0, 9999, n), rng.integers(0, 9999, n), r
            parts.append(h)
            pad = np.zeros((cap, 4), dtype=np.int32)
            pad[:n] = h.view(np.int32).reshape(-1, 4)
            recv.append(torch.from_numpy(pad))
        off, hits, owner = D.merge_hits(counts, parts)
        toff, thits, towner = D.merge_hits_torch(torch.from_numpy(counts.astype(np.int64)), recv, counts.sum(axis=1).tolist())
        assert toff.tolist() == off.tolist()
        assert np.array_equal(thits.numpy().reshape(-1).view(HIT_DTYPE), hits)
        assert towner.tolist() == owner.tolist()


def test_shard_covers_every_genome_once():
    from prairiedog_b200 import distributed as D
    for n in (0, 1, 5, 8, 100):
        for w in (1, 2, 3, 8):
            items = list(range(n))
            parts = [D.shard(items, r, w) for r in range(w)]
            assert sum(parts, []) == items
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


def test_owner_is_a_monotone_partition():
    """range partition of csrc/comm.cu:range_owner: monotone in the key (owned sets are ranges), balanced
    on uniform codes, opaque keys on the last rank; mix64 keeps its known answers (edge checksums use it)"""
    from prairiedog_b200 import distributed as D
    rng = np.random.default_rng(3)
    for k in (5, 8, 11, 16, 31):
        codes = np.sort(rng.integers(0, 1 << (2 * k), size=1 << 14, dtype=np.uint64))
        for w in (1, 2, 3, 4, 8):
            o = D.owner(codes, w, k)
            assert o.min() >= 0 and o.max() < w
            assert (np.diff(o) >= 0).all()
            if k >= 8:
                assert np.bincount(o, minlength=w).min() > codes.size // (2 * w)
    opaque = np.array([(1 << 63) | 5, 7], dtype=np.uint64)
    assert D.owner(opaque, 4, 11).tolist() == [3, 0]
    assert int(D.mix64(np.array([0], dtype=np.uint64))[0]) == 0
    assert int(D.mix64(np.array([1], dtype=np.uint64))[0]) == 0xb456bcfc34c2cb2c


@pytest.mark.timeout(300)
def test_world_size_2_gloo(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 11, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").exists()


def test_bind_host_to_device_is_best_effort():
    """without NVML (this container) nothing is bound and nothing raises; the affinity stays as it was"""
    import os
    from prairiedog_b200 import distributed as D
    before = os.sched_getaffinity(0)
    n = D.bind_host_to_device(0)
    assert n == 0 or n == len(os.sched_getaffinity(0))
    if n == 0:
        assert os.sched_getaffinity(0) == before
