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
