"""`-m gpu`, needs >= 2 GPUs (skipped otherwise; run with `gpurun --gpus 2`): the N > 1 build and the N > 1 query.
Genomes shard by sample, every rank builds the index of its own genomes, distinct node keys are exchanged inside
pdg_finalize (csrc/comm.cu: OR of presence bitmaps for k <= 15, otherwise a RANGE-partitioned all-to-all of each
rank's sorted table).  Every rank must end with the node table of the single-process graph; edges never move, so the
per-rank edge lists are disjoint and their union is the single-process edge list; a path query over the ranks
(B200Graph.path_many / query_many / path / connected, collective) returns the single-process answer."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT, data_path

pytestmark = pytest.mark.gpu
FILES = ["syn101.fasta", "syn102.fasta", "syn201.fasta", "syn202.fasta", "syn301.fasta", "syn302.fasta"]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, k, flags, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    from prairiedog_b200 import B200Graph, distributed as D
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        g = B200Graph(k=k, device=rank)
        assert D.attach(g, dev) == (rank, world)
        mine = D.shard(FILES, rank, world)
        files = [(f, np.fromfile(data_path(f), dtype=np.uint8), None) for f in mine]
        g.add_fasta_batch(files)
        g._flush()
        g.engine.finalize(flags)
        nodes = g.engine.export_nodes()
        src, tgt, ty, val = g.engine.export_edges()
        labels = [g._type_string(t) for t in ty.tolist()]
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), nodes=nodes, src=src, tgt=tgt, val=val,
                 labels=np.array(labels, dtype=object), counts=np.array(g.engine.counts(), dtype=np.uint64))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("k,flags", [(11, 0), (11, 1), (31, 0), (5, 0)], ids=["k11-bitmap", "k11-hash", "k31-hash", "k5"])
def test_two_ranks_equal_one(tmp_path, lib_built, k, flags):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from prairiedog_b200 import B200Graph
    mp.spawn(_worker, args=(2, _free_port(), k, flags, str(tmp_path)), nprocs=2, join=True)
    one = B200Graph(k=k)
    one.add_fasta_batch([(f, np.fromfile(data_path(f), dtype=np.uint8), None) for f in FILES])
    one.save()
    want_nodes = one.engine.export_nodes()
    s, t, ty, v = one.engine.export_edges()
    want_edges = sorted(zip(s.tolist(), t.tolist(), [one._type_string(x) for x in ty.tolist()], v.tolist()))
    got_edges = []
    for r in range(2):
        z = np.load(os.path.join(str(tmp_path), "rank%d.npz" % r), allow_pickle=True)
        assert np.array_equal(z["nodes"], want_nodes)                    # same global node table on every rank
        assert int(z["counts"][0]) == want_nodes.size
        got_edges += list(zip(z["src"].tolist(), z["tgt"].tolist(), z["labels"].tolist(), z["val"].tolist()))
    assert len(got_edges) == len(set(got_edges))
    assert sorted(got_edges) == want_edges


def _query_worker(rank, world, port, k, pairs, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import pickle
    import torch
    import torch.distributed as dist
    from prairiedog_b200 import B200Graph, distributed as D
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        g = B200Graph(k=k, device=rank)
        assert D.attach(g, dev) == (rank, world)
        mine = D.shard(FILES, rank, world)
        g.add_fasta_batch([(f, np.fromfile(data_path(f), dtype=np.uint8), None) for f in mine])
        g.save()
        a, off, hits = g.path_many(pairs, limit=300)               # collective: hits of the whole pangenome
        owner = g.hit_owner
        rows = g.query_many(pairs, limit=300)                      # collective: labels + strings
        paths, metas = g.path(*pairs[0])
        conn = g.connected_many(pairs[:6])
        with open(os.path.join(out_dir, "q%d.pkl" % rank), "wb") as f:
            pickle.dump({"off": off, "hits": hits, "owner": owner, "rows": rows,
                         "paths": [[n.value for n in p] for p in paths], "metas": list(metas),
                         "conn": [(ok, sorted((e.edge_type, e.edge_value) for e in es)) for ok, es in conn]}, f)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("k", [11, 31])
def test_two_rank_query_equals_one(tmp_path, lib_built, k):
    """SURVEY.md 8(e) "Queries" / prairiedog/prairiedog.py:17-33: a query is answered over the whole pangenome."""
    import pickle
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from prairiedog_b200 import B200Graph
    one = B200Graph(k=k)
    one.add_fasta_batch([(f, np.fromfile(data_path(f), dtype=np.uint8), None) for f in FILES])
    one.save()
    n_res = int(one.engine.stats()["n_residues"])
    seq = one.engine.fetch_residues(0, n_res)
    rng = np.random.default_rng(k)
    pairs = []
    for _ in range(50):                     # spread over the whole store: both ranks own some of the positives
        p, d = int(rng.integers(0, n_res - k - 120)), int(rng.integers(1, 100))
        pairs.append((seq[p:p + k], seq[p + d:p + d + k]))
    a, off, hits = one.path_many(pairs, limit=300)
    assert hits.size > 20
    want_rows = one.query_many(pairs, limit=300)
    want_paths, want_metas = one.path(*pairs[0])
    want_conn = [(ok, sorted((e.edge_type, e.edge_value) for e in es)) for ok, es in one.connected_many(pairs[:6])]
    mp.spawn(_query_worker, args=(2, _free_port(), k, pairs, str(tmp_path)), nprocs=2, join=True)
    seen_owner = set()
    for r in range(2):
        z = pickle.load(open(os.path.join(str(tmp_path), "q%d.pkl" % r), "rb"))
        assert z["off"].tolist() == off.tolist()
        for f in ("type_id", "i", "j"):                 # global type ids = the ids of the single-process graph
            assert z["hits"][f].tolist() == hits[f].tolist(), f
        assert z["rows"] == want_rows                   # edge_type labels and strings, pair by pair, in order
        assert z["paths"] == [[n.value for n in p] for p in want_paths] and z["metas"] == list(want_metas)
        assert z["conn"] == want_conn
        seen_owner |= set(z["owner"].tolist())
    assert seen_owner == {0, 1}
