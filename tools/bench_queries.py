#!/usr/bin/env python
"""Config 5 of BASELINE.json: batched path queries (all paths between two k-mers) over a pangenome of synthetic
genomes, on 1 GPU or -- under torchrun -- over the N-rank graph (every rank holds the index of its block of genomes,
B200Graph.path_many is collective and returns the hits of the whole pangenome).

Pairs follow SURVEY.md 8(d): 50 % positives (same contig, b = a + d, d in [1, 1000]), 25 % same genome / another
place, 25 % different genomes; L = 1000; seed 20261104.  Checks, on every line:
  * every positive pair's own hit (type, i, i + d - 1) is among the returned hits (all positives);
  * --check N (default 2000) sampled pairs are compared HIT FOR HIT with oracle/pdg_oracle.c.  A hit lies inside one
    contig, so the hits of a pair partition by genome: the oracle holds --oracle-genomes of the genomes (spread over
    all ranks) and the comparison is exact for every hit whose edge type belongs to one of them.

  python tools/bench_queries.py [--genomes G] [--k K] [--queries Q] [--check N] [--strings]
  python -m torch.distributed.run --nproc-per-node N ... tools/bench_queries.py --genomes 1000 --k 31
Prints one JSON line (rank 0).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# ASCII letter index of synth (A C G T) -> 2-bit code of the library (A=0 T=1 C=2 G=3)
DIGIT_TO_CODE = np.array([0, 2, 3, 1], dtype=np.uint64)


def codes_at(digits, pos, k):
    """k-mer codes starting at the given positions of one genome's digit array."""
    out = np.zeros(pos.size, dtype=np.uint64)
    for j in range(k):
        out = (out << np.uint64(2)) | DIGIT_TO_CODE[digits[pos + j]]
    return out


def make_pairs(a, src, digits, cuts, type_base, rng):
    """Q pairs drawn from the genomes in ``src`` (digits / cuts / type_base are dicts keyed by genome)."""
    def pick(gi, n):
        c = rng.integers(0, cuts[gi].size - 1, size=n)
        ln = cuts[gi][c + 1] - cuts[gi][c]
        i = (rng.random(n) * (ln - a.k)).astype(np.int64)
        return c, i

    Q = a.queries
    kind = rng.choice(3, size=Q, p=[0.5, 0.25, 0.25])
    qa, qb = np.zeros(Q, dtype=np.uint64), np.zeros(Q, dtype=np.uint64)
    want = np.full((Q, 3), -1, dtype=np.int64)
    ga = src[rng.integers(0, src.size, size=Q)]
    for gi in src.tolist():
        idx = np.nonzero(ga == gi)[0]
        if idx.size == 0:
            continue
        c, i = pick(gi, idx.size)
        ln = cuts[gi][c + 1] - cuts[gi][c]
        qa[idx] = codes_at(digits[gi], cuts[gi][c] + i, a.k)
        # positives: b = a + d inside the same contig
        pos = kind[idx] == 0
        d = rng.integers(1, a.limit + 1, size=idx.size)
        d = np.minimum(d, np.maximum(ln - a.k - i, 1))
        ok = pos & (i + d + a.k <= ln)
        jb = np.where(ok, i + d, i)
        bpos = codes_at(digits[gi], cuts[gi][c] + jb, a.k)
        qb[idx[ok]] = bpos[ok]
        want[idx[ok], 0] = type_base[gi] + c[ok]
        want[idx[ok], 1] = i[ok]
        want[idx[ok], 2] = (i + d - 1)[ok]
        # same genome, another random place; a different genome
        rest = ~ok
        same = rest & (kind[idx] != 2)
        c2, i2 = pick(gi, idx.size)
        qb[idx[same]] = codes_at(digits[gi], cuts[gi][c2] + i2, a.k)[same]
        other = rest & (kind[idx] == 2)
        g2 = int(src[(int(np.nonzero(src == gi)[0][0]) + 1 + int(rng.integers(0, max(src.size - 1, 1)))) % src.size])
        c3, i3 = pick(g2, idx.size)
        qb[idx[other]] = codes_at(digits[g2], cuts[g2][c3] + i3, a.k)[other]
    return qa, qb, want, ga


def oracle_check(a, rng, qa, qb, ga, off, hits, orc_genomes, bufs, type_base, n_check):
    """n_check sampled pairs hit for hit against oracle/pdg_oracle.c holding ``orc_genomes``: GPU hits whose type id
    belongs to one of those genomes == the oracle's hits (type ids mapped back to the global numbering)."""
    from oracle.c_oracle import COracle
    o = COracle(a.k)
    glob_of_local = []                           # oracle type id -> global type id
    for gi in orc_genomes:
        base, n = o.add_fasta_bytes(bufs[gi])
        glob_of_local += list(range(type_base[gi], type_base[gi] + n))
    o.build()
    glob_of_local = np.array(glob_of_local, dtype=np.int64)
    held = np.zeros(max(type_base.values()) + 100000, dtype=bool)
    held[glob_of_local] = True
    checked = compared_hits = 0
    tid = hits["type_id"].astype(np.int64)
    # half of the sample from pairs whose own genome the oracle holds (their positives must be found there)
    own = np.nonzero(np.isin(ga, np.array(orc_genomes)))[0]
    n_own = min(own.size, n_check // 2)
    sample = np.concatenate([rng.choice(own, size=n_own, replace=False),
                             rng.choice(qa.size, size=min(n_check, qa.size) - n_own, replace=False)])
    for q in sample.tolist():
        w = o.query(int(qa[q]), int(qb[q]), a.limit)
        want = sorted((int(glob_of_local[t]), int(i), int(j)) for t, i, j in w.tolist())
        lo, hi = int(off[q]), int(off[q + 1])
        sel = held[tid[lo:hi]]
        got = sorted(zip(tid[lo:hi][sel].tolist(), hits["i"][lo:hi][sel].tolist(), hits["j"][lo:hi][sel].tolist()))
        assert got == want, ("pair %d differs from the oracle" % q, got[:3], want[:3])
        checked += 1
        compared_hits += len(want)
    return checked, compared_hits


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genomes", type=int, default=100, help="genomes of the whole pangenome (all ranks)")
    ap.add_argument("--bases", type=int, default=5_000_000)
    ap.add_argument("--contigs", type=int, default=200)
    ap.add_argument("--k", type=int, default=11)
    ap.add_argument("--queries", type=int, default=100_000)
    ap.add_argument("--limit", type=int, default=1000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", type=int, default=2000, help="pairs compared hit for hit with the C oracle (0: none)")
    ap.add_argument("--oracle-genomes", type=int, default=6)
    ap.add_argument("--source-genomes", type=int, default=48, help="genomes the pairs are drawn from (spread over all)")
    ap.add_argument("--strings", action="store_true", help="also spell every hit (pdg_query_strings) and time it")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from prairiedog_b200 import B200Graph, synth, distributed as D
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    g = B200Graph(k=a.k, device=local)
    D.attach(g, dev)
    per = -(-a.genomes // world)
    first, count = min(rank * per, a.genomes), min(per, a.genomes - min(rank * per, a.genomes))
    # genomes the pairs come from / the oracle holds: spread evenly, so every rank owns some of each
    src = np.unique(np.linspace(0, a.genomes - 1, min(a.source_genomes, a.genomes)).astype(np.int64))
    orc = src[np.unique(np.linspace(0, src.size - 1, min(a.oracle_genomes, src.size)).astype(np.int64))].tolist()
    base = synth.base_genome(a.bases)
    digits, cuts, bufs, type_base = {}, {}, {}, {}
    for gi in src.tolist():            # every synthetic genome has exactly --contigs contigs (checked below)
        d = synth.genome_digits(gi, base)
        buf, lens = synth.fasta_bytes(gi, d, n_contigs=a.contigs)
        assert lens.size == a.contigs
        digits[gi], cuts[gi], type_base[gi] = d, np.concatenate([[0], np.cumsum(lens)]).astype(np.int64), a.contigs * gi
        if gi in orc:
            bufs[gi] = buf
    t0 = time.time()
    files = []
    for name, buf, lens in synth.genomes(count, n_bases=a.bases, n_contigs=a.contigs, first=first):
        assert lens.size == a.contigs
        files.append((name, buf, torch.from_numpy(buf).to(dev)))
    g.add_fasta_batch(files)
    g.save()
    torch.cuda.synchronize()
    t_build = time.time() - t0
    del files

    rng = np.random.Generator(np.random.PCG64(20261104))           # the same pairs on every rank
    qa, qb, want, ga = make_pairs(a, src, digits, cuts, type_base, rng)
    Q = a.queries
    limit = a.limit

    def run_local():
        return g.engine.query_paths(qa, qb, limit)

    def run_global():
        if world == 1:
            return run_local()
        off_t, hits_t = g.engine.query_paths_device(qa, qb, limit)     # what B200Graph.path_many does at world > 1
        goff, ghits, _ = D.exchange_hits_device(off_t, hits_t, int(g._type_bases[rank]))
        return goff, ghits

    if world > 1:
        g._type_bases = D.type_bases(g._next_type, dev)
        assert int(g._type_bases[rank]) == a.contigs * first           # global ids = contigs x genome index

    def timed(fn):
        fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / a.reps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    ms_local, (loff, lhits) = timed(run_local)
    ms, (off, hits) = timed(run_global) if world > 1 else (ms_local, (loff, lhits))
    line = {}
    if a.strings:            # spelling the hits of this rank's shard (they live where the residues are)
        import ctypes
        run_local()
        g.engine.query_strings()                 # warm-up: allocates the pinned result buffer once
        torch.cuda.synchronize()
        t1 = time.time()
        soff, sbytes = g.engine.query_strings()
        dt = time.time() - t1
        # the same without the device->host copy of the characters: sizing + decode kernel + offsets only
        tot = ctypes.c_uint64()
        t2 = time.time()
        g.engine.lib.pdg_query_strings_size(g.engine.h, ctypes.byref(tot))
        g.engine.lib.pdg_query_strings_fetch(g.engine.h, soff.ctypes.data, None)
        dt_dev = time.time() - t2
        lens = lhits["j"].astype(np.int64) - lhits["i"].astype(np.int64) + 1 + a.k
        x = int(rng.integers(0, max(lhits.size, 1)))
        if lhits.size:       # one spot check against the contig slice
            h = lhits[x]
            s = sbytes[int(soff[x]):int(soff[x]) + int(lens[x])].tobytes().decode("ascii")
            tid = int(h["type_id"])
            fl = g._files[tid // a.contigs]
            begin = fl.residue_base + int(fl.seq_begin[tid % a.contigs])
            assert s == g.engine.fetch_residues(begin + int(h["i"]), int(lens[x]))
        sms = torch.tensor([dt * 1e3], device=dev)
        if world > 1:
            dist.all_reduce(sms, op=dist.ReduceOp.MAX)
        line["strings"] = {"ms_per_batch_rank_max": float(sms.item()), "bytes_rank0": int(sbytes.size),
                           "hits_rank0": int(lhits.size), "gb_per_s_rank0": sbytes.size / dt / 1e9,
                           "vs_hits_only": float(sms.item()) / ms_local,
                           "ms_without_d2h_of_characters_rank0": dt_dev * 1e3,
                           "bound": "device->host copy of the characters over PCIe: bytes / ~50 GB/s",
                           "what": "pdg_query_strings_size + _fetch on the hits of the rank's shard: sizing, decode "
                                   "kernel and the device->host copy of the ASCII into pinned memory"}
    if rank == 0:
        off = off.astype(np.int64)
        pos_idx = np.nonzero(want[:, 0] >= 0)[0]
        hkey = (hits["type_id"].astype(np.int64) << 42) | (hits["i"].astype(np.int64) << 21) | hits["j"].astype(np.int64)
        qof = np.repeat(np.arange(Q), np.diff(off))
        wkey = (want[:, 0] << 42) | (want[:, 1] << 21) | want[:, 2]
        found = np.zeros(Q, dtype=bool)
        m = hkey == wkey[qof]
        found[qof[m]] = True
        missing = int((~found[pos_idx]).sum())
        checked = compared = 0
        if a.check:
            checked, compared = oracle_check(a, rng, qa, qb, ga, off, hits, orc, bufs, type_base, a.check)
        line.update({"metric": "batched_path_queries_per_sec", "value": Q / (ms * 1e-3), "unit": "queries/s",
                     "n_gpus": world, "ms_per_batch": ms, "ms_per_batch_local_kernels": ms_local,
                     "ms_per_batch_local_kernels_note": "this rank's shard only, through Engine.query_paths: kernels + the "
                                                        "copy of its hits into a PAGEABLE host array (the collective form "
                                                        "keeps them on the device and copies the merged answer out once, "
                                                        "through pinned memory -- it can be the faster of the two)", "queries": Q,
                     "hits": int(hits.size), "positives": int(pos_idx.size), "positives_missing_own_hit": missing,
                     "oracle_checked": checked, "oracle_compared_hits": compared, "oracle_genomes": orc,
                     "k": a.k, "limit": a.limit, "genomes": a.genomes, "genomes_per_gpu": per,
                     "nodes": g.engine.counts()[0], "edges_rank0": g.engine.counts()[1],
                     "build_s_incl_python": t_build,
                     "includes": "host->device copy of the pairs, count + scan + fill kernels, device->host copy of "
                                 "offsets and hits" + ("; N > 1: plus the all-gather of every rank's counts and hit "
                                                        "records and the merge into pair order" if world > 1 else "")})
        print(json.dumps(line))
        assert missing == 0
        assert checked >= min(a.check, Q)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
