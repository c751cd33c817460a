#!/usr/bin/env python
"""Config 5 of BASELINE.json: batched path queries (all paths between two k-mers) over a pangenome of
synthetic genomes.  Pairs follow SURVEY.md 8(d): 50 % positives (same contig, b = a + d, d in [1, 1000]),
25 % same genome / different contig, 25 % different genomes; L = 1000.  Every positive pair's own hit
(type, i, i + d - 1) must be among the returned hits (checked for all of them); a sample of queries is
checked hit for hit against the C oracle when it is available (--check N genomes only: the oracle sorts
on one core).  Prints one JSON line.

  python tools/bench_queries.py [--genomes G] [--k K] [--queries Q] [--check]
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genomes", type=int, default=100)
    ap.add_argument("--bases", type=int, default=5_000_000)
    ap.add_argument("--contigs", type=int, default=200)
    ap.add_argument("--k", type=int, default=11)
    ap.add_argument("--queries", type=int, default=100_000)
    ap.add_argument("--limit", type=int, default=1000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", action="store_true", help="compare a sample of queries with the C oracle")
    a = ap.parse_args()
    import torch
    from prairiedog_b200 import B200Graph, synth
    rng = np.random.Generator(np.random.PCG64(20261104))
    g = B200Graph(k=a.k)
    base = synth.base_genome(a.bases)
    files, digits, cuts = [], [], []
    for gi in range(a.genomes):
        d = synth.genome_digits(gi, base)
        buf, lens = synth.fasta_bytes(gi, d, n_contigs=a.contigs)
        files.append(("syn%04d.fasta" % gi, buf, torch.from_numpy(buf).cuda()))
        digits.append(d)
        cuts.append(np.concatenate([[0], np.cumsum(lens)]).astype(np.int64))
    t0 = time.time()
    g.add_fasta_batch(files)
    g.save()
    torch.cuda.synchronize()
    t_build = time.time() - t0
    type_base = np.concatenate([[0], np.cumsum([c.size - 1 for c in cuts])]).astype(np.int64)

    def pick(gi, n):
        """n random (contig, in-contig index) pairs of genome gi with room for k-mers"""
        c = rng.integers(0, cuts[gi].size - 1, size=n)
        ln = cuts[gi][c + 1] - cuts[gi][c]
        i = (rng.random(n) * (ln - a.k)).astype(np.int64)
        return c, i

    Q = a.queries
    kind = rng.choice(3, size=Q, p=[0.5, 0.25, 0.25])
    qa, qb = np.zeros(Q, dtype=np.uint64), np.zeros(Q, dtype=np.uint64)
    want = np.full((Q, 3), -1, dtype=np.int64)
    ga = rng.integers(0, a.genomes, size=Q)
    for gi in range(a.genomes):
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
        # same genome, another random place; different genome
        rest = ~ok
        same = rest & (kind[idx] != 2)
        c2, i2 = pick(gi, idx.size)
        qb[idx[same]] = codes_at(digits[gi], cuts[gi][c2] + i2, a.k)[same]
        other = rest & (kind[idx] == 2)
        g2 = (gi + 1 + rng.integers(0, max(a.genomes - 1, 1))) % a.genomes
        c3, i3 = pick(g2, idx.size)
        qb[idx[other]] = codes_at(digits[g2], cuts[g2][c3] + i3, a.k)[other]

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    off, hits = g.engine.query_paths(qa, qb, a.limit)       # warm-up
    torch.cuda.synchronize()
    e0.record()
    for _ in range(a.reps):
        off, hits = g.engine.query_paths(qa, qb, a.limit)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.reps
    # every positive pair's own hit must be present
    off = off.astype(np.int64)
    missing = 0
    pos_idx = np.nonzero(want[:, 0] >= 0)[0]
    hkey = (hits["type_id"].astype(np.int64) << 42) | (hits["i"].astype(np.int64) << 21) | hits["j"].astype(np.int64)
    qof = np.repeat(np.arange(Q), np.diff(off))
    wkey = (want[:, 0] << 42) | (want[:, 1] << 21) | want[:, 2]
    found = np.zeros(Q, dtype=bool)
    m = hkey == wkey[qof]
    found[qof[m]] = True
    missing = int((~found[pos_idx]).sum())
    checked = 0
    if a.check:
        from oracle.c_oracle import COracle
        o = COracle(a.k)
        for name, buf, _ in files:
            o.add_fasta_bytes(buf)
        o.build()
        for q in rng.choice(Q, size=min(2000, Q), replace=False).tolist():
            w = sorted(map(tuple, o.query(int(qa[q]), int(qb[q]), a.limit).tolist()))
            h = hits[off[q]:off[q + 1]]
            got = sorted((int(x["type_id"]), int(x["i"]), int(x["j"])) for x in h)
            assert got == w, (q, got[:3], w[:3])
            checked += 1
    print(json.dumps({"metric": "batched_path_queries_per_sec", "value": Q / (ms * 1e-3), "unit": "queries/s",
                      "ms_per_batch": ms, "queries": Q, "hits": int(hits.size), "positives": int(pos_idx.size),
                      "positives_missing_own_hit": missing, "oracle_checked": checked, "k": a.k, "limit": a.limit,
                      "genomes": a.genomes, "nodes": g.engine.counts()[0], "edges": g.engine.counts()[1],
                      "build_s_incl_python": t_build,
                      "includes": "host->device copy of the pairs, count + scan + fill kernels, device->host "
                                  "copy of offsets and hits"}))
    assert missing == 0


if __name__ == "__main__":
    main()
