#!/usr/bin/env python
"""Writes tests/golden/bench_checksums.json: what bench.py must find at the BASELINE.json configs.

TEST INFRASTRUCTURE.  Everything here comes from oracle/pdg_oracle.c (the CPU restatement that is pinned to
the unmodified reference by tests/test_oracle_pinning.py) through its streaming forms -- orc_stream_sums,
orc_mark_codes, orc_emit_codes / orc_sort_unique / orc_merge_unique -- which tests/test_oracle_pinning.py
checks against orc_build / orc_checksum.  Nothing from prairiedog_b200's CUDA side is involved; the only product
module used is `synth`, the deterministic input generator bench.py itself uses.

Layouts (a rank numbers its edge types from 0, so an edge checksum belongs to a rank BLOCK of genomes):
  weak     config 3: rank r holds genomes [100 r, 100 r + 100), r = 0..7, for k = 11 and k = 31
  config4  1000 genomes, k = 31, rank r of N holds genomes [per r, per r + per), per = ceil(1000 / N), N = 2, 4, 8
Node tables are global: the union over the first 100 N genomes (weak, N = 1, 2, 4, 8) or over all 1000.

  python tests/golden/make_bench_checksums.py [--workers 8]          (about 10 minutes on 8 cores, < 40 GB)
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden", "bench_checksums.json")
BASES, CONTIGS, TOTAL, BATCH = 5_000_000, 200, 1000, 25
M64 = (1 << 64) - 1
# (layout name, block size, genomes covered)
LAYOUTS = [("weak", 100, 800), ("config4_n8", 125, 1000), ("config4_n4", 250, 1000), ("config4_n2", 500, 1000)]
NODE_PREFIXES = [100, 200, 400, 800, 1000]


def batch_job(args):
    """Genomes [g0, g0 + BATCH): per genome and k the k-mer / edge counts and the edge checksum under every
    layout's type numbering; per batch the k = 11 presence bitmap and the sorted distinct k = 31 codes."""
    g0, tmp = args
    from oracle import c_oracle
    from oracle.c_oracle import COracle
    from prairiedog_b200 import synth
    rows = []
    bitmap = np.zeros((4 ** 11) // 64, dtype=np.uint64)
    codes31 = []
    for name, buf, lens in synth.genomes(BATCH, n_bases=BASES, n_contigs=CONTIGS, first=g0):
        g = g0 + len(rows)
        row = {"g": g}
        for k in (11, 31):
            o = COracle(k)
            _, n_headers = o.add_fasta_bytes(buf)
            assert n_headers == lens.size
            # type id of contig c of genome g inside a block that starts at genome b: headers of [b, g) + c.
            # every synthetic genome has exactly `n_headers` = CONTIGS contigs (asserted by the caller)
            offs = [n_headers * (g % size) for _, size, _ in LAYOUTS]
            nk, ne, sums = o.stream_sums(offs)
            row["k%d" % k] = {"n_headers": int(n_headers), "n_kmers": nk, "n_edges": ne, "sums": sums}
            if k == 11:
                o.mark_codes(bitmap)
            else:
                codes31.append(o.distinct_codes())
            del o
        rows.append(row)
    allc = np.concatenate(codes31)
    del codes31
    uniq = c_oracle.sort_unique(allc, 62)
    np.save(os.path.join(tmp, "codes31_%04d.npy" % g0), uniq)
    np.save(os.path.join(tmp, "bitmap11_%04d.npy" % g0), bitmap)
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=min(8, os.cpu_count() or 1))
    ap.add_argument("--total", type=int, default=TOTAL, help="genomes (smaller values only for a dry run)")
    a = ap.parse_args()
    from oracle import c_oracle
    c_oracle.build()
    t0 = time.time()
    tmp = tempfile.mkdtemp(prefix="pdg_golden_")
    starts = list(range(0, a.total, BATCH))
    with mp.Pool(a.workers) as pool:
        per_batch = pool.map(batch_job, [(g0, tmp) for g0 in starts], chunksize=1)
    rows = [r for b in per_batch for r in b]
    assert [r["g"] for r in rows] == list(range(len(rows)))
    assert all(r["k11"]["n_headers"] == CONTIGS for r in rows), "a genome with != %d contigs breaks the type offsets" % CONTIGS
    print("per-genome sums done in %.0f s" % (time.time() - t0), flush=True)

    out = {"generator": "tests/golden/make_bench_checksums.py (oracle/pdg_oracle.c streaming forms)",
           "synth": {"bases": BASES, "contigs": CONTIGS, "seed": 20260000, "related": True},
           "checksum": "pdg_checksum / orc_checksum: nodes = sum mix64(code + golden ratio); edges = sum of the chained "
                       "mix64 over (src, tgt, type << 32 | value), type ids numbered from 0 inside a rank block",
           "blocks": {}, "nodes": {}}
    for li, (lname, size, cover) in enumerate(LAYOUTS):
        for k in (11, 31):
            if lname != "weak" and k != 31:
                continue
            blocks = []
            for b0 in range(0, min(cover, len(rows)), size):
                part = rows[b0:b0 + size]
                if len(part) < size:
                    break
                blocks.append({"first_genome": b0, "genomes": size,
                               "n_kmers": sum(r["k%d" % k]["n_kmers"] for r in part),
                               "n_edges": sum(r["k%d" % k]["n_edges"] for r in part),
                               "edges_sum": sum(r["k%d" % k]["sums"][li] for r in part) & M64})
            out["blocks"]["%s_k%d" % (lname, k)] = blocks

    # ---- global node tables of the prefixes ----
    bitmap = np.zeros((4 ** 11) // 64, dtype=np.uint64)
    union = np.zeros(0, dtype=np.uint64)
    done = 0
    for g0 in starts:
        bitmap |= np.load(os.path.join(tmp, "bitmap11_%04d.npy" % g0))
        union = c_oracle.merge_unique(union, np.load(os.path.join(tmp, "codes31_%04d.npy" % g0)))
        os.remove(os.path.join(tmp, "codes31_%04d.npy" % g0))
        os.remove(os.path.join(tmp, "bitmap11_%04d.npy" % g0))
        done = g0 + BATCH
        if done in NODE_PREFIXES:
            n11, s11 = c_oracle.bitmap_sum(bitmap)
            out["nodes"]["first_%d_k11" % done] = {"n_nodes": n11, "nodes_sum": s11}
            out["nodes"]["first_%d_k31" % done] = {"n_nodes": int(union.size), "nodes_sum": c_oracle.codes_sum(union)}
            print("first %d genomes: k11 %d nodes, k31 %d nodes (%.0f s)" % (done, n11, union.size, time.time() - t0),
                  flush=True)
    os.rmdir(tmp)
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote %s in %.0f s" % (OUT, time.time() - t0))


if __name__ == "__main__":
    main()
