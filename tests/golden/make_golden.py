#!/usr/bin/env python
"""Generates tests/golden/*.json by running the UNMODIFIED reference in the build container.

Needs /root/reference and oracle/_ref (python oracle/build_ref.py).  The GPU box has neither, so
the vectors are committed.  Run from the repo root:  python tests/golden/make_golden.py

Golden files:
  fixtures_k11.json       the 4 FASTA fixtures of the reference (tests/*.fa*, copied to tests/data)
                          each alone, the 3 SHORTENED together (== tests/pangenome.lemongraph) and
                          all 4 together (BASELINE config 2): Kmers stats, canonical nodes/edges,
                          query answers through LGGraph.connected/path
  prebuilt_pangenome.json canonical dump of the reference's golden DB tests/pangenome.lemongraph
  synthetic.json          seeded synthetic genomes (several k, multi-contig, repeats) through the
                          reference: stats, canonical graph checksum material, queries
  corner_cases.json       the FASTA corner cases of SURVEY.md A.1 (k=4): emission or exception
"""
import json
import os
import random
import shutil
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import live_reference as live  # noqa: E402

DATA = os.path.join(ROOT, "tests", "data")
SHORT = ["SRR1060582_SHORTENED.fasta", "SRR1106609_SHORTENED.fasta",
         "GCA_900015695.1_ED647_contigs_genomic_SHORTENED.fasta"]


def dump_graph(paths, k, queries):
    g, stats = live.build_graph(paths, k)
    nodes, edges = live.canonical(g)
    qs = []
    for a, b in queries:
        try:
            c, n_start, hits = live.query(g, a, b)
        except RecursionError:   # the reference crashes on long paths
            print("RecursionError in the reference for", a, b, [os.path.basename(p) for p in paths])
            qs.append({"a": a, "b": b, "raises": "RecursionError"})
            continue
        qs.append({"a": a, "b": b, "connected": c, "n_start": n_start, "hits": hits})
    return {"k": k, "files": [os.path.basename(p) for p in paths], "stats": stats,
            "nodes": nodes, "edges": [list(e) for e in edges], "queries": qs}


def synth_fasta(path, seed, contig_lens, repeat=None, width=60):
    rnd = random.Random(seed)
    with open(path, "w") as f:
        for c, n in enumerate(contig_lens):
            seq = "".join(rnd.choice("ACGT") for _ in range(n))
            if repeat and c == repeat[0]:
                unit = seq[:repeat[1]]
                seq = (unit * (n // len(unit) + 1))[:n]
            f.write(">syn%d|contig_%d_len_%d\n" % (seed, c, n))
            for i in range(0, n, width):
                f.write(seq[i:i + width] + "\n")


def pick_queries(paths, k, rnd, n):
    """positive pairs at various distances, a different-contig pair, an absent k-mer"""
    from oracle import ref_spec
    qs = []
    contigs = []
    for p in paths:
        h, s = ref_spec.load_fasta(p)
        contigs += [x for x in s if len(x) > k + 1]
    for _ in range(n):
        s = rnd.choice(contigs)
        i = rnd.randrange(0, len(s) - k)
        d = rnd.choice([1, 2, 5, 30, 100])
        j = min(i + d, len(s) - k)
        if j > i:
            qs.append((s[i:i + k], s[j:j + k]))
    a, b = contigs[0], contigs[-1]
    qs.append((a[0:k], b[len(b) - k:]))
    qs.append((a[1:k + 1], a[0:k]))          # backwards: not connected unless repeated
    qs.append(("A" * k, a[0:k]))
    return qs


def main():
    out = {}
    fx = {}
    pre_q = [("CCGGAAGAAAA", "CGGAAGAAAAA"), ("ATACGACGCCA", "CGTCCGGACGT"), ("GCTGGATACGT", "CGTCCGGACGT")]
    for name in ["15.fa"] + SHORT:
        fx[name] = dump_graph([os.path.join(DATA, name)], 11, pre_q)
    fx["shortened3"] = dump_graph([os.path.join(DATA, n) for n in SHORT], 11, pre_q)
    fx["all4"] = dump_graph([os.path.join(DATA, n) for n in ["15.fa"] + SHORT], 11, pre_q)
    json.dump(fx, open(os.path.join(HERE, "fixtures_k11.json"), "w"), indent=0, sort_keys=True)

    # the reference's own golden DB (tests/pangenome.lemongraph), copied because LMDB writes a lock
    ns = live.load()
    tmp = tempfile.mkdtemp()
    db = os.path.join(tmp, "pangenome.lemongraph")
    shutil.copy2(os.path.join(live.REF, "tests", "pangenome.lemongraph"), db)
    g = ns.LGGraph(db, delete_on_exit=False)
    nodes, edges = live.canonical(g)
    qs = []
    for a, b in pre_q:
        c, n_start, hits = live.query(g, a, b)
        qs.append({"a": a, "b": b, "connected": c, "n_start": n_start, "hits": hits})
    json.dump({"nodes": nodes, "edges": [list(e) for e in edges], "queries": qs},
              open(os.path.join(HERE, "prebuilt_pangenome.json"), "w"), indent=0, sort_keys=True)

    # synthetic genomes through the reference (small: LemonGraph does ~20 kbp/s)
    syn = {}
    sdir = os.path.join(ROOT, "tests", "data")
    cases = {
        "syn_k11": (11, [(101, [700, 45, 1200, 12, 300], None), (102, [900, 400], None)]),
        "syn_k5_repeats": (5, [(201, [150, 230], (0, 7)), (202, [200], None)]),   # ln < 239: SURVEY.md R2
        "syn_k16": (16, [(301, [500, 260], None)]),
        "syn_k17": (17, [(302, [500, 260], None)]),
        "syn_k31": (31, [(401, [800, 32, 500], None), (402, [640], (0, 40))]),
        "syn_k3": (3, [(501, [60, 4, 3, 50], None)]),
    }
    rnd = random.Random(7)
    for name, (k, files) in cases.items():
        paths = []
        for seed, lens, rep in files:
            p = os.path.join(sdir, "syn%d.fasta" % seed)
            synth_fasta(p, seed, lens, rep)
            paths.append(p)
        syn[name] = dump_graph(paths, k, pick_queries(paths, k, rnd, 12))
    json.dump(syn, open(os.path.join(HERE, "synthetic.json"), "w"), indent=0, sort_keys=True)

    # corner cases of Kmers, k = 4
    corner = {}
    texts = {
        "crlf_blank_multiline": ">a\r\nACGTA\r\n\r\nCGT\r\n>b\r\nTTTTACG\r\n",
        "short_middle": ">a\nACGTAC\n>b\nAC\n>c\nGGGTTTA\n",
        "short_first_eqk": ">a\nACGT\n>b\nGGGTTTA\n",
        "two_short": ">a\nACGTAC\n>b\nAC\n>c\nGG\n>d\nTTTTTT\n",
        "last_short": ">a\nACGTAC\n>b\nAC\n",
        "empty_record": ">a\n>b\nACGTA\n>c\nGGGGGG\n",
        "leading_text": "ACGTACGT\n>a\nGGGGTTTT\n",
        "lower_and_n": ">a\nacgtNACGT\n",
        "no_trailing_newline": ">a\nACGTACGT",
        "dup_header": ">a\nACGTAC\n>a\nACGTTT\n",
        "trailing_blanks": ">a  \nACGTA  \nCC\t\n",
    }
    cdir = os.path.join(sdir, "corner")
    os.makedirs(cdir, exist_ok=True)
    for name, text in texts.items():
        p = os.path.join(cdir, name + ".fasta")
        with open(p, "w", newline="") as f:
            f.write(text)
        try:
            km = ns.Kmers(p, 4)
            em = []
            while km.has_next:
                em.append(list(km.next()))
            km.reset()
            rec = {"headers": km.headers, "sequences": km.sequences, "len": len(km),
                   "unique": km.unique_kmers, "emitted": em}
            g, stats = live.build_graph([p], 4)
            nodes, edges = live.canonical(g)
            rec.update({"nodes": nodes, "edges": [list(e) for e in edges], "returned": stats[0]["returned"]})
        except Exception as e:  # noqa: BLE001 - the exception type is the golden value
            rec = {"raises": type(e).__name__}
        corner[name] = rec
    json.dump(corner, open(os.path.join(HERE, "corner_cases.json"), "w"), indent=0, sort_keys=True)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
