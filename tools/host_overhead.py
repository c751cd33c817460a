#!/usr/bin/env python
"""Wall-clock split of one e2e step (host images) into library calls and Python bookkeeping."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    from prairiedog_b200 import B200Graph, synth
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    g = B200Graph(k=11)
    files = []
    for name, buf, _ in synth.genomes(n):
        p = torch.from_numpy(buf).pin_memory()
        files.append((name, p.numpy(), None))
    eng = g.engine
    orig_ingest, orig_contigs, orig_fin = eng.ingest_batch, eng.add_contigs, eng.finalize
    T = {}

    def timed(name, fn):
        def w(*a, **k):
            t0 = time.perf_counter()
            r = fn(*a, **k)
            T[name] = T.get(name, 0.0) + time.perf_counter() - t0
            return r
        return w
    eng.ingest_batch = timed("ingest_batch", orig_ingest)
    eng.add_contigs = timed("add_contigs", orig_contigs)
    eng.finalize = timed("finalize", orig_fin)
    for it in range(6):
        T.clear()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g.clear()
        t1 = time.perf_counter()
        g.add_fasta_batch(files)
        t2 = time.perf_counter()
        g.save()
        t3 = time.perf_counter()
        c = g.engine.counts()
        torch.cuda.synchronize()
        t4 = time.perf_counter()
        print("step %.2f ms: clear %.2f | add_fasta_batch %.2f (ingest_batch %.2f, add_contigs %.2f, python %.2f) | "
              "save %.2f (finalize %.2f) | counts %.2f" % (
                  1e3 * (t4 - t0), 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * T["ingest_batch"], 1e3 * T["add_contigs"],
                  1e3 * (t2 - t1 - T["ingest_batch"] - T["add_contigs"]), 1e3 * (t3 - t2), 1e3 * T["finalize"],
                  1e3 * (t4 - t3)))


if __name__ == "__main__":
    main()
