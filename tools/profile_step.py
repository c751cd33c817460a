#!/usr/bin/env python
"""One warm-up build and one profiled build of G synthetic genomes (for ncu; not a benchmark)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from prairiedog_b200 import B200Graph, synth
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 11
    g = B200Graph(k=k)
    files = [(name, buf, torch.from_numpy(buf).cuda()) for name, buf, _ in synth.genomes(n)]
    for it in range(2):
        g.clear()
        g.add_fasta_batch(files)
        g.save()
        torch.cuda.synchronize()
    print(g.engine.counts(), g.engine.stats())


if __name__ == "__main__":
    main()
