"""Deterministic synthetic E. coli-like FASTA genomes for benchmarks and tests.

Genome g is a base genome (i.i.d. uniform ACGT, PCG64(seed)) with a per-base substitution rate and
a few kbp-sized insertions / deletions drawn from PCG64(seed + 1 + g), cut into contigs at sorted
uniform cut points (minimum length enforced), written as 80-column FASTA with headers
``>syn{g:04d}|NODE_{c}_length_{len}`` (unique, no ' in ').  ``related=False`` makes every genome
independent (the worst case for deduplication).
"""
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
SEED = 20260000


def base_genome(n_bases, seed=SEED):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.integers(0, 4, size=n_bases, dtype=np.uint8)


def genome_digits(g, base, sub_rate=0.01, n_indels=20, related=True, seed=SEED):
    rng = np.random.Generator(np.random.PCG64(seed + 1 + g))
    if not related:
        return rng.integers(0, 4, size=base.size, dtype=np.uint8)
    d = base.copy()
    n_sub = rng.binomial(d.size, sub_rate)
    idx = rng.integers(0, d.size, size=n_sub)
    d[idx] = (d[idx] + rng.integers(1, 4, size=n_sub, dtype=np.uint8)) & 3
    for _ in range(n_indels):
        at = int(rng.integers(0, max(d.size - 6000, 1)))
        ln = int(rng.integers(1000, 5001)) if d.size > 20000 else int(rng.integers(5, 50))
        if rng.integers(0, 2):
            d = np.concatenate([d[:at], d[at + ln:]])
        else:
            d = np.concatenate([d[:at], rng.integers(0, 4, size=ln, dtype=np.uint8), d[at:]])
    return d


def cut_points(n, n_contigs, min_len, rng):
    """Sorted cut points giving n_contigs pieces of at least min_len bases."""
    n_contigs = max(1, min(n_contigs, n // max(min_len, 1)))
    free = n - n_contigs * min_len
    c = np.sort(rng.integers(0, free + 1, size=n_contigs - 1))
    return np.concatenate([[0], c + np.arange(1, n_contigs) * min_len, [n]]).astype(np.int64)


def fasta_bytes(g, digits, n_contigs=200, min_len=500, width=80, seed=SEED):
    """FASTA image (uint8 array) of one genome, plus its contig lengths."""
    rng = np.random.Generator(np.random.PCG64(seed + 500000 + g))
    cuts = cut_points(digits.size, n_contigs, min_len, rng)
    letters = ACGT[digits]
    parts = []
    lens = np.diff(cuts)
    for c in range(lens.size):
        n = int(lens[c])
        parts.append(np.frombuffer((">syn%04d|NODE_%d_length_%d\n" % (g, c, n)).encode(), dtype=np.uint8))
        seq = letters[cuts[c]:cuts[c + 1]]
        full, rem = divmod(n, width)
        body = np.empty(n + full + (1 if rem else 0), dtype=np.uint8)
        if full:
            blk = body[:full * (width + 1)].reshape(full, width + 1)
            blk[:, :width] = seq[:full * width].reshape(full, width)
            blk[:, width] = 10
        if rem:
            body[full * (width + 1):-1] = seq[full * width:]
            body[-1] = 10
        parts.append(body)
    return np.concatenate(parts), lens


def genomes(n_genomes, n_bases=5_000_000, n_contigs=200, related=True, seed=SEED, first=0):
    """Yields (name, fasta uint8 array, contig lengths) for genomes first .. first+n_genomes-1."""
    base = base_genome(n_bases, seed) if related else np.zeros(n_bases, dtype=np.uint8)
    for g in range(first, first + n_genomes):
        d = genome_digits(g, base, related=related, seed=seed)
        buf, lens = fasta_bytes(g, d, n_contigs=n_contigs, seed=seed)
        yield "syn%04d.fasta" % g, buf, lens
