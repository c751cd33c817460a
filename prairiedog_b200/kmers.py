"""Kmers: FASTA loader + sliding-window k-mer iterator -- API mirror of prairiedog/kmers.py:27-150.

Same constructor, attributes (filepath, headers, sequences, li, pi, k, unique_kmers), iterator
protocol (has_next / contig_has_next / next / reset), ``len()`` and ``str()`` as the reference, but
the file is parsed and 2-bit packed by the CUDA library (pdg_ingest_fasta) and the total / distinct
k-mer counts come from pdg_kmer_stats instead of a Python set of 5 million strings.
"""
import itertools
import logging
import os
import typing

import numpy as np

from . import fasta
from .engine import Engine

log = logging.getLogger("prairiedog")


def possible_kmers(k: int = 11) -> typing.Generator:
    """All k-mers in the order of their integer codes (prairiedog/kmers.py:17-24)."""
    return (''.join(x) for x in itertools.product('ATCG', repeat=k))


class Kmers:
    def __init__(self, filepath, k=11, device=0):
        self.filepath = filepath
        self.li = 0  # current sequence
        self.pi = 0  # current position in the sequence
        self.k = k
        self.device = device
        self._sequences = None
        self._load()
        self._n = 0
        self.unique_kmers = self._count_unique()

    def __str__(self):
        return os.path.basename(self.filepath)

    def __len__(self):
        return self._n

    # ---- loading (kmers.py:51-77) ----
    def _load(self):
        self._host = np.fromfile(self.filepath, dtype=np.uint8)
        self._engine = Engine(self.k, self.device)
        self._d_bytes = self._engine.to_device(self._host)
        info, table = self._engine.ingest(self._d_bytes)
        self._info, self._table = info, table
        self.dirty = int(info.n_dirty) != 0
        if self.dirty:
            # bytes the 2-bit store cannot hold (N, IUPAC, lower case): exact host escape path
            self.headers, self._sequences = fasta.host_load(self.filepath)
            self.layout = None
            return
        self.headers = fasta.header_strings(self._host, table)
        self.layout = fasta.Layout(table, int(info.n_residues))

    @property
    def sequences(self):
        """The merged sequences (kmers.py:72-75), decoded from the device store on first use."""
        if self._sequences is None:
            lay = self.layout
            whole = self._engine.fetch_residues(0, lay.n_residues) if lay.n_residues else ""
            self._sequences = [whole[b:b + n] for b, n in zip(lay.seq_begin.tolist(), lay.seq_len.tolist())]
        return self._sequences

    @property
    def well_formed(self) -> bool:
        """True when the bulk device path reproduces the reference's emission exactly: clean
        alphabet, headers aligned with sequences, short contigs only where the reference's
        iterator simply skips them."""
        return (not self.dirty) and self.layout.benign(self.k)

    # ---- iterator protocol (kmers.py:79-132) ----
    def _end_of_kmers(self) -> bool:
        return (self.pi + self.k) > len(self.sequences[self.li])

    @property
    def has_next(self) -> bool:
        is_last_sequence = self.li == (len(self.sequences) - 1)
        return not (is_last_sequence and self._end_of_kmers())

    @property
    def contig_has_next(self) -> bool:
        return not self._end_of_kmers()

    def next(self) -> typing.Tuple[str, str]:
        if not self.has_next:
            return "", ""
        if not self.contig_has_next:
            self.li += 1
            self.pi = 0
        if self.k > len(self.sequences[self.li]) - 1:
            log.warning("Contig {} in file {} was shorter than K of {}, skipping...".format(
                self.headers[self.li], self.filepath, self.k))
            self.li += 1
            self.pi = 0
        header = self.headers[self.li]
        sl = self.sequences[self.li][self.pi: self.pi + self.k]
        self.pi += 1
        return header, sl

    def reset(self):
        self.li = 0
        self.pi = 0

    # ---- counts (kmers.py:134-150) ----
    def _count_unique(self) -> int:
        if self.well_formed:
            n, u = self._engine.kmer_stats(self.layout.seq_begin, self.layout.seq_len)
            self._n = n
            return u
        # irregular layouts (SURVEY.md A.1 corner cases) and dirty alphabets: walk the iterator
        # like the reference so that counts -- and the IndexError cases -- match it exactly
        st = set()
        c = 0
        while self.has_next:
            _, kmer = self.next()
            st.add(kmer)
            c += 1
        self.reset()
        self._n = c
        return len(st)
