"""Host-side bookkeeping around the device FASTA parser (pdg_ingest_fasta).

The device returns one row {byte_begin, byte_end, residue_index} per '>' line.  This module turns
those rows into the reference's ``headers`` / ``sequences`` lists exactly as Kmers._load builds
them (prairiedog/kmers.py:61-75): a sequence is flushed when the next header arrives only if it
is non-empty, and once more, unconditionally, at end of file.
"""
import numpy as np


class Layout(object):
    """Contig layout of one parsed file, in residue coordinates relative to the file."""

    def __init__(self, table, n_residues):
        table = np.asarray(table, dtype=np.uint64).reshape(-1, 3)
        self.n_headers = int(table.shape[0])
        self.n_residues = int(n_residues)
        cuts = np.concatenate([[0], table[:, 2].astype(np.int64), [self.n_residues]]).astype(np.int64)
        begin, end = cuts[:-1], cuts[1:]
        keep = end > begin
        keep[-1] = True                              # kmers.py:75: the last sequence is always appended
        self.seq_begin = begin[keep]
        self.seq_len = (end - begin)[keep]
        # headers[i] labels sequences[i] only if nothing precedes the first header and no record is empty
        self.aligned = bool(self.n_headers > 0 and int(table[0, 2]) == 0
                            and bool((end[1:] > begin[1:]).all()))

    def benign(self, k):
        """True when the reference's iterator simply skips every contig with len <= k: the layout
        is aligned, the last contig is long and no two short contigs are adjacent."""
        if not self.aligned:
            return False
        short = self.seq_len <= k
        if short[-1]:
            return False
        return not bool((short[1:] & short[:-1]).any())


def header_strings(host_bytes, table):
    """Header lines (verbatim, '>' included, right-stripped like kmers.py:63) as str."""
    buf = host_bytes if isinstance(host_bytes, (bytes, bytearray)) else bytes(memoryview(host_bytes))
    out = []
    for b, e, _ in np.asarray(table, dtype=np.uint64).reshape(-1, 3).tolist():
        out.append(buf[b:e].decode("utf-8", "replace").rstrip())
    return out


def host_load(path):
    """Escape path for files the 2-bit store cannot hold (bytes outside ACGT): the loader of
    prairiedog/kmers.py:61-75 on the host.  Only used when pdg_ingest_fasta reports n_dirty != 0."""
    headers, sequences, seq = [], [], []
    with open(path) as f:
        for line in f:
            ln = line.rstrip()
            if ln.startswith(">"):
                if seq:
                    sequences.append("".join(seq))
                    seq = []
                headers.append(ln)
            elif ln:
                seq.append(ln)
    sequences.append("".join(seq))
    return headers, sequences
