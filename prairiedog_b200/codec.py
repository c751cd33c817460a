"""k-mer string <-> integer code.

Encoding (include/pdgpu.h): code = sum(base[j] << 2*(k-1-j)) with A=0, T=1, C=2, G=3 -- the
enumeration order of itertools.product('ATCG') used by prairiedog/kmers.py:17-24 (possible_kmers)
and by the reference's kmer_map fixture (tests/conftest.py:195-203), so code + 1 is that fixture's
node id.  Only upper-case ACGT strings of length exactly k have a code; every other node value is
an opaque key interned by the host (bit 63 set).
"""
import numpy as np

ALPHABET = "ATCG"
OPAQUE_BIT = 1 << 63
_LUT = np.full(256, 255, dtype=np.uint8)
for _i, _c in enumerate(ALPHABET):
    _LUT[ord(_c)] = _i
_CHR = np.frombuffer(ALPHABET.encode(), dtype=np.uint8)


def encode(value, k):
    """Code of a clean k-mer, or None."""
    if not isinstance(value, str) or len(value) != k or k > 31:
        return None
    try:
        b = np.frombuffer(value.encode("ascii"), dtype=np.uint8)
    except UnicodeEncodeError:
        return None
    d = _LUT[b]
    if (d == 255).any():
        return None
    code = 0
    for x in d.tolist():
        code = (code << 2) | x
    return code


def decode(code, k):
    return "".join(ALPHABET[(code >> (2 * (k - 1 - j))) & 3] for j in range(k))


def decode_many(codes, k):
    """List of k-mer strings for a uint64 array of codes."""
    codes = np.asarray(codes, dtype=np.uint64)
    if codes.size == 0:
        return []
    shifts = (2 * (k - 1 - np.arange(k, dtype=np.uint64))).astype(np.uint64)
    digits = ((codes[:, None] >> shifts[None, :]) & np.uint64(3)).astype(np.intp)
    txt = _CHR[digits]                       # (n, k) uint8
    flat = txt.tobytes().decode("ascii")
    return [flat[i * k:(i + 1) * k] for i in range(codes.size)]


def windows_to_codes(seq, k):
    """Codes of every k-mer of an upper-case ACGT string (host helper for small API-path inputs)."""
    d = _LUT[np.frombuffer(seq.encode("ascii"), dtype=np.uint8)].astype(np.uint64)
    n = len(seq) - k + 1
    if n <= 0:
        return np.zeros(0, dtype=np.uint64)
    out = np.zeros(n, dtype=np.uint64)
    for j in range(k):
        out = (out << np.uint64(2)) | d[j:j + n]
    return out
