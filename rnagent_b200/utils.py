"""Host-side helpers that keep the reference's signatures (rlif/rna/utils.py)."""
import numpy as np


def hamming_distance(seq1, seq2):
    """rlif/rna/utils.py:67-75: zip-truncating Hamming distance -> (hd, mismatch indices).

    Host-side twin of the distance the fold kernel fuses; used for single strings only.
    """
    a = np.frombuffer(seq1.encode() if isinstance(seq1, str) else bytes(seq1), dtype=np.uint8)
    b = np.frombuffer(seq2.encode() if isinstance(seq2, str) else bytes(seq2), dtype=np.uint8)
    n = min(len(a), len(b))
    mism = a[:n] != b[:n]
    return int(mism.sum()), np.where(mism)[0]
