"""Synthetic workloads named by BASELINE.json `configs` and the work model.

Host-side numpy only; nothing here folds.  The generators are deterministic in their seeds so the
GPU run, the CPU baseline and the parity tests see the same bytes.
"""
import numpy as np

ALPHABET = np.frombuffer(b"ACGU", dtype=np.uint8)

# Lengths of the Eterna100 targets (rlif/data/eterna/1.rna .. 100.rna, one dot-bracket line each):
# min 12, median 104, mean 159, max 400.  Config C5 resamples sequence lengths from this multiset.
ETERNA100_LENGTHS = (
    16, 116, 36, 192, 101, 379, 284, 12, 258, 45, 36, 213, 67, 98, 30, 105, 105, 67, 104, 42, 105, 400, 17, 104,
    62, 26, 57, 378, 105, 31, 102, 122, 45, 96, 123, 151, 103, 316, 174, 38, 35, 116, 85, 63, 63, 97, 50, 102,
    101, 105, 338, 80, 400, 92, 102, 387, 36, 111, 68, 105, 67, 53, 400, 110, 40, 36, 61, 282, 200, 184, 88, 73,
    370, 380, 75, 393, 105, 284, 101, 397, 212, 214, 119, 104, 389, 355, 97, 34, 97, 400, 392, 100, 116, 389, 82,
    358, 400, 202, 364, 382)


def eterna_mixed_lengths(count, seed=1999):
    """C5: `count` lengths drawn with replacement from the Eterna100 length multiset."""
    rng = np.random.default_rng(seed)
    return np.asarray(ETERNA100_LENGTHS, dtype=np.int64)[rng.integers(0, 100, size=count)]


def random_sequences(count, n, seed):
    """`count` i.i.d. uniform ACGU sequences of length n -> (uint8 concatenation, int32 CSR offsets)."""
    rng = np.random.default_rng(seed)
    buf = ALPHABET[rng.integers(0, 4, size=(count, n), dtype=np.uint8)].reshape(-1)
    off = (np.arange(count + 1, dtype=np.int64) * n).astype(np.int32)
    return np.ascontiguousarray(buf), off


def random_sequences_mixed(lengths, seed):
    """Uniform ACGU sequences with the given lengths (CSR)."""
    rng = np.random.default_rng(seed)
    lengths = np.asarray(lengths, dtype=np.int64)
    off = np.zeros(len(lengths) + 1, dtype=np.int32)
    off[1:] = np.cumsum(lengths)
    buf = ALPHABET[rng.integers(0, 4, size=int(off[-1]), dtype=np.uint8)]
    return np.ascontiguousarray(buf), off


def random_structure(n, rng, p_open=0.35, min_loop=3):
    """A random valid dot-bracket string of length n (hairpin loops >= min_loop unpaired bases)."""
    out = np.full(n, ord("."), dtype=np.uint8)
    stack = []
    last_open = -10
    for k in range(n):
        r = rng.random()
        remaining = n - k
        if stack and (remaining <= len(stack) or (r < 0.3 and k - last_open > min_loop)):
            if k - last_open > min_loop:
                out[k] = ord(")")
                stack.pop()
                continue
        if r > 1.0 - p_open and remaining > len(stack) + min_loop + 2:
            out[k] = ord("(")
            stack.append(k)
            last_open = k
    for k in stack:  # unmatched openers become unpaired
        out[k] = ord(".")
    return out


def random_targets(off, seed):
    """One random dot-bracket target per CSR slot (same offsets as the sequences)."""
    rng = np.random.default_rng(seed)
    total = int(off[-1])
    out = np.empty(total, dtype=np.uint8)
    # a pool of structures per distinct length keeps generation O(#lengths), not O(B) Python loops
    lens = np.diff(off)
    pools = {}
    for n in np.unique(lens):
        pools[int(n)] = [random_structure(int(n), rng) for _ in range(64)]
    pick = rng.integers(0, 64, size=len(lens))
    for b, n in enumerate(lens):
        out[off[b]:off[b + 1]] = pools[int(n)][pick[b]]
    return out


def interior_candidates(d, maxloop=30, turn=3):
    """I(d): number of (n1, n2) with n1 + n2 <= MAXLOOP and an inner span q - p >= TURN + 1."""
    lim = min(maxloop, d - 2 - (turn + 1))
    return 0 if lim < 0 else (lim + 1) * (lim + 2) // 2


def work_candidates(n, maxloop=30, turn=3):
    """W(n) of SURVEY.md section 8d: sequence-independent count of min-plus candidates of one fold."""
    w = 0
    for d in range(turn + 1, n):
        w += (n - d) * (interior_candidates(d, maxloop, turn) + max(0, d - 2 * turn - 2) + 3)
    w += sum(j - (turn + 1) for j in range(turn + 2, n + 1))
    return w


def algorithmic_hbm_bytes(n):
    """Bytes one fold must move through HBM: sequence, target, structure (n each) + energy + HD."""
    return 3 * n + 8
