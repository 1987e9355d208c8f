"""Batched drop-in for the reference's local search, `Solution.permute` (rlif/rna/solution.py:157-219),
as `Solution.evaluate(permute=True)` reaches it (rlif/rna/solution.py:133-141).

The reference runs, per solution, up to `permutation_budget` sequential rounds of "mutate around the
mismatches, RNA.fold, keep if the Hamming distance improves".  Here every round folds the candidates of
ALL solutions in one batched GPU fold (`rnafold_permute_batch`, csrc/permute.cu); there is no CPU path.

Python's `random` stream cannot be reproduced on a GPU, so the draws are a counter-based hash of
(seed, sequence number, round, position, kind) -- `draw()` below mirrors the device function bit for bit and
is what tools/make_permute_golden.py patches into the reference's own `permute` to record golden vectors.
"""
import collections
import ctypes

import numpy as np

from . import _lib
from .vienna import mfe_float, pack, unpack

_M = (1 << 64) - 1

# rlif/config/RnaDesign.yml:34-38 (training) and rlif/config/Testing.yml:4-8 (inference)
DEFAULTS = dict(permutation_budget=50, permutation_radius=1, permutation_threshold=5, allow_gu_permutations=False,
                mutation_probability=0.5)
TESTING = dict(permutation_budget=15, permutation_radius=5, permutation_threshold=15, allow_gu_permutations=True,
               mutation_probability=0.5)


def _mix64(x):
    x = (x + 0x9E3779B97F4A7C15) & _M
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M
    return x ^ (x >> 31)


def draw(seed, index, step, pos, kind):
    """64-bit draw of round `step` at position `pos` of sequence number `index`
    (kind 0: random.random(), 1: random.randint(0, 3), 2: random.randint(0, 5)); == rnafold_permute_draw."""
    key = _mix64((seed + index) & _M)
    return _mix64(key ^ ((step & 0xFFFFFFFF) << 32 | (pos & 0xFFFFFFFF) << 4 | kind))


def draw_random(seed, index, step, pos):
    return (draw(seed, index, step, pos, 0) >> 11) * (1.0 / 9007199254740992.0)


def draw_randint(seed, index, step, pos, hi):
    """random.randint(0, hi) for hi in (3, 5)."""
    return (draw(seed, index, step, pos, 1 if hi == 3 else 2) >> 33) % (hi + 1)


PermuteResult = collections.namedtuple("PermuteResult", "sequences structures mfe hd reward rounds folds")


def permute_batch(seqs, targets, config=None, seed=0, reward_exp=9.0):
    """evaluate(permute=True) over a batch: fold every sequence; those with 0 < HD <= permutation_threshold are
    improved by the local search.  -> PermuteResult(best sequences, their structures, mfe float32[B], hd int32[B],
    reward float64[B], rounds used int32[B], total number of folds)."""
    cfg = dict(DEFAULTS)
    cfg.update({k: v for k, v in (config or {}).items() if k in DEFAULTS})
    buf, off = pack(seqs)
    tbuf, toff = pack(targets)
    if not np.array_equal(off, toff):
        raise ValueError("every target must have the length of its sequence")
    B = len(off) - 1
    db = np.zeros_like(buf)
    mfe = np.zeros(B, dtype=np.int32)
    hd = np.zeros(B, dtype=np.int32)
    reward = np.zeros(B, dtype=np.float64)
    rounds = np.zeros(B, dtype=np.int32)
    folds = ctypes.c_int32(0)
    L = _lib.lib()
    _lib.check(L.rnafold_permute_batch(
        buf.ctypes.data, off.ctypes.data, B, tbuf.ctypes.data, int(cfg["permutation_budget"]),
        int(cfg["permutation_radius"]), int(cfg["permutation_threshold"]), float(cfg["mutation_probability"]),
        int(bool(cfg["allow_gu_permutations"])), int(seed) & _M, float(reward_exp), db.ctypes.data, mfe.ctypes.data,
        hd.ctypes.data, reward.ctypes.data, rounds.ctypes.data, ctypes.byref(folds)))
    return PermuteResult(unpack(buf, off), unpack(db, off), mfe_float(mfe), hd, reward, rounds, folds.value)
