"""Drop-in for the reference's fold wrapper and its third-party engine.

Mirrors, name for name:
  * `RNA.fold(seq) -> (structure, mfe)`           bound at rlif/settings.py:63-64 as ConfigManager.fold_fn
  * `RNA.read_parameter_file(path)`               rlif/rna/vienna.py:18
  * `RNA.cvar.<temperature|dangles|noGU|no_closingGU|uniq_ML>`   rlif/rna/vienna.py:21-26
  * `set_vienna_params(n)`, `config_vienna(**kw)`, `fold(sequence, worker)`   rlif/rna/vienna.py:7-47
plus the batched entry points the vectorised rollout uses (`fold_batch`, `fold_hamming_batch`).
Everything executes on the B200 through librnafold_b200.so; nothing here folds on the CPU.
"""
import os

import numpy as np

from . import _lib

PARAMETERS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
# same numbering as rlif/rna/vienna.py:11-15
PARAM_FILES = {1: "turner2004_37.par", 2: "turner1999_37.par", 3: "andronescu2007_37.par", 4: "langdon2018_37.par"}


def mfe_float(dcal):
    """ViennaRNA returns `(float)(e / 100.)`; Python sees that float32 widened to a double."""
    return np.asarray(dcal, dtype=np.float64).__truediv__(100.0).astype(np.float32)


def read_parameter_file(path):
    """RNA.read_parameter_file: process-global, takes effect on the next fold."""
    _lib.check(_lib.lib().rnafold_read_parameter_file(os.fsencode(path)))


def set_vienna_params(params):
    """rlif/rna/vienna.py:7-19 (with its NameError fixed): 1..4 -> parameter file; returns the path."""
    path = os.path.join(PARAMETERS, PARAM_FILES[params])
    read_parameter_file(path)
    return path


class _CVar:
    """RNA.cvar look-alike.  Only the reference's default model (rlif/settings.py:35-39) exists on the
    device; assigning anything else raises NotImplementedError instead of folding with a wrong model."""
    _names = ("temperature", "dangles", "noGU", "no_closingGU", "uniq_ML")

    def __setattr__(self, name, value):
        if name not in self._names:
            raise AttributeError(f"RNA.cvar has no model detail {name!r} on this engine")
        _lib.check(_lib.lib().rnafold_set_md(name.encode(), float(value)))

    def __getattr__(self, name):
        if name not in self._names:
            raise AttributeError(name)
        import ctypes
        v = ctypes.c_double(0)
        _lib.check(_lib.lib().rnafold_get_md(name.encode(), ctypes.byref(v)))
        return int(v.value) if name != "temperature" else v.value


cvar = _CVar()


def config_vienna(**kwargs):
    """rlif/rna/vienna.py:21-26."""
    for arg, val in kwargs.items():
        setattr(cvar, arg, val)


def pack(seqs):
    """list[str|bytes] -> (uint8 concatenation, int32 CSR offsets)."""
    bs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
    off = np.zeros(len(bs) + 1, dtype=np.int32)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs], dtype=np.int64)
    buf = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, dtype=np.uint8)
    return buf, off


def unpack(buf, off):
    raw = buf.tobytes().decode("ascii")
    return [raw[off[k]:off[k + 1]] for k in range(len(off) - 1)]


def fold_batch(seqs):
    """Fold a list of sequences: -> (list[str] structures, float32[B] mfe kcal/mol)."""
    buf, off = pack(seqs)
    B = len(off) - 1
    db = np.zeros_like(buf)
    mfe = np.zeros(B, dtype=np.int32)
    _lib.fold_hamming_raw(buf, off, B, db=db, mfe=mfe)
    return unpack(db, off), mfe_float(mfe)


def fold_hamming_batch(seqs, targets, reward_exp=9.0):
    """Fold + Hamming distance to the targets + reward (rlif/rna/solution.py:134-148 over a batch).

    -> (structures, mfe float32[B], hd int32[B], mismatch_indices list[int64 array], reward float64[B])
    """
    buf, off = pack(seqs)
    tbuf, toff = pack(targets)
    if not np.array_equal(off, toff):
        raise ValueError("every target must have the length of its sequence")
    B = len(off) - 1
    db = np.zeros_like(buf)
    mfe = np.zeros(B, dtype=np.int32)
    hd = np.zeros(B, dtype=np.int32)
    reward = np.zeros(B, dtype=np.float64)
    mism = np.zeros_like(buf)
    _lib.fold_hamming_raw(buf, off, B, target=tbuf, db=db, mfe=mfe, hd=hd, reward=reward, mismatch=mism,
                          reward_exp=reward_exp)
    idx = [np.nonzero(mism[off[k]:off[k + 1]])[0].astype(np.int64) for k in range(B)]
    return unpack(db, off), mfe_float(mfe), hd, idx, reward


def fold(sequence, worker=None):
    """`fold_fn(string) -> (structure, mfe)`; `worker` is the unused second argument of the reference's
    Windows variant (rlif/rna/vienna.py:28)."""
    structs, mfe = fold_batch([sequence])
    return structs[0], float(mfe[0])
