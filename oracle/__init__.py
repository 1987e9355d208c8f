"""ctypes binding of the CPU oracle (oracle/rnafold_oracle.c).  TEST INFRASTRUCTURE ONLY.

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package (rnagent_b200) never imports this module.

Restates RNA.fold of ViennaRNA 2.4.14 as bound by the reference at rlif/settings.py:63-64;
parity pinned by tests/golden/vienna_logged_rows.jsonl (rows of rlif/logs/*/solutions*.csv).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "librnafold_oracle.so")
_lib = None


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("rnafold_oracle.c", "rnafold_pf.c", "rnafold_oracle.h", "Makefile")]
    if (not force and os.path.exists(_SO)
            and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in src)):
        return _SO
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.orc_params_new.restype = ctypes.c_void_p
        L.orc_params_free.argtypes = [ctypes.c_void_p]
        L.orc_params_load.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        L.orc_params_export.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.orc_fold.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p,
                               ctypes.POINTER(ctypes.c_int)]
        L.orc_last_fold_candidates.restype = ctypes.c_longlong
        L.orc_eval_structure.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p,
                                         ctypes.POINTER(ctypes.c_int)]
        L.orc_bruteforce_mfe.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int,
                                         ctypes.POINTER(ctypes.c_int)]
        L.orc_fold_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                     ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_int]
        L.orc_pf.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int, ctypes.c_int,
                             ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p]
        L.orc_dist_mountain.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_int]
        L.orc_dist_mountain.restype = ctypes.c_double
        L.orc_mea.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_char_p, ctypes.c_void_p]
        L.orc_pf_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                   ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
        _lib = L
    return _lib


def mfe_float(dcal):
    """ViennaRNA returns (float)(e/100.): float32 of the double quotient, widened by Python."""
    return float(np.float32(np.float64(dcal) / 100.0))


class Params:
    """One loaded parameter set (RNA.read_parameter_file equivalent for the oracle)."""

    def __init__(self, path):
        self._L = lib()
        self._p = self._L.orc_params_new()
        rc = self._L.orc_params_load(self._p, os.fsencode(path))
        if rc != 0:
            raise RuntimeError(f"oracle: cannot load parameter file {path!r} (rc={rc})")
        self.path = path

    def __del__(self):
        try:
            self._L.orc_params_free(self._p)
        except Exception:
            pass

    def export(self):
        n = self._L.orc_params_export_size()
        out = np.zeros(n, dtype=np.int32)
        k = self._L.orc_params_export(self._p, out.ctypes.data)
        assert k == n, (k, n)
        return out

    def fold_dcal(self, seq):
        b = seq.encode()
        db = ctypes.create_string_buffer(len(b) + 1)
        e = ctypes.c_int(0)
        rc = self._L.orc_fold(self._p, b, len(b), db, ctypes.byref(e))
        if rc != 0:
            raise RuntimeError(f"oracle: backtrack failed rc={rc} for {seq}")
        return db.value.decode(), e.value

    def fold(self, seq):
        s, e = self.fold_dcal(seq)
        return s, mfe_float(e)

    def last_candidates(self):
        return int(self._L.orc_last_fold_candidates())

    PF_STATS = ("ensemble_energy", "ensemble_defect", "probability", "ensemble_diversity", "centroid_dist")

    def pf(self, seq, structure, e_dcal, smooth=True, bpp=False):
        """Ensemble statistics of Solution.compute_statistics (rlif/rna/solution.py:221-238) for `seq`, with `structure`
        (its MFE fold) and that structure's energy in dcal/mol -> dict (+ 'centroid_structure', + 'bpp' n x n if asked)."""
        n = len(seq)
        stats = np.zeros(5, dtype=np.float64)
        cen = ctypes.create_string_buffer(n + 1)
        mat = np.zeros((n, n), dtype=np.float64) if bpp else None
        rc = self._L.orc_pf(self._p, seq.encode(), n, structure.encode(), int(e_dcal), 1 if smooth else 0, stats.ctypes.data,
                            cen, mat.ctypes.data if bpp else None)
        if rc:
            raise RuntimeError("orc_pf failed")
        out = dict(zip(self.PF_STATS, (float(v) for v in stats)))
        out["centroid_structure"] = cen.value.decode()
        if bpp:
            out["bpp"] = mat
        return out

    def statistics(self, seq, structure, e_dcal, gamma=1.0):
        """Everything Solution.compute_statistics sets (rlif/rna/solution.py:221-238): pf() plus the centroid's energy, the
        MEA structure (fc.MEA(), gamma = 1), its expected accuracy and its energy (kcal/mol as fc.eval_structure returns)."""
        out = self.pf(seq, structure, e_dcal, bpp=True)
        s, v = mea(out["bpp"], gamma)
        out["MEA_structure"], out["MEA"] = s, v
        out["centroid_en"] = self.eval_structure(seq, out["centroid_structure"]) / 100.0
        out["MEA_en"] = self.eval_structure(seq, s) / 100.0
        return out

    def pf_batch(self, seqs_u8, off, structs_u8, mfe_dcal, smooth=True, nthreads=0):
        B = len(off) - 1
        out = np.zeros((B, 5), dtype=np.float64)
        seqs_u8, structs_u8 = np.ascontiguousarray(seqs_u8), np.ascontiguousarray(structs_u8)
        off = np.ascontiguousarray(off, dtype=np.int32)
        e = np.ascontiguousarray(mfe_dcal, dtype=np.int32)
        rc = self._L.orc_pf_batch(self._p, seqs_u8.ctypes.data, off.ctypes.data, B, structs_u8.ctypes.data, e.ctypes.data,
                                  1 if smooth else 0, out.ctypes.data, int(nthreads))
        if rc:
            raise RuntimeError("orc_pf_batch failed")
        return out

    def eval_structure(self, seq, db):
        e = ctypes.c_int(0)
        rc = self._L.orc_eval_structure(self._p, seq.encode(), len(seq), db.encode(), ctypes.byref(e))
        if rc != 0:
            raise ValueError(f"oracle: invalid structure rc={rc}")
        return e.value

    def bruteforce_mfe(self, seq):
        e = ctypes.c_int(0)
        rc = self._L.orc_bruteforce_mfe(self._p, seq.encode(), len(seq), ctypes.byref(e))
        if rc != 0:
            raise ValueError("oracle: bruteforce needs n <= 20")
        return e.value

    def fold_batch(self, seqs_u8, off, targets_u8=None, nthreads=0):
        """seqs_u8: uint8 ASCII concatenation, off: int32[B+1].  Returns (db_u8, mfe_dcal, hd)."""
        seqs_u8 = np.ascontiguousarray(seqs_u8, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int32)
        B = len(off) - 1
        db = np.zeros_like(seqs_u8)
        mfe = np.zeros(B, dtype=np.int32)
        hd = np.zeros(B, dtype=np.int32)
        tp = None
        if targets_u8 is not None:
            targets_u8 = np.ascontiguousarray(targets_u8, dtype=np.uint8)
            tp = targets_u8.ctypes.data
        rc = self._L.orc_fold_batch(self._p, seqs_u8.ctypes.data, off.ctypes.data, B, tp,
                                    db.ctypes.data, mfe.ctypes.data, hd.ctypes.data, int(nthreads))
        if rc != 0:
            raise RuntimeError(f"oracle: batch fold failed rc={rc}")
        return db, mfe, hd


def hamming_distance(a, b):
    """rlif/rna/utils.py:67-75 restated: zip-truncating; returns (hd, mismatch indices)."""
    m = [x != y for x, y in zip(a, b)]
    return sum(m), np.where(m)[0]


def mea(bpp, gamma=1.0):
    """fc.MEA(gamma) from a base-pair probability matrix (n x n, upper triangle) -> (structure, expected accuracy)."""
    mat = np.ascontiguousarray(bpp, dtype=np.float64)
    n = mat.shape[0]
    buf = ctypes.create_string_buffer(n + 1)
    val = np.zeros(1, dtype=np.float64)
    rc = lib().orc_mea(mat.ctypes.data, n, float(gamma), buf, val.ctypes.data)
    if rc:
        raise RuntimeError("orc_mea: backtracking failed")
    return buf.value.decode(), float(val[0])


def dist_mountain(a, b, weighted=True):
    """RNA.dist_mountain(a, b) as used by Solution.compute_statistics (rlif/rna/solution.py:238)."""
    if len(a) != len(b):
        raise ValueError("structures of unequal length")
    return float(lib().orc_dist_mountain(a.encode(), b.encode(), len(a), 1 if weighted else 0))
