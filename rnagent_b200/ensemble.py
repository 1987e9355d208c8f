"""Ensemble statistics of the reference's `Solution.compute_statistics` (rlif/rna/solution.py:221-238) for batches,
on the GPU (rnafold_ensemble_batch -> csrc/ensemble.cu): partition function, base-pair probabilities and what the
reference derives from them.

    reference attribute (solution.py)                      here
    partition_prob, partition_fn = fc.pf()                 propensity[k], ensemble_energy[k]
    positional_entropy = fc.positional_entropy()           positional_entropy[k]
    centroid_structure, centroid_dist = fc.centroid()      centroid_structure[k], centroid_dist[k]
    probability = fc.pr_structure(fold)                    probability[k]
    ensemble_diversity = fc.mean_bp_distance()             ensemble_diversity[k]
    ensemble_defect = fc.ensemble_defect(fold)             ensemble_defect[k]   (the `ed` column of rlif/logs/*)
    md = RNA.dist_mountain(target.seq, fold)               md[k]                (the `md` column)
    centroid_en = fc.eval_structure(centroid)              centroid_en[k]       (kcal/mol)
    MEA_structure, MEA = fc.MEA()                          MEA_structure[k], MEA[k]   (gamma = 1; parity unpinned: no logged column)
    MEA_en = fc.eval_structure(MEA_structure)              MEA_en[k]

`eval_structure_batch` is fc.eval_structure for any batch of (sequence, structure) pairs (rnafold_eval_structure_batch).
"""
import numpy as np

from . import _lib
from .vienna import pack, unpack

STATS = ("ensemble_energy", "ensemble_defect", "probability", "ensemble_diversity", "centroid_dist", "md", "MEA", "centroid_en", "MEA_en")


def propensity_string(down, up):
    """ViennaRNA's pair-propensity string (vrna_db_from_probs / bppm_symbol) from P(paired downstream) and P(paired upstream)."""
    out = []
    for p1, p2 in zip(down, up):
        p0 = 1.0 - p1 - p2
        if p0 > 0.667:
            c = '.'
        elif p1 > 0.667:
            c = '('
        elif p2 > 0.667:
            c = ')'
        elif (p1 + p2) > p0:
            c = '{' if p1 / (p1 + p2) > 0.667 else ('}' if p2 / (p1 + p2) > 0.667 else '|')
        elif p0 > (p1 + p2):
            c = ','
        else:
            c = ':'
        out.append(c)
    return ''.join(out)


class EnsembleResult(object):
    def __init__(self, seqs, structures, mfe, stats, centroids, pos, off, bpp=None, mea=None):
        self.sequences, self.structures, self.mfe = seqs, structures, mfe
        for k, name in enumerate(STATS):
            setattr(self, name, stats[:, k].copy())
        self.centroid_structure = centroids
        self.MEA_structure = mea
        self.positional_entropy = [pos[off[k]:off[k + 1], 2].copy() for k in range(len(seqs))]
        self.propensity = [propensity_string(pos[off[k]:off[k + 1], 0], pos[off[k]:off[k + 1], 1]) for k in range(len(seqs))]
        self.bpp = bpp

    def __len__(self):
        return len(self.sequences)


def ensemble_batch(seqs, targets=None, structures=None, mfe_dcal=None, bpp=False, gamma=1.0):
    """Statistics of every sequence's ensemble.  `structures` / `mfe_dcal` default to the sequences' own folds (fc.mfe());
    `targets` (dot-bracket strings) are only needed for the mountain distance; `gamma` is fc.MEA()'s pair weight."""
    L = _lib.lib()
    seqs = list(seqs)
    B = len(seqs)
    buf, off = pack(seqs)
    if structures is None or mfe_dcal is None:
        db = np.zeros_like(buf)
        mfe = np.zeros(B, dtype=np.int32)
        _lib.fold_hamming_raw(buf, off, B, db=db, mfe=mfe)
        structures = unpack(db, off)
    else:
        structures = list(structures)
        if [len(s) for s in structures] != [len(s) for s in seqs]:
            raise ValueError("structures and sequences differ in length")
        db = np.frombuffer(''.join(structures).encode(), dtype=np.uint8).copy()
        mfe = np.ascontiguousarray(mfe_dcal, dtype=np.int32)
    tgt = None
    if targets is not None:
        targets = list(targets)
        if [len(t) for t in targets] != [len(s) for s in seqs]:
            raise ValueError("targets and sequences differ in length")
        tgt = np.frombuffer(''.join(targets).encode(), dtype=np.uint8).copy()
    stats = np.zeros((B, len(STATS)), dtype=np.float64)
    cen = np.zeros_like(buf)
    mea = np.zeros_like(buf)
    pos = np.zeros((len(buf), 3), dtype=np.float64)
    mat = None
    if bpp:
        if B != 1:
            raise ValueError("bpp=True needs a single sequence")
        mat = np.zeros((len(seqs[0]), len(seqs[0])), dtype=np.float64)
    _lib.check(L.rnafold_ensemble_batch(_lib._ptr(buf), _lib._ptr(off), B, _lib._ptr(db), _lib._ptr(mfe), _lib._ptr(tgt),
                                        _lib._ptr(stats), _lib._ptr(cen), _lib._ptr(pos), _lib._ptr(mat),
                                        float(gamma), _lib._ptr(mea)))
    if targets is None:
        stats[:, 5] = np.nan
    return EnsembleResult(seqs, structures, mfe, stats, unpack(cen, off), pos, off, mat, unpack(mea, off))


def eval_structure_batch(seqs, structures, strict=True):
    """fc.eval_structure(structure) for every (sequence, structure) pair -> int32 energies in dcal/mol (x / 100 = kcal/mol)
    on the GPU.  strict: raise ValueError on unbalanced brackets, bracket pairs that cannot pair or hairpins under 3 bases;
    otherwise return (energies, codes)."""
    L = _lib.lib()
    seqs, structures = list(seqs), list(structures)
    if [len(s) for s in structures] != [len(s) for s in seqs]:
        raise ValueError("structures and sequences differ in length")
    B = len(seqs)
    buf, off = pack(seqs)
    db = np.frombuffer(''.join(structures).encode(), dtype=np.uint8).copy()
    e = np.zeros(B, dtype=np.int32)
    codes = np.zeros(B, dtype=np.int32)
    _lib.check(L.rnafold_eval_structure_batch(_lib._ptr(buf), _lib._ptr(off), B, _lib._ptr(db), _lib._ptr(e), _lib._ptr(codes)))
    if strict:
        bad = np.flatnonzero(codes)
        if len(bad):
            raise ValueError("eval_structure: structure %d is invalid (code %d)" % (int(bad[0]), int(codes[bad[0]])))
        return e
    return e, codes
