"""Pins the CPU oracle against every known answer the reference ships for RNA.fold.

The reference has no assertions of its own; its committed logs
rlif/logs/*/solutions*.csv hold (sequence, HD-to-target, MFE) rows written by
rlif/utils/comparison.py:225-233 from `RNA.fold`.  tools/make_golden.py extracted them.
"""
import collections

import numpy as np
import pytest

import oracle


def _fe_matches(e_dcal, fe_text):
    # logged with '{:.3f}' then possibly shortened by a spreadsheet ("-22" for "-22.000")
    return abs(round(oracle.mfe_float(e_dcal), 3) - float(fe_text)) < 5e-4


def test_worked_examples(oracle_params):
    # hand-computed from the .par text, valid under both Turner sets
    for name in ("turner2004", "turner1999"):
        P = oracle_params(name)
        assert P.fold_dcal("GCAAAAAAAAAUUUGC") == ("(((((......)))))", -220)
        assert P.fold_dcal("GCAGAAAAAAAUCUGC") == ("(((((......)))))", -490)
    assert oracle_params("turner2004").fold("GCAAAAAAAAAUUUGC")[1] == float(np.float32(-2.2))


def test_every_logged_energy(golden_rows, golden_folds):
    bad = [(r["file"], r["line"]) for r, (_, e) in zip(golden_rows, golden_folds) if not _fe_matches(e, r["fe"])]
    assert len(golden_rows) == 1726
    assert not bad, bad[:10]


def test_logged_structures_and_hd(golden_rows, golden_folds):
    """HD of our fold to the target equals the logged HD; for HD=0 rows the whole structure is pinned
    (this is what pins the backtrack's tie-breaking order)."""
    stats = collections.Counter()
    bad = []
    for r, (s, _) in zip(golden_rows, golden_folds):
        hd = sum(a != b for a, b in zip(s, r["target"]))
        ok = hd == r["hd"]
        stats[(r["params"], ok)] += 1
        if not ok:
            bad.append(r)
    # every Turner 1999 row (700, all sources, lengths 34..451) matches
    assert stats[("turner1999", False)] == 0 and stats[("turner1999", True)] == 700
    # Turner 2004: the only misses are 29 rows from sources whose HD is not reproducible under any
    # parameter set (see conftest.HD_UNPINNED_SOURCES); every RNAinverse/MODENA/NUPACK/antaRNA/rnaMCTS
    # row matches.
    assert all(r["source"] in ("rlfold", "mrlfold", "LEARNA") for r in bad), [r["source"] for r in bad]
    assert len(bad) == 29
    full = sum(1 for r in golden_rows if r["hd"] == 0 and r not in bad)
    assert full >= 1090


def test_eval_of_own_structure_equals_mfe(golden_rows, golden_folds, oracle_params):
    for r, (s, e) in zip(golden_rows, golden_folds):
        P = oracle_params(r["params"])
        assert P.eval_structure(r["seq"], s) == e


def test_bruteforce_small(oracle_params):
    rng = np.random.default_rng(7)
    for name in ("turner2004", "turner1999"):
        P = oracle_params(name)
        for _ in range(60):
            n = int(rng.integers(5, 15))
            seq = "".join(rng.choice(list("ACGU"), n))
            s, e = P.fold_dcal(seq)
            assert P.bruteforce_mfe(seq) == e, seq
            assert P.eval_structure(seq, s) == e


def test_edge_cases(oracle_params):
    P = oracle_params("turner2004")
    assert P.fold_dcal("") == ("", 0)
    assert P.fold_dcal("ACGU") == ("....", 0)
    assert P.fold_dcal("NNNNNNNNNNNN") == ("." * 12, 0)
    s, e = P.fold_dcal("GGGGAAAACCCC")
    assert s == "((((....))))" and e < 0
    assert P.fold_dcal("ggggaaaacccc") == (s, e)       # case-insensitive encoding
    assert P.fold_dcal("GGGGAAAACCCC".replace("U", "T"))[1] == e


def test_batch_matches_single(oracle_params):
    P = oracle_params("turner2004")
    rng = np.random.default_rng(3)
    seqs = ["".join(rng.choice(list("ACGU"), int(n))) for n in rng.integers(0, 90, 40)]
    buf = np.frombuffer("".join(seqs).encode(), dtype=np.uint8)
    off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int32)
    db, mfe, hd = P.fold_batch(buf, off, targets_u8=np.full_like(buf, ord(".")), nthreads=2)
    for k, s in enumerate(seqs):
        st, e = P.fold_dcal(s)
        assert db[off[k]:off[k + 1]].tobytes().decode() == st and mfe[k] == e
        assert hd[k] == len(s) - st.count(".")


def test_hd_unpinned_rows_fixture(golden_rows, golden_folds, oracle_params):
    """tests/golden/hd_unpinned_rows.json (tools/make_hd_unpinned.py): the 29 rows whose logged HD the oracle does not
    reproduce, with the outcome under every parameter set, kept for a future live RNA.fold to settle.  What the fixture
    shows today: every one comes from an in-process source (rlfold / mrlfold / LEARNA); the logged ENERGY always matches
    the log file's parameter set; in 22 rows the target structure is not even co-optimal under that set, i.e. the logged
    (HD, energy) pair cannot stem from one fold -- so these rows cannot pin a tie-break either way."""
    import json
    import os
    from conftest import GOLDEN
    rows = json.load(open(os.path.join(GOLDEN, "hd_unpinned_rows.json")))
    assert len(rows) == 29 and {r["source"] for r in rows} <= {"rlfold", "mrlfold", "LEARNA"}
    keys = {(r["file"], r["line"]) for r in rows}
    miss = {(r["file"], r["line"]) for r, (s, _) in zip(golden_rows, golden_folds)
            if sum(a != b for a, b in zip(s, r["target"])) != r["hd"]}
    assert keys == miss
    for r in rows:
        own = r["per_set"][r["params"]]
        assert own["fe_matches"] and own["hd"] != r["hd"]
        for name, v in r["per_set"].items():
            s, e = oracle_params(name).fold_dcal(r["seq"])
            assert (s, e) == (v["structure"], v["dcal"])
        assert not any(v["hd"] == r["hd"] and v["fe_matches"] for v in r["per_set"].values())
    assert sum(not r["per_set"][r["params"]]["target_is_cooptimal"] for r in rows) == 22


def test_logged_ensemble_defect_and_mountain_distance(golden_rows, golden_folds, oracle_params):
    """Row N3 (Solution.compute_statistics, rlif/rna/solution.py:221-238): the `ed` and `md` columns of the reference's
    logs (rlif/utils/comparison.py:225-233, '{:.3f}') against the partition-function oracle.  Every row whose HD reproduces
    (1,697 of 1,726; the others carry an HD from another evaluation, see test_hd_unpinned_rows_fixture) matches to 5e-4
    on both columns, 12..451 nt, both parameter sets."""
    import oracle
    checked = 0
    for name in ("turner2004", "turner1999"):
        P = oracle_params(name)
        idx = [k for k, r in enumerate(golden_rows) if r["params"] == name
               and sum(a != b for a, b in zip(golden_folds[k][0], r["target"])) == r["hd"]]
        seqs = [golden_rows[k]["seq"] for k in idx]
        buf = np.frombuffer("".join(seqs).encode(), dtype=np.uint8)
        dbs = np.frombuffer("".join(golden_folds[k][0] for k in idx).encode(), dtype=np.uint8)
        off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int32)
        mfe = np.array([golden_folds[k][1] for k in idx], dtype=np.int32)
        stats = P.pf_batch(buf, off, dbs, mfe, smooth=True, nthreads=0)
        for n_, k in enumerate(idx):
            r = golden_rows[k]
            assert abs(stats[n_][1] - float(r["ed"])) <= 0.00051, (r["file"], r["line"], r["ed"], stats[n_][1])
            md = oracle.dist_mountain(r["target"], golden_folds[k][0])
            assert abs(md - float(r["md"])) <= 0.00051, (r["file"], r["line"], r["md"], md)
            assert stats[n_][0] <= mfe[n_] / 100.0 + 1e-9 and 0 < stats[n_][2] <= 1.0 + 1e-9     # F <= MFE, 0 < p(mfe) <= 1
            checked += 1
    assert checked == 1697


def test_partition_function_against_enumeration(oracle_params):
    """Z and every base-pair probability against an explicit sum over all structures (n <= 15), on the integer energies of
    the independent loop evaluator (smooth=False: the same energy model as the MFE recursions)."""
    import math
    P = oracle_params("turner2004")
    kT = (37 + 273.15) * 1.98717
    pairs = {"AU", "UA", "CG", "GC", "GU", "UG"}

    def structures(n):
        out, cur, stack = [], [], []

        def rec(i):
            if i == n:
                if not stack:
                    out.append("".join(cur))
                return
            cur.append("."); rec(i + 1); cur.pop()
            cur.append("("); stack.append(i); rec(i + 1); stack.pop(); cur.pop()
            if stack and i - stack[-1] > 3:
                o = stack.pop(); cur.append(")"); rec(i + 1); cur.pop(); stack.append(o)
        rec(0)
        return out
    rng = np.random.default_rng(11)
    for trial in range(8):
        n = int(rng.integers(10, 16))
        seq = "".join(rng.choice(list("ACGU" if trial % 2 else "GCGCAU"), n))
        Z, pij = 0.0, np.zeros((n, n))
        for st in structures(n):
            stack, pl, ok = [], [], True
            for k, ch in enumerate(st):
                if ch == "(":
                    stack.append(k)
                elif ch == ")":
                    o = stack.pop()
                    ok = ok and (seq[o] + seq[k]) in pairs
                    pl.append((o, k))
            if not ok:
                continue
            w = math.exp(-P.eval_structure(seq, st) * 10 / kT)
            Z += w
            for o, k in pl:
                pij[o, k] += w
        s, e = P.fold_dcal(seq)
        r = P.pf(seq, s, e, smooth=False, bpp=True)
        assert abs(r["ensemble_energy"] + kT * math.log(Z) / 1000) < 1e-9
        assert np.abs(r["bpp"] - pij / Z).max() < 1e-12
        assert abs(r["probability"] - math.exp(-e * 10 / kT) / Z) < 1e-12


def test_mea_against_its_definition(oracle_params):
    """fc.MEA() (rlif/rna/solution.py:233) has no logged column to pin it: the restatement is checked against the definition --
    the structure maximising sum over pairs 2 gamma p(i,j) + sum over unpaired pu(i) -- by enumeration, and the returned
    structure must reach the returned value; the evaluator takes loops above MAXLOOP (fc.eval_structure extrapolates)."""
    import math
    P = oracle_params("turner2004")

    def structures(n):
        out, cur, stack = [], [], []

        def rec(i):
            if i == n:
                if not stack:
                    out.append("".join(cur))
                return
            cur.append("."); rec(i + 1); cur.pop()
            cur.append("("); stack.append(i); rec(i + 1); stack.pop(); cur.pop()
            if stack and i - stack[-1] > 3:
                o = stack.pop(); cur.append(")"); rec(i + 1); cur.pop(); stack.append(o)
        rec(0)
        return out

    def accuracy(st, bpp, pu, gamma):
        stack, acc = [], 0.0
        for k, ch in enumerate(st):
            if ch == "(":
                stack.append(k)
            elif ch == ")":
                acc += 2 * gamma * bpp[stack.pop(), k]
            else:
                acc += pu[k]
        return acc
    rng = np.random.default_rng(5)
    for trial in range(10):
        n = int(rng.integers(9, 15))
        seq = "".join(rng.choice(list("GCGCAU" if trial % 2 else "ACGU"), n))
        s, e = P.fold_dcal(seq)
        for gamma in (1.0, 3.0):
            if trial < 5:
                bpp = P.pf(seq, s, e, bpp=True)["bpp"]
            else:   # any consistent matrix will do: random probabilities with row sums below 1
                bpp = np.triu(rng.random((n, n)) ** 3, 4) / n * 2.5
            cut = 1e-4 / (1 + gamma)
            kept = np.where(bpp >= cut, bpp, 0.0)
            pu = 1.0 - kept.sum(0) - kept.sum(1)
            best = max(accuracy(st, kept, pu, gamma) for st in structures(n))
            st, val = oracle.mea(bpp, gamma)
            assert abs(val - best) < 1e-12
            assert abs(accuracy(st, kept, pu, gamma) - val) < 1e-12
    # the evaluator beyond MAXLOOP: one 40-nt bulge costs bulge[30] + (int)(lxc log(40/30)) + the two stacks' helices
    inner = "GGGGAAAACCCC"
    seq = "GGGG" + "A" * 40 + inner + "CCCC"
    st = "((((" + "." * 40 + "((((....))))" + "))))"
    e_long = P.eval_structure(seq, st)
    seq30 = "GGGG" + "A" * 30 + inner + "CCCC"
    st30 = "((((" + "." * 30 + "((((....))))" + "))))"
    assert e_long - P.eval_structure(seq30, st30) == int(107.856 * math.log(40 / 30.))
    full = P.statistics("UGUUGAAGGCCUAACA", *P.fold_dcal("UGUUGAAGGCCUAACA"))
    assert full["MEA_en"] == P.eval_structure("UGUUGAAGGCCUAACA", full["MEA_structure"]) / 100.0 and full["MEA"] > 0
