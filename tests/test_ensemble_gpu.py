"""GPU: ensemble statistics (Solution.compute_statistics, rlif/rna/solution.py:221-238) through rnafold_ensemble_batch against
the reference's logged `ed` / `md` columns and against the CPU oracle (oracle/rnafold_pf.c).  Floating point: fp64 on both
sides, different summation orders -> relative tolerance 1e-9 against the oracle, 5e-4 (the logs' three decimals) against the logs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-9


def test_logged_ed_and_md_columns(golden_rows):
    """All logged rows whose HD reproduces (1,697): ensemble defect and mountain distance to 5e-4, 12..451 nt, both sets."""
    import rnagent_b200 as rb
    checked = 0
    for setno, name in ((1, "turner2004"), (2, "turner1999")):
        rb.set_vienna_params(setno)
        rows = [r for r in golden_rows if r["params"] == name]
        res = rb.ensemble_batch([r["seq"] for r in rows], targets=[r["target"] for r in rows])
        for k, r in enumerate(rows):
            if rb.hamming_distance(r["target"], res.structures[k])[0] != r["hd"]:
                continue
            assert abs(res.ensemble_defect[k] - float(r["ed"])) <= 0.00051, (r["file"], r["line"], r["ed"], res.ensemble_defect[k])
            assert abs(res.md[k] - float(r["md"])) <= 0.00051, (r["file"], r["line"], r["md"], res.md[k])
            checked += 1
    rb.set_vienna_params(1)
    assert checked == 1697


@pytest.mark.parametrize("setno,name", [(1, "turner2004"), (2, "turner1999"), (3, "andronescu2007")])
def test_against_the_oracle(oracle_params, setno, name):
    import oracle
    import rnagent_b200 as rb
    rb.set_vienna_params(setno)
    P = oracle_params(name)
    rng = np.random.default_rng(40 + setno)
    lens = [0, 1, 4, 5, 8, 12] + [int(v) for v in rng.integers(13, 140, 90)] + [200, 257, 330]
    seqs = ["".join(rng.choice(list("ACGU"), n)) for n in lens]
    seqs[7] = "GGGGAAAACCCC".lower().replace("u", "t")            # case / T handling as the fold
    seqs[8] = "GCGCNNNNGCGC"                                       # unknown letters never pair
    from rnagent_b200 import synth
    targets = [synth.random_structure(len(s), rng).tobytes().decode() if len(s) else "" for s in seqs]
    res = rb.ensemble_batch(seqs, targets=targets)
    for k, s in enumerate(seqs):
        st, e = P.fold_dcal(s)
        assert (res.structures[k], int(res.mfe[k])) == (st, e)
        if not s:
            continue
        want = P.pf(s, st, e, smooth=True, bpp=True)
        for key in ("ensemble_energy", "ensemble_defect", "probability", "ensemble_diversity", "centroid_dist"):
            got = getattr(res, key)[k]
            assert abs(got - want[key]) <= RTOL * max(1.0, abs(want[key])), (k, len(s), key, got, want[key])
        assert res.centroid_structure[k] == want["centroid_structure"]
        # fc.MEA() and the two fc.eval_structure calls (solution.py:232-234): value to 1e-9, structures and energies exact
        mea_s, mea_v = oracle.mea(want["bpp"], 1.0)
        assert abs(res.MEA[k] - mea_v) <= RTOL * max(1.0, mea_v), (k, len(s), res.MEA[k], mea_v)
        assert res.MEA_structure[k] == mea_s, (k, len(s))
        assert res.centroid_en[k] == P.eval_structure(s, want["centroid_structure"]) / 100.0
        assert res.MEA_en[k] == P.eval_structure(s, mea_s) / 100.0
        assert abs(res.md[k] - oracle.dist_mountain(targets[k], st)) < 1e-12
        full = want["bpp"] + want["bpp"].T
        down = np.triu(want["bpp"]).sum(axis=1)
        assert res.propensity[k] == rb.ensemble.propensity_string(down, full.sum(axis=1) - down)
        pu = 1.0 - full.sum(axis=1)
        with np.errstate(divide="ignore", invalid="ignore"):
            ent = -(np.where(full > 0, full * np.log(full), 0.0).sum(axis=1) + np.where(pu > 0, pu * np.log(pu), 0.0)) / np.log(2.0)
        assert np.abs(res.positional_entropy[k] - ent).max() < 1e-9
    wide = rb.ensemble_batch(seqs[20:40], gamma=3.0)          # a larger pair weight lists and keeps more pairs
    for k, s in enumerate(seqs[20:40]):
        st, e = P.fold_dcal(s)
        mea_s, mea_v = oracle.mea(P.pf(s, st, e, bpp=True)["bpp"], 3.0)
        assert wide.MEA_structure[k] == mea_s and abs(wide.MEA[k] - mea_v) <= RTOL * mea_v
    one = rb.ensemble_batch([seqs[40]], bpp=True)
    s40, e40 = P.fold_dcal(seqs[40])
    assert np.abs(one.bpp - P.pf(seqs[40], s40, e40, bpp=True)["bpp"]).max() < 1e-12 and np.isnan(one.md[0])
    rb.set_vienna_params(1)


def test_solution_compute_statistics_and_limits():
    import rnagent_b200 as rb
    from rnagent_b200 import _lib
    rb.set_vienna_params(1)
    cfg = dict(kernel_size=25, reward_exp=9, use_nucleotides=True, encoding_type=2)
    sol = rb.Solution(target="(((((......)))))", config=cfg, string="UGUUGAAGGCCUAACA")   # first row of rlif/logs/eterna/solutions.csv
    assert (sol.hd, sol.fe) == (0, float(np.float32(-0.8)))
    assert round(sol.ensemble_defect, 3) == 0.121 and sol.md == 0.0                          # ... and its logged ed / md
    assert sol.centroid_structure == ".((((......))))." and 0 < sol.probability < 0.5       # p(1,16) = 0.47: not in the centroid
    assert len(sol.partition_prob) == 16 and sol.partition_fn < -0.8 and len(sol.positional_entropy) == 16
    # 2 p(1,16) = 0.94 < pu(1) + pu(16): the outer pair is no MEA candidate either, the MEA structure is the centroid
    assert sol.MEA_structure == ".((((......))))." and sol.MEA_en == sol.centroid_en and 8.0 < sol.MEA < 16.0
    assert sol.centroid_en == rb.eval_structure_batch([sol.string], [sol.centroid_structure])[0] / 100.0
    with pytest.raises(ValueError):
        rb.ensemble_batch(["ACGU"], gamma=0.0)
    with pytest.raises(ValueError):
        rb.ensemble_batch(["A" * 1001])
    assert len(rb.ensemble_batch([])) == 0


def test_comparison_writer_regenerates_the_reference_log(tmp_path):
    """Row N4: `solutions.csv` / `<n>.log` of rlif/utils/comparison.py:150-255 from GPU results, against the first block of the
    reference's own rlif/logs/eterna/solutions.csv (tests/golden/eterna_solutions_head.csv).  The committed log went through a
    spreadsheet (trailing zeros and the log-name field stripped, padding commas and a per-method tally appended): the block's
    fields are compared one by one -- strings exactly, numbers as the values our '{:.3f}' text denotes."""
    import os
    import rnagent_b200 as rb
    from conftest import GOLDEN
    from rnagent_b200 import comparison
    rb.set_vienna_params(1)
    ref = [l.rstrip("\n").split(",") for l in open(os.path.join(GOLDEN, "eterna_solutions_head.csv"))]
    target, limit, length = ref[0][0], ref[0][1], ref[0][2]
    rows = [r for r in ref[2:] if r[0] and set(r[0]) <= set("ACGU")]
    assert (target, limit, length) == ("(((((......)))))", "60", "16") and len(rows) == 8 and ref[1][:8] == comparison.COLUMNS
    logged_t = {r[1]: float(r[2]) for r in rows}
    t = [0.0]

    def make(seq_out, name):
        def fn(seq, time_limit=10):
            t[0] += logged_t[name]       # the method's run time is the logged one (a fake clock): everything else is computed
            return seq_out
        return fn
    names = [r[1] for r in rows]
    out, log = tmp_path / "solutions.csv", tmp_path / "1.log"
    solved, times = comparison.mass_fold(target, [make(r[0], r[1]) for r in rows], names, time_limit=int(limit), file=str(out),
                                         log=str(log), clock=lambda: t[0])
    assert solved.tolist() == [1.0] * 8
    got = [l.rstrip("\n").split(",") for l in open(out)]
    assert got[0] == [target, limit, length, "1"] and got[1] == comparison.COLUMNS and got[-1] == [""]
    for g, r in zip(got[2:], rows):
        assert g[0] == r[0] and g[1] == r[1] and g[3] == r[3] and g[7] == r[7]                   # seq, source, hd, attempts
        assert all(len(g[c].split(".")[1]) == 3 for c in (2, 4, 5, 6))                           # '{:.3f}'
        assert [float(g[c]) for c in (2, 4, 5, 6)] == [float(r[c]) for c in (2, 4, 5, 6)], (g, r)   # t, md, fe, ed
    assert [l.split() for l in open(log)] == [[r[0], "%.3f" % float(r[2]), r[1]] for r in rows]
    assert comparison.timeout_row("NUPACK", 60.0) == "---, NUPACK, 60.000\n"


@pytest.mark.parametrize("setno,name", [(1, "turner2004"), (2, "turner1999"), (3, "andronescu2007")])
def test_eval_structure_batch(oracle_params, setno, name):
    """fc.eval_structure (solution.py:232,234) on the GPU against the oracle's independent loop evaluator: MFE folds (energy ==
    mfe), random structures threaded onto sequences that can form them (every loop kind, loops far above MAXLOOP), and the
    error codes."""
    import rnagent_b200 as rb
    from rnagent_b200 import synth
    rb.set_vienna_params(setno)
    P = oracle_params(name)
    rng = np.random.default_rng(70 + setno)
    seqs, dbs = [], []
    for n in [0, 1, 5, 9] + [int(v) for v in rng.integers(10, 400, 300)] + [1000, 2500, 4000]:
        db = synth.random_structure(n, rng).tobytes().decode() if n else ""
        s = list(rng.choice(list("ACGU"), n))
        stack = []
        for k, ch in enumerate(db):        # make every bracket pair a pairable one (GU pairs included)
            if ch == "(":
                stack.append(k)
            elif ch == ")":
                o = stack.pop()
                s[o], s[k] = [("G", "C"), ("C", "G"), ("A", "U"), ("U", "A"), ("G", "U"), ("U", "G")][int(rng.integers(0, 6))]
        seqs.append("".join(s))
        dbs.append(db)
    seqs.append("GGGG" + "A" * 40 + "GGGGAAAACCCC" + "A" * 35 + "CCCC")      # a 75-nt interior loop
    dbs.append("((((" + "." * 40 + "((((....))))" + "." * 35 + "))))")
    got = rb.eval_structure_batch(seqs, dbs)
    for k, (s, d) in enumerate(zip(seqs, dbs)):
        assert int(got[k]) == P.eval_structure(s, d), (k, len(s))
    structs, mfes = rb.fold_batch(seqs[4:200])
    e2 = rb.eval_structure_batch(seqs[4:200], list(structs))
    assert [int(v) for v in e2] == [int(round(float(m) * 100)) for m in mfes]
    e, codes = rb.eval_structure_batch(["GGGAAACCC", "GGGAAACCC", "AAAAAAAAA", "GGGAACCC", "GGGAAACCC"],
                                       ["(((...)))", "((....)))", "(((...)))", "(((..)))", "((.....)."], strict=False)
    assert list(codes) == [0, 3, 1, 1, 3] and int(e[0]) == P.eval_structure("GGGAAACCC", "(((...)))") and list(e[1:]) == [0, 0, 0, 0]
    with pytest.raises(ValueError):
        rb.eval_structure_batch(["GGGAAACCC"], ["(((...))."])
    rb.set_vienna_params(1)
