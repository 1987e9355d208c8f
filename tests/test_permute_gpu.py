"""GPU parity of the batched local search (Solution.permute, rlif/rna/solution.py:157-219) against golden vectors
recorded from the reference's own permute() code (tools/make_permute_golden.py), plus size-independent properties."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _cases():
    g = json.load(open(os.path.join(GOLDEN, "permute_cases.json")))
    return g["configs"], g["cases"]


@pytest.mark.parametrize("cname", ["training", "testing", "wide"])
def test_replays_reference_permute_bit_for_bit(cname):
    import rnagent_b200 as rb
    rb.set_vienna_params(1)
    cfgs, cases = _cases()
    cs = [c for c in cases if c["config"] == cname]
    assert [c["index"] for c in cs] == list(range(len(cs))) and len(set(c["seed"] for c in cs)) == 1
    res = rb.permute_batch([c["seq"] for c in cs], [c["target"] for c in cs], cfgs[cname], seed=cs[0]["seed"], reward_exp=9)
    checked = 0
    for k, c in enumerate(cs):
        # every case, including the four whose result depends on the order in which the reference walks its Python set
        # of candidate positions (order_safe False): the device replays CPython's set order (rnafold_permute_set_order)
        got = (res.sequences[k], int(res.hd[k]), res.structures[k], float(res.mfe[k]))
        want = (c["out_seq"], c["out_hd"], c["out_folded"], float(np.float32(c["out_fe"])))
        assert got == want, (k, c["seq"], got, want)
        assert abs(res.reward[k] - c["out_r"]) < 1e-12
        searched = 0 < c["hd0"] <= cfgs[cname]["permutation_threshold"]
        assert int(res.rounds[k]) == (c["folds"] - 2 if searched else 0)   # reference: 1 fold + rounds + 1 re-fold
        checked += 1
    assert checked == len(cs)
    assert res.folds == len(cs) + int(res.rounds.sum())


def test_properties_at_scale(oracle_params):
    """4,096 near misses of n = 100: HD never gets worse, untouched outside (0, threshold], the reported structure /
    energy / HD are those of the returned sequence (oracle re-fold), pairs of the target stay complementary."""
    import rnagent_b200 as rb
    rb.set_vienna_params(1)
    rng = np.random.default_rng(99)
    B, n = 4096, 100
    base = ["".join(rng.choice(list("ACGU"), n)) for _ in range(B)]
    targets, _ = rb.fold_batch(base)
    seqs = []
    for s in base:
        s = list(s)
        for _ in range(int(rng.integers(0, 4))):
            s[int(rng.integers(0, n))] = "ACGU"[int(rng.integers(0, 4))]
        seqs.append("".join(s))
    _, _, hd0, _, _ = rb.fold_hamming_batch(seqs, targets)
    cfg = dict(rb.permute.TESTING)
    res = rb.permute_batch(seqs, targets, cfg, seed=7)
    assert np.all(res.hd <= hd0)
    same = (hd0 == 0) | (hd0 > cfg["permutation_threshold"])
    assert all(res.sequences[k] == seqs[k] and res.rounds[k] == 0 for k in np.nonzero(same)[0])
    searched = ~same
    assert searched.sum() > 500 and (res.hd[searched] < hd0[searched]).sum() > 50
    assert np.all(res.rounds[searched] >= 1) and np.all(res.rounds <= cfg["permutation_budget"])
    assert np.all((res.rounds[searched] < cfg["permutation_budget"]) <= (res.hd[searched] == 0))  # early stop only at HD 0
    # oracle re-fold of the returned sequences
    P = oracle_params("turner2004")
    buf = np.frombuffer("".join(res.sequences).encode(), dtype=np.uint8)
    off = (np.arange(B + 1) * n).astype(np.int32)
    tb = np.frombuffer("".join(targets).encode(), dtype=np.uint8)
    db, mfe, hd = P.fold_batch(buf, off, targets_u8=tb, nthreads=0)
    raw = db.tobytes().decode()
    assert [raw[off[k]:off[k + 1]] for k in range(B)] == res.structures
    assert np.array_equal(hd, res.hd) and np.array_equal(mfe, np.round(res.mfe.astype(np.float64) * 100).astype(np.int32))
    # same seed -> same result; another seed -> another search
    again = rb.permute_batch(seqs, targets, cfg, seed=7)
    assert again.sequences == res.sequences and np.array_equal(again.rounds, res.rounds)
    other = rb.permute_batch(seqs, targets, cfg, seed=8)
    assert other.sequences != res.sequences


def test_edge_cases():
    import rnagent_b200 as rb
    rb.set_vienna_params(1)
    r = rb.permute_batch([], [], None)
    assert r.sequences == [] and r.folds == 0
    r = rb.permute_batch(["ACGU", ""], ["....", ""], None)
    assert r.sequences == ["ACGU", ""] and list(r.hd) == [0, 0] and list(r.rounds) == [0, 0]
    # budget 0: evaluation only
    r = rb.permute_batch(["GGGGAAAACCCC"], ["(((......)))"], dict(permutation_budget=0))
    assert r.sequences == ["GGGGAAAACCCC"] and r.hd[0] == 2 and r.rounds[0] == 0 and r.folds == 1
    with pytest.raises(ValueError):
        rb.permute_batch(["ACGU"], ["..."], None)
    with pytest.raises(ValueError):
        rb.permute_batch(["ACGU"], ["...."], dict(permutation_radius=0))


def test_environment_with_permute():
    """RnaDesign.permute (RLIFenv.py:13,69-73): same actions with and without the local search -- the searched
    environments end with an HD that is never worse, the others with identical solutions."""
    import rnagent_b200 as rb
    from rnagent_b200 import synth
    rb.set_vienna_params(1)
    seq, off = synth.random_sequences(16, 40, 21)
    structs, _ = rb.fold_batch([seq[off[k]:off[k + 1]].tobytes().decode() for k in range(16)])
    base = dict(kernel_size=25, reward_exp=9, encoding_type=2, use_nucleotides=True, permutation_budget=15,
                permutation_radius=5, permutation_threshold=15, allow_gu_permutations=True, mutation_probability=0.5)
    N = 256
    runs = {}
    for permute in (False, True):
        env = rb.RnaDesignBatch(dict(base, permute=permute), N, targets=structs, auto_reset=False, seed=5)
        rng = np.random.default_rng(3)
        env.reset()
        finished = np.zeros(N, dtype=bool)
        rewards = np.zeros(N)
        for _ in range(45):
            acts = rng.integers(0, 4, N).astype(np.int32)
            _, r, done, _ = env.step(acts)
            new = done & ~finished
            rewards[new] = r[new]
            finished |= done
            if finished.all():
                break
        assert finished.all()
        runs[permute] = (env.get_attr("prev_solution"), rewards.copy())
        env.close()
    (plain, r0), (srch, r1) = runs[False], runs[True]
    better = 0
    for e in range(N):
        a, b = plain[e], srch[e]
        assert b.hd <= a.hd and b.target.seq == a.target.seq
        if a.hd == 0 or a.hd > 15:
            assert b.string == a.string and b.folded_structure == a.folded_structure
        assert rb.fold(b.string) == (b.folded_structure, b.fe)
        assert abs(r1[e] - b.r) < 1e-12 and abs(r0[e] - a.r) < 1e-12
        better += b.hd < a.hd
    assert better > 10


def test_environment_search_is_keyed_on_environment_and_episode():
    """The local search of an environment draws from (seed, environment index, episode): it does not depend on which other
    environments finish in the same step.  The first 40 environments of an 80-environment batch (round-robin targets,
    the same actions) end with exactly the solutions they reach in a 40-environment batch."""
    import rnagent_b200 as rb
    from rnagent_b200 import synth
    rb.set_vienna_params(1)
    seq, off = synth.random_sequences(5, 36, 77)
    structs, _ = rb.fold_batch([seq[off[k]:off[k + 1]].tobytes().decode() for k in range(5)])
    cfg = dict(kernel_size=25, reward_exp=9, encoding_type=2, use_nucleotides=True, permute=True, permutation_budget=15,
               permutation_radius=5, permutation_threshold=15, allow_gu_permutations=True, mutation_probability=0.5)
    acts = np.random.default_rng(8).integers(0, 4, (120, 80)).astype(np.int32)
    out = {}
    for N in (40, 80):
        env = rb.RnaDesignBatch(cfg, N, targets=structs, auto_reset=True, seed=11)
        env.reset()
        for t in range(120):                      # several episodes each: the episode counter enters the key
            env.step(np.ascontiguousarray(acts[t, :N]))
        out[N] = [(p.string, p.folded_structure, p.hd) for p in env.get_attr("prev_solution")]
        assert env.stats().episodes >= 2 * N
        env.close()
    assert out[80][:40] == out[40]
    assert len({s for s, _, _ in out[80]}) > 40   # ... and the environments are not all alike
