#!/usr/bin/env python3
"""Golden vectors for the local search: run the REFERENCE's own Solution.evaluate(permute=True) /
Solution.permute (rlif/rna/solution.py:125-219, imported unmodified from /root/reference).

Same stubbing as tools/make_env_golden.py (RNA.fold is served by the CPU oracle).  The only other patch:
the module-level `random` the reference's solution.py uses is replaced by a stand-in whose
random() / randint() return the counter-based draws of rnagent_b200.permute.draw(), keyed by the
(round `step`, position `mismatch`) found in the calling frame of permute() -- so the reference's code
makes exactly the decisions the device kernels make for the same seed.

The reference visits the candidate positions in Python-set order; the device visits them in ascending
order.  The order matters only if two mutated positions of one round write the same nucleotide; each
case records `order_safe` (True when every round of the reference run iterated in ascending order or
had no write conflicts), and the tests require bit-equality on those.

Output: tests/golden/permute_cases.json.  Run in the build container only (needs /root/reference).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_env_golden as meg  # noqa: E402  (stubs + oracle parameters)
from rnagent_b200 import permute as pm  # noqa: E402  (host mirror of the device draw function; no GPU needed)


class FakeRandom(object):
    """Stand-in for the `random` module inside rlif.rna.solution."""

    def __init__(self):
        self.seed, self.index = 0, 0
        self.order_safe = True
        self.rounds = {}

    def _ctx(self):
        f = sys._getframe(2)
        while f is not None and f.f_code.co_name != "permute":
            f = f.f_back
        loc = f.f_locals
        step, m = int(loc["step"]), int(loc["mismatch"])
        # record, per round, the set order and the positions written so far, to flag order-sensitive rounds
        r = self.rounds.setdefault(step, {"order": list(loc["best_mismatch_indices"]), "writes": []})
        return step, m, r, loc

    def random(self):
        step, m, r, loc = self._ctx()
        return pm.draw_random(self.seed, self.index, step, m)

    def randint(self, a, b):
        step, m, r, loc = self._ctx()
        assert a == 0 and b in (3, 5)
        if b == 3:  # first draw of a mutated position: note what this mutation writes
            sol = loc["self"]
            w = [m]
            if sol.target.seq[m] != ".":
                pr = sol.target.base_pair_indices.get(m)
                if pr is None:
                    pr = sol.target.rev_base_pair_indices.get(m)
                if pr is not None:
                    w.append(int(pr))
            r["writes"].append(w)
        return pm.draw_randint(self.seed, self.index, step, m, b)

    def finish(self):
        """True if ascending-order iteration would have produced the same candidates in every round."""
        safe = True
        for r in self.rounds.values():
            flat = [p for w in r["writes"] for p in w]
            conflict = len(flat) != len(set(flat))
            ascending = r["order"] == sorted(r["order"])
            if conflict and not ascending:
                safe = False
        self.rounds = {}
        return safe


def main():
    meg.install_stubs()
    from rlif.rna import DotBracket          # noqa: E402  (reference code)
    import rlif.rna.solution as refsol        # noqa: E402  (reference code)
    fake = FakeRandom()
    refsol.random = fake
    nfolds = [0]
    real_fold = refsol.fold_fn

    def counting_fold(s):
        nfolds[0] += 1
        return real_fold(s)
    refsol.fold_fn = counting_fold

    rng = np.random.default_rng(157)
    cfgs = {"training": dict(pm.DEFAULTS), "testing": dict(pm.TESTING),
            "wide": dict(permutation_budget=8, permutation_radius=3, permutation_threshold=30, allow_gu_permutations=False,
                         mutation_probability=0.8)}
    cases = []
    for cname, pcfg in cfgs.items():
        cfg = dict(permute=True, verbose=False, encoding_type=2, use_nucleotides=True, kernel_size=25, reward_exp=9, **pcfg)
        made = 0
        tries = 0
        seed = int(rng.integers(0, 2**31))  # one seed per configuration: its cases form one batch, index = position
        while made < 24 and tries < 4000:
            tries += 1
            n = int(rng.integers(20, 110))
            base = "".join(rng.choice(list("ACGU"), n))
            target, _ = meg.P2004.fold(base)
            if target.count("(") < 3:
                continue
            # a near miss: the solving sequence with a few point mutations
            s = list(base)
            for _ in range(int(rng.integers(0, 5))):
                s[int(rng.integers(0, n))] = "ACGU"[int(rng.integers(0, 4))]
            seq = "".join(s)
            folded, _ = meg.P2004.fold(seq)
            hd0 = sum(a != b for a, b in zip(folded, target))
            # keep mostly searchable cases, plus a few that must pass through untouched (hd 0 / above threshold)
            searchable = 0 < hd0 <= pcfg["permutation_threshold"]
            if not searchable and made % 6 != 5:
                continue
            fake.seed, fake.index = seed, made
            tgt = DotBracket(target, file_id=0, file_nr=0, encoding_type=2)
            sol = refsol.Solution(target=tgt, config=cfg)
            sol.str = list(seq)
            nfolds[0] = 0
            r, hd, mism = sol.evaluate(permute=True, reward=True)
            safe = fake.finish()
            cases.append(dict(config=cname, target=target, seq=seq, seed=seed, index=made, hd0=int(hd0),
                              out_seq="".join(sol.str), out_hd=int(hd), out_folded=sol.folded_structure,
                              out_fe=float(sol.fe), out_r=float(r), folds=int(nfolds[0]), order_safe=bool(safe)))
            made += 1
    out = os.path.join(ROOT, "tests", "golden", "permute_cases.json")
    with open(out, "w") as f:
        json.dump({"configs": cfgs, "cases": cases}, f, indent=0)
    improved = sum(c["out_hd"] < c["hd0"] for c in cases)
    solved = sum(c["out_hd"] == 0 and c["hd0"] > 0 for c in cases)
    print(f"{len(cases)} cases ({improved} improved, {solved} solved, {sum(not c['order_safe'] for c in cases)} order-sensitive) "
          f"-> {out} ({os.path.getsize(out)} bytes)", file=sys.stderr)


if __name__ == "__main__":
    main()
