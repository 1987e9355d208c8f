#!/usr/bin/env python3
"""Generate golden episodes by running the REFERENCE's own environment code.

Imports rlif/rna/{dotbracket,solution,utils}.py and rlif/environments/RLIFenv.py from /root/reference
unmodified.  The third-party modules they import but which are absent here are stubbed:
  * RNA       -> `fold` is served by the CPU oracle (oracle/), so the reference's Solution.evaluate
                 runs its real code path (fold -> hamming_distance -> reward)
  * forgi     -> only used for boosting / graph helpers, which the driven path never calls
  * gym       -> bare Env / spaces classes
  * yaml.load -> PyYAML 6 needs an explicit Loader (the reference calls yaml.load(f))
`rlif/__init__.py` is not executed (it imports the TF1 learning stack); a namespace stand-in is used.

Each episode drives RnaDesign.step with seeded random actions, exactly like the reference's smoke
loops (rlif/utils/env_test.py:20-32), and records every observation, reward and done flag plus the
final sequence / folded structure / HD / MFE.  Output: tests/golden/env_episodes.json.

Run in the build container only (needs /root/reference); the fixture travels with the repo.
"""
import json
import os
import random
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("RNAGENT_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

P2004 = oracle.Params(os.path.join(ROOT, "rnagent_b200", "data", "turner2004_37.par"))


def install_stubs():
    import yaml
    _load = yaml.load
    yaml.load = lambda f, Loader=None: _load(f, Loader=Loader or yaml.SafeLoader)

    rna = types.ModuleType("RNA")
    rna.fold = lambda s: P2004.fold(s)
    rna.read_parameter_file = lambda p: None
    rna.cvar = types.SimpleNamespace()
    sys.modules["RNA"] = rna

    forgi = types.ModuleType("forgi")
    forgi.load_rna = lambda *a, **k: (None,)
    sys.modules["forgi"] = forgi

    gym = types.ModuleType("gym")
    gym.Env = object
    spaces = types.ModuleType("gym.spaces")
    spaces.Box = lambda **k: ("Box", k)
    spaces.Discrete = lambda n: ("Discrete", n)
    gym.spaces = spaces
    sys.modules["gym"] = gym
    sys.modules["gym.spaces"] = spaces

    pkg = types.ModuleType("rlif")
    pkg.__path__ = [os.path.join(REF, "rlif")]
    sys.modules["rlif"] = pkg


def main():
    install_stubs()
    from rlif.rna import DotBracket                      # noqa: E402  (reference code)
    from rlif.environments.RLIFenv import RnaDesign      # noqa: E402  (reference code)

    cfg = dict(permute=False, verbose=False, seq_len=[0, 500], seq_count=10, encoding_type=2, use_nucleotides=True,
               kernel_size=25, reward_exp=9, permutation_budget=0, permutation_radius=1, permutation_threshold=5,
               allow_gu_permutations=False, mutation_probability=0.5)

    targets = []
    for k in (1, 3, 8, 10, 15, 20, 23, 26, 30, 45, 57, 65):  # Eterna100 targets of 12..67 nt
        targets.append(open(os.path.join(REF, "rlif", "data", "eterna", f"{k}.rna")).read().strip())
    targets += ["((((....))))", "....", "(((...)))..", "..((..((...))..((...))..))", ".(((((...)))))", "((...))((...))",
                "(((((((..((((........)))).(((((.......))))).....(((((.......))))))))))))....", "(.(...).)" + "." * 40 + "((....))"]
    # long Eterna100 targets: 100, 151, 200 nt (wave kernel, 128- and 256-thread classes), 258 and 400 nt (cluster kernel);
    # every episode crosses the observation window (2k = 50) many times over
    for k in (92, 36, 69, 9, 22):
        targets.append(open(os.path.join(REF, "rlif", "data", "eterna", f"{k}.rna")).read().strip())

    class FakeData:
        def __init__(self, seqs):
            self.sequences = [DotBracket(s, file_id=i, file_nr=i, encoding_type=2) for i, s in enumerate(seqs)]
            self.n_seqs = len(self.sequences)

    RnaDesign.load_data = lambda self: FakeData(targets)
    env = RnaDesign(cfg, rank=0)
    env.randomize = False          # round-robin over the targets (RLIFenv.py:87-90)
    env.testing_mode = False       # no partition-function statistics (trainer.py:69)
    env.current_sequence = -1
    env.next_target_structure()

    rng = random.Random(2024)
    episodes = []
    for ep in range(2 * len(targets)):
        obs = env.reset()
        tgt = env.target_structure
        rec = {"target": tgt.seq, "motifs": tgt.struct_motifs, "pairs": sorted([int(a), int(b)] for a, b in tgt.base_pair_indices.items()),
               "actions": [], "obs": [np.packbits(obs.astype(np.uint8).reshape(-1)).tobytes().hex()], "rewards": [], "dones": []}
        done = False
        while not done:
            a = rng.randint(0, 3)
            obs, r, done, _ = env.step(a)
            rec["actions"].append(a)
            rec["obs"].append(np.packbits(obs.astype(np.uint8).reshape(-1)).tobytes().hex())
            rec["rewards"].append(float(r))
            rec["dones"].append(bool(done))
        sol = env.solution
        rec.update(sequence=sol.string, folded=sol.folded_structure, hd=int(sol.hd), fe=float(sol.fe),
                   mismatch=[int(v) for v in sol.mismatch_indices], obs_shape=list(obs.shape))
        episodes.append(rec)
    out = os.path.join(ROOT, "tests", "golden", "env_episodes.json")
    with open(out, "w") as f:
        json.dump({"config": cfg, "episodes": episodes}, f)
    steps = sum(len(e["actions"]) for e in episodes)
    print(f"{len(episodes)} episodes, {steps} steps -> {out} ({os.path.getsize(out)} bytes)", file=sys.stderr)


if __name__ == "__main__":
    main()
