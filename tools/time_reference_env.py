#!/usr/bin/env python3
"""Time the REFERENCE's own environment loop (rlif/environments/RLIFenv.py:59-81, unmodified, imported from
/root/reference with the stubs of tools/make_env_golden.py; RNA.fold served by the oracle) on the C4 workload shape:
n = 100 targets (folds of random sequences), uniform random actions, one process.  Build container only."""
import os, sys, time, random
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import make_env_golden as meg

meg.install_stubs()
from rlif.rna import DotBracket                      # noqa: E402  (reference code)
from rlif.environments.RLIFenv import RnaDesign      # noqa: E402  (reference code)
from rnagent_b200 import synth                       # noqa: E402

seq, off = synth.random_sequences(64, 100, 4100)
targets = [meg.P2004.fold(seq[off[k]:off[k + 1]].tobytes().decode())[0] for k in range(64)]
cfg = dict(permute=False, verbose=False, seq_len=[0, 500], seq_count=10, encoding_type=2, use_nucleotides=True, kernel_size=25,
           reward_exp=9, permutation_budget=0, permutation_radius=1, permutation_threshold=5, allow_gu_permutations=False,
           mutation_probability=0.5)


class FakeData:
    def __init__(self, seqs):
        self.sequences = [DotBracket(s, file_id=i, file_nr=i, encoding_type=2) for i, s in enumerate(seqs)]
        self.n_seqs = len(self.sequences)


RnaDesign.load_data = lambda self: FakeData(targets)
env = RnaDesign(cfg, rank=0)
for testing_mode in (False,):
    env.testing_mode = testing_mode
    rng = random.Random(1)
    steps = episodes = 0
    env.reset()
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < float(sys.argv[1]) if len(sys.argv) > 1 else time.perf_counter() - t0 < 10.0:
        _, _, done, _ = env.step(rng.randint(0, 3))
        steps += 1
        if done:
            episodes += 1
            env.reset()
    dt = time.perf_counter() - t0
    print(f"reference RnaDesign.step, n=100 targets, one process: {steps / dt:.0f} env steps/s, {episodes / dt:.1f} episodes/s "
          f"({1e6 * dt / steps:.1f} us per step incl. {1e3 * dt / max(episodes, 1):.2f} ms per episode)")
