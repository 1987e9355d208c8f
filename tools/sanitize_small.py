#!/usr/bin/env python3
"""A small pass over every kernel family for compute-sanitizer (memcheck / racecheck / initcheck are slow: tiny inputs).

    compute-sanitizer --tool memcheck python tools/sanitize_small.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import rnagent_b200 as rb  # noqa: E402
from rnagent_b200 import synth  # noqa: E402

which = set(sys.argv[1:]) or {"wave", "cluster", "ensemble", "eval", "env", "permute", "noalloc"}
rng = np.random.default_rng(1)
rb.set_vienna_params(1)
rnd = lambda n: "".join(rng.choice(list("ACGU"), n))
if "wave" in which:
    seqs = [rnd(n) for n in (0, 5, 12, 30, 64, 100, 100, 128, 129, 200, 256, 290)]
    s, e = rb.fold_batch(seqs)
    print("wave ok", [float(v) for v in e][:6])
    s, e = rb.fold_batch([rnd(100) for _ in range(200)])      # throughput instance (more sequences than SMs)
    print("wave throughput ok", float(np.sum(e)))
if "cluster" in which:
    s, e = rb.fold_batch([rnd(300), rnd(400), rnd(451)])
    print("cluster ok", [float(v) for v in e])
if "ensemble" in which:
    res = rb.ensemble_batch([rnd(16), rnd(40), rnd(77)], targets=["." * 16, "." * 40, "." * 77])
    print("ensemble ok", res.ensemble_defect, res.MEA)
if "eval" in which:
    st, _ = rb.fold_batch([rnd(50), rnd(120)])
    print("eval ok", rb.eval_structure_batch([rnd(50), rnd(120)], ["." * 50, "." * 120]))
if "env" in which:
    dev = torch.device("cuda:0")
    targets = [synth.random_structure(int(n), rng).tobytes().decode() for n in (14, 30, 45)]
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
    env = rb.RnaDesignBatch(cfg, 48, targets=targets, auto_reset=True, randomize=True, seed=1)
    env.reset()
    acts = torch.randint(0, 4, (80, 48), dtype=torch.int32, device=dev)
    obs = torch.zeros((48, 10, 50), dtype=torch.uint8, device=dev)
    rew = torch.zeros((80, 48), dtype=torch.float64, device=dev)
    done = torch.zeros((80, 48), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    for t in range(40):
        env.step_device(acts[t], obs, rew[t], done[t])
    env.set_reward_delay(8)
    for t in range(40, 80):
        env.step_device(acts[t], obs, rew[t], done[t])
    env.flush()
    torch.cuda.synchronize()
    print("env ok", env.stats().summary())
    env.close()
if "permute" in which:
    tg = synth.random_structure(40, rng).tobytes().decode()
    res = rb.permute_batch([rnd(40) for _ in range(6)], [tg] * 6, dict(permutation_budget=3, permutation_radius=2, permutation_threshold=40), seed=0)
    print("permute ok", res.hd)
print("done")
