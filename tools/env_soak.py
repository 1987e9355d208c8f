#!/usr/bin/env python3
"""Soak of the deferred-reward path at bench scale: N environments, T steps, rewards deferred by DELAY steps against
step-by-step folding -- rewards [T, N], dones, final statistics and last solutions must be identical.

    python tools/env_soak.py [N=4096] [T=1500] [DELAY=32] [mixed]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import rnagent_b200 as rb  # noqa: E402
from rnagent_b200 import synth  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1500
DELAY = int(sys.argv[3]) if len(sys.argv) > 3 else 32
mixed = len(sys.argv) > 4 and sys.argv[4] == "mixed"
dev = torch.device("cuda:0")
rb.set_vienna_params(1)
rng = np.random.default_rng(21)
lens = list(synth.eterna_mixed_lengths(64, 5)) if mixed else [100] * 64
targets = [synth.random_structure(int(n), rng).tobytes().decode() for n in lens]
cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
g = torch.Generator(device=dev)
g.manual_seed(3)
acts = torch.randint(0, 4, (T, N), dtype=torch.int32, device=dev, generator=g)
out = []
for delay in (1, DELAY):
    env = rb.RnaDesignBatch(cfg, N, targets=targets, auto_reset=True, randomize=True, seed=2)
    env.reset()
    env.set_reward_delay(delay)
    obs = torch.zeros((N, 10, 50), dtype=torch.uint8, device=dev)
    rew = torch.full((T, N), -1.0, dtype=torch.float64, device=dev)
    done = torch.zeros((T, N), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in range(T):
        env.step_device(acts[t], obs, rew[t], done[t])
    env.flush()
    torch.cuda.synchronize()
    secs = time.perf_counter() - t0
    out.append((rew.cpu().numpy(), done.cpu().numpy(), env.stats().as_vector(), env.snapshot(), obs.cpu().numpy()))
    print(f"delay {delay:2d}: {N * T / secs / 1e6:.2f} M env steps/s, {int(out[-1][1].sum())} episodes")
    env.close()
a, b = out
ok = (a[0] == b[0]).all() and (a[1] == b[1]).all() and np.allclose(a[2], b[2]) and (a[4] == b[4]).all()
for key in ("last_seq", "last_db", "last_hd", "last_mfe", "last_t", "seq", "idx", "cur", "ep"):
    ok = ok and bool((a[3][key] == b[3][key]).all())
print("identical" if ok else "DIFFERENT", "(rewards, dones, statistics, last solutions, state, last observation)")
sys.exit(0 if ok else 1)
