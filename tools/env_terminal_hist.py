#!/usr/bin/env python3
"""Distribution of terminal environments per vector step in bench.py's C4 configuration after SETTLE untimed steps
(argv[1], default 300), and the step time over the counted window."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import rnagent_b200
from rnagent_b200 import synth, vienna
import time
SETTLE = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda:0")
vienna.set_vienna_params(1)
for rank in (0, 2):
    sseq, soff = synth.random_sequences(256, 100, 4100 + rank)
    targets, _ = rnagent_b200.fold_batch([sseq[soff[k]:soff[k + 1]].tobytes().decode() for k in range(256)])
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
    N = 4096
    env = rnagent_b200.RnaDesignBatch(cfg, N, targets=targets, rank=rank, auto_reset=True, randomize=True, seed=rank)
    g = torch.Generator(device=dev); g.manual_seed(rank)
    acts = torch.randint(0, 4, (SETTLE + 400, N), dtype=torch.int32, device=dev, generator=g)
    d_obs = torch.empty((N, 10, 50), dtype=torch.uint8, device=dev); d_r = torch.empty(N, dtype=torch.float64, device=dev)
    d_done = torch.empty(N, dtype=torch.uint8, device=dev)
    env.reset()
    counts = []
    for k in range(SETTLE + 400):
        env.step_device(acts[k], d_obs, d_r, d_done)
        if k >= SETTLE:
            counts.append(int(d_done.sum().item()))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(400):
        env.step_device(acts[SETTLE + k], d_obs, d_r, d_done)
    torch.cuda.synchronize()
    us = (time.perf_counter() - t0) / 400 * 1e6
    c = np.array(counts)
    lens = [t.count("(") + t.count(".") for t in targets]
    print(f"rank {rank}: terminal per step mean {c.mean():.1f} median {np.median(c):.0f} p90 {np.percentile(c, 90):.0f} max {c.max()} "
          f"steps with > 148: {(c > 148).mean() * 100:.1f} %; episode steps {min(lens)}..{max(lens)}; {us:.1f} us per vector step after {SETTLE + 400} steps")
    env.close()
