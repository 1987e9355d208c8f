#!/usr/bin/env python3
"""CPU cost of enqueueing one environment step (no synchronisation in the loop) next to the GPU-bound step time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rnagent_b200
from rnagent_b200 import synth, vienna

dev = torch.device("cuda:0")
vienna.set_vienna_params(1)
sseq, soff = synth.random_sequences(256, 100, 4100)
targets, _ = rnagent_b200.fold_batch([sseq[soff[k]:soff[k + 1]].tobytes().decode() for k in range(256)])
cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
N = 4096
env = rnagent_b200.RnaDesignBatch(cfg, N, targets=targets, rank=0, auto_reset=True, randomize=True, seed=0)
acts = torch.randint(0, 4, (700, N), dtype=torch.int32, device=dev)
d_obs = torch.empty((N, 10, 50), dtype=torch.uint8, device=dev)
d_r = torch.empty(N, dtype=torch.float64, device=dev)
d_done = torch.empty(N, dtype=torch.uint8, device=dev)
env.reset()
for k in range(300):
    env.step_device(acts[k], d_obs, d_r, d_done)
torch.cuda.synchronize()
for chunk in (50, 50, 200):
    t0 = time.perf_counter()
    for k in range(chunk):
        env.step_device(acts[300 + k], d_obs, d_r, d_done)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"{chunk} steps: enqueue {1e6 * (t1 - t0) / chunk:.1f} us/step on the CPU, {1e6 * (t2 - t0) / chunk:.1f} us/step until the GPU is done")
