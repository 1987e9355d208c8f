#!/usr/bin/env python3
"""Environment stepping rate: 4096 RnaDesign envs, n=100 targets, random actions on the device.
ENV_SETTLE=<steps> untimed steps first (default 20: episode cohorts still in step; 1500: drifted apart),
ENV_DELAY=<k> deferred rewards (one reward row per step, the terminal folds of k steps in one call)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rnagent_b200
from rnagent_b200 import synth, vienna

dev = torch.device("cuda:0")
vienna.set_vienna_params(1)
mixed = len(sys.argv) > 1 and sys.argv[1] in ("mixed", "eterna")
if mixed:
    lens = [60 + 4 * (k % 40) for k in range(256)] if sys.argv[1] == "mixed" else list(synth.eterna_mixed_lengths(256, 77))
    sseq, soff = synth.random_sequences_mixed(lens, 4100)
else:
    sseq, soff = synth.random_sequences(256, 100, 4100)
targets, _ = rnagent_b200.fold_batch([sseq[soff[k]:soff[k + 1]].tobytes().decode() for k in range(256)])
cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
N, steps = 4096, (600 if not (mixed and sys.argv[1] == "eterna") else 120)
settle = 400 if (mixed and sys.argv[1] == "eterna") else 20   # long targets: episodes of up to ~400 steps must de-synchronise first
settle = int(os.environ.get("ENV_SETTLE", settle))
delay = int(os.environ.get("ENV_DELAY", 1))
env = rnagent_b200.RnaDesignBatch(cfg, N, targets=targets, rank=0, auto_reset=True, randomize=True, seed=0)
acts = torch.randint(0, 4, (steps + settle, N), dtype=torch.int32, device=dev)
d_obs = torch.empty((N, 10, 50), dtype=torch.uint8, device=dev)
d_r = torch.empty(N, dtype=torch.float64, device=dev)
d_rT = torch.empty((steps, N), dtype=torch.float64, device=dev) if delay > 1 else None
d_done = torch.empty(N, dtype=torch.uint8, device=dev)
env.reset()
for k in range(settle):
    env.step_device(acts[k], d_obs, d_r, d_done)
env.stats(clear=True)
env.set_reward_delay(delay)
torch.cuda.synchronize()
t0 = time.perf_counter()
for k in range(steps):
    env.step_device(acts[settle + k], d_obs, d_r if delay <= 1 else d_rT[k], d_done)
env.flush()
torch.cuda.synchronize()
secs = time.perf_counter() - t0
st = env.stats()
print(f"{sys.argv[1] if mixed else 'n100'}: {st.steps / secs / 1e6:.2f} M env steps/s, {st.episodes / secs / 1e3:.1f} k episodes/s, {secs / steps * 1e6:.1f} us per vector step, mean HD {st.sum_hd / max(st.episodes, 1):.2f} (settle {settle}, reward delay {delay})")
