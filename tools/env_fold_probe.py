#!/usr/bin/env python3
"""Time the fold of 4096 environment-made sequences against 4096 random ones (device-resident buffers)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import rnagent_b200
from rnagent_b200 import synth, vienna, _lib

dev = torch.device("cuda:0")
vienna.set_vienna_params(1)
sseq, soff = synth.random_sequences(256, 100, 4100)
targets, _ = rnagent_b200.fold_batch([sseq[soff[k]:soff[k + 1]].tobytes().decode() for k in range(256)])
cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
N = 4096
env = rnagent_b200.RnaDesignBatch(cfg, N, targets=targets, rank=0, auto_reset=True, randomize=True, seed=0)
acts = torch.randint(0, 4, (120, N), dtype=torch.int32, device=dev)
env.reset()
for k in range(110):
    env.step_device(acts[k])
snap = env.snapshot()
seqs = [bytes(r[:100]).decode() for r in snap['last_seq']]
print("example env sequence:", seqs[0])


def timed(label, seq_u8, off):
    B = len(off) - 1
    d_seq = torch.from_numpy(seq_u8).to(dev)
    d_db = torch.empty(len(seq_u8), dtype=torch.uint8, device=dev)
    d_mfe = torch.empty(B, dtype=torch.int32, device=dev)
    for _ in range(2):
        _lib.fold_hamming_raw(d_seq, off, B, db=d_db, mfe=d_mfe)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        _lib.fold_hamming_raw(d_seq, off, B, db=d_db, mfe=d_mfe)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    print(f"{label}: {dt * 1e3:.2f} ms per {B} folds = {B / dt / 1e6:.2f} M folds/s, launch info {_lib.last_launch_info()}")


buf = np.frombuffer("".join(seqs).encode(), dtype=np.uint8).copy()
off = (np.arange(N + 1) * 100).astype(np.int32)
timed("env-made", buf, off)
r, roff = synth.random_sequences(N, 100, 7)
timed("random  ", r, roff)
