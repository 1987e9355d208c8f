#!/usr/bin/env python3
"""Latency of the per-call drop-in `fold(seq)` (what a reference call site sees after `fold_fn = rnagent_b200.fold`)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rnagent_b200 as rb

rb.set_vienna_params(1)
rng = np.random.default_rng(1)
for n in (30, 100, 200, 400):
    seqs = ["".join(rng.choice(list("ACGU"), n)) for _ in range(200)]
    for s in seqs[:20]:
        rb.fold(s)
    t0 = time.perf_counter()
    for s in seqs:
        rb.fold(s)
    dt = (time.perf_counter() - t0) / len(seqs)
    print(f"n={n}: {dt * 1e6:.0f} us per fold() call")
