#!/usr/bin/env python3
"""Latency of one fold call against the batch size (n = 100, device buffers): the curve an environment step rides on."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rnagent_b200 import _lib, synth, vienna
vienna.set_vienna_params(1)
dev = torch.device("cuda:0")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
for count in (64, 148, 149, 200, 296, 444, 592, 740):
    seq, off = synth.random_sequences(count, n, 5)
    tgt = synth.random_targets(off, 6)
    d = [torch.from_numpy(x).to(dev) for x in (seq, off, tgt)]
    db = torch.empty_like(d[0]); mfe = torch.empty(count, dtype=torch.int32, device=dev); hd = torch.empty_like(mfe)
    st = torch.cuda.Stream(device=dev)
    def run():
        _lib.fold_hamming_raw(d[0], d[1], count, target=d[2], db=db, mfe=mfe, hd=hd, stream=st.cuda_stream)
    for _ in range(3): run()
    st.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        run(); st.synchronize()
    print(f"count {count:5d}: {1e6 * (time.perf_counter() - t0) / 20:7.1f} us per call  ({_lib.last_launch_info()})")
