#!/usr/bin/env python3
"""Debug helper (GPU box): fold random batches on the GPU and with the oracle, print the first differences."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
import rnagent_b200
from rnagent_b200 import _lib
from rnagent_b200.vienna import pack, unpack

def main():
    nlo, nhi, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    alphabet = sys.argv[4] if len(sys.argv) > 4 else "ACGU"
    rng = np.random.default_rng(1)
    seqs = ["".join(rng.choice(list(alphabet), int(n))) for n in rng.integers(nlo, nhi + 1, B)]
    P = oracle.Params(os.path.join(os.path.dirname(rnagent_b200.__file__), "data", "turner2004_37.par"))
    buf, off = pack(seqs)
    db = np.zeros_like(buf); mfe = np.zeros(B, dtype=np.int32)
    try:
        _lib.fold_hamming_raw(buf, off, B, db=db, mfe=mfe)
    except Exception as e:
        print("ERR", e)
    got = unpack(db, off)
    wdb, wmfe, _ = P.fold_batch(buf, off, nthreads=0)
    want = unpack(wdb, off)
    bad = [k for k in range(B) if got[k] != want[k] or mfe[k] != wmfe[k]]
    print(f"{len(bad)} / {B} differ; launches={_lib.last_launch_info()}")
    for k in bad[:5]:
        print(len(seqs[k]), seqs[k]); print(" got ", got[k], mfe[k]); print(" want", want[k], wmfe[k])

main()
