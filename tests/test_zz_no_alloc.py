"""GPU: the allocation-free entry point rnafold_fold_hamming_batch_noalloc (static __device__ scratch, no cudaMalloc)
against the existing rnafold_fold_hamming_batch on the same inputs, bit for bit, on two calls in a row.

Collected last on purpose (file name): a failure here must not hide other tests.
No counterpart in the reference.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_batch(torch, dev, count, lo, hi, seed, hairpins=()):
    from rnagent_b200 import synth
    rng = np.random.default_rng(seed)
    lengths = rng.integers(lo, hi + 1, count).tolist()
    for b in hairpins:
        lengths[b] = 200
    seq, off = synth.random_sequences_mixed(lengths, seed=seed + 1)
    seq = seq.copy()
    for b in hairpins:   # a perfect 98-pair GC helix: below the int16 tables' checked range -> int32 re-run path
        seq[off[b]:off[b + 1]] = np.frombuffer(b"G" * 98 + b"AAAA" + b"C" * 98, dtype=np.uint8)
    tgt = synth.random_targets(off, seed=seed + 2)
    return seq, off, tgt, [torch.from_numpy(x).to(dev) for x in (seq, off, tgt)]


def _outputs(torch, dev, total, count):
    return (torch.full((total,), 7, dtype=torch.uint8, device=dev), torch.full((count,), 7, dtype=torch.int32, device=dev),
            torch.full((count,), 7, dtype=torch.int32, device=dev), torch.full((count,), 7.0, dtype=torch.float64, device=dev))


def test_noalloc_entry_point_matches_the_allocating_one_twice():
    import torch
    import rnagent_b200
    from rnagent_b200 import _lib
    L = _lib.lib()
    rnagent_b200.set_vienna_params(1)
    dev = torch.device("cuda", _lib.device_index())
    st = torch.cuda.Stream(device=dev)
    # two different batches back to back: the second call sees the first call's leftovers in the shared static array
    # (queue counters beyond the batch size, range-check flags); GC hairpins in the first batch force the int32
    # re-run path, so their flags are set when the second call starts
    for call, (count, lo, hi, nmax, seed) in enumerate([(3000, 0, 200, 200, 41), (2500, 60, 100, 100, 42)]):
        seq, off, tgt, (d_seq, d_off, d_tgt) = _device_batch(torch, dev, count, lo, hi, seed, (5, 17, 2999) if call == 0 else ())
        total = int(off[-1])
        want = _outputs(torch, dev, total, count)
        _lib.fold_hamming_raw(d_seq, d_off, count, target=d_tgt, db=want[0], mfe=want[1], hd=want[2], reward=want[3],
                              reward_exp=9.0, stream=st.cuda_stream)
        st.synchronize()
        got = _outputs(torch, dev, total, count)
        rc = L.rnafold_fold_hamming_batch_noalloc(d_seq.data_ptr(), d_off.data_ptr(), count, nmax, d_tgt.data_ptr(),
                                                  got[0].data_ptr(), got[1].data_ptr(), got[2].data_ptr(), got[3].data_ptr(),
                                                  9.0, st.cuda_stream)
        _lib.check(rc)
        st.synchronize()
        assert _lib.last_status() == 0
        for w, g, name in zip(want, got, ("structure", "mfe", "hd", "reward")):
            assert torch.equal(w, g), (call, name)          # integer / byte outputs bit-exact; fp64 reward bit-exact too
        if call == 0:
            assert int(got[1][5]) < -25000 and int(got[1][5]) == int(got[1][2999])   # the re-run path produced them
        assert int((got[1] != 7).sum()) > count // 2 and set(np.unique(got[0].cpu().numpy()).tolist()) <= {40, 41, 46}


def test_noalloc_limits_are_rejected_with_a_clear_error():
    import torch
    import rnagent_b200
    from rnagent_b200 import _lib
    L = _lib.lib()
    rnagent_b200.set_vienna_params(1)
    dev = torch.device("cuda", _lib.device_index())
    seq, off, tgt, (d_seq, d_off, d_tgt) = _device_batch(torch, dev, 64, 20, 40, 7)
    d_db, d_mfe, d_hd, d_rw = _outputs(torch, dev, int(off[-1]), 64)

    def call(B, nmax, seq_ptr=None, off_ptr=None):
        return L.rnafold_fold_hamming_batch_noalloc(seq_ptr or d_seq.data_ptr(), off_ptr or d_off.data_ptr(), B, nmax,
                                                    d_tgt.data_ptr(), d_db.data_ptr(), d_mfe.data_ptr(), d_hd.data_ptr(),
                                                    d_rw.data_ptr(), 9.0, None)
    assert call(65537, 40) == -1 and b"RNAFOLD_NOALLOC_MAX_BATCH" in L.rnafold_last_error()
    assert call(64, 201) == -1 and b"RNAFOLD_NOALLOC_MAX_LEN" in L.rnafold_last_error()
    assert call(64, 40, seq_ptr=seq.ctypes.data) == -1 and b"device pointer" in L.rnafold_last_error()
    # a sequence longer than the stated nmax: nothing is folded (outputs untouched), the status word says so
    assert call(64, 30) == 0
    with pytest.raises(_lib.RnaFoldError):
        _lib.last_status()
    assert int((d_mfe != 7).sum()) == 0
    assert _lib.last_status() == 0                      # cleared by the failed query
    assert call(64, 40) == 0 and _lib.last_status() == 0
    torch.cuda.synchronize()
    assert int((d_mfe == 7).sum()) < 8
