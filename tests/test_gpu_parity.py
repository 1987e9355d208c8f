"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle (bit-exact)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rand_batch(rng, B, nlo, nhi, alphabet="ACGU"):
    lens = rng.integers(nlo, nhi + 1, B)
    return ["".join(rng.choice(list(alphabet), int(n))) for n in lens]


def _oracle_batch(P, seqs, targets=None):
    buf = np.frombuffer("".join(seqs).encode(), dtype=np.uint8)
    off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int32)
    t = np.frombuffer("".join(targets).encode(), dtype=np.uint8) if targets else None
    db, mfe, hd = P.fold_batch(buf, off, targets_u8=t, nthreads=0)
    raw = db.tobytes().decode()
    return [raw[off[k]:off[k + 1]] for k in range(len(seqs))], mfe, hd


@pytest.fixture(scope="module")
def engine():
    import rnagent_b200
    from rnagent_b200 import _lib
    _lib.lib()
    yield rnagent_b200
    rnagent_b200.set_vienna_params(1)


def _check(engine, P, seqs):
    from rnagent_b200 import _lib
    from rnagent_b200.vienna import pack, unpack
    buf, off = pack(seqs)
    db = np.zeros_like(buf)
    mfe = np.zeros(len(seqs), dtype=np.int32)
    _lib.fold_hamming_raw(buf, off, len(seqs), db=db, mfe=mfe)
    got = unpack(db, off)
    want, wmfe, _ = _oracle_batch(P, seqs)
    bad = [k for k in range(len(seqs)) if got[k] != want[k] or mfe[k] != wmfe[k]]
    assert not bad, (len(bad), seqs[bad[0]], got[bad[0]], want[bad[0]], int(mfe[bad[0]]), int(wmfe[bad[0]]))


def test_golden_rows_both_parameter_sets(engine, golden_rows, golden_folds):
    from rnagent_b200 import _lib
    from rnagent_b200.vienna import pack, unpack
    for setno, name in ((1, "turner2004"), (2, "turner1999")):
        engine.set_vienna_params(setno)
        idx = [k for k, r in enumerate(golden_rows) if r["params"] == name]
        seqs = [golden_rows[k]["seq"] for k in idx]
        tgts = [golden_rows[k]["target"] for k in idx]
        structs, mfe, hd, mism, reward = engine.fold_hamming_batch(seqs, tgts, reward_exp=9)
        for n, k in enumerate(idx):
            s, e = golden_folds[k]
            assert structs[n] == s and int(round(float(mfe[n]) * 100)) == e, golden_rows[k]["line"]
            want_hd = sum(a != b for a, b in zip(s, tgts[n]))
            assert hd[n] == want_hd and len(mism[n]) == want_hd
            r = (1 - want_hd / len(s)) ** 9 + (0.25 if want_hd == 0 else 0)
            assert abs(reward[n] - r) < 1e-12
        launches, routed = _lib.last_launch_info()
        assert launches >= 3 and routed[0] > 0 and routed[1] > 0   # single-CTA and cluster kernels both exercised
    engine.set_vienna_params(1)


@pytest.mark.parametrize("setno,name", [(1, "turner2004"), (2, "turner1999"), (3, "andronescu2007"), (4, "langdon2018")])
def test_random_short(engine, oracle_params, setno, name):
    engine.set_vienna_params(setno)
    rng = np.random.default_rng(100 + setno)
    _check(engine, oracle_params(name), _rand_batch(rng, 3000, 0, 120))


def test_random_n100_block(engine, oracle_params):
    engine.set_vienna_params(1)
    rng = np.random.default_rng(100)
    _check(engine, oracle_params("turner2004"), _rand_batch(rng, 4096, 100, 100))


def test_random_long_cluster_kernel(engine, oracle_params):
    """257..464 nt: 2-CTA cluster kernel (DSMEM row exchange) + warp-per-sequence f5/backtrack kernel."""
    from rnagent_b200 import _lib
    engine.set_vienna_params(1)
    rng = np.random.default_rng(400)
    _check(engine, oracle_params("turner2004"), _rand_batch(rng, 160, 257, 432))
    assert _lib.last_launch_info()[1][1] == 160
    _check(engine, oracle_params("turner2004"), _rand_batch(rng, 64, 400, 400))
    _check(engine, oracle_params("turner2004"), _rand_batch(rng, 48, 433, 464) + _rand_batch(rng, 8, 464, 464))
    assert _lib.last_launch_info()[1][1] == 56   # the longest reference targets (451 nt) stay on the cluster path
    engine.set_vienna_params(2)
    _check(engine, oracle_params("turner1999"), _rand_batch(rng, 64, 300, 420, alphabet="GGCCAU"))
    engine.set_vienna_params(1)


def test_beyond_cluster_reach_uses_global_tables(engine, oracle_params):
    from rnagent_b200 import _lib
    engine.set_vienna_params(1)
    rng = np.random.default_rng(600)
    _check(engine, oracle_params("turner2004"), _rand_batch(rng, 12, 470, 620))
    assert _lib.last_launch_info()[1][2] == 12


def test_long_gc_rich_overflows_int16_in_cluster_path(engine, oracle_params):
    engine.set_vienna_params(1)
    rng = np.random.default_rng(17)
    _check(engine, oracle_params("turner2004"), _rand_batch(rng, 24, 300, 440, alphabet="GC") + _rand_batch(rng, 24, 300, 440))


def test_biased_composition_and_unknown_letters(engine, oracle_params):
    engine.set_vienna_params(1)
    rng = np.random.default_rng(5)
    P = oracle_params("turner2004")
    _check(engine, P, _rand_batch(rng, 1500, 20, 140, alphabet="GC"))      # deep stems, many ties
    _check(engine, P, _rand_batch(rng, 1500, 20, 140, alphabet="GGCCAU"))
    _check(engine, P, _rand_batch(rng, 1500, 20, 140, alphabet="ACGUN"))   # unpairable letters
    _check(engine, P, _rand_batch(rng, 500, 20, 140, alphabet="acgut"))


def test_int16_range_overflow_reruns_in_int32(engine, oracle_params):
    """GC-only sequences of 200-256 nt reach energies below the int16 tables' checked range: the fast
    kernel flags them on the device and the int32 kernel re-folds them; results stay bit-exact."""
    engine.set_vienna_params(1)
    rng = np.random.default_rng(16)
    P = oracle_params("turner2004")
    seqs = _rand_batch(rng, 64, 200, 256, alphabet="GC") + ["GC" * 128, "G" * 120 + "AAAA" + "C" * 120]
    seqs += _rand_batch(rng, 64, 129, 256)
    _check(engine, P, seqs)
    want, wmfe, _ = _oracle_batch(P, seqs[:66])
    assert wmfe.min() < -25000   # the case is really exercised


def test_every_length_up_to_70(engine, oracle_params):
    engine.set_vienna_params(1)
    rng = np.random.default_rng(70)
    seqs = ["".join(rng.choice(list("ACGU"), n)) for n in range(0, 71) for _ in range(12)]
    _check(engine, oracle_params("turner2004"), seqs)


def test_small_mixed_batches_take_one_launch(engine, oracle_params):
    """Fewer sequences than SMs are folded by ONE launch of the longest sequence's class (the RL step's case):
    short sequences then run in the wide-CTA, windowed instances as well."""
    from rnagent_b200 import _lib
    engine.set_vienna_params(1)
    rng = np.random.default_rng(4242)
    P = oracle_params("turner2004")
    for hi in (256, 224, 128, 100, 40):
        lens = [hi] + [int(v) for v in rng.integers(0, hi + 1, 90)]
        seqs = ["".join(rng.choice(list("ACGU"), n)) for n in lens]
        _check(engine, P, seqs)
        assert _lib.last_launch_info()[0] <= 2  # fold launch + (empty) int32 re-run


def test_fold_signature_and_edge_cases(engine):
    engine.set_vienna_params(1)
    s, e = engine.fold("GCAAAAAAAAAUUUGC")
    assert s == "(((((......)))))" and e == float(np.float32(-2.2))
    assert engine.fold("") == ("", 0.0)
    assert engine.fold("ACGU") == ("....", 0.0)
    structs, mfe = engine.fold_batch([])
    assert structs == [] and len(mfe) == 0
    with pytest.raises(NotImplementedError):
        engine.cvar.dangles = 0
    engine.cvar.dangles = 2
    assert engine.cvar.dangles == 2
    with pytest.raises(ValueError):
        engine.fold("A" * 5000)


def test_device_buffers_async(engine, oracle_params):
    import torch
    from rnagent_b200 import _lib
    from rnagent_b200.vienna import pack, unpack
    rng = np.random.default_rng(11)
    seqs = _rand_batch(rng, 2048, 60, 100)
    buf, off = pack(seqs)
    dev = torch.device("cuda", _lib.device_index())
    d_seq = torch.from_numpy(buf).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    d_db = torch.zeros_like(d_seq)
    d_mfe = torch.zeros(len(seqs), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    st = torch.cuda.Stream(device=dev)
    _lib.fold_hamming_raw(d_seq, d_off, len(seqs), db=d_db, mfe=d_mfe, stream=st.cuda_stream)
    st.synchronize()
    want, wmfe, _ = _oracle_batch(oracle_params("turner2004"), seqs)
    assert unpack(d_db.cpu().numpy(), off) == want
    assert np.array_equal(d_mfe.cpu().numpy(), wmfe)


def _full_size_checks(engine, P, name, count, n, seed, subset, rng_seed):
    """BASELINE.json full-size workload through the C ABI: size-independent properties on every sequence
    (energy of the returned structure re-evaluated by the oracle's independent loop-decomposition evaluator ==
    returned MFE; Hamming distance == byte comparison; idempotence) + bit-equality with the oracle on a subset."""
    from rnagent_b200 import _lib, synth
    seq, off = synth.random_sequences(count, n, seed)
    tgt = synth.random_targets(off, seed + 1)
    out = []
    for _ in range(2):
        db = np.zeros_like(seq)
        mfe = np.zeros(count, dtype=np.int32)
        hd = np.zeros(count, dtype=np.int32)
        _lib.fold_hamming_raw(seq, off, count, target=tgt, db=db, mfe=mfe, hd=hd, reward_exp=9.0)
        out.append((db, mfe, hd))
    (db, mfe, hd), (db2, mfe2, hd2) = out
    assert np.array_equal(db, db2) and np.array_equal(mfe, mfe2) and np.array_equal(hd, hd2)   # idempotent
    assert set(np.unique(db).tolist()) <= {ord("("), ord(")"), ord(".")}
    assert np.array_equal(hd, (db.reshape(count, n) != tgt.reshape(count, n)).sum(axis=1))
    sraw, draw = seq.tobytes().decode(), db.tobytes().decode()
    bad = 0
    for k in range(count):
        s, d = sraw[k * n:(k + 1) * n], draw[k * n:(k + 1) * n]
        bad += P.eval_structure(s, d) != int(mfe[k])
    assert bad == 0, bad
    assert (mfe <= 0).all() and mfe.min() < -1000
    pick = np.sort(np.random.default_rng(rng_seed).choice(count, subset, replace=False))
    sub = np.concatenate([seq[off[k]:off[k + 1]] for k in pick])
    soff = (np.arange(subset + 1) * n).astype(np.int32)
    wdb, wmfe, _ = P.fold_batch(sub, soff, nthreads=0)
    assert np.array_equal(wmfe, mfe[pick])
    assert np.array_equal(wdb.reshape(subset, n), db.reshape(count, n)[pick])
    return int(mfe.astype(np.int64).sum())


def test_full_size_c2_65536_x_n100(engine, oracle_params):
    engine.set_vienna_params(1)
    chk = _full_size_checks(engine, oracle_params("turner2004"), "c2", 65536, 100, 100, 4096, 1)
    assert chk == -151306730   # bench.py's checksum_mfe_dcal of the same workload


def test_full_size_c3_8192_x_n400(engine, oracle_params):
    from rnagent_b200 import _lib
    engine.set_vienna_params(1)
    _full_size_checks(engine, oracle_params("turner2004"), "c3", 8192, 400, 400, 192, 2)
    assert _lib.last_launch_info()[1][1] == 8192   # all through the cluster kernel


def test_full_size_c5_16384_mixed_turner1999(engine, oracle_params):
    """BASELINE.json configs[4] at full size, bench.py's own workload (16,384 sequences, Eterna100 length mix 12..400,
    Turner 1999): one call that spans every wave class and the cluster classes.  Size-independent properties on every
    sequence + bit-equality with the oracle on a subset drawn across all lengths."""
    from rnagent_b200 import _lib, synth
    engine.set_vienna_params(2)
    try:
        P = oracle_params("turner1999")
        count, seed = 16384, 1999
        seq, off = synth.random_sequences_mixed(synth.eterna_mixed_lengths(count, seed), seed)
        tgt = synth.random_targets(off, seed + 1)
        out = []
        for _ in range(2):
            db = np.zeros_like(seq)
            mfe = np.zeros(count, dtype=np.int32)
            hd = np.zeros(count, dtype=np.int32)
            _lib.fold_hamming_raw(seq, off, count, target=tgt, db=db, mfe=mfe, hd=hd, reward_exp=9.0)
            out.append((db, mfe, hd))
        launches, routed = _lib.last_launch_info()
        assert routed[0] > 0 and routed[1] > 0 and routed[0] + routed[1] + routed[2] == count and launches >= 10
        (db, mfe, hd), (db2, mfe2, hd2) = out
        assert np.array_equal(db, db2) and np.array_equal(mfe, mfe2) and np.array_equal(hd, hd2)   # idempotent
        assert set(np.unique(db).tolist()) <= {ord("("), ord(")"), ord(".")}
        mism = (db != tgt).astype(np.int64)
        assert np.array_equal(hd, np.add.reduceat(mism, off[:-1].astype(np.int64)))                  # HD == byte comparison
        sraw, draw = seq.tobytes().decode(), db.tobytes().decode()
        bad = sum(P.eval_structure(sraw[off[k]:off[k + 1]], draw[off[k]:off[k + 1]]) != int(mfe[k]) for k in range(count))
        assert bad == 0, bad                                   # energy of the returned structure == returned MFE
        assert (mfe <= 0).all()
        pick = np.sort(np.random.default_rng(5).choice(count, 768, replace=False))
        sub = np.concatenate([seq[off[k]:off[k + 1]] for k in pick])
        soff = np.concatenate([[0], np.cumsum([off[k + 1] - off[k] for k in pick])]).astype(np.int32)
        wdb, wmfe, _ = P.fold_batch(sub, soff, nthreads=0)
        assert np.array_equal(wmfe, mfe[pick])
        assert np.array_equal(wdb, np.concatenate([db[off[k]:off[k + 1]] for k in pick]))
        assert len(set((off[k + 1] - off[k]) for k in pick)) > 60        # the subset spans the length mix
    finally:
        engine.set_vienna_params(1)


def test_load_fasta_folds_all_records_in_one_batch(engine, tmp_path):
    engine.set_vienna_params(1)
    fa = tmp_path / "x.fasta"
    fa.write_text(">a\nGGGGAAAA\nCCCC\n>b\nGCAAAAAAAAATTTGC\n>c\nACGT\n")
    sols = engine.load_fasta(str(fa))
    assert [(s["name"], s["folded_structure"], s["fe"]) for s in sols] == [
        ("a", "((((....))))", float(np.float32(-5.4))), ("b", "(((((......)))))", float(np.float32(-2.2))), ("c", "....", 0.0)]
    assert sols[1]["string"] == "GCAAAAAAAAAUUUGC" and sols[1]["target"].base_pair_indices[0] == 15
    quirk = engine.load_fasta(str(fa), reference_quirks=True)   # the reference's loop: next header's name, last record dropped
    assert [(s["name"], s["string"]) for s in quirk] == [("b", "GGGGAAAACCCC"), ("c", "GCAAAAAAAAAUUUGC")]


@pytest.mark.parametrize("setno,name", [(2, "turner1999"), (3, "andronescu2007"), (4, "langdon2018")])
def test_other_parameter_sets_on_every_kernel_family(engine, oracle_params, setno, name):
    """Wide-CTA wave instances (129..256 nt), cluster kernel (257..464) and the global-table kernel under the
    reference's other parameter files (rlif/rna/vienna.py:11-15)."""
    engine.set_vienna_params(setno)
    rng = np.random.default_rng(900 + setno)
    seqs = _rand_batch(rng, 200, 129, 256) + _rand_batch(rng, 40, 257, 464) + _rand_batch(rng, 4, 470, 520)
    _check(engine, oracle_params(name), seqs)
    engine.set_vienna_params(1)


def test_int32_tables_when_a_parameter_set_exceeds_the_int16_margin(engine, tmp_path):
    """A parameter file whose loop terms are too large for the int16 fast path (wave_int16_ok) routes every
    sequence through the int32 instances of the wave kernel / the int32 CTA kernels; results stay bit-exact."""
    import oracle
    from conftest import param_file
    from rnagent_b200 import _lib
    text = open(param_file("turner2004")).read()
    head, rest = text.split("# bulge\n", 1)
    body, tail = rest.split("\n\n", 1)
    vals = body.split()
    vals[-1] = "2950"          # bulge[30]: beyond the 2700 dcal margin of the int16 tables
    big = tmp_path / "big_bulge.par"
    big.write_text(head + "# bulge\n" + " ".join(vals) + "\n\n" + tail)
    engine.read_parameter_file(str(big))
    P = oracle.Params(str(big))
    assert P.export().max() >= 2950
    rng = np.random.default_rng(3232)
    try:
        _check(engine, P, _rand_batch(rng, 1500, 0, 128))
        _check(engine, P, _rand_batch(rng, 120, 129, 256) + _rand_batch(rng, 6, 257, 330))
        assert _lib.last_launch_info()[1][1] == 0    # no int16 cluster kernel under this parameter set
    finally:
        engine.set_vienna_params(1)
