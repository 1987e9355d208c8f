"""CPU-only checks: the C-ABI library loads and exports every declared symbol, the host-side target
parser matches the reference's DotBracket on the golden episodes, the parameter loader matches the
oracle's, and the rollout statistics reduce over a 2-rank gloo group."""
import json
import os
import re
import subprocess
import sys

import numpy as np

from conftest import GOLDEN, ROOT, param_file


def test_library_exports_every_declared_symbol():
    from rnagent_b200 import _lib
    L = _lib.load_library()
    hdr = open(os.path.join(ROOT, "include", "rnafold_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(rnafold_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 25
    for n in sorted(names):
        assert hasattr(L, n), n
    assert names <= set(_lib.EXPORTS) | {"rnafold_version", "rnafold_last_error"}, names - set(_lib.EXPORTS)


def test_version_and_errors_without_device():
    from rnagent_b200 import _lib
    L = _lib.load_library()
    assert b"sm_100a" in L.rnafold_version()
    assert L.rnafold_current_device() == -1 or L.rnafold_current_device() >= 0
    assert L.rnafold_wave_smem_bytes(100, 2) > 0 and L.rnafold_wave_smem_bytes(100, 3) < 0


def test_parameter_loader_matches_oracle(oracle_params):
    from rnagent_b200 import _lib
    for name in ("turner2004", "turner1999", "andronescu2007", "langdon2018"):
        ours = _lib.parse_params(param_file(name))
        want = oracle_params(name).export()
        assert np.array_equal(ours, want), name
    assert np.array_equal(_lib.parse_params(None), oracle_params("turner2004").export())  # built-in default


def test_loader_reads_the_reference_original_parameter_files(oracle_params):
    """In the build container (where the reference checkout is mounted): the C++ loader parses the reference's own
    `rlif/utils/parameters/rna_*.par` -- enthalpy sections, C comments, INF / DEF tokens, sections in either order --
    to exactly the tables of the shipped 37 C re-serialisations.  Skipped where the reference is absent (GPU box)."""
    import pytest
    from rnagent_b200 import _lib
    ref = os.path.join(os.environ.get("RNAGENT_REFERENCE", "/root/reference"), "rlif", "utils", "parameters")
    if not os.path.isdir(ref):
        pytest.skip("reference checkout not present")
    for name in ("turner2004", "turner1999", "andronescu2007", "langdon2018"):
        ours = _lib.parse_params(os.path.join(ref, f"rna_{name}.par"))
        assert np.array_equal(ours, _lib.parse_params(param_file(name))), name
        assert np.array_equal(ours, oracle_params(name).export()), name


def test_target_parser_matches_reference_dotbracket():
    from rnagent_b200 import DotBracket
    eps = json.load(open(os.path.join(GOLDEN, "env_episodes.json")))["episodes"]
    for e in eps:
        t = DotBracket(e["target"])
        assert t.struct_motifs == e["motifs"]
        assert sorted([a, b] for a, b in t.base_pair_indices.items()) == e["pairs"]
        assert t.structure_encoding.shape == (6, len(e["target"])) and t.structure_encoding.sum() == len(e["target"])
    t = DotBracket("..((..((...))..((...))..))")
    assert t.struct_motifs == "EEOOIIOOHHHCCMMOOHHHCCIICC"
    assert t.rev_base_pair_indices[25] == 2


def test_permute_draws_mirror_the_library_and_golden_cases_are_consistent(oracle_params):
    """The Python mirror of the local search's counter-based draws equals the library's (host-callable) function,
    and the golden cases recorded from the reference's permute() agree with the oracle on what they claim."""
    import ctypes
    from rnagent_b200 import _lib, permute as pm
    L = _lib.load_library()
    rng = np.random.default_rng(1)
    for _ in range(2000):
        seed, idx, step, pos, kind = (int(rng.integers(0, 2**62)), int(rng.integers(0, 70000)), int(rng.integers(0, 60)),
                                      int(rng.integers(0, 4000)), int(rng.integers(0, 3)))
        assert L.rnafold_permute_draw(ctypes.c_uint64(seed), idx, step, pos, kind) == pm.draw(seed, idx, step, pos, kind)
    u = [pm.draw_random(5, 0, s, p) for s in range(40) for p in range(50)]
    assert 0.0 <= min(u) and max(u) < 1.0 and 0.45 < float(np.mean(u)) < 0.55
    g = json.load(open(os.path.join(GOLDEN, "permute_cases.json")))
    P = oracle_params("turner2004")
    assert len(g["cases"]) >= 60 and sum(c["out_hd"] < c["hd0"] for c in g["cases"]) >= 20
    for c in g["cases"]:
        s, e = P.fold(c["out_seq"])
        assert s == c["out_folded"] and sum(a != b for a, b in zip(s, c["target"])) == c["out_hd"] <= c["hd0"]
        thr = g["configs"][c["config"]]["permutation_threshold"]
        if not 0 < c["hd0"] <= thr:
            assert c["out_seq"] == c["seq"] and c["folds"] == 1


def test_rna_and_fasta_formats(tmp_path):
    """.rna targets (rlif/rna/utils.py:115-129) and FASTA records (utils.py:9-63): host-side parsing only."""
    import rnagent_b200 as rb
    d = tmp_path / "eterna"
    d.mkdir()
    (d / "1.rna").write_text("((((....))))\n")
    (d / "2.rna").write_text("..((...))..\nsecond line ignored\n")
    t = rb.load_sequence(1, dataset="eterna", data_dir=str(tmp_path))
    assert t.seq == "((((....))))" and t.file_nr == 1 and t.base_pair_indices[0] == 11
    assert rb.load_sequence(7, dataset="eterna", data_dir=str(tmp_path)) == []
    ds = rb.load_dataset("eterna", start=1, n_seqs=5, data_dir=str(tmp_path))
    assert [x.seq for x in ds] == ["((((....))))", "..((...)).."]
    fa = tmp_path / "x.fasta"
    fa.write_text(">a\nGGGGAAAA\nCCCC\n>b desc\nACGT\n")
    assert rb.read_fasta_records(str(fa)) == [("a", "GGGGAAAACCCC"), ("b desc", "ACGU")]
    out = tmp_path / "o.fa"
    rb.write_fasta(str(out), [dict(string="GGGGAAAACCCC", target="((((....))))", name="s1")], write_dot_brackets=True, linebreak=8)
    assert out.read_text() == ">s1\nGGGGAAAA\nCCCC\n((((....\n))))\n"
    assert rb.read_fasta_records(str(out))[0] == ("s1", "GGGGAAAACCCC((((....))))")


def test_solution_state_machine_replays_reference_episodes():
    """Host-side `Solution` (insert_nucleotide / find_next_unfilled / get_state, rlif/rna/solution.py:84-123) and the
    RnaDesign.step control flow (RLIFenv.py:59-81) against the episodes recorded from the reference's own code:
    every observation up to the terminal step, bit for bit, without a GPU."""
    from rnagent_b200 import DotBracket, Solution
    g = json.load(open(os.path.join(GOLDEN, "env_episodes.json")))
    cfg = g["config"]
    steps = 0
    for ep in g["episodes"]:
        tgt = DotBracket(ep["target"])
        sol = Solution(target=tgt, config=cfg)
        unpack = lambda hx: np.unpackbits(np.frombuffer(bytes.fromhex(hx), dtype=np.uint8))[:np.prod(ep["obs_shape"])].reshape(ep["obs_shape"])
        assert np.array_equal(sol.get_state().astype(np.uint8), unpack(ep["obs"][0]))
        for k, a in enumerate(ep["actions"]):
            sol.insert_nucleotide(a)
            done = not (sol.index < tgt.len - 1)
            if not done:
                sol.find_next_unfilled()
            assert done == ep["dones"][k]
            assert np.array_equal(sol.get_state().astype(np.uint8), unpack(ep["obs"][k + 1])), (ep["target"], k)
            steps += 1
        assert sol.string == ep["sequence"]
    assert steps >= 900


def test_hamming_distance_signature():
    from rnagent_b200 import hamming_distance
    hd, idx = hamming_distance("((..))", "(....)")
    assert hd == 2 and list(idx) == [1, 4]
    assert hamming_distance("((..))", "((")[0] == 0          # zip-truncating like the reference


def test_work_model_and_workloads():
    from rnagent_b200 import synth
    assert synth.work_candidates(100) == 1538430 and synth.work_candidates(400) == 45149380
    seq, off = synth.random_sequences(16, 100, 100)
    assert seq.shape == (1600,) and off[-1] == 1600 and set(seq.tobytes()) <= set(b"ACGU")
    lens = synth.eterna_mixed_lengths(1000)
    assert lens.min() >= 12 and lens.max() <= 400
    tg = synth.random_targets(off, 1)
    for k in range(16):
        s = tg[off[k]:off[k + 1]].tobytes().decode()
        depth = 0
        for c in s:
            depth += (c == "(") - (c == ")")
            assert depth >= 0
        assert depth == 0


_GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import torch.distributed as dist
from rnagent_b200 import RolloutStats, shard_sequences, shard_targets
from rnagent_b200.synth import eterna_mixed_lengths, work_candidates
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=rank, world_size=world)
targets = ["t%d" % k for k in range(7)]
mine = shard_targets(targets, rank, world)
local = RolloutStats(episodes=len(mine), solved=rank, sum_hd=10 * (rank + 1), steps=100 + rank, sum_reward=0.5 * (rank + 1), sum_mfe=-1.5 * (rank + 1))
tot = local.all_reduce()
assert tot.episodes == 7 and tot.solved == 1 and tot.sum_hd == 30 and tot.steps == 201, vars(tot)
assert abs(tot.sum_reward - 1.5) < 1e-12 and abs(tot.sum_mfe + 4.5) < 1e-12
assert sorted(mine + shard_targets(targets, 1 - rank, world)) == sorted(targets)
# fold batches: contiguous shards balanced on the work sum, tiling the batch exactly
lens = eterna_mixed_lengths(600, 5)
b, e = shard_sequences(lens, rank, world)
ob, oe = shard_sequences(lens, 1 - rank, world)
assert (b, oe) == (0, 600) and e == ob if rank == 0 else (ob, e) == (0, 600) and oe == b
t = torch_t = __import__("torch").tensor([float(sum(work_candidates(int(n)) for n in lens[b:e]))], dtype=__import__("torch").float64)
parts = [t.clone(), t.clone()]
dist.all_gather(parts, t)
w = [float(p.item()) for p in parts]
assert max(w) / sum(w) < 0.56, w     # within one long sequence of an even split
dist.destroy_process_group()
print("ok", rank)
"""


def test_rollout_stats_all_reduce_gloo_world2(tmp_path):
    port = 29500 + os.getpid() % 400
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER.format(root=ROOT, port=port))
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    for p in procs:
        out, _ = p.communicate(timeout=120)
        assert p.returncode == 0, out.decode()


def test_permute_candidate_order_is_cpythons_set_order():
    """The local search visits its candidate positions in the order the reference's `for mismatch in
    best_mismatch_indices` walks a Python set (rlif/rna/solution.py:166-189).  The library replays CPython's set
    insertion; here it is compared with real `set` objects built by the reference's own get_surrounding arithmetic
    (this interpreter is CPython >= 3.7: mode 0), and mode 2 with the ascending order."""
    import random
    import sys
    from rnagent_b200 import _lib
    assert sys.implementation.name == "cpython" and sys.version_info >= (3, 7)
    L = _lib.load_library()
    rng = random.Random(5)
    for trial in range(3000):
        n = rng.randint(1, 460)
        radius = rng.choice([1, 1, 2, 3, 5, 5, 8])
        mm = np.zeros(n, dtype=np.uint8)
        for _ in range(rng.randint(0, min(n, 40))):
            mm[rng.randrange(n)] = 1
        all_indices = []
        for index in np.where(mm)[0]:
            for i in range(1, radius + 1):
                if index - i >= 0:
                    all_indices.append(int(index - i))
                if index + i <= n - 1:
                    all_indices.append(int(index + i))
        want = list(set(all_indices))
        out = np.zeros(n + 1, dtype=np.int32)
        k = L.rnafold_permute_candidate_order(mm.ctypes.data, n, radius, 0, out.ctypes.data)
        assert out[:k].tolist() == want, (n, radius, np.where(mm)[0].tolist())
        k2 = L.rnafold_permute_candidate_order(mm.ctypes.data, n, radius, 2, out.ctypes.data)
        assert out[:k2].tolist() == sorted(want)
        k1 = L.rnafold_permute_candidate_order(mm.ctypes.data, n, radius, 1, out.ctypes.data)   # CPython 3.6 growth rule
        assert sorted(out[:k1].tolist()) == sorted(want)


def test_tester_loop_and_writers_on_a_scripted_env(tmp_path):
    """rlif/learning/tester.py:43-196 on the host: the evaluation loop's attribute round trip, bookkeeping and the
    three result files, driven by a scripted VecEnv (no GPU): episodes of three steps, every second one solved."""
    from rnagent_b200 import Dataset, PolicyModel, Tester
    from rnagent_b200.environments import _PrevSolution
    targets = ["((((....))))", "..((...))..", "....."]
    data = Dataset("scripted", sequences=targets)
    assert [t.file_nr for t in data.sequences] == [1, 2, 3] and data.n_seqs == 3

    class FakeEnv(object):
        n_envs = 1

        def __init__(self):
            self.attrs, self.t, self.ep, self.cur, self.prev = {}, 0, 0, -1, None

        def set_attr(self, k, v):
            self.attrs[k] = v
            if k == 'dataset':
                self.cur = 0
            if k == 'current_sequence':
                self.cur = v

        def get_attr(self, k):
            if k == 'prev_solution':
                return [self.prev]
            if k == 'next_target_structure':
                return [None]
            return [self.attrs[k]]

        def reset(self):
            self.cur = (self.cur + 1) % 3
            self.t = 0
            return np.zeros((1, 10, 50, 1), dtype=np.uint8)

        def step(self, a):
            self.t += 1
            done = self.t == 3
            if done:
                tgt = data[self.cur]
                hd = self.ep % 2
                self.prev = _PrevSolution(tgt, "A" * tgt.len, tgt.seq if hd == 0 else "." * tgt.len, hd, -1.5, 9)
                self.ep += 1
                self.reset()
            return np.zeros((1, 10, 50, 1), dtype=np.uint8), np.zeros(1), np.array([done]), [{}]

    class Wrapper(object):
        config = {"environment": dict(encoding_type=2, permute=False, permutation_budget=50, permutation_threshold=5, permutation_radius=1)}
        _model_path, current_checkpoint, env = "m", 3, None
    w = Wrapper()
    w.test_env = FakeEnv()
    w.model = PolicyModel(lambda s: np.zeros(1, dtype=np.int32))
    t = Tester(w, results_dir=str(tmp_path))
    solved, summary = t.timed_evaluation(dataset="scripted", time_limit=1e6, permute=False, verbose=False, test_set=data, max_episodes=6)
    assert w.test_env.attrs['randomize'] is False and w.test_env.attrs['meta_learning'] is True and w.test_env.attrs['permute'] is False
    assert t.attempts.tolist() == [2, 2, 2] and [s[0] for s in solved] == [1, 0, 2]      # walk starts at index 1 (tester.py:63)
    assert t.min_hd.tolist() == [0, 0, 0] and summary[1] == 3
    log = open(t.last_log).read()
    assert "solved 3/3 sequences with 1000000.0s eval budget.\n" + "=" * 100 + "\nPermutation strategy: False, budget: 50, threshold: 5, radius: 1\n" in log
    assert "\n\n\nSeq Nr.     2, Len:  11, DBr: 0.64, HD:   0, Reward: 1.25000  ||   G:0.00 | C:0.00 | A:1.00 | U:0.00 |\nS:  AAAAAAAAAAA\nF:  ..((...))..\nT:  ..((...))..\nD2; EEOOHHHCCEE\nFE: -1.500\nSolved in: 1 attempts, " in log
    rows = open(tmp_path / "scripted.csv").read().split("\n")
    assert rows[0] == "Date, Solved, Time(s), Model, Checkpoint, Dataset, seq1, seq2, seq3, seq1(t), seq2(t), seq3(t)"
    assert rows[1].split(", ")[1:9] == ["3", "1000000.0", "m", "3", "scripted", "0", "0", "0"]
    t.write_csv([solved], str(tmp_path / "tt.csv"), summary, 1e6)
    t.write_csv([solved], str(tmp_path / "tt.csv"), summary, 1e6)
    lines = open(tmp_path / "tt.csv").read().split("\n")
    assert lines[0].startswith("Date, Solved, Time(s), Model, Checkpoint, learn_test") and lines[1].endswith(", 3, 1000000.0, m, 3, 3")
