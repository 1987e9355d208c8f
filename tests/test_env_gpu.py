"""GPU: the batched environment against golden episodes recorded from the REFERENCE's own RnaDesign /
Solution code (tools/make_env_golden.py), and its VecEnv behaviour (auto-reset, statistics)."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(GOLDEN, "env_episodes.json")))


def _unpack(hexs, shape):
    bits = np.unpackbits(np.frombuffer(bytes.fromhex(hexs), dtype=np.uint8))
    return bits[:int(np.prod(shape))].reshape(shape)


def test_replays_reference_episodes_bit_for_bit(golden):
    import rnagent_b200
    rnagent_b200.set_vienna_params(1)
    eps = golden["episodes"]
    targets = []
    for e in eps:                       # the recording walked its target list round-robin, starting at index 1
        if e["target"] not in targets:
            targets.append(e["target"])
    targets = targets[-1:] + targets[:-1]      # the constructor takes target 0, the first reset() moves on to 1
    n_envs = 3
    env = rnagent_b200.RnaDesignBatch(golden["config"], n_envs, targets=targets, auto_reset=False)
    for e in eps:
        obs = env.reset()
        shape = tuple(e["obs_shape"])
        assert obs.shape == (n_envs,) + shape
        for k in range(n_envs):
            assert np.array_equal(obs[k], _unpack(e["obs"][0], shape)), e["target"]
        for t, a in enumerate(e["actions"]):
            obs, r, done, infos = env.step(np.full(n_envs, a, dtype=np.int32))
            want = _unpack(e["obs"][t + 1], shape)
            for k in range(n_envs):
                assert np.array_equal(obs[k], want), (e["target"], t)
                assert r[k] == pytest.approx(e["rewards"][t], abs=1e-12) and bool(done[k]) == e["dones"][t]
        sol = env.get_attr("prev_solution")
        for k in range(n_envs):
            assert sol[k].string == e["sequence"] and sol[k].folded_structure == e["folded"]
            assert sol[k].hd == e["hd"] and sol[k].fe == pytest.approx(e["fe"], abs=1e-6)
            assert sol[k].target.seq == e["target"]
    st = env.stats()
    assert st.episodes == n_envs * len(eps) and st.steps == n_envs * sum(len(e["actions"]) for e in eps)
    assert st.solved == n_envs * sum(1 for e in eps if e["hd"] == 0)
    env.close()


def test_auto_reset_statistics_and_device_step():
    import torch
    import rnagent_b200
    from rnagent_b200 import _lib, synth
    rnagent_b200.set_vienna_params(1)
    seq, off = synth.random_sequences(24, 60, 5)
    structs, _ = rnagent_b200.fold_batch([seq[off[k]:off[k + 1]].tobytes().decode() for k in range(24)])
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
    N = 512
    env = rnagent_b200.RnaDesignBatch(cfg, N, targets=structs, auto_reset=True, randomize=True, seed=3)
    rng = np.random.default_rng(0)
    obs = env.reset()
    assert obs.shape == (N, 10, 50, 1) and obs.dtype == np.uint8
    episodes = 0
    for _ in range(200):
        obs, r, done, _ = env.step(rng.integers(0, 4, N).astype(np.int32))
        episodes += int(done.sum())
        assert np.all((r > 0) == done)          # reward only on the terminal step
        assert np.all(r[done] <= 1.25 + 1e-12)
    st = env.stats()
    assert st.episodes == episodes and st.steps == 200 * N and episodes > N
    assert 0 < st.summary()["mean_reward"] < 1.25 and st.summary()["mean_mfe"] < 0
    # torch tensors on the device go through the same entry point without host copies
    dev = torch.device("cuda", _lib.device_index())
    a = torch.randint(0, 4, (N,), dtype=torch.int32, device=dev)
    d_obs = torch.zeros((N, 10, 50), dtype=torch.uint8, device=dev)
    d_r = torch.zeros(N, dtype=torch.float64, device=dev)
    d_done = torch.zeros(N, dtype=torch.uint8, device=dev)
    env.step_device(a, d_obs, d_r, d_done)
    assert int(d_obs.sum()) > 0 and env.stats().steps == 201 * N
    with pytest.raises(NotImplementedError):
        env.set_attr("boosting", True)
    env.close()


def test_solution_object_evaluates_like_the_reference(golden):
    """Per-object drop-in (`Solution.evaluate`, rlif/rna/solution.py:125-155) on the recorded episodes' final
    sequences: folded structure, HD, mismatch indices, MFE and reward; with permute=True never worse."""
    import rnagent_b200 as rb
    rb.set_vienna_params(1)
    cfg = dict(golden["config"], permutation_budget=15, permutation_radius=5, permutation_threshold=15,
               allow_gu_permutations=True)
    for ep in golden["episodes"][:16]:
        sol = rb.Solution(target=ep["target"], config=cfg)
        sol.str = list(ep["sequence"])
        r, hd, mism = sol.evaluate(reward=True)
        assert (sol.folded_structure, hd, [int(v) for v in mism]) == (ep["folded"], ep["hd"], ep["mismatch"])
        assert sol.fe == float(np.float32(ep["fe"])) and abs(r - ep["rewards"][-1]) < 1e-12
        better = rb.Solution(target=ep["target"], config=cfg, string=ep["sequence"])
        assert (better.hd, better.folded_structure) == (ep["hd"], ep["folded"])
        r2, hd2, _ = better.evaluate(permute=True, reward=True)
        assert hd2 <= hd and rb.fold(better.string)[0] == better.folded_structure


def _pair_policy(obs):
    """G for an opening bracket (its partner gets C), A elsewhere: row 0 of the window's centre column is motif 'O'."""
    k = obs.shape[2] // 2
    return np.where(obs[:, 0, k, 0] == 1, 2, 0).astype(np.int32)


def test_gc_penalty_reward_shaping(golden):
    """config gc_penalty / gc (rlif/rna/solution.py:150-151): r * 0.85 + clip(U fraction, 0, .15) on the batched path,
    against the reference formula evaluated on the recorded episodes' own sequences."""
    import rnagent_b200
    rnagent_b200.set_vienna_params(1)
    eps = [e for e in golden["episodes"] if len(e["target"]) <= 100][:12]
    for key in ("gc_penalty", "gc"):
        for e in eps[:6] if key == "gc" else eps:
            env = rnagent_b200.RnaDesignBatch(dict(golden["config"], **{key: True}), 2, targets=[e["target"]], auto_reset=False)
            env.reset()
            for a in e["actions"]:
                obs, r, done, _ = env.step(np.full(2, a, dtype=np.int32))
            assert done.all()
            n = len(e["sequence"])
            u, inc = 0.0, 1. / n
            for ch in e["sequence"]:
                if ch == "U":
                    u += inc
            want = e["rewards"][-1] * 0.85 + float(np.clip(u, 0, .15))
            assert abs(r[0] - want) < 1e-12 and r[0] == r[1]
            assert abs(env.get_attr("prev_solution")[0].r - want) < 1e-12
            env.close()


def test_create_env_rollout_and_timed_evaluation(tmp_path):
    """The rollout surface of rlif/learning/trainer.py:523-565, :307-313 and rlif/learning/tester.py:43-128 on the
    batched environment: create_env / rollout / set_attr('dataset' | 'randomize' | 'current_sequence' | 'permute') /
    get_attr('prev_solution'), the tester's bookkeeping and the files it writes."""
    import re
    import rnagent_b200 as rb
    from rnagent_b200 import Dataset, PolicyModel, Tester
    rb.set_vienna_params(1)
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True, permutation_budget=15,
               permutation_radius=5, permutation_threshold=15, allow_gu_permutations=True, mutation_probability=0.5)
    targets = ["((((....))))", "..((((...))))..", "(((..(((...)))..)))", "....", "((((....))))..((((....))))", ".((((((....)))))).",
               "(((((.....)))))"]
    # -- trainer.py:523-565 / :307-313
    env = rb.create_env("RnaDesign", cfg, 16, targets=targets)
    stats, total_r = rb.rollout(env, _pair_policy, 60)
    assert stats.steps == 16 * 60 and stats.episodes > 16 and stats.solved > 0 and total_r > 0
    prev = env.get_attr("prev_solution")
    assert all(p is not None and rb.fold(p.string)[0] == p.folded_structure for p in prev)
    assert all(p.target.seq in targets and p.hd == rb.hamming_distance(p.target.seq, p.folded_structure)[0] for p in prev)
    with pytest.raises(NotImplementedError):
        rb.create_env("RankedRnaDesign", cfg, 4)
    # -- tester.py:43-128 with the reference's attribute round trip
    class Wrapper(object):
        config = {"environment": cfg}
        _model_path, current_checkpoint = "models/x", 7
    w = Wrapper()
    w.test_env = rb.create_env("RnaDesign", cfg, 1, targets=["...."])        # replaced by set_attr('dataset', ...)
    w.env = env
    w.model = PolicyModel(_pair_policy, env)
    tester = Tester(w, results_dir=str(tmp_path))
    data = Dataset("toyset", sequences=targets)
    solved, summary = tester.timed_evaluation(dataset="toyset", time_limit=1e9, permute=True, verbose=False, test_set=data,
                                              max_episodes=21)
    assert w.model.env is env                                   # training environment restored (tester.py:121)
    assert w.test_env.dataset == targets and w.test_env.permute is True and w.test_env.randomize is False
    # current_sequence = 0 then reset(): the walk starts at target index 1 (file_nr 2) and goes round-robin
    assert tester.attempts.tolist() == [3] * 7 and int(tester.attempts.sum()) == 21
    assert len(solved) >= 12 and summary[1] == len(solved) and summary[3:] == ["models/x", 7, "toyset"]
    for target_id, sol, att, mhd, secs, limit in solved:
        assert sol.hd == 0 and mhd == 0 and sol.target.file_nr - 1 == target_id and sol.target.seq == targets[target_id]
        assert rb.fold(sol.string)[0] == sol.target.seq and 1 <= att <= 3 and limit == 1e9
    assert all(tester.min_hd[t] == 0 for t in {s[0] for s in solved})
    # -- the files (tester.py:130-196)
    log = open(tester.last_log).read().split("\n")
    assert re.fullmatch(r"Dataset: toyset, date: \d\d-\d\d_\d\d-\d\d, solved %d/7 sequences with 1000000000.0s eval budget\." % len(solved), log[0])
    assert log[1] == "=" * 100 and log[2] == "Permutation strategy: False, budget: 15, threshold: 15, radius: 5"
    body = "\n".join(log[3:])
    first = solved[0][1]
    assert ("\n\nSeq Nr. {:5}, Len: {:3}, DBr: {:.2f}, HD: {:3}, Reward: {:.5f}".format(first.target.file_nr, first.target.len,
            first.target.percent_unpaired, 0, first.r)) in "\n" + body
    assert body.count("S:  ") == len(solved) and body.count("Solved in: ") == len(solved)
    assert "S:  " + first.string + "\nF:  " + first.folded_structure + "\nT:  " + first.target.seq + "\nD2; " + first.target.struct_motifs in body
    rows = open(tmp_path / "toyset.csv").read().split("\n")
    assert rows[0] == "Date, Solved, Time(s), Model, Checkpoint, Dataset" + "".join(", seq%d" % (i + 1) for i in range(7)) + \
        "".join(", seq%d(t)" % (i + 1) for i in range(7))
    cells = rows[1].split(", ")
    assert cells[1:6] == [str(len(solved)), "1000000000.0", "models/x", "7", "toyset"]
    assert [int(c) for c in cells[6:13]] == tester.min_hd.tolist() and all(re.fullmatch(r"\(\d+\.\d+\)", c) for c in cells[13:])
    tester.evaluate(time_limit=1e9, datasets=("toyset",), test_set=data, max_episodes=7)
    assert open(tmp_path / "training_tests.csv").read() == \
        "Date, Solved, Time(s), Model, Checkpoint, learn_test, taneda, learn_validation, eterna\n"   # first call: header only (quirk)
    env.close()
    w.test_env.close()


@pytest.mark.parametrize("gc_penalty", [False, True])
def test_deferred_rewards_equal_step_by_step(gc_penalty):
    """rnafold_env_config_delay: folding the terminal sequences of up to 8 steps together changes WHEN a reward is written,
    never its value or its place -- rewards [T, N], dones, every observation, the rollout statistics and the last solutions
    equal those of step-by-step folding; short targets cap the window at their episode length; stats() in the middle of a
    window delivers what is pending."""
    import torch
    import rnagent_b200 as rb
    from rnagent_b200 import synth
    rb.set_vienna_params(1)
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(3)
    for lens, delay in (([30, 45, 60, 100], 8), ([12, 14, 40], 16), ([100], 5)):
        targets = [synth.random_structure(int(n), rng).tobytes().decode() for n in lens for _ in range(3)]
        cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True, gc_penalty=gc_penalty)
        N, T = 384, 230
        envs = [rb.RnaDesignBatch(cfg, N, targets=targets, auto_reset=True, randomize=True, seed=11) for _ in range(2)]
        g = torch.Generator(device=dev)
        g.manual_seed(5)
        acts = torch.randint(0, 4, (T, N), dtype=torch.int32, device=dev, generator=g)
        out = []
        for which, env in enumerate(envs):
            env.reset()
            env.set_reward_delay(delay if which else 1)
            obs = torch.zeros((T, N, 10, 50), dtype=torch.uint8, device=dev)
            rew = torch.full((T, N), -7.0, dtype=torch.float64, device=dev)
            done = torch.zeros((T, N), dtype=torch.uint8, device=dev)
            mid = None
            for t in range(T):
                env.step_device(acts[t], obs[t], rew[t], done[t])
                if t == 101:
                    mid = env.stats()            # closes the open window first
            env.flush()
            torch.cuda.synchronize()
            st, snap = env.stats(), env.snapshot()
            out.append((obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy(), st.as_vector(), mid.as_vector(), snap))
            env.close()
        a, b = out
        assert (a[2] == b[2]).all() and a[2].sum() > N                 # dones, several episodes per environment
        assert (a[0] == b[0]).all()                                     # observations
        assert (a[1] == b[1]).all() and (a[1] != -7.0).all()            # rewards: same values at the same (step, env)
        assert (a[1][a[2] == 0] == 0.0).all() and (a[1][a[2] == 1] > 0.0).mean() > 0.9
        assert np.allclose(a[3], b[3]) and np.allclose(a[4], b[4])     # statistics at the end and in the middle
        for key in ("last_seq", "last_db", "last_hd", "last_mfe", "last_t", "seq", "idx", "cur", "ep"):
            assert (a[5][key] == b[5][key]).all(), key


def test_rollout_device_matches_host_rollout():
    import torch
    import rnagent_b200 as rb
    from rnagent_b200 import synth
    rb.set_vienna_params(1)
    rng = np.random.default_rng(8)
    targets = [synth.random_structure(40, rng).tobytes().decode() for _ in range(5)]
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
    N, T = 64, 120

    def policy_np(state):                    # a deterministic function of the observation
        return (state.reshape(N, -1).astype(np.int64).sum(axis=1) % 4).astype(np.int32)

    def policy_t(obs):
        return (obs.reshape(N, -1).to(torch.int64).sum(dim=1) % 4).to(torch.int32)
    env = rb.RnaDesignBatch(cfg, N, targets=targets, auto_reset=True, randomize=False, seed=0)
    stats_h, total_h = rb.rollout(env, policy_np, T)
    env.close()
    env = rb.RnaDesignBatch(cfg, N, targets=targets, auto_reset=True, randomize=False, seed=0)
    rew, done, _, stats_d = rb.rollout_device(env, policy_t, T, reward_delay=8)
    env.close()
    assert np.allclose(stats_h.as_vector(), stats_d.as_vector()) and stats_d.episodes > N
    assert abs(float(rew.sum().item()) - total_h) < 1e-9 and int(done.sum().item()) == stats_d.episodes


def test_deferred_rewards_interleaved_with_other_calls():
    """Whatever reads or changes the environments' state closes the open window first: a host step(), a change of the delay,
    reset(), new targets and a stream switch in the middle of a window leave the same rewards as step-by-step folding."""
    import torch
    import rnagent_b200 as rb
    from rnagent_b200 import synth
    rb.set_vienna_params(1)
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(17)
    targets = [synth.random_structure(int(n), rng).tobytes().decode() for n in (24, 33, 47, 60)]
    targets2 = [synth.random_structure(int(n), rng).tobytes().decode() for n in (18, 52)]
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
    N, T = 96, 150
    g = torch.Generator(device=dev)
    g.manual_seed(9)
    acts = torch.randint(0, 4, (T, N), dtype=torch.int32, device=dev, generator=g)
    side = torch.cuda.Stream(device=dev)
    results = []
    for deferred in (False, True):
        env = rb.RnaDesignBatch(cfg, N, targets=targets, auto_reset=True, randomize=True, seed=4)
        env.reset()
        rew = torch.full((T, N), -3.0, dtype=torch.float64, device=dev)
        done = torch.zeros((T, N), dtype=torch.uint8, device=dev)
        obs = torch.zeros((N, 10, 50), dtype=torch.uint8, device=dev)
        host_r = []
        torch.cuda.synchronize()
        if deferred:
            env.set_reward_delay(12)
        for t in range(T):
            if t == 20 or t == 21:                      # a host step in the middle of a window
                _, r, d, _ = env.step(acts[t].cpu().numpy())
                host_r.append((r.copy(), d.copy()))
                rew[t] = torch.from_numpy(r).to(dev)
                done[t] = torch.from_numpy(d.astype(np.uint8)).to(dev)
                continue
            if t == 45 and deferred:
                env.set_reward_delay(5)                 # closes the window, then a shorter one
            if t == 70:
                env.reset()                             # pending rewards are delivered before the reset
            if t == 100:
                env.set_data(targets2)                  # ... and before new targets
                env.reset()
            if t == 120:
                env.use_torch_stream(True)
            if 120 <= t < 135:
                with torch.cuda.stream(side):           # steps ordered on a torch side stream
                    side.wait_stream(torch.cuda.current_stream())
                    env.step_device(acts[t], obs, rew[t], done[t])
                torch.cuda.current_stream().wait_stream(side)
            else:
                env.step_device(acts[t], obs, rew[t], done[t])
        env.flush()
        torch.cuda.synchronize()
        results.append((rew.cpu().numpy(), done.cpu().numpy(), env.stats().as_vector(), host_r))
        env.close()
    a, b = results
    assert (a[1] == b[1]).all() and a[1].sum() > 2 * N
    assert (a[0] == b[0]).all() and (a[0] != -3.0).all()
    assert np.allclose(a[2], b[2])
    for (r0, d0), (r1, d1) in zip(a[3], b[3]):
        assert (r0 == r1).all() and (d0 == d1).all()
