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
            assert sol[k].target == e["target"]
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
