"""Batched drop-in for the reference's environment layer, running on the B200.

Reference                                               here
------------------------------------------------------  -------------------------------------------
RnaDesign(config, rank).reset()/step(a)                 RnaDesignBatch(config, n_envs).reset()/step(actions)
  rlif/environments/RLIFenv.py:6-114                      N environments in lock-step on one GPU
create_env(...) -> SubprocVecEnv / DummyVecEnv          the same VecEnv-shaped surface: step/reset/get_attr/set_attr,
  rlif/learning/trainer.py:523-565                        finished environments re-armed inside step (auto_reset)
while not done: a = model.predict(s); s,_,done,_ =      rollout(env, policy, n_steps)
  env.step(a)   trainer.py:307-309, tester.py:85-87
Solution.evaluate reward                                fused into the fold kernels (fold + Hamming + reward)

Every per-step operation is a CUDA kernel behind the C ABI (rnafold_env_*); there is no CPU path.
"""
import ctypes

import numpy as np

from . import _lib
from .dotbracket import DotBracket

ROWS = 10


class _PrevSolution(object):
    """What callers read from env.get_attr('prev_solution') (rlif/learning/trainer.py:311-313)."""

    def __init__(self, target, string, folded, hd, fe, reward_exp, gc_penalty=False):
        # target: the DotBracket of the episode (file_nr indexes the tester's per-target arrays, tester.py:91)
        self.target, self.string, self.folded_structure, self.hd, self.fe = target, string, folded, hd, fe
        self.str = list(string)
        self.time = None
        self.mismatch_indices = np.where([a != b for a, b in zip(target.seq, folded)])[0]
        n = max(1, len(string))
        self.r = (1 - float(hd) / float(n)) ** reward_exp + (0.25 if hd == 0 else 0.0)
        if gc_penalty:                      # solution.py:150-151, gcau_content accumulates one increment per U
            u, inc = 0.0, 1. / n
            for ch in string:
                if ch == 'U':
                    u += inc
            self.r = self.r * 0.85 + float(np.clip(u, 0, .15))
        self.source = 'rlif'  # the local search's 'rlif*' tag (solution.py:215) is not tracked per environment

    def gcau_content(self):
        from .solution import gcau_content
        return gcau_content(self.str)

    def summary(self, print_output=False, colorize=False):
        from .solution import summary_lines
        lines = summary_lines(self)
        if print_output:
            print('\n'.join(lines))
        return lines

    def __repr__(self):
        return f"<solution hd={self.hd} fe={self.fe:.2f} {self.string}>"


class RnaDesignBatch(object):
    """N RnaDesign environments stepped in lock-step on one GPU.

    config keys read (same names as rlif/config/RnaDesign.yml `environment`): kernel_size, reward_exp, permute,
    permutation_budget / _radius / _threshold, mutation_probability, allow_gu_permutations, encoding_type,
    use_nucleotides, gc_penalty / gc.  With permute=True the terminal evaluation runs the batched local search
    (Solution.permute, rlif/rna/solution.py:157-219 -> rnafold_permute_batch).  boosting / testing-mode
    statistics are not part of this path and raise NotImplementedError instead of silently returning something else.

    step() / reset() return fresh arrays like the reference's VecEnvs (copy=False hands out the environment's own
    buffers, which the next call overwrites).
    """

    def __init__(self, config, n_envs, targets=None, rank=0, auto_reset=True, randomize=False, seed=0, copy=True):
        self.config = dict(config)
        if self.config.get('encoding_type', 2) != 2:
            raise NotImplementedError("only encoding_type=2 (rlif/config/RnaDesign.yml:26) is implemented")
        if not self.config.get('use_nucleotides'):
            # config_check('use_nucleotides') is False when the key is absent (solution.py:51-59, :71): the reference
            # then allocates 6 rows and insert_nucleotide indexes row 6+ -> IndexError on the first step
            raise NotImplementedError("use_nucleotides must be true (rlif/config/RnaDesign.yml:27); without it the "
                                      "reference's insert_nucleotide fails (solution.py:92)")
        self.copy = bool(copy)
        self.n_envs = int(n_envs)
        self.env_id = rank
        self.kernel_size = int(self.config.get('kernel_size', 25))
        self.reward_exp = float(self.config.get('reward_exp', 9))
        self.randomize = bool(randomize)
        self.meta_learning = True
        self.permute = bool(self.config.get('permute', False))
        self.testing_mode = False
        self.verbose = False
        self.seed = int(seed) + 7919 * int(rank)
        self.auto_reset = bool(auto_reset)
        self.L = _lib.lib()
        h = ctypes.c_int32(-1)
        _lib.check(self.L.rnafold_env_create(self.n_envs, self.kernel_size, ctypes.byref(h)))
        self._torch_stream, self._stream = False, None
        self._h = h.value
        self._configure()
        self.observation_shape = (ROWS, 2 * self.kernel_size, 1)
        self.dataset = []
        self._obs = np.zeros((self.n_envs, ROWS, 2 * self.kernel_size, 1), dtype=np.uint8)
        self._rew = np.zeros(self.n_envs, dtype=np.float64)
        self._done = np.zeros(self.n_envs, dtype=np.uint8)
        self._infos = [{} for _ in range(self.n_envs)]
        if targets is not None:
            self.set_data(targets)

    def _configure(self):
        from .permute import DEFAULTS
        _lib.check(self.L.rnafold_env_config(self._h, int(self.auto_reset), int(self.randomize), self.seed))
        self.gc_penalty = bool(self.config.get('gc_penalty') or self.config.get('gc'))   # solution.py:150
        _lib.check(self.L.rnafold_env_config_reward(self._h, int(self.gc_penalty)))
        c = {k: self.config.get(k, v) for k, v in DEFAULTS.items()}
        _lib.check(self.L.rnafold_env_config_permute(
            self._h, int(bool(self.permute)), int(c['permutation_budget']), int(c['permutation_radius']),
            int(c['permutation_threshold']), float(c['mutation_probability']), int(bool(c['allow_gu_permutations']))))

    # ---- RnaDesign.set_data / set_sequence (RLIFenv.py:106-113)
    def set_data(self, data):
        items = list(data.sequences if hasattr(data, 'sequences') else data)
        seqs = [d.seq if hasattr(d, 'seq') else str(d) for d in items]
        if not seqs:
            raise ValueError("empty data set")
        self.dataset = seqs
        self._data = data            # what get_attr('dataset') hands back (tester.py:66)
        self._dots = [d if isinstance(d, DotBracket) else DotBracket(s, None, k + 1) for k, (d, s) in enumerate(zip(items, seqs))]
        raw = np.frombuffer(''.join(seqs).encode(), dtype=np.uint8).copy()
        off = np.zeros(len(seqs) + 1, dtype=np.int32)
        off[1:] = np.cumsum([len(s) for s in seqs])
        self._configure()
        _lib.check(self.L.rnafold_env_set_targets(self._h, raw.ctypes.data, off.ctypes.data, len(seqs)))

    def set_sequence(self, sequence):
        self.set_data([sequence])

    # ---- VecEnv surface
    def reset(self):
        _lib.check(self.L.rnafold_env_reset(self._h, self._obs.ctypes.data))
        return self._obs.copy() if self.copy else self._obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int32)
        if a.shape != (self.n_envs,):
            raise ValueError(f"expected {self.n_envs} actions")
        _lib.check(self.L.rnafold_env_step(self._h, a.ctypes.data, self.reward_exp, self._obs.ctypes.data,
                                           self._rew.ctypes.data, self._done.ctypes.data))
        if self.copy:
            return self._obs.copy(), self._rew.copy(), self._done.view(np.bool_).copy(), [{} for _ in range(self.n_envs)]
        return self._obs, self._rew, self._done.view(np.bool_), self._infos

    def step_device(self, actions, obs=None, reward=None, done=None):
        """Same step with torch CUDA tensors (int32 actions; uint8 obs [N,10,2k], float64 reward, uint8 done)."""
        p = lambda t: None if t is None else t.data_ptr()
        if self._torch_stream:      # order the step against the caller's torch work (policy before, consumers after)
            import torch
            s = torch.cuda.current_stream().cuda_stream
            if s != self._stream:
                _lib.check(self.L.rnafold_env_set_stream(self._h, s, 0))
                self._stream = s
        _lib.check(self.L.rnafold_env_step(self._h, actions.data_ptr(), self.reward_exp, p(obs), p(reward), p(done)))

    def use_torch_stream(self, enabled=True):
        """step_device on torch's current CUDA stream (True: the step is stream-ordered with the tensors' producers and
        consumers) or on the environments' own non-blocking stream (False, default: synchronise yourself)."""
        self._torch_stream = bool(enabled)
        if not enabled and self._stream is not None:
            _lib.check(self.L.rnafold_env_set_stream(self._h, None, 1))
            self._stream = None

    def set_reward_delay(self, steps):
        """Deferred rewards for step_device (rnafold_env_config_delay): the terminal sequences of up to `steps` consecutive
        steps are folded together, and each reward lands in the `reward` tensor that was passed to ITS step -- at the latest
        after `steps` steps, on flush(), or before stats() / snapshot() / reset().  Pass a different reward tensor to every
        step of a window (a [T, N] rollout buffer).  steps=1: every step folds its own terminals (default)."""
        _lib.check(self.L.rnafold_env_config_delay(self._h, int(steps)))

    def flush(self):
        """Deliver the rewards that are still deferred (stream-ordered on the environments' stream, no host wait)."""
        _lib.check(self.L.rnafold_env_flush(self._h))

    def stats(self, clear=False):
        """Rollout statistics accumulated on the device since the last clear."""
        ist = np.zeros(4, dtype=np.uint64)
        fst = np.zeros(2, dtype=np.float64)
        _lib.check(self.L.rnafold_env_stats(self._h, ist.ctypes.data, fst.ctypes.data, int(clear)))
        return RolloutStats(int(ist[0]), int(ist[1]), int(ist[2]), int(ist[3]), float(fst[0]), float(fst[1]))

    def snapshot(self):
        N = self.n_envs
        lmax = max(len(s) for s in self.dataset)
        buf = {k: np.zeros((N, lmax), dtype=np.uint8) for k in ('seq', 'last_seq', 'last_db')}
        ints = {k: np.zeros(N, dtype=np.int32) for k in ('idx', 'cur', 'ep', 'last_hd', 'last_mfe', 'last_t')}
        lm = ctypes.c_int32(0)
        _lib.check(self.L.rnafold_env_snapshot(self._h, buf['seq'].ctypes.data, ints['idx'].ctypes.data, ints['cur'].ctypes.data,
                                               ints['ep'].ctypes.data, buf['last_seq'].ctypes.data, buf['last_db'].ctypes.data,
                                               ints['last_hd'].ctypes.data, ints['last_mfe'].ctypes.data,
                                               ints['last_t'].ctypes.data, ctypes.byref(lm)))
        buf.update(ints)
        return buf

    def get_attr(self, name):
        if name in ('ep', 'current_sequence'):
            return list(self.snapshot()['ep' if name == 'ep' else 'cur'])
        if name == 'prev_solution':
            s = self.snapshot()
            out = []
            for e in range(self.n_envs):
                t = int(s['last_t'][e])
                if t < 0:
                    out.append(None)
                    continue
                n = len(self.dataset[t])
                out.append(_PrevSolution(self._dots[t], s['last_seq'][e, :n].tobytes().decode(),
                                         s['last_db'][e, :n].tobytes().decode(), int(s['last_hd'][e]),
                                         float(np.float32(s['last_mfe'][e] / 100.0)), self.reward_exp,
                                         self.gc_penalty))
            return out
        if name == 'target_structure':
            return [self._dots[int(t)] for t in self.snapshot()['cur']]
        if name == 'dataset':
            return [self._data] * self.n_envs
        if name == 'next_target_structure':      # fetched, never called, by tester.py:67
            return [self.reset] * self.n_envs
        return [getattr(self, name)] * self.n_envs

    def set_attr(self, name, value):
        if name == 'dataset':
            return self.set_data(value)
        if name == 'current_sequence':       # tester.py:63; the next reset moves on to (value + 1) % len(dataset)
            return _lib.check(self.L.rnafold_env_set_current(self._h, int(value)))
        if name == 'meta_learning' and not value:
            raise NotImplementedError("meta_learning=False (same target every episode): use a one-target data set")
        if name == 'boosting' and value:
            raise NotImplementedError("boosting=True is not on the batched path")
        if name == 'testing_mode' and value:
            raise NotImplementedError("testing_mode=True computes the ensemble statistics of every terminal solution inside step(); "
                                      "call rnagent_b200.ensemble_batch on env.get_attr('prev_solution') instead")
        setattr(self, name, value)
        if name in ('randomize', 'auto_reset', 'seed', 'permute', 'config'):
            self._configure()

    def close(self):
        if self._h is not None and self._h >= 0:
            self.L.rnafold_env_destroy(self._h)
            self._h = -1

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RolloutStats(object):
    """Per-GPU counters of a rollout; `all_reduce` sums them over the ranks of a torch.distributed group
    (NCCL on GPUs, gloo in the CPU tests) -- the only collective on this path."""
    FIELDS = ('episodes', 'solved', 'sum_hd', 'steps', 'sum_reward', 'sum_mfe')

    def __init__(self, episodes=0, solved=0, sum_hd=0, steps=0, sum_reward=0.0, sum_mfe=0.0):
        self.episodes, self.solved, self.sum_hd, self.steps = episodes, solved, sum_hd, steps
        self.sum_reward, self.sum_mfe = sum_reward, sum_mfe

    def as_vector(self):
        return [float(getattr(self, f)) for f in self.FIELDS]

    def all_reduce(self, device=None, group=None):
        import torch
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            return self
        t = torch.tensor(self.as_vector(), dtype=torch.float64, device=device or 'cpu')
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        v = t.tolist()
        return RolloutStats(int(v[0]), int(v[1]), int(v[2]), int(v[3]), v[4], v[5])

    def summary(self):
        e = max(1, self.episodes)
        return dict(episodes=self.episodes, steps=self.steps, solved_frac=self.solved / e, mean_hd=self.sum_hd / e,
                    mean_reward=self.sum_reward / e, mean_mfe=self.sum_mfe / e)


def shard_targets(targets, rank, world):
    """Environment shard of one rank: target k lives on rank k mod world."""
    return list(targets[rank::world]) or list(targets)


def shard_sequences(lengths, rank, world):
    """Fold-batch shard of one rank: contiguous index ranges balanced on the work
    sum W(n_i) (cubic in the length), not on the count -- lengths of 12..400 nt differ 3e4-fold in cost.
    -> (begin, end) into the batch; the ranges of all ranks tile [0, len(lengths))."""
    from .synth import work_candidates
    import numpy as np
    lengths = np.asarray(lengths, dtype=np.int64)
    if world <= 1 or len(lengths) == 0:
        return (0, len(lengths)) if rank == 0 else (len(lengths), len(lengths))
    uniq = {int(n): float(work_candidates(int(n))) + 1.0 for n in np.unique(lengths)}
    cum = np.cumsum([uniq[int(n)] for n in lengths])
    cuts = [int(np.searchsorted(cum, cum[-1] * r / world, side='left')) for r in range(world + 1)]
    cuts[0], cuts[-1] = 0, len(lengths)
    return cuts[rank], cuts[rank + 1]


def create_env(env_name, config, n_workers=1, targets=None, **kw):
    """rlif/learning/trainer.py:523-565 shaped factory: one batched environment replaces the n_workers processes."""
    if env_name != 'RnaDesign':
        raise NotImplementedError(env_name)
    return RnaDesignBatch(config, n_workers, targets=targets, **kw)


def rollout(env, policy, n_steps):
    """The reference's vectorised loop (trainer.py:307-309): a, _ = policy(state); state, r, done, _ = env.step(a)."""
    state = env.reset()
    total_r = 0.0
    for _ in range(n_steps):
        actions = policy(state)
        state, r, done, _ = env.step(actions)
        total_r += float(r.sum())
    return env.stats(), total_r


def rollout_device(env, policy, n_steps, reward_delay=8, obs=None):
    """The vectorised rollout loop with everything on the device (trainer.py:307-309 keeps (state, r, done) per step for PPO):
    `policy(obs_uint8[N, 10, 2k]) -> int32 actions[N]` on CUDA tensors.  Returns (rewards[T, N] float64, dones[T, N] uint8,
    last observation, RolloutStats).  Rewards are deferred by up to `reward_delay` steps inside the loop -- the policy never
    sees them -- and complete when this function returns; values equal those of step-by-step folding."""
    import torch
    dev = torch.device("cuda", torch.cuda.current_device())
    N = env.n_envs
    if obs is None:
        obs = torch.empty((N, 10, 2 * env.kernel_size), dtype=torch.uint8, device=dev)
        obs.copy_(torch.from_numpy(env.reset()).reshape(N, 10, -1))
    rewards = torch.zeros((n_steps, N), dtype=torch.float64, device=dev)
    dones = torch.zeros((n_steps, N), dtype=torch.uint8, device=dev)
    env.set_reward_delay(reward_delay)
    env.use_torch_stream(True)
    try:
        for t in range(n_steps):
            actions = policy(obs)
            env.step_device(actions, obs, rewards[t], dones[t])
        env.flush()
    finally:
        env.set_reward_delay(1)
        torch.cuda.synchronize()
        env.use_torch_stream(False)
    return rewards, dones, obs, env.stats()
