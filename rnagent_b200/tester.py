"""Timed evaluation over benchmark target sets and the result files it writes -- the reference's
rlif/learning/tester.py:11-196 with the batched environment underneath.

Kept from the reference: the attribute round trip on the (vectorised) environment (`set_attr('dataset' | 'randomize' |
'meta_learning' | 'current_sequence' | 'permute')`, `get_attr('dataset' | 'next_target_structure' | 'target_structure' |
'prev_solution')`, tester.py:60-67, :80, :89), the `while not done: action, _ = model.predict(state); state, _, done, _ =
env.step(action)` loop (tester.py:84-87), the per-target bookkeeping (attempts, min_hd, t_per_sequence, solved list with
`[target_id, solution, attempts, min_hd, seconds, time_limit]` entries, tester.py:69-115) and the three writers with
their exact line formats (tester.py:130-196).

Different on purpose: the test environment may hold N > 1 lock-step environments (the reference's is a single
DummyVecEnv): every environment that finishes on a vector step is one episode of the loop above, in environment
order.  With N = 1 the control flow is the reference's.  Results directories are arguments (`results_dir`), not
rlif.settings constants.  `model` is anything with `predict(state) -> (actions, _)`; the wrapper needs `.model`,
`.test_env`, `.config['environment']`, and optionally `._model_path`, `.current_checkpoint`, `.env`.
"""
import datetime
import os
import time

import numpy as np

from .dataset import Dataset


class PolicyModel(object):
    """Minimal stand-in for the stable_baselines model surface the tester touches: predict / env / set_env."""

    def __init__(self, policy, env=None):
        self.policy, self.env = policy, env

    def set_env(self, env):
        self.env = env

    def predict(self, state):
        return self.policy(state), None


class Tester(object):
    def __init__(self, wrapper, budget=20, results_dir='results', data_dir=None):
        self.wrapper = wrapper
        self.budget = budget
        self.results_dir = results_dir
        self.data_dir = data_dir
        self.test_state = None
        self.done = None

    # ---- tester.py:24-41
    def evaluate(self, time_limit=60, verbose=False, permute=False, datasets=('rfam_runge', 'rfam_modena', 'eterna'), **kw):
        results, desc = [None for _ in datasets], None
        for i, dataset in enumerate(datasets):
            results[i], desc = self.timed_evaluation(dataset=dataset, time_limit=time_limit, verbose=verbose, permute=permute, **kw)
        path = os.path.join(self.results_dir, 'training_tests.csv')
        self.write_csv(results, path, desc, time_limit)
        return results

    # ---- tester.py:43-128
    def timed_evaluation(self, dataset='rfam_runge', time_limit=60, permute=False, verbose=True, test_set=None,
                         max_episodes=None):
        model = self.wrapper.model
        model.set_env(self.wrapper.test_env)
        n_seqs = 29 if dataset == 'rfam_modena' else 100
        if test_set is None:
            test_set = Dataset(dataset=dataset, start=1, n_seqs=n_seqs,
                               encoding_type=self.wrapper.config['environment']['encoding_type'], data_dir=self.data_dir)
        n_seqs = test_set.n_seqs
        env = model.env
        env.set_attr('dataset', test_set)
        env.set_attr('randomize', False)
        env.set_attr('meta_learning', True)
        env.set_attr('current_sequence', 0)
        env.set_attr('permute', permute)
        data = env.get_attr('dataset')[0]
        get_next_target = env.get_attr('next_target_structure')  # noqa: F841  (fetched and unused, as in the reference)
        assert data is test_set

        solved = []
        t_total = 0
        attempts = np.zeros([n_seqs], dtype=np.uint8)
        min_hd = np.full([n_seqs], 500, dtype=np.uint16)   # np.ones(uint8) * 500 promotes to uint16 under the reference's numpy
        t_per_sequence = np.zeros([n_seqs])
        test_state = env.reset()
        n_envs = getattr(env, 'n_envs', 1)
        ep_start = np.full(n_envs, time.time())
        episodes = 0
        try:
            while t_total <= time_limit and (max_episodes is None or episodes < max_episodes):
                done = np.zeros(n_envs, dtype=bool)
                while not done.any():
                    action, _ = model.predict(test_state)
                    test_state, _, done, _ = env.step(action)
                    done = np.asarray(done, dtype=bool)
                now = time.time()
                solutions = env.get_attr('prev_solution')
                for e in np.where(done)[0]:
                    solution = solutions[e]
                    target_id = solution.target.file_nr - 1
                    attempts[target_id] += 1
                    t_episode = now - ep_start[e]
                    ep_start[e] = now
                    t_total += t_episode / n_envs        # wall clock: N environments share it
                    t_per_sequence[target_id] += t_episode
                    episodes += 1
                    if solution.hd < min_hd[target_id]:
                        min_hd[target_id] = solution.hd
                    if solution.hd <= 0:
                        if verbose:
                            solution.summary(True)
                            print('({}/{}) Solved sequence: {} in {} iterations, {:.2f} seconds...\n'.format(
                                len(solved), n_seqs, target_id, attempts[target_id], t_per_sequence[target_id]))
                        solved.append([target_id, solution, attempts[target_id], min_hd[target_id],
                                       round(t_per_sequence[target_id], 2), time_limit])
        except KeyboardInterrupt:
            pass
        if verbose:
            print('Solved {}/{}'.format(len(solved), n_seqs))
        self.date = date = str(datetime.datetime.now().strftime("%m-%d %H:%M"))
        if getattr(self.wrapper, 'env', None) is not None:
            model.set_env(self.wrapper.env)  # restore the training environment
        test_summary = [date, len(solved), time_limit, getattr(self.wrapper, '_model_path', None),
                        getattr(self.wrapper, 'current_checkpoint', None), dataset]
        self.attempts, self.min_hd, self.t_per_sequence = attempts, min_hd, t_per_sequence
        self.write_test_results(solved, test_set, time_limit)
        self.write_detailed_csv(test_summary, min_hd, t_per_sequence)
        return solved, test_summary

    # ---- tester.py:130-160: <results>/<dataset>/<date>_<solved>.log
    def write_test_results(self, results, dataset, time_limit):
        date = datetime.datetime.now().strftime("%m-%d_%H-%M")
        directory = os.path.join(self.results_dir, dataset.dataset)
        config = self.wrapper.config['environment']
        if not os.path.isdir(directory):
            os.makedirs(directory)
        filename = os.path.join(directory, '{}_{}.log'.format(date, len(results)))
        with open(filename, 'w') as f:
            msg = 'Dataset: {}, date: {}, solved {}/{} sequences with {}s eval budget.\n'.format(
                dataset.dataset, date, len(results), dataset.n_seqs, time_limit)
            msg += ''.join(['='] * 100) + '\n'
            msg += 'Permutation strategy: {}, budget: {}, threshold: {}, radius: {}\n'.format(
                config['permute'], config['permutation_budget'], config['permutation_threshold'], config['permutation_radius'])
            f.write(msg)
            for result in results:
                for line in result[1].summary():
                    f.write(line + '\n')
                f.write('Solved in: {} attempts, {}s\n'.format(result[2], result[4]))
        self.last_log = filename
        return filename

    # ---- tester.py:162-174 (quirk kept: the call that creates the file writes only the header)
    def write_csv(self, results, path, test_summary, time_limit):
        header = 'Date, Solved, Time(s), Model, Checkpoint, learn_test, taneda, learn_validation, eterna\n'
        data = ', '.join([str(x) for x in test_summary[:-1]])
        data += ', ' + ', '.join([str(len(x)) for x in results]) + '\n'
        os.makedirs(os.path.dirname(path) or '.', exist_ok=True)
        if not os.path.isfile(path):
            with open(path, 'w') as f:
                f.write(header)
        else:
            with open(path, 'a') as f:
                f.write(data)

    # ---- tester.py:176-196: <results>/<dataset>.csv, one row per evaluation with per-target min HD and seconds
    def write_detailed_csv(self, test_summary, hd, t_per_sequence):
        header = 'Date, Solved, Time(s), Model, Checkpoint, Dataset'
        sequence_data = ''
        for i in range(len(hd)):
            sequence_data += ', ' + str(hd[i])
            header += ', seq{}'.format(i + 1)
        for i in range(len(t_per_sequence)):
            sequence_data += ', ({})'.format(round(t_per_sequence[i], 2))
            header += ', seq{}(t)'.format(i + 1)
        header += '\n'
        data_entry = ', '.join([str(x) for x in test_summary]) + sequence_data + '\n'
        os.makedirs(self.results_dir, exist_ok=True)
        path = os.path.join(self.results_dir, test_summary[-1] + '.csv')
        if not os.path.isfile(path):
            with open(path, 'w') as f:
                f.write(header)
        with open(path, 'a') as f:
            f.write(data_entry)
        return path
