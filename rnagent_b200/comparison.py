"""Result files of the reference's method comparison (rlif/utils/comparison.py:150-255): `solutions.csv` -- one block per
target: `<target>,<time limit>,<length>,<log name>`, a column line, one row per method -- and `<n>.log`.
The reference's logs under rlif/logs/* (the golden vectors of this repository) were written by this code.

`mass_fold` keeps the reference's control flow: every method is a callable `fn(target, time_limit=...)` returning a
sequence, a list of sequences or a list of Solution objects; candidates are evaluated by `Solution(string=...)`
(fold + Hamming distance + ensemble statistics on the GPU) until HD 0 or the time limit; the best candidate's row is
written with '{},{},{:.3f},{},{:.3f},{:.3f},{:.3f},{}'.  The external tools themselves (RNAinverse, MODENA, NUPACK,
antaRNA, MCTS, LEARNA; hard-coded paths in comparison.py:20-137) are the caller's business.
"""
import time

import numpy as np

from .dotbracket import DotBracket
from .solution import Solution

COLUMNS = ['seq', 'source', 't', 'hd', 'md', 'fe', 'ed', 'attempts']
# the configuration comparison.py:153-167 evaluates candidates with
CONFIG = {'encoding_type': 2, 'use_nucleotides': True, 'use_padding': True, 'full_state': True, 'kernel_size': 45, 'permute': False,
          'permutation_budget': 50, 'permutation_radius': 1, 'permutation_threshold': 5, 'allow_gu_permutations': False,
          'mutation_probability': 0.5, 'detailed_comparison': True, 'reward_exp': 9, 'write_threshold': 0}


def header_lines(seq, time_limit, log_name):
    """comparison.py:177-178."""
    return [','.join([seq, str(time_limit), str(len(seq)), log_name]) + '\n', ','.join(COLUMNS) + '\n']


def result_row(solution, name, t, attempts):
    """comparison.py:225-233."""
    return '{},{},{:.3f},{},{:.3f},{:.3f},{:.3f},{}\n'.format(solution.string, name, t, solution.hd, solution.md, solution.fe,
                                                              solution.ensemble_defect, attempts)


def timeout_row(name, t_total):
    """comparison.py:246."""
    return '{}, {}, {:.3f}\n'.format('---', name, t_total)


def log_line(to_log, t_total, name):
    """comparison.py:251."""
    return '{} {:.3f} {}\n'.format(to_log, t_total, name)


def mass_fold(seq, fns, names, time_limit=10, file=None, log=None, config=None, clock=time.time):
    """comparison.py:150-255 for one target.  Returns (solved per method, seconds per method)."""
    config = dict(CONFIG if config is None else config)
    target = DotBracket(seq)
    fh = open(file, 'a+') if file is not None else None
    lh = open(log, 'a+') if log is not None else None
    if fh is not None:
        for line in header_lines(seq, time_limit, (log or '').split('/')[-1].split('.')[0]):
            fh.write(line)
    solved, times = np.zeros(len(names)), np.zeros(len(names))
    for i, fn in enumerate(fns):
        t_total, hd, attempts = 0, 1000, 0
        best, lhd, to_log, result_string = None, 500, '---', None
        while hd != 0 and t_total < time_limit:
            t0 = clock()
            result = fn(seq, time_limit=time_limit - t_total)
            t_eval = clock() - t0
            if result is not None:
                if isinstance(result, str):
                    solutions = [Solution(target=target, config=config, string=result, time=t_eval, source=names[i])]
                elif isinstance(result, list) and result and isinstance(result[0], str):
                    solutions = [Solution(target=target, config=config, string=s, time=t_eval, source=names[i]) for s in result]
                else:
                    solutions = list(result)
                for solution in solutions:
                    attempts += 1
                    hd = solution.hd
                    if hd <= lhd:
                        lhd, best, to_log = hd, solution, solution.string
                        result_string = result_row(best, names[i], t_total + t_eval / len(solutions), attempts)
                    if hd == 0:
                        solved[i] += 1
                        break
            t_total += t_eval
        if best is None:
            result_string = timeout_row(names[i], t_total)
        times[i] += t_total
        if fh is not None:
            fh.write(result_string)
        if lh is not None:
            lh.write(log_line(to_log, t_total, names[i]))
    if fh is not None:
        fh.write('\n')
        fh.close()
    if lh is not None:
        lh.close()
    return solved, times
