"""Per-object drop-in for the reference's `Solution` (rlif/rna/solution.py:11-155): the same attributes and
methods the environment, the tester and the CLI touch on ONE candidate sequence.

The state machine (`create_encoding`, `insert_nucleotide`, `find_next_unfilled`, `get_state`, solution.py:62-123)
is plain host code, kept quirk for quirk (the last slot may be overwritten; one-hot bits accumulate).  The
arithmetic -- `evaluate` = fold + Hamming distance + reward, optionally the local search `permute`
(solution.py:125-219) -- runs on the GPU through the batched C-ABI calls with a batch of one; there is no CPU fold.
For many candidates use `fold_hamming_batch` / `permute_batch` / `RnaDesignBatch` instead: same results, one launch.
"""
import time as t

import numpy as np

from .dotbracket import DotBracket
from .permute import DEFAULTS as _PERMUTE_DEFAULTS
from .permute import permute_batch
from .vienna import fold_hamming_batch


def gcau_content(chars):
    """rlif/rna/solution.py:257-268: proportions of G, C, A, U, accumulated one increment per nucleotide."""
    gcau = dict(G=0, C=0, A=0, U=0)
    increment = 1. / max(1, len(chars))
    for nucleotide in chars:
        if nucleotide in gcau:
            gcau[nucleotide] += increment
    return gcau


def summary_lines(sol):
    """Solution.summary (rlif/rna/solution.py:329-345) without the console colouring: the six lines the result
    logs hold per solution (header with sequence number / length / unpaired fraction / HD / reward / nucleotide
    proportions, S:, F:, T:, D2; and FE:)."""
    gcau = sol.gcau_content()
    t = sol.target
    nr = t.file_nr if t.file_nr is not None else 0
    header = '\n\nSeq Nr. {:5}, Len: {:3}, DBr: {:.2f}, HD: {:3}, Reward: {:.5f}'.format(nr, t.len, t.percent_unpaired, sol.hd, sol.r)
    header += '  ||   G:{:.2f} | C:{:.2f} | A:{:.2f} | U:{:.2f} |'.format(gcau['G'], gcau['C'], gcau['A'], gcau['U'])
    return [header, 'S:  ' + sol.string, 'F:  ' + sol.folded_structure, 'T:  ' + t.seq, 'D2; ' + t.struct_motifs,
            'FE: {:.3f}'.format(sol.fe)]


class Solution(object):
    def __init__(self, target, config=None, string=None, time=None, source='rlif'):
        self.config = dict(config or {})
        self.target = target if isinstance(target, DotBracket) else DotBracket(str(target))
        self.time = time
        self.start = None
        self.mapping = {0: 'AU', 1: 'CG', 2: 'GC', 3: 'UA', 4: 'GU', 5: 'UG'}
        self.reverse_action = {0: 3, 1: 2, 2: 1, 3: 0, 4: 5, 5: 4}
        self.source = source
        self.str = ['-'] * self.target.len
        self.hd = 100
        self.md = 0
        self.fe = 0
        self.r = 0
        self.mismatch_indices = None
        self.folded_structure = ''
        self.reward_exp = self.config.get('reward_exp', 9)
        self.kernel_size = self.config.get('kernel_size', 25)
        self.index = 0
        self.permute_seed = int(self.config.get('permute_seed', 0))
        self.init_vars()
        self.create_encoding()
        if self.config_check('boosting'):
            raise NotImplementedError("boosting (forgi graph heuristics) is not part of this path")
        if string is not None:
            self.str = [x for x in string]
            self.evaluate(string, permute=False, compute_statistics=True)   # as the reference (solution.py:44-46)
        if time is None:
            self.start = t.time()

    def config_check(self, parameter):
        return self.config[parameter] if self.config.get(parameter) is not None else False

    # ---- solution.py:62-75
    def create_encoding(self):
        k = self.kernel_size
        self.rows, length = self.target.structure_encoding.shape
        structure_encoding = np.zeros([self.rows, length + k * 2])
        structure_encoding[:, k:k + length] = self.target.structure_encoding
        if self.config_check('use_nucleotides'):   # absent -> False, as the reference (solution.py:51-59, :71)
            self.encoding = np.vstack([structure_encoding, np.zeros([4, length + k * 2])])
        else:
            self.encoding = structure_encoding

    @property
    def string(self):
        return ''.join(self.str)

    # ---- solution.py:84-123
    def insert_nucleotide(self, action):
        i, k = self.index, self.kernel_size
        self.encoding[self.rows + action, i + k] = 1
        self.str[i] = self.mapping[action][0]
        if self.target.seq[i] == '(':
            pair_index = self.target.base_pair_indices.get(i)
            if pair_index is not None:
                self.encoding[self.rows + self.reverse_action[action], pair_index + k] = 1
                self.str[pair_index] = self.mapping[action][1]

    def find_next_unfilled(self):
        index = self.index
        while index < self.target.len - 1:
            if self.str[index] == '-':
                break
            index += 1
        self.index = index

    def get_state(self, reshape=False):
        i, k = self.index, self.kernel_size
        state = self.encoding[:, i:i + 2 * k]
        return state.flatten() if reshape else np.expand_dims(state, axis=2)

    # ---- solution.py:125-155
    def evaluate(self, string=None, permute=False, compute_statistics=False, boost=False, reward=False, verbose=False):
        if boost:
            raise NotImplementedError("boosting is not part of this path")
        if string is None:
            string = self.string
        tgt = self.target.seq
        if permute:
            cfg = {k: self.config.get(k, v) for k, v in _PERMUTE_DEFAULTS.items()}
            res = permute_batch([string], [tgt], cfg, seed=self.permute_seed, reward_exp=self.reward_exp)
            if res.rounds[0] > 0:
                self.str = list(res.sequences[0])
                if res.hd[0] == 0:
                    self.source = 'rlif*'
            self.folded_structure, self.fe, self.hd = res.structures[0], float(res.mfe[0]), int(res.hd[0])
            self.mismatch_indices = np.where([a != b for a, b in zip(tgt, self.folded_structure)])[0]
        else:
            structs, mfe, hd, mism, _ = fold_hamming_batch([string], [tgt], reward_exp=self.reward_exp)
            self.folded_structure, self.fe, self.hd, self.mismatch_indices = structs[0], float(mfe[0]), int(hd[0]), mism[0]
        if compute_statistics:
            self.compute_statistics()
        if reward:
            self.r = (1 - float(self.hd) / float(self.target.len)) ** self.reward_exp
            gcau = self.gcau_content()
            if self.hd == 0:
                self.r += .25
            if self.config_check('gc_penalty') or self.config_check('gc'):
                self.r = self.r * 0.85 + np.clip(gcau['U'], 0, .15)
        if self.start is not None:
            self.time = t.time() - self.start
        return self.r, self.hd, self.mismatch_indices

    def permute(self, original_mismatch_indices=None, verbose=True):
        """solution.py:157-219 on the GPU: -> (best sequence as a list of characters, its Hamming distance)."""
        cfg = {k: self.config.get(k, v) for k, v in _PERMUTE_DEFAULTS.items()}
        cfg['permutation_threshold'] = max(int(cfg['permutation_threshold']), self.target.len)   # permute() itself has no gate
        res = permute_batch([self.string], [self.target.seq], cfg, seed=self.permute_seed, reward_exp=self.reward_exp)
        if res.hd[0] == 0 and res.rounds[0] > 0:
            self.source = 'rlif*'
        return list(res.sequences[0]), int(res.hd[0])

    # ---- solution.py:221-238 (and init_vars, :240-255)
    def init_vars(self):
        for name in ('probability', 'positional_entropy', 'partition_prob', 'partition_fn', 'centroid_structure', 'centroid_dist',
                     'centroid_en', 'MEA_structure', 'MEA', 'MEA_en', 'ensemble_diversity', 'ensemble_defect'):
            setattr(self, name, None)

    def compute_statistics(self):
        """Partition function, centroid, ensemble defect / diversity, mountain distance of the current string on the GPU
        (rnafold_ensemble_batch), with the centroid's and the MEA structure's energies (fc.eval_structure) in kcal/mol."""
        from .ensemble import ensemble_batch
        res = ensemble_batch([self.string], targets=[self.target.seq])
        self.partition_prob, self.partition_fn = res.propensity[0], float(res.ensemble_energy[0])
        self.positional_entropy = res.positional_entropy[0]
        self.centroid_structure, self.centroid_dist = res.centroid_structure[0], float(res.centroid_dist[0])
        self.probability = float(res.probability[0])
        self.centroid_en = float(res.centroid_en[0])
        self.MEA_structure, self.MEA, self.MEA_en = res.MEA_structure[0], float(res.MEA[0]), float(res.MEA_en[0])
        self.ensemble_diversity = float(res.ensemble_diversity[0])
        self.ensemble_defect = float(res.ensemble_defect[0])
        self.md = float(res.md[0])

    def gcau_content(self):
        return gcau_content(self.str)

    def summary(self, print_output=False, colorize=False):
        lines = summary_lines(self)
        if print_output:
            print('\n'.join(lines))
        return lines

    def __repr__(self):
        return "<Solution hd={} fe={} {}>".format(self.hd, self.fe, self.string)
