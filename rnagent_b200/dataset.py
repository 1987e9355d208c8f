"""Benchmark target sets as the reference's rollout code handles them (rlif/rna/dataset.py:8-52).

`Dataset(dataset, start=1, n_seqs=100)` loads the consecutive one-line dot-bracket files `<data_dir>/<dataset>/<n>.rna`
(rlif/rna/utils.py:115-129) into `DotBracket` targets that carry their file number (`file_nr`, what
rlif/learning/tester.py:91 uses to index its per-target statistics); `Dataset(name, sequences=[...])` wraps targets
loaded elsewhere (dataset.py:44-45).  Length-indexed metadata and FASTA-backed sets are not part of this path.
"""
import os

from .dotbracket import DotBracket
from .utils import load_sequence


class Dataset(object):
    def __init__(self, dataset, length=None, start=1, n_seqs=100, sequences=None, encoding_type=2, data_dir=None):
        if length is not None:
            raise NotImplementedError("length-indexed metadata (rlif/data/metadata) is not part of this path")
        self.dataset = dataset
        self.path = os.path.join(data_dir or os.environ.get("RLIF_DATA", "."), dataset)
        self.sequence_length = length
        if sequences is not None:
            self.sequences = [s if isinstance(s, DotBracket) else DotBracket(str(s), None, k + 1, encoding_type=encoding_type)
                              for k, s in enumerate(sequences)]
        else:
            loaded = [load_sequence(k, dataset=dataset, encoding_type=encoding_type, data_dir=data_dir)
                      for k in range(start, start + n_seqs)]
            self.sequences = [t for t in loaded if isinstance(t, DotBracket)]
        self.n_seqs = len(self.sequences)

    def __getitem__(self, index):
        return self.sequences[index]

    def __len__(self):
        return self.n_seqs

    def statistics(self):
        lens = [t.len for t in self.sequences] or [0]
        return "Dataset {}: {} sequences, length {}..{}".format(self.dataset, self.n_seqs, min(lens), max(lens))

    __repr__ = statistics
