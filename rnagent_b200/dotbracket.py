"""Target structures as the reference's environments see them (rlif/rna/dotbracket.py:4-211).

`DotBracket` keeps the attributes the hot path reads -- seq, len, base_pair_indices,
rev_base_pair_indices, struct_motifs, structure_encoding (encoding_type 2: one-hot rows O C I H M E),
percent_unpaired -- computed by the library's own parser (rnafold_parse_target, host-only C++), the
same code the batched environment uses on upload.  The forgi graph helpers are out of scope.
"""
import numpy as np

from . import _lib

MOTIFS = "OCIHME"


class DotBracket(object):
    def __init__(self, sequence, file_id=None, file_nr=None, encoding_type=2):
        if encoding_type != 2:
            raise NotImplementedError("only encoding_type 2 (the reference's RnaDesign.yml setting) is implemented")
        self.encoding_type = encoding_type
        self.file_nr = file_nr
        self.file_id = file_id
        self.seq = sequence
        self.len = len(sequence)
        self.name = ''
        raw = np.frombuffer(sequence.encode(), dtype=np.uint8)
        pair = np.zeros(self.len, dtype=np.int16)
        mot = np.zeros(self.len, dtype=np.uint8)
        _lib.check(_lib.load_library().rnafold_parse_target(raw.ctypes.data, self.len, pair.ctypes.data, mot.ctypes.data))
        self.base_pair_indices = {int(i): int(p) for i, p in enumerate(pair) if p >= 0}
        self.rev_base_pair_indices = {v: k for k, v in self.base_pair_indices.items()}
        self.struct_motifs = ''.join(MOTIFS[m] for m in mot)
        self.structure_encoding = np.zeros([6, self.len])
        self.structure_encoding[mot, np.arange(self.len)] = 1
        self.percent_unpaired = float(sequence.count('.')) / self.len if self.len else 0.0

    def __repr__(self):
        return self.seq
