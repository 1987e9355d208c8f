"""Host-side helpers that keep the reference's signatures (rlif/rna/utils.py)."""
import numpy as np


def hamming_distance(seq1, seq2):
    """rlif/rna/utils.py:67-75: zip-truncating Hamming distance -> (hd, mismatch indices).

    Host-side twin of the distance the fold kernel fuses; used for single strings only.
    """
    a = np.frombuffer(seq1.encode() if isinstance(seq1, str) else bytes(seq1), dtype=np.uint8)
    b = np.frombuffer(seq2.encode() if isinstance(seq2, str) else bytes(seq2), dtype=np.uint8)
    n = min(len(a), len(b))
    mism = a[:n] != b[:n]
    return int(mism.sum()), np.where(mism)[0]


# ---- data formats on either side of the path ----------------------------

def load_sequence(num, dataset='rfam_learn', encoding_type=2, data_dir=None):
    """rlif/rna/utils.py:115-129: `<data_dir>/<dataset>/<num>.rna` (first line = dot-bracket target) -> DotBracket.

    The reference strips the last character of the first line (`readline()[:-1]`); files there end with a newline.
    A missing file returns [] like the reference's except-branch intends.  `data_dir` replaces settings.DATA.
    """
    import os
    from .dotbracket import DotBracket
    filename = os.path.join(data_dir or os.environ.get("RLIF_DATA", "."), dataset, '{}.rna'.format(num))
    try:
        with open(filename, 'r') as f:
            line = f.readline()
    except OSError:
        return []
    seq = line[:-1] if line.endswith('\n') else line
    return DotBracket(seq.strip('\r'), filename, num, encoding_type=encoding_type)


def load_dataset(dataset, start=1, n_seqs=100, encoding_type=2, data_dir=None):
    """Dataset(dataset, start=, n_seqs=) of rlif/rna/dataset.py:46-49: consecutive `<num>.rna` targets."""
    out = [load_sequence(k, dataset=dataset, encoding_type=encoding_type, data_dir=data_dir) for k in range(start, start + n_seqs)]
    return [t for t in out if not isinstance(t, list)]


def read_fasta_records(filename):
    """All (name, sequence) records of a FASTA file, T -> U as the reference does before folding."""
    records, name, seq = [], None, []
    with open(filename, 'r') as f:
        for line in f:
            line = line.rstrip('\r\n')
            if line.startswith('>'):
                if name is not None:
                    records.append((name, ''.join(seq).replace('T', 'U')))
                name, seq = line[1:], []
            elif line:
                seq.append(line)
    if name is not None:
        records.append((name, ''.join(seq).replace('T', 'U')))
    return records


def load_fasta(filename, config=None, reference_quirks=False):
    """rlif/rna/utils.py:9-35: FASTA -> one evaluated solution per record; the target of each record is the fold
    of its own sequence.  All records are folded by ONE batched GPU call instead of one RNA.fold each.

    Returns a list of dicts(name, string, target (DotBracket), folded_structure, fe, hd=0).
    reference_quirks=True reproduces the reference loop literally: a record is emitted only when the NEXT header
    is read, carries that next header's name, and the last record of the file is dropped.
    """
    from .dotbracket import DotBracket
    from .vienna import fold_batch
    recs = read_fasta_records(filename)
    if reference_quirks:
        recs = [(recs[k + 1][0], recs[k][1]) for k in range(len(recs) - 1) if recs[k][1] != '']
    if not recs:
        return []
    structs, fe = fold_batch([s for _, s in recs])
    out = []
    for (name, s), db, e in zip(recs, structs, fe):
        target = DotBracket(db)
        target.name = name
        target.nucleotides = s
        out.append(dict(name=name, string=s, target=target, folded_structure=db, fe=float(e), hd=0))
    return out


def write_fasta(filename, sequences, write_dot_brackets=False, linebreak=80):
    """rlif/rna/utils.py:38-63 as documented there (`>Name`, the nucleotides in lines of `linebreak`, optionally the
    target dot-bracket = Vienna format); the reference body itself raises (`readlines().count()`, undefined name).
    `sequences`: objects or dicts with `string` and `target` (DotBracket or str; `name` optional)."""
    def get(o, k, default=None):
        return o.get(k, default) if isinstance(o, dict) else getattr(o, k, default)
    with open(filename, 'a+') as fasta:
        fasta.seek(0)
        if len(fasta.readlines()) > 1:
            fasta.write('\n')
        for n, sol in enumerate(sequences):
            target = get(sol, 'target')
            name = get(sol, 'name') or getattr(target, 'name', '') or 'solution_{}'.format(n)
            string = get(sol, 'string')
            fasta.write('>' + name + '\n')
            for i in range(0, len(string), linebreak):
                fasta.write(string[i:i + linebreak] + '\n')
            if write_dot_brackets:
                db = target.seq if hasattr(target, 'seq') else str(target)
                for i in range(0, len(db), linebreak):
                    fasta.write(db[i:i + linebreak] + '\n')
