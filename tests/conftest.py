import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
DATA = os.path.join(ROOT, "rnagent_b200", "data")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def param_file(name):
    return os.path.join(DATA, f"{name}_37.par")


# Which parameter set was live when each reference log was written (established by folding every
# row under all four sets with the oracle: the split is clean per file; see DESIGN.md "Parity").
LOG_PARAMS = {
    "rlif/logs/eterna/solutions.csv": "turner2004",
    "rlif/logs/rfam_modena/solutions.csv": "turner2004",
    "rlif/logs/eterna/solutions2.csv": "turner1999",
    "rlif/logs/rfam_runge/solutions.csv": "turner1999",
}
# Sources whose logged HD was not produced by the committed Solution(string=...) path under the
# file's parameter set (RL agents switch parameter sets per model; LEARNA runs in-process):
# their energies are pinned, their HD is not reproducible under ANY single parameter set.
HD_UNPINNED_SOURCES = {"rlfold", "mrlfold", "LEARNA"}


def load_golden_rows():
    rows = []
    with open(os.path.join(GOLDEN, "vienna_logged_rows.jsonl")) as f:
        for line in f:
            r = json.loads(line)
            r["params"] = LOG_PARAMS[r["file"]]
            r["source"] = r["source"].strip()
            rows.append(r)
    return rows


@pytest.fixture(scope="session")
def golden_rows():
    return load_golden_rows()


@pytest.fixture(scope="session")
def oracle_params():
    import oracle
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = oracle.Params(param_file(name))
        return cache[name]
    return get


@pytest.fixture(scope="session")
def golden_folds(golden_rows, oracle_params):
    """Oracle (structure, dcal) of every golden row, folded once per session (multi-threaded)."""
    import numpy as np
    out = [None] * len(golden_rows)
    for name in sorted(set(r["params"] for r in golden_rows)):
        idx = [k for k, r in enumerate(golden_rows) if r["params"] == name]
        seqs = [golden_rows[k]["seq"] for k in idx]
        buf = np.frombuffer("".join(seqs).encode(), dtype=np.uint8)
        off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int32)
        db, mfe, _ = oracle_params(name).fold_batch(buf, off, nthreads=0)
        raw = db.tobytes().decode()
        for n, k in enumerate(idx):
            out[k] = (raw[off[n]:off[n + 1]], int(mfe[n]))
    return out
