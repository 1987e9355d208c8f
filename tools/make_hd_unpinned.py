#!/usr/bin/env python3
"""Record the golden rows whose logged Hamming distance the oracle does not reproduce under the parameter set of
their log file, with the outcome under EVERY parameter set -> tests/golden/hd_unpinned_rows.json.

These are the rows a live ViennaRNA 2.4.14 `RNA.fold` would have to settle: either the logged HD came from another
parameter set / code path (the sources are the RL agents and LEARNA, which ran in-process with their own settings),
or the oracle's backtrack breaks a tie differently.  The fixture keeps, per row and per set: the oracle's structure,
energy (dcal/mol), its HD to the target, whether the logged energy matches, and the energy of the TARGET structure
itself on that sequence (equal to the MFE <=> the target is a co-optimal structure the backtrack did not pick).
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
from conftest import load_golden_rows, param_file  # noqa: E402

SETS = ("turner2004", "turner1999", "andronescu2007", "langdon2018")


def main():
    P = {s: oracle.Params(param_file(s)) for s in SETS}
    out = []
    for r in load_golden_rows():
        s0, _ = P[r["params"]].fold_dcal(r["seq"])
        if sum(a != b for a, b in zip(s0, r["target"])) == r["hd"]:
            continue
        rec = {k: r[k] for k in ("file", "line", "source", "target", "seq", "hd", "fe", "params")}
        rec["per_set"] = {}
        for s in SETS:
            st, e = P[s].fold_dcal(r["seq"])
            try:
                et = P[s].eval_structure(r["seq"], r["target"])
            except Exception:
                et = None
            rec["per_set"][s] = {"structure": st, "dcal": e, "hd": sum(a != b for a, b in zip(st, r["target"])),
                                 "fe_matches": abs(e / 100.0 - float(r["fe"])) < 0.0051, "target_dcal": et,
                                 "target_is_cooptimal": et == e}
        out.append(rec)
    path = os.path.join(ROOT, "tests", "golden", "hd_unpinned_rows.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=0)
    any_set = sum(any(v["hd"] == r["hd"] and v["fe_matches"] for v in r["per_set"].values()) for r in out)
    coopt = sum(r["per_set"][r["params"]]["target_is_cooptimal"] for r in out)
    print(f"{len(out)} rows; {any_set} reproduce HD and energy under some other set; {coopt} have a co-optimal target -> {path}", file=sys.stderr)


if __name__ == "__main__":
    main()
