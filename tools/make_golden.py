#!/usr/bin/env python3
"""Extract the known-answer rows for RNA.fold that the reference ships in its logs.

The reference has no test-suite; the only artefacts that pin results of its fold
call (`RNA.fold` of ViennaRNA 2.4.14, bound at rlif/settings.py:63-64) are the CSV
logs written by rlif/utils/comparison.py:177-233 -- each row is
`seq,source,t,hd,md,fe,ed,attempts` under a header row that carries the target
dot-bracket.  `hd` is hamming_distance(target, RNA.fold(seq)[0]) and `fe` is
RNA.fold(seq)[1] printed with {:.3f} (then damaged by a spreadsheet round trip:
trailing commas, "-22" for "-22.000", extra 0/1 columns).

Run in the build container (needs /root/reference); writes
tests/golden/vienna_logged_rows.jsonl which travels with the repo.
"""
import json
import os
import sys

REF = os.environ.get("RNAGENT_REFERENCE", "/root/reference")
FILES = [
    "rlif/logs/eterna/solutions.csv",
    "rlif/logs/eterna/solutions2.csv",
    "rlif/logs/rfam_runge/solutions.csv",
    "rlif/logs/rfam_modena/solutions.csv",
]
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden",
                   "vienna_logged_rows.jsonl")


def is_db(s):
    return len(s) > 0 and set(s) <= set("().")


def is_seq(s):
    return len(s) > 0 and set(s) <= set("ACGU")


def main():
    rows = []
    for rel in FILES:
        target = None
        with open(os.path.join(REF, rel)) as f:
            for lineno, line in enumerate(f, 1):
                cols = [c.strip() for c in line.rstrip("\n").split(",")]
                if not cols or not cols[0]:
                    continue
                if is_db(cols[0]):
                    target = cols[0]
                    continue
                if cols[0] == "seq" or not is_seq(cols[0]):
                    continue
                if len(cols) < 6 or cols[3] == "" or cols[5] == "":
                    continue
                seq = cols[0]
                if target is None or len(target) != len(seq):
                    continue
                rows.append({
                    "file": rel, "line": lineno, "source": cols[1], "target": target,
                    "seq": seq, "hd": int(float(cols[3])), "fe": cols[5],
                    # mountain distance and ensemble defect of the row's solution (Solution.compute_statistics,
                    # rlif/rna/solution.py:221-238, written by rlif/utils/comparison.py:225-233): the fixture for row N3
                    "md": cols[4], "ed": cols[6] if len(cols) > 6 else "",
                })
    with open(OUT, "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")
    n0 = sum(1 for r in rows if r["hd"] == 0)
    print(f"{len(rows)} rows ({n0} with hd=0) -> {os.path.normpath(OUT)}", file=sys.stderr)


if __name__ == "__main__":
    main()
