#!/usr/bin/env python3
"""Re-serialise the reference's energy parameter files into the 37 C-only files the package ships.

Input : /root/reference/rlif/utils/parameters/rna_*.par ("RNAfold parameter file v2.0", the data that
        rlif/rna/vienna.py:7-19 hands to RNA.read_parameter_file).
Output: rnagent_b200/data/<name>_37.par -- same v2.0 grammar, but only the free-energy sections
        (every *_enthalpies section is unused at 37 C and is dropped, as are all comments), values
        re-flowed 10 per line, special-loop enthalpies written as 0.  The product's C++ loader and
        the oracle's C loader both read this grammar, so one parser path serves shipped and
        user-supplied files.

Run in the build container only (needs /root/reference).
"""
import os
import re
import sys

REF = os.environ.get("RNAGENT_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "rlif", "utils", "parameters")
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "rnagent_b200", "data")
ORDER = ["stack", "mismatch_hairpin", "mismatch_interior", "mismatch_interior_1n", "mismatch_interior_23",
         "mismatch_multi", "mismatch_exterior", "dangle5", "dangle3", "int11", "int21", "int22",
         "hairpin", "bulge", "interior", "ML_params", "NINIO", "Misc", "Triloops", "Tetraloops", "Hexaloops"]


def parse(path):
    text = re.sub(r"/\*.*?\*/", " ", open(path).read(), flags=re.S)
    sections, name = {}, None
    for line in text.splitlines():
        if line.startswith("#"):
            name = line.lstrip("#").strip().split(" ")[0] if not line.startswith("##") else None
            if name == "END":
                break
            if name:
                sections[name] = []
        elif name:
            sections[name].extend(line.split())
    return sections


def main():
    os.makedirs(DST, exist_ok=True)
    for fn in sorted(os.listdir(SRC)):
        if not fn.endswith(".par"):
            continue
        sec = parse(os.path.join(SRC, fn))
        out = ["## RNAfold parameter file v2.0", ""]
        for name in ORDER:
            toks = sec.get(name)
            if toks is None:
                continue
            out.append("# " + name)
            if name.endswith("loops"):
                for k in range(0, len(toks) - 2, 3):
                    out.append(f"{toks[k]} {toks[k + 1]} 0")
            else:
                for k in range(0, len(toks), 10):
                    out.append(" ".join(toks[k:k + 10]))
            out.append("")
        out.append("#END")
        dst = os.path.join(DST, fn.replace("rna_", "").replace(".par", "_37.par"))
        with open(dst, "w") as f:
            f.write("\n".join(out) + "\n")
        print(f"{fn} -> {os.path.normpath(dst)} ({os.path.getsize(dst)} bytes)", file=sys.stderr)


if __name__ == "__main__":
    main()
