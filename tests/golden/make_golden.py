#!/usr/bin/env python3
"""Generates the golden fixtures of tests/golden/ by running the UNMODIFIED reference binaries
(oracle/_ref/lara_scalar = default CMake semantics, oracle/_ref/lara_simd = -mavx2 / SEQAN_SIMD_ENABLED build,
both built from /root/reference by oracle/build_ref.sh) on small seeded synthetic inputs.

Run once in the build container (needs /root/reference to have been compiled); the inputs and the reference's
outputs are committed, so the tests never read /root/reference.

    python tests/golden/make_golden.py
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import gen_inputs  # noqa: E402

REF = {"scalar": os.path.join(ROOT, "oracle", "_ref", "lara_scalar"),
       "simd": os.path.join(ROOT, "oracle", "_ref", "lara_simd")}

# name -> (generator arguments, list of cases); a case = (tag, build, extra lara options)
INPUTS = {
    "dbn8x70": dict(gen=dict(seed=101, n=8, length=60, fmt="dbn", lenmax=80, related=False),
                    cases=[("default", "scalar", []),
                           ("pairs", "scalar", ["-o", "pairs"]),
                           ("fasta", "scalar", ["-o", "fasta"]),
                           ("simd", "simd", []),
                           ("m3", "scalar", ["-m", "3", "-n", "120"])]),
    "ebp7x60": dict(gen=dict(seed=102, n=7, length=50, fmt="ebpseq", lenmax=75),
                    cases=[("n80", "scalar", ["-n", "80", "-a", "10"]),
                           ("n80_simd", "simd", ["-n", "80", "-a", "10"]),
                           ("n60_pairs", "scalar", ["-n", "60", "-o", "pairs"]),
                           ("lin", "scalar", ["-n", "60", "-l", "100", "900"]),
                           ("opts", "scalar", ["-n", "60", "-b", "0.7", "-c", "1.5", "-f", "1.7", "-u", "25",
                                               "-x", "-1.5", "-y", "-5"])]),
    "ebp5x130": dict(gen=dict(seed=103, n=5, length=120, fmt="ebpseq", lenmax=140),
                     cases=[("default", "scalar", []),
                            ("simd", "simd", ["-n", "150"])]),
    "ebp4unrel": dict(gen=dict(seed=104, n=4, length=40, fmt="ebpseq", lenmax=110, related=False),
                      cases=[("n100", "scalar", ["-n", "100"])]),
    "dp4x90": dict(gen=dict(seed=105, n=4, length=80, fmt="dp", lenmax=100),
                   cases=[("n100", "scalar", ["-n", "100"]),
                          ("n100_pairs_simd", "simd", ["-n", "100", "-o", "pairs"])]),
    # fixed structures in the two other SeqAn structure formats LaRA can read a sequence from
    "bpseq5x60": dict(gen=dict(seed=106, n=5, length=50, fmt="bpseq", lenmax=70, related=False),
                      cases=[("default", "scalar", []),
                             # -r: a second input file whose records are appended (src/io.hpp:78-79)
                             ("ref_ct", "scalar", ["-n", "100"], "ct4x80/input.ct")]),
    "ct4x80": dict(gen=dict(seed=107, n=4, length=70, fmt="ct", lenmax=90, related=False),
                   cases=[("default", "scalar", []),
                          ("simd_pairs", "simd", ["-o", "pairs"])]),
    # gap costs outside the common case: open == extend (SeqAn switches to Needleman-Wunsch, linear gap costs) and
    # open cheaper than extend
    "ebp6gaps": dict(gen=dict(seed=109, n=6, length=50, fmt="ebpseq", lenmax=75),
                     cases=[("linear", "scalar", ["-n", "80", "-x", "-3", "-y", "-3"]),
                            ("linear_simd", "simd", ["-n", "80", "-x", "-3", "-y", "-3"]),
                            ("open_cheaper", "scalar", ["-n", "80", "-x", "-4", "-y", "-2"]),
                            ("open_cheaper_simd_pairs", "simd", ["-n", "80", "-x", "-4", "-y", "-2", "-o", "pairs"])]),
    # -s: score matrix file read by SeqAn's loadScoreMatrix (src/parameters.hpp:306-323)
    "ebp6matrix": dict(gen=dict(seed=108, n=6, length=55, fmt="ebpseq", lenmax=70),
                       cases=[("s_matrix", "scalar", ["-n", "80", "-s", "@golden/matrix_test.txt"]),
                              ("s_matrix_simd", "simd", ["-n", "80", "-s", "@golden/matrix_test.txt", "-c", "1.0"])]),
}

MATRIX_TEST = """# test matrix for -s (SeqAn score matrix file format): header lines, alphabet line, one row per letter
#  asymmetric on purpose, N row short on purpose (missing cells stay 0)
   A     C     G     U     N
A  2.05 -1.50 -0.75 -1.25 -1.00
C -1.40  1.85 -2.00 -0.60 -1.00
G -0.80 -2.10  1.60 -1.10 -1.00
U -1.30 -0.55 -1.05  1.95 -1.00
N -1.00 -1.00 -1.00
"""


def main():
    for b in REF.values():
        if not os.path.exists(b):
            raise SystemExit("missing %s -- run oracle/build_ref.sh (needs /root/reference)" % b)
    manifest = []
    with open(os.path.join(HERE, "matrix_test.txt"), "w") as f:
        f.write(MATRIX_TEST)
    generated = {}
    for name, spec in INPUTS.items():
        d = os.path.join(HERE, name)
        if os.path.isdir(d):
            shutil.rmtree(d)
        g = spec["gen"]
        generated[name] = gen_inputs.generate(d, g["seed"], g["n"], g["length"], g["fmt"], lenmax=g.get("lenmax", 0),
                                              related=g.get("related", True))
    for name, spec in INPUTS.items():
        d = os.path.join(HERE, name)
        g = spec["gen"]
        files = generated[name]
        rel = [os.path.relpath(f, HERE) for f in files]
        for case in spec["cases"]:
            tag, build, opts = case[:3]
            ref_file = case[3] if len(case) > 3 else None
            out = os.path.join(d, "ref_%s.out" % tag)
            if g["fmt"] == "dp":
                inargs = []
                for f in files:
                    inargs += ["-d", f]
            else:
                inargs = ["-i", files[0]]
            if ref_file:
                inargs += ["-r", os.path.join(HERE, ref_file)]
            run_opts = [os.path.join(HERE, o[len("@golden/"):]) if o.startswith("@golden/") else o for o in opts]
            cmd = [REF[build]] + inargs + run_opts + ["-j", "4", "-w", out]
            cp = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            if cp.returncode != 0:
                raise SystemExit("reference failed: %s\n%s" % (" ".join(cmd), cp.stderr.decode()[-800:]))
            entry = {"input": name, "files": rel, "format": g["fmt"], "tag": tag, "build": build,
                     "options": opts, "output": os.path.relpath(out, HERE),
                     "bytes": os.path.getsize(out)}
            if ref_file:
                entry["ref_file"] = ref_file
            manifest.append(entry)
            print("%-10s %-16s %-6s %7d bytes" % (name, tag, build, os.path.getsize(out)))
    with open(os.path.join(HERE, "manifest.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py",
                   "reference": "seqan/lara @ f94301fd built by oracle/build_ref.sh (g++ -O3, scalar and -mavx2)",
                   "cases": manifest}, f, indent=1)
    print("wrote manifest with %d cases" % len(manifest))


if __name__ == "__main__":
    main()
