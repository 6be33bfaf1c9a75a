#!/usr/bin/env python3
"""Developer tool (GPU box): the COMPLETE BASELINE config C5 (2,000 synthetic RNAs of 50..600 nt, 1,999,000 pairs, `-o pairs`)
through lara_b200_cli on one GPU, timed; then the size-independent check: the records of the pairs among the first 300
sequences, cut out of the complete output in order, must be byte for byte the output of the reference binary on those
300 sequences (tests/golden/configs.json, case c5_300_pairs: pairs are independent, the writer prints them in (i, j)
order).  usage: tools/c5_full.py [nseq]   (0 / none = all 2,000)"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_config_md5 as mk  # noqa: E402


def names_of(path, n):
    out = []
    with open(path) as f:
        for ln in f:
            if ln.startswith("## S"):                   # extended bpseq header: "## S<k>: <name>"
                out.append(ln.split(":", 1)[1].strip())
                if len(out) == n:
                    break
    return out


def main():
    nseq = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    cases = json.load(open(os.path.join(ROOT, "tests", "golden", "configs.json")))["cases"]
    gold = cases["c5_300_pairs"]
    d = tempfile.mkdtemp(prefix="lara_c5full_")
    argv, files, n = mk.make_inputs({"config": "c5", "nseq": nseq}, d)
    out = os.path.join(d, "gpu.pairs")
    cmd = [os.path.join(ROOT, "lara_b200", "lara_b200_cli")] + argv + ["-o", "pairs", "-w", out, "-v", "1"]
    t0 = time.time()
    cp = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=sys.stderr)     # stderr streams (LARA_B200_VERBOSE=1 shows progress)
    wall = time.time() - t0
    if cp.returncode != 0:
        print(json.dumps({"error": "cli rc %d" % cp.returncode}))
        return 1
    pairs = n * (n - 1) // 2
    # the first 300 sequences of the generated file are the 300 the reference was run on
    sub_argv, sub_files, _ = mk.make_inputs({"config": "c5", "nseq": 300}, tempfile.mkdtemp(prefix="lara_c5sub_"))
    assert mk.input_digest(sub_files) == gold["input_sha256"], "generator drift"
    first = set(names_of(sub_files[0], 300))
    assert len(first) == 300
    h = hashlib.sha256()
    kept = 0
    nbytes = 0
    total_records = 0
    with open(out, "rb") as f:
        while True:
            head = f.readline()
            if not head:
                break
            a = f.readline()
            b = f.readline()
            total_records += 1
            txt = head.decode()
            left, rest = txt[1:].split(" && ", 1)
            right = rest.rsplit(" (score", 1)[0]
            if left in first and right in first:
                for part in (head, a, b):
                    h.update(part)
                    nbytes += len(part)
                kept += 1
    res = {"workload": "c5 complete" if not nseq else "c5 first %d" % nseq, "n_sequences": n, "pairs": pairs,
           "records_written": total_records, "output_bytes": os.path.getsize(out), "cli_wall_s": round(wall, 2),
           "pairs_per_s_wall": round(pairs / wall, 1),
           "subset_check": {"pairs_among_first_300": kept, "bytes": nbytes, "sha256": h.hexdigest(),
                            "reference_sha256": gold["output_sha256"], "reference_bytes": gold["output_bytes"],
                            "identical": h.hexdigest() == gold["output_sha256"] and nbytes == gold["output_bytes"]}}
    print(json.dumps(res))
    os.remove(out)
    return 0 if res["subset_check"]["identical"] and total_records == pairs else 2


if __name__ == "__main__":
    sys.exit(main())
