#!/usr/bin/env python3
"""Developer tool (GPU box): one BASELINE.json config (or its first NSEQ records) through the GPU CLI
(lara_b200/lara_b200_cli) and through the unmodified reference binary (oracle/_ref/lara_scalar), same files, same
options; compares the written outputs byte for byte and prints both wall times.

usage: config_check.py <c1..c5> [nseq] [--no-ref] [extra lara options ...]
"""
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_inputs  # noqa: E402
import bench  # noqa: E402


def main():
    cfg = sys.argv[1]
    rest = sys.argv[2:]
    nseq = 0
    if rest and rest[0].isdigit():
        nseq = int(rest[0]); rest = rest[1:]
    no_ref = "--no-ref" in rest
    extra = [x for x in rest if x != "--no-ref"]
    c = gen_inputs.CONFIGS[cfg]
    d = tempfile.mkdtemp(prefix="lara_cfg_%s_" % cfg)
    files = gen_inputs.generate(d, c["seed"], c["n"], c["length"], c["fmt"], skew=(cfg == "c5"))
    n = c["n"]
    if c["fmt"] == "dp":
        if nseq:
            files = files[:nseq]; n = nseq
        inargs = []
        for f in files:
            inargs += ["-d", f]
    else:
        f = files[0]
        if nseq and nseq < n:
            f = bench.subset_file(f, nseq, os.path.join(d, "subset." + c["fmt"])); n = nseq
        inargs = ["-i", f]
    fmt = ["-o", "pairs"] if cfg == "c5" else []
    out_gpu, out_ref = os.path.join(d, "gpu.out"), os.path.join(d, "ref.out")
    t0 = time.perf_counter()
    cp = subprocess.run([os.path.join(ROOT, "lara_b200", "lara_b200_cli")] + inargs + fmt + extra + ["-w", out_gpu, "-v", "1"],
                        stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    tg = time.perf_counter() - t0
    print("config %s: %d sequences, %d pairs" % (cfg, n, n * (n - 1) // 2))
    print("GPU CLI rc=%d wall %.2f s, %.0f pairs/s" % (cp.returncode, tg, n * (n - 1) / 2 / tg))
    err = cp.stderr.decode()
    keep = 100000 if os.environ.get("LARA_B200_VERBOSE") else 8          # verbose runs: the whole stage listing
    print("\n".join("   " + ln for ln in err.strip().split("\n")[-keep:]))
    if cp.returncode != 0 or no_ref:
        return cp.returncode
    ref = os.path.join(ROOT, "oracle", "_ref", "lara_scalar")
    t0 = time.perf_counter()
    cr = subprocess.run([ref] + inargs + fmt + extra + ["-w", out_ref, "-j", str(os.cpu_count())],
                        stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    tr = time.perf_counter() - t0
    print("reference rc=%d wall %.2f s (-j %d), %.1f pairs/s" % (cr.returncode, tr, os.cpu_count(), n * (n - 1) / 2 / tr))
    if cr.returncode != 0:
        print(cr.stderr.decode()[-1000:])
        return 2
    A, B = open(out_gpu, "rb").read(), open(out_ref, "rb").read()
    print("outputs %s (%d vs %d bytes)" % ("IDENTICAL" if A == B else "DIFFER", len(A), len(B)))
    if A != B:
        la, lb = A.decode().splitlines(), B.decode().splitlines()
        k = 0
        for i, (x, y) in enumerate(zip(la, lb)):
            if x != y:
                print("  line %d: gpu %r ref %r" % (i + 1, x[:70], y[:70]))
                k += 1
                if k > 6:
                    break
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
