#!/usr/bin/env python3
"""Developer tool: per-source-line stall samples / executed instructions of one kernel.

Joins `ncu -i X.ncu-rep --page source --csv` (SASS rows, in address order) with the line table of the matching
cubin (`nvdisasm -g -c`), by instruction order inside the function.

usage: ncu_lines.py <source_page.csv> <cubin> <function substring> [top N]
"""
import csv
import collections
import re
import subprocess
import sys


def line_table(cubin, func):
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], stdout=subprocess.PIPE, text=True).stdout.split("\n")
    out = []
    inside = False
    cur = None
    for ln in dis:
        if ln.startswith(".text."):
            inside = func in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            out.append((cur, ln.split("*/", 1)[1].strip()))
    return out


def main():
    src_csv, cubin, func = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src_csv)))
    hdr = None
    inst = []
    cur = None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = [] if func in r[1] or func.split("IL")[0] in r[1] else None
            if cur is not None:
                inst.append(cur)
            continue
        if r and r[0] == "Address":
            hdr = r
            continue
        if cur is not None and hdr and len(r) >= len(hdr) - 2:
            cur.append(r)
    data = inst[0]
    table = line_table(cubin, func)
    print("sass rows in profile: %d, in cubin: %d" % (len(data), len(table)))
    si = hdr.index("# Samples")
    ie = hdr.index("Instructions Executed")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    for k, r in enumerate(data):
        key = table[k][0] if k < len(table) else None
        a = agg[key]
        a[0] += int(r[si])
        a[1] += int(r[ie])
        for h in stalls:
            v = int(r[hdr.index(h)])
            if v:
                a[2][h[6:]] += v
    tot_s = sum(a[0] for a in agg.values())
    tot_i = sum(a[1] for a in agg.values())
    print("total samples %d, warp instructions %d" % (tot_s, tot_i))
    srcs = {}
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        text = ""
        if key:
            f = key[0]
            if f not in srcs:
                try:
                    srcs[f] = open("/root/repo/lara_b200/csrc/" + f).read().split("\n")
                except Exception:
                    srcs[f] = []
            if key[1] - 1 < len(srcs[f]):
                text = srcs[f][key[1] - 1].strip()[:70]
        print("%-24s samples %6d (%4.1f%%) inst %9d (%4.1f%%) %s | %s" % (
            "%s:%d" % key if key else "?", a[0], 100.0 * a[0] / tot_s, a[1], 100.0 * a[1] / tot_i,
            dict(a[2].most_common(3)), text))


if __name__ == "__main__":
    main()
