#!/usr/bin/env python3
"""Benchmark of the hot path: LaRA 2's all-pairs Lagrangian subgradient loop on B200.

    python bench.py --gpus N --steps K --warmup W            (N=1: plain python; N>1: under torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W
    python bench.py --workload c4|c5|c2|c1 ...               (the other BASELINE.json configs; the default line is C3)

One "step" = one pass of the hot path (lara::solve, src/subgradient_solver.hpp:110) over the whole workload.  Default
workload: BASELINE.json configs[2] (C3, the config the north-star target is quoted on): 500 synthetic RNAs x 300 nt,
`.ebpseq` base pair weights, all 124,750 pairs, reference default options (500 iterations).  With N ranks the fixed
pair set is dealt round-robin from the cost-sorted list to the ranks (no collective on the data path; strong scaling).

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  value            pairs/s with the inputs resident in HBM when the timed region starts (device CUDA-event span)
  e2e              the same metric through the C ABI with HOST buffers (lara_gpu_set_sequences + lara_gpu_solve_pairs
                   + results copied to host memory), wall clock, every one of the --steps steps
  gcups_effective  the metric's GCUPS: sum over pairs and iterations of lenA*lenB / time; `gcups` is
                   the same number under its old name
  gcups_computed   DP cells the kernels really update per second: the band kernels fill only the diagonals that hold
                   active edges (exact: the rest of the matrix is closed forms), a tenth of the matrix on C3
  roofline         DP fill kernel gotoh_band_kernel (one launch per iteration step of a batch), HBM bound:
                   achieved = algorithmic bytes of the launches (4.5 B per band cell + 16 B closed-form record per row
                   and column) / summed CUDA-event duration of the launches, against MEASURED_PEAKS.json hbm_gbs;
                   `isolated` = the same from one extra single-worker step (no overlap with the other worker's kernels);
                   traffic = DRAM bytes per launch: ratio dram/algorithmic of one launch captured by ncu in this very run
                   (a short sub-process, see traffic_probe) x this run's algorithmic bytes per launch;
                   int32_issue = 9 algorithmic ops per band cell against the measured issue rate
  cpu_baseline     the UNMODIFIED reference binaries (oracle/_ref/lara_scalar and lara_simd, built from /root/reference
                   by oracle/build_ref.sh) on this box's host cores, on a bounded sample of the same workload
"""
import argparse
import csv
import io
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

OPS_PER_CELL = 9.0          # 5 IADD + 4 tie-aware maxima

# BASELINE.json configs; sample = what the CPU legs (cpu_baseline, --impl reference) run per step
WORKLOADS = {
    "c1": dict(seed=1, n=10, length=120, fmt="dbn", cfg=0, sample=10),
    "c2": dict(seed=2, n=100, length=150, fmt="ebpseq", cfg=1, sample=48),
    "c3": dict(seed=3, n=500, length=300, fmt="ebpseq", cfg=2, sample=48),
    "c4": dict(seed=4, n=50, length=1500, fmt="dp", cfg=3, sample=6),       # 50 dense RNAfold dot plots
    "c5": dict(seed=5, n=2000, length=0, fmt="ebpseq", cfg=4, sample=48),   # lengths 50..600, lognormal; -o pairs
}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def make_workload(name, nseq=0):
    import gen_inputs
    w = dict(WORKLOADS[name])
    d = tempfile.mkdtemp(prefix="lara_bench_%s_" % name)
    files = gen_inputs.generate(d, w["seed"], w["n"], w["length"], w["fmt"], skew=(name == "c5"))
    if nseq and nseq < w["n"]:
        files = first_records(files, w["fmt"], nseq, d)
        w["n"] = nseq
    return w, files


def first_records(files, fmt, m, d):
    if fmt == "dp":
        return files[:m]
    return [subset_file(files[0], m, os.path.join(d, "subset%d.%s" % (m, fmt)))]


def subset_file(path, m, out):
    """First m records of a generated .ebpseq / .dbn file, byte-identical per record."""
    lines = open(path).read().split("\n")
    res = []
    if path.endswith(".dbn"):
        res = lines[:3 * m]
    else:
        rec = 0
        for ln in lines:
            if ln.startswith("## S"):
                k = int(ln[4:ln.index(":")])
                if k <= m:
                    res.append(ln)
            elif ln.startswith("## "):
                res.append(ln)
            elif ln.startswith("# S"):
                rec = int(ln[3:])
                if rec <= m:
                    res.append(ln)
            elif rec <= m:
                res.append(ln)
    with open(out, "w") as f:
        f.write("\n".join(res).rstrip("\n") + "\n")
    return out


def input_args(files, fmt):
    if fmt == "dp":
        a = []
        for f in files:
            a += ["-d", f]
        return a
    return ["-i", files[0]]


def load_store(files, fmt):
    import lara_b200 as lb
    if fmt == "dp":
        return lb.InputStorage.from_files(dotplots=files)
    return lb.InputStorage.from_files(in_file=files[0])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([x.strip() for x in ln.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
            except Exception:
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for nm, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def run_reference_binary(in_args, threads, fmt="lib", out=None, binary=None):
    from oracle import binding as ob
    binary = binary or ob.REF_SCALAR
    t0 = time.perf_counter()
    cp = ob.run_reference(binary, list(in_args) + ["-j", str(threads), "-o", fmt, "-w", out])
    dt = time.perf_counter() - t0
    if cp.returncode != 0:
        raise RuntimeError("reference binary failed: " + cp.stderr.decode()[-500:])
    return dt, out


def cpu_sample(args, w, files):
    """-> (input arguments of the sample, number of sequences, pairs, description)"""
    d = os.path.dirname(files[0])
    m = min(w["sample"], w["n"])
    sample = first_records(files, w["fmt"], m, d) if m < w["n"] else files
    return input_args(sample, w["fmt"]), m, m * (m - 1) // 2


def out_format(workload):
    return "pairs" if workload == "c5" else "lib"


def reference_arm(args, rank, world):
    """Times the unmodified reference binary on the host cores: rank 0 only, bounded sample per step."""
    if rank != 0:
        return 0
    from oracle import binding as ob
    w, files = make_workload(args.workload, args.nseq)
    in_args, m, n_pairs = cpu_sample(args, w, files)
    cores = os.cpu_count()
    if not os.path.exists(ob.REF_SCALAR):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/lara_scalar missing (build on a box with /root/reference)"}))
        return 0
    out = os.path.join(os.path.dirname(files[0]), "ref.out")
    for _ in range(args.warmup):
        run_reference_binary(in_args, cores, out_format(args.workload), out)
    times = []
    for _ in range(args.steps):
        dt, _ = run_reference_binary(in_args, cores, out_format(args.workload), out)
        times.append(dt)
    t = sum(times)
    value = n_pairs * args.steps / t
    line = {
        "impl": "reference", "metric": "aligned pairs/s", "value": value, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": workload_config(args, w, world),
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": cores, "kind": "reference",
                         "sample": "first %d of the %d sequences (%d pairs, all iterations), oracle/_ref/lara_scalar -j %d, process wall "
                                   "clock incl. parse+write" % (m, w["n"], n_pairs, cores)},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_config(args, w, world):
    length = ("%d nt" % w["length"]) if w["length"] else "50..600 nt (lognormal)"
    return {"workload": "%s: %d synthetic RNAs x %s %s, all %d pairs, 500 iterations (BASELINE.json configs[%d])"
                        % (args.workload, w["n"], length, w["fmt"], w["n"] * (w["n"] - 1) // 2, w["cfg"]),
            "n_sequences": w["n"], "length": w["length"], "pairs": w["n"] * (w["n"] - 1) // 2,
            "num_iterations": 500, "partition": "round-robin over the cost-sorted pair list, %d rank(s)" % world,
            "l2": "per-batch working set (band scores, trace, interaction records: > 10 GB on C3) is larger than L2; no explicit flush"}


def traffic_probe(args):
    """One DP launch of this workload (first batch, every pair live) under ncu in a sub-process: DRAM bytes read +
    written against the algorithmic bytes of that same launch.  None when ncu is not available."""
    ncu = None
    for c in ("ncu", "/usr/local/cuda/bin/ncu"):
        try:
            if subprocess.run([c, "--version"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode == 0:
                ncu = c
                break
        except Exception:
            continue
    if not ncu:
        return None
    nseq = args.nseq or {"c3": 250, "c5": 340, "c2": 0, "c1": 0, "c4": 0}[args.workload]
    cmd = [ncu, "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum", "--clock-control", "none",
           "-k", "regex:gotoh_band_kernel", "-s", "4", "-c", "1", "--csv",
           sys.executable, os.path.abspath(__file__), "--probe", "--workload", args.workload, "--nseq", str(nseq)]
    try:
        cp = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600,
                            env=dict(os.environ, LARA_B200_WORKERS="1"))
    except Exception as ex:
        log("traffic probe failed: %r" % (ex,))
        return None
    probe = None
    rows = []
    for ln in cp.stdout.split("\n"):
        if ln.startswith("PROBE "):
            probe = json.loads(ln[6:])
        elif ln.startswith('"'):
            rows.append(ln)
    if not probe or len(rows) < 2:
        log("traffic probe: no capture (rc %d) %s" % (cp.returncode, cp.stderr[-300:]))
        return None
    tab = list(csv.reader(io.StringIO("\n".join(rows))))
    hdr = tab[0]
    vals = {}
    for r in tab[1:]:
        if len(r) != len(hdr):
            continue
        rec = dict(zip(hdr, r))
        try:
            v = float(rec["Metric Value"].replace(",", ""))
        except Exception:
            continue
        unit = rec.get("Metric Unit", "")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0,
                 "nsecond": 1e-3, "msecond": 1e3}.get(unit, 1.0)
        vals[rec["Metric Name"]] = v * scale
    if "dram__bytes_read.sum" not in vals:
        return None
    dram = vals["dram__bytes_read.sum"] + vals.get("dram__bytes_write.sum", 0.0)
    return {"dram_bytes": dram, "algorithmic_bytes": probe["launch_bytes"], "pairs_in_launch": probe["launch_pairs"],
            "ratio": dram / probe["launch_bytes"] if probe["launch_bytes"] else None,
            "duration_us_under_ncu": vals.get("gpu__time_duration.sum"),
            "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum -k gotoh_band_kernel -s 4 -c 1 over a one-step run of "
                   "the first %s sequences of the workload, one worker: the launch of iteration 5 of the first batch" % (nseq or "all")}


def probe_main(args):
    """Sub-process of traffic_probe (runs under ncu): one step, prints the algorithmic bytes of a full-live DP launch."""
    import lara_b200 as lb
    import numpy as np
    from lara_b200 import partition
    w, files = make_workload(args.workload, args.nseq)
    store = load_store(files, w["fmt"])
    ia, ib = partition.all_pairs(len(store))
    order = partition.cost_order(store.lengths(), ia, ib)
    p = lb.Parameters()
    p.set(num_iterations=8)
    ctx = lb.Context(p, device_ids=[0])
    ctx.set_sequences(store.flat())
    ctx.prepare_pairs(ia[order].astype(np.int32), ib[order].astype(np.int32))
    ctx.run_prepared()
    st = ctx.stats()
    print("PROBE " + json.dumps({"launch_bytes": int(st.reserved[10]), "launch_pairs": int(st.reserved[11])}), flush=True)
    ctx.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--nseq", type=int, default=0, help="developer switch: only the first NSEQ sequences")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-isolated", action="store_true", help="skip the extra single-worker step of the roofline")
    ap.add_argument("--no-traffic", action="store_true", help="skip the ncu sub-process that measures the DRAM bytes of one DP launch")
    ap.add_argument("--probe", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--emulate-world", type=int, default=0,
                    help="developer switch: run only shard 0 of K of the pair list on this GPU (the work one rank of a "
                         "K-GPU run does); value and gcups then count that shard only")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        return reference_arm(args, rank, world)
    if args.probe:
        return probe_main(args)

    import numpy as np
    import torch
    import lara_b200 as lb

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the hot path has no CPU fallback")
    multi_proc = world > 1
    if multi_proc:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        device_ids = [local_rank]
    else:
        torch.cuda.set_device(0)
        device_ids = list(range(max(1, args.gpus)))       # one process driving N GPUs through the C ABI (-j N)

    def barrier():
        if multi_proc:
            dist.barrier()
        for d in device_ids:
            torch.cuda.synchronize(d)

    def reduce_max(x):
        if not multi_proc:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x):
        if not multi_proc:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    w, files = make_workload(args.workload, args.nseq)
    store = load_store(files, w["fmt"])
    flat = store.flat()
    n = len(store)
    lens = store.lengths()
    from lara_b200 import partition
    ia, ib = partition.all_pairs(n)
    # cost-sorted (longest first) global order, dealt round-robin to the ranks
    order = partition.cost_order(lens, ia, ib)
    mine = partition.shard(order, rank, world) if multi_proc else order
    if args.emulate_world > 1 and not multi_proc:
        mine = partition.shard(order, 0, args.emulate_world)
    my_a, my_b = ia[mine].astype(np.int32), ib[mine].astype(np.int32)
    n_pairs_total = len(mine) if (args.emulate_world > 1 and not multi_proc) else len(ia)

    params = lb.Parameters()
    ctx = lb.Context(params, device_ids=device_ids)
    ctx.set_sequences(flat)
    ctx.prepare_pairs(my_a, my_b)

    for _ in range(args.warmup):
        ctx.run_prepared()

    sampler = ClockSampler(device_ids[0])
    barrier()
    if rank == 0:
        sampler.start()
    dev_ms = 0.0
    dp_ms = post_ms = setup_ms = 0.0
    cells = 0
    dp_bytes = dp_cells_band = dp_pair_iters = 0
    launches = dp_launches = iter_steps = slots = live_slots = refilled = 0
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        ctx.run_prepared()
        st = ctx.stats()
        dev_ms += st.ms_device
        dp_ms += st.ms_dp
        post_ms += st.ms_bounds
        setup_ms += st.ms_setup
        cells += st.dp_cells
        launches += st.kernel_launches
        dp_launches += st.dp_launches
        dp_bytes += st.reserved[4]
        dp_cells_band += st.reserved[5]
        dp_pair_iters += st.sum_iterations
        iter_steps += st.reserved[7]
        slots += st.reserved[8]
        live_slots += st.reserved[9]
        refilled += st.reserved[12]
        log("step: %.1f ms device, dp %.1f post %.1f setup %.1f (summed over %d concurrent workers), band pairs %d (Q16 %d, mean width %.0f) of %d, "
            "mean iterations %.1f, E/pair %.0f, D/pair %.0f, %d batches, %d iteration steps, live lanes %.3f, refilled %d"
            % (st.ms_device, st.ms_dp, st.ms_bounds, st.ms_setup, st.reserved[0], st.reserved[1], st.reserved[2], st.reserved[3] / max(st.reserved[1], 1), st.n_pairs,
               st.sum_iterations / max(st.n_pairs, 1), st.n_active_edges / max(st.n_pairs, 1), st.n_interactions / max(st.n_pairs, 1),
               st.reserved[6], st.reserved[7], st.reserved[9] / max(st.reserved[8], 1), st.reserved[12]))
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - wall0)
    clocks = sampler.stop() if rank == 0 else None
    t_ms = reduce_max(dev_ms)
    wall_ms = reduce_max(wall_ms)
    cells_all = reduce_sum(float(cells))
    band_cells_all = reduce_sum(float(dp_cells_band))
    launches_all = int(reduce_sum(float(launches)))
    value = n_pairs_total * args.steps / (t_ms * 1e-3)

    # ---- end to end through the C ABI with host buffers, every step ----------------------------
    e2e = None
    if not args.no_e2e:
        # one untimed call first: the page-locked result buffers are allocated (and cached by the library) on first use
        ctx.set_sequences(flat)
        ctx.free(ctx.solve_pairs(my_a, my_b))
        barrier()
        t0 = time.perf_counter()
        h2d = d2h = 0
        per_step = []
        for _ in range(args.steps):
            t1 = time.perf_counter()
            ctx.set_sequences(flat)
            res = ctx.solve_pairs(my_a, my_b)
            st = ctx.stats()
            h2d += st.h2d_bytes + 8 * len(my_a)
            d2h += st.d2h_bytes
            ctx.free(res)
            per_step.append(time.perf_counter() - t1)
        barrier()
        e2e_s = reduce_max(time.perf_counter() - t0)
        e2e = {"value": n_pairs_total * args.steps / e2e_s, "unit": "pairs/s",
               "h2d_bytes_per_step": int(reduce_sum(float(h2d)) / args.steps),
               "d2h_bytes_per_step": int(reduce_sum(float(d2h)) / args.steps), "steps": args.steps,
               "ms_per_step_min_max_rank0": [1e3 * min(per_step), 1e3 * max(per_step)],
               "timing": "host wall clock around lara_gpu_set_sequences + lara_gpu_solve_pairs (results in host memory), all steps, "
                         "after one untimed call"}

    if rank != 0:
        ctx.close()
        if multi_proc:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the metric's kernel (DP fill) -------------------------------------------------
    # bound: HBM.  achieved = algorithmic bytes of the DP launches (per band cell 4 B score + 0.5 B trace nibble, plus one
    # 16 B closed-form record per row and column; lara_stats.reserved[4]) / summed CUDA-event durations of the DP
    # launches (one launch per iteration step: lara_stats.reserved[7]).  With two workers the launches of one batch overlap
    # the post kernel of the other, so the in-region figure is a lower bound; `isolated` repeats one step with a single worker.
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    int_rates = {k: ctx.int_issue_rate(k) for k in (0, 1, 2)}
    iso = None
    if not multi_proc and args.gpus == 1 and not args.no_isolated:
        prev_workers = os.environ.get("LARA_B200_WORKERS")
        os.environ["LARA_B200_WORKERS"] = "1"
        try:
            ctx1 = lb.Context(params, device_ids=device_ids)
            ctx1.set_sequences(flat)
            ctx1.prepare_pairs(my_a, my_b)
            ctx1.run_prepared()
            s1 = ctx1.stats()
            iso = {"dp_ms": s1.ms_dp, "post_ms": s1.ms_bounds, "setup_ms": s1.ms_setup, "device_ms": s1.ms_device,
                   "iter_steps": s1.reserved[7], "dp_bytes": s1.reserved[4], "dp_cells_band": s1.reserved[5],
                   "dp_cells": s1.dp_cells}
            ctx1.close()
        except Exception as ex:          # the extra pass is informative only; the timed region above stands
            log("isolated roofline pass failed: %r" % (ex,))
            iso = None
        finally:
            if prev_workers is None:
                os.environ.pop("LARA_B200_WORKERS", None)
            else:
                os.environ["LARA_B200_WORKERS"] = prev_workers
    achieved = dp_bytes / (dp_ms * 1e-3) / 1e9 if dp_ms > 0 else 0.0
    roofline = {
        "kernel": "gotoh_band_kernel", "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
        "frac": achieved / hbm_peak if hbm_peak else None,
        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
        "algorithmic_bytes": "4.5 B per band cell (4 B score read + 4-bit trace write) + 16 B closed-form record per row and column",
        "timing": "sum of the CUDA-event spans of the %d DP launches of the timed region, one per iteration step of a batch (%d concurrent "
                  "workers: spans include time shared with the other worker's kernels)" % (iter_steps, int(st.reserved[0])),
        "launches": iter_steps,
        "dp_ms_per_launch": dp_ms / iter_steps if iter_steps else None,
        "algorithmic_bytes_per_launch": dp_bytes / iter_steps if iter_steps else None,
        "live_lane_fraction": live_slots / slots if slots else None,
        "dp_gcups_kernel_only": float(cells) / (dp_ms * 1e-3) / 1e9 if dp_ms > 0 else None,
        "dp_gcups_computed_kernel_only": float(dp_cells_band) / (dp_ms * 1e-3) / 1e9 if dp_ms > 0 else None,
        "band_cells_per_matrix_cell": dp_cells_band / float(cells) if cells else None,
        "traffic": None,
        "int32_issue": {"ops_per_band_cell": OPS_PER_CELL, "achieved_gops": dp_cells_band * OPS_PER_CELL / (dp_ms * 1e-3) / 1e9 if dp_ms > 0 else None,
                        "peak_gops": max(int_rates.values()),
                        "peak_source": "lara_gpu_int_issue_rate on this GPU (IADD %.0f, LOP3+VIMNMX %.0f, DP mix %.0f Gop/s)"
                                       % (int_rates[0], int_rates[1], int_rates[2])},
        "share_of_step": {"dp": dp_ms / dev_ms if dev_ms else None, "post": post_ms / dev_ms if dev_ms else None,
                          "setup": setup_ms / dev_ms if dev_ms else None},
    }
    if iso:
        a1 = iso["dp_bytes"] / (iso["dp_ms"] * 1e-3) / 1e9
        roofline["isolated"] = {
            "achieved": a1, "frac": a1 / hbm_peak, "dp_ms_per_launch": iso["dp_ms"] / max(iso["iter_steps"], 1),
            "algorithmic_bytes_per_launch": iso["dp_bytes"] / max(iso["iter_steps"], 1),
            "dp_gcups_kernel_only": iso["dp_cells"] / (iso["dp_ms"] * 1e-3) / 1e9,
            "pairs_per_s": n_pairs_total / (iso["device_ms"] * 1e-3),
            "share_of_step": {"dp": iso["dp_ms"] / iso["device_ms"], "post": iso["post_ms"] / iso["device_ms"],
                              "setup": iso["setup_ms"] / iso["device_ms"]},
            "timing": "one extra step with a single worker: launches back to back on one stream"}
    if not args.no_traffic and not multi_proc and args.gpus == 1:
        tp = traffic_probe(args)
        if tp and tp["ratio"] and iter_steps:
            roofline["traffic"] = tp["ratio"] * dp_bytes / iter_steps
            roofline["traffic_probe"] = tp
    if roofline["traffic"] is None:
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof) and iter_steps:
            try:
                t = json.load(open(prof))["gotoh_band_kernel"]
                roofline["traffic"] = t["dram_bytes"] / t["algorithmic_bytes"] * dp_bytes / iter_steps
                roofline["traffic_source"] = "no ncu capture in this run; ratio from " + t.get("source", "profiles/traffic.json")
            except Exception:
                pass

    # ---- CPU baseline: the unmodified reference binaries on a bounded sample -----------------------------
    cpu = None
    if not args.no_cpu_baseline and not multi_proc and args.gpus == 1:
        from oracle import binding as ob
        in_args, m, mp = cpu_sample(args, w, files)
        cores = os.cpu_count()
        fmt = out_format(args.workload)
        if os.path.exists(ob.REF_SCALAR):
            ref_out = os.path.join(os.path.dirname(files[0]), "ref_sample.out")
            dt, _ = run_reference_binary(in_args, cores, fmt, ref_out)
            # live parity check on the same sample: GPU command line program vs reference binary, output byte for byte
            mine_out = os.path.join(os.path.dirname(files[0]), "gpu_sample.out")
            cp = subprocess.run([os.path.join(ROOT, "lara_b200", "lara_b200_cli")] + in_args + ["-o", fmt, "-w", mine_out],
                                stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            same = cp.returncode == 0 and open(mine_out, "rb").read() == open(ref_out, "rb").read()
            cpu = {"value": mp / dt, "unit": "pairs/s", "cores": cores, "kind": "reference",
                   "sample": "first %d of the %d sequences (%d pairs, all 500 iterations), oracle/_ref/lara_scalar -j %d, "
                             "process wall clock %.1f s incl. parse+write" % (m, n, mp, cores, dt),
                   "output_identical_to_gpu_on_sample": bool(same)}
            if os.path.exists(ob.REF_SIMD):
                # the -mavx2 build (SEQAN_SIMD_ENABLED: inter-sequence SIMD DP, src/subgradient_solver_simd.hpp) beside it
                dts, _ = run_reference_binary(in_args, cores, fmt, ref_out + ".simd", binary=ob.REF_SIMD)
                cpu["simd_build"] = {"value": mp / dts, "unit": "pairs/s",
                                     "sample": "same sample, oracle/_ref/lara_simd (-mavx2) -j %d, %.1f s" % (cores, dts)}
        else:
            cpu = {"value": None, "unit": "pairs/s", "cores": cores, "kind": "reference",
                   "sample": "oracle/_ref/lara_scalar not present on this box"}

    gc_eff = cells_all / (t_ms * 1e-3) / 1e9
    line = {
        "metric": "aligned pairs/s", "value": value, "unit": "pairs/s", "n_gpus": max(world, args.gpus),
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": dict(workload_config(args, w, max(world, args.gpus)),
                       **({"emulate_world": args.emulate_world, "pairs_run": n_pairs_total} if args.emulate_world > 1 else {})),
        "gcups": gc_eff, "gcups_effective": gc_eff, "gcups_computed": band_cells_all / (t_ms * 1e-3) / 1e9,
        "gcups_note": "effective = lenA*lenB of every iteration run (the metric's definition); computed = cells the band kernels update",
        "wall_ms_per_step": wall_ms / args.steps,
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches_all, "refilled_pairs_per_step": refilled / args.steps,
        "roofline": roofline, "cpu_baseline": cpu,
    }
    print(json.dumps(line))
    ctx.close()
    if multi_proc:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
