#!/usr/bin/env python
"""bench.py — scored cell x locus nonzeros / second per EM iteration (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2]

A step is ONE EM iteration (compute_new_excluded, cellector/src/main.rs:308-347: alpha/beta tables, per-cell
scoring, IQR threshold, exclusion, per-locus column pass, -80 locus filter) over the synthetic workload, driven
through the C ABI of libcellector_cuda.so.  N=1 runs BASELINE.json configs[1] (C2: 10k cells x 100k loci, 2 %
foreign cells); N>1 is weak scaling: every rank holds a C2-sized cell shard (N x 10k cells), per-locus arrays are
allreduced with NCCL and the threshold is global.  L2 is flushed between timed steps.  One JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from cellector_b200 import synth  # noqa: E402

METRIC = "scored cell x locus nonzeros per second per EM iteration"
UNIT = "nnz/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=list(synth.CONFIGS))
    ap.add_argument("--cpu-sample-cells", type=int, default=5000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--flags", type=int, default=0, help="extra clx_create flags (e.g. 4 = v0 scoring kernel)")
    return ap.parse_args()


def workload_config(name, world):
    base = synth.CONFIGS[name]
    cfg = synth.SynthConfig(**{**base.__dict__})
    cfg.n_cells = base.n_cells * world  # weak scaling: fixed cells per GPU
    return cfg


def config_dict(name, cfg, world, extra=None):
    d = {"workload": f"{name}: synthetic vartrix ref/alt, {cfg.n_cells // world} cells x {cfg.n_loci} loci per GPU, "
                     f"density {cfg.density}, {cfg.minority_frac:.0%} minority genotype, min_alt=min_ref=4",
         "cells_per_gpu": cfg.n_cells // world, "loci": cfg.n_loci, "density": cfg.density, "seed": cfg.seed,
         "l2": "flushed between timed steps (512 MiB overwrite, then 256 MiB of it read back so the resident lines are clean)", "sharding": f"cells x{world}"}
    if extra:
        d.update(extra)
    return d


class ClockSampler:
    """nvidia-smi clocks line of B200_PROFILING.md sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def pinned_copy(a):
    import torch
    t = torch.empty(a.shape, dtype=torch.int32, pin_memory=True)
    v = t.numpy().view(np.uint32)
    v[...] = a
    return t, v


def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path (C++ restatement in oracle/: the Rust binary cannot be
    built here), single thread like the reference, on a bounded cell slice of the same workload."""
    if rank != 0:
        return
    from tests.oracle import Oracle
    cfg = workload_config(args.workload, 1)
    data = synth.generate(cfg)
    n_s = min(args.cpu_sample_cells, data.n_cells)
    keep = data.cell < n_s
    orc = Oracle()
    coo = [a[keep] for a in (data.locus, data.cell, data.ref, data.alt)]
    _, nnz_used = orc.load_cell_data(*coo, data.n_loci, n_s)
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        st = orc.compute_new_excluded(5.0)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
        if not st.any_change:
            orc.reset()
    ms = 1e3 * sum(times) / len(times)
    value = nnz_used / (ms * 1e-3)
    sample = f"one EM iteration per step over the first {n_s} cells x all {data.n_loci} loci of {args.workload} ({nnz_used} scored nonzeros)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config_dict(args.workload, cfg, 1),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample,
                             "host_cores": os.cpu_count()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference = oracle/ C++ restatement of cellector (Rust toolchain absent); the reference is single-threaded (rayon declared, never used), so cores=1"}
    ctx_all = all_cores_context(coo, data.n_loci, n_s)
    if ctx_all:
        line["cpu_baseline"]["all_cores_context"] = ctx_all
    print(json.dumps(line), flush=True)


def all_cores_context(coo, n_loci, n_s):
    """Context only: the oracle port on disjoint contiguous cell slices, one thread per host core (ctypes releases
    the GIL).  The reference itself is single-threaded."""
    from tests.oracle import Oracle
    n_thr = max(1, min(os.cpu_count() or 1, 64))
    if n_thr < 2 or n_s < 250 * n_thr:  # too few cells per slice: the per-slice locus filter would remove most loci
        return None
    cuts = np.linspace(0, n_s, n_thr + 1).astype(np.int64)
    parts = []
    for k in range(n_thr):
        m = (coo[1] >= cuts[k]) & (coo[1] < cuts[k + 1])
        o = Oracle()
        _, nz = o.load_cell_data(coo[0][m], coo[1][m] - np.uint32(cuts[k]), coo[2][m], coo[3][m], n_loci, int(cuts[k + 1] - cuts[k]))
        parts.append((o, nz))
    best = None
    for rep in range(3):
        th = [threading.Thread(target=lambda o=o: o.compute_new_excluded(5.0)) for o, _ in parts]
        t0 = time.perf_counter()
        for x in th:
            x.start()
        for x in th:
            x.join()
        dt = time.perf_counter() - t0
        if rep > 0:
            best = dt if best is None else min(best, dt)
    nz_all = sum(nz for _, nz in parts)
    for o, _ in parts:
        o.close()
    return {"value": nz_all / best, "unit": UNIT, "cores": n_thr,
            "how": "the oracle port on disjoint cell slices, one thread per host core, per-slice locus filter; not the reference's behaviour (it is single-threaded)"}


def cpu_baseline(args, data):
    """Oracle port timed on the host: ~10 s of single-thread work on the whole per-GPU workload when it is small enough
    (C2: all 10,000 cells), else on its first cells; plus, for context only, the same iteration run over disjoint cell
    slices on every host core at once (the reference itself is single-threaded)."""
    from tests.oracle import Oracle
    n_s = min(args.cpu_sample_cells * 2, data.n_cells)
    keep = data.cell < n_s
    coo = [a[keep] for a in (data.locus, data.cell, data.ref, data.alt)]
    orc = Oracle()
    _, nnz_used = orc.load_cell_data(*coo, data.n_loci, n_s)
    ts, spent = [], 0.0
    orc.compute_new_excluded(5.0)  # warm-up
    while len(ts) < 3 or (spent < 10.0 and len(ts) < 12):
        t0 = time.perf_counter()
        st = orc.compute_new_excluded(5.0)
        ts.append(time.perf_counter() - t0)
        spent += ts[-1]
        if not st.any_change:
            orc.reset()
    orc.close()
    t = sum(ts) / len(ts)
    out = {"value": nnz_used / t, "unit": UNIT, "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
           "sample": f"{len(ts)} EM iterations (mean; faithful cost: expected-LL per entry and the per-entry records) over the "
                     f"first {n_s} cells x all loci of {args.workload} ({nnz_used} scored nonzeros, {t:.2f} s each, {spent:.1f} s timed)"}
    ctx_all = all_cores_context(coo, data.n_loci, n_s)
    if ctx_all:
        out["all_cores_context"] = ctx_all
    return out


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return 0
    import torch
    import torch.distributed as dist
    from cellector_b200 import Cellector, api
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    cfg = workload_config(args.workload, world)
    data = synth.generate(cfg, shard=rank, n_shards=world)
    pins = [pinned_copy(a) for a in (data.locus, data.cell, data.ref, data.alt)]
    host = [p[1] for p in pins]

    lib = api.load_library()
    ctx = Cellector(local_rank, args.flags, lib)
    if world > 1:
        uid = [Cellector.nccl_unique_id(lib) if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(rank, world, uid[0])
    n_loci_used, nnz_used = ctx.load_cell_data(*host, data.n_loci, data.n_cells)
    nnz_total = sum_over_ranks(float(nnz_used))

    no_flush = os.environ.get("CLX_BENCH_NO_FLUSH") == "1"  # diagnostics only: a reported number always flushes

    def step():
        if not no_flush:
            ctx.flush_l2()
        st = ctx.compute_new_excluded(5.0)
        if not st.any_change:
            ctx.reset()
        return st

    def timed_loop():
        for _ in range(max(args.warmup, 3)):
            step()
        ctx.reset()
        ctx.reset_phase_ms()
        l0 = ctx.launch_count()
        barrier()
        ms, w0 = [], time.perf_counter()
        with ClockSampler(local_rank) as clk:
            for _ in range(args.steps):
                ms.append(step().device_ms)
            barrier()
        wall = 1e3 * (time.perf_counter() - w0) / args.steps
        return ms, wall, ctx.launch_count() - l0, clk

    # pass 1 — the metric: K EM iterations, no per-kernel events inside the iteration graph
    dev_ms, wall_ms, launches, clocks = timed_loop()
    ms_per_step = max_over_ranks(sum(dev_ms) / len(dev_ms))
    value = nnz_total / (ms_per_step * 1e-3)
    # pass 2 — the same K iterations with a CUDA event pair around every kernel phase (they are extra graph nodes and
    # cost tens of microseconds per iteration, so they stay out of pass 1): per-kernel durations for the roofline
    ctx.set_profiling(True)
    dev_ms_prof, _, _, _ = timed_loop()
    phases = ctx.phase_ms()
    ctx.set_profiling(False)
    ms_per_step_prof = max_over_ranks(sum(dev_ms_prof) / len(dev_ms_prof))

    # roofline of the dominant kernel (per-cell scoring), timed with CUDA events on the library's stream
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    sc_ms, sc_n = phases["score"]
    # ALGORITHMIC bytes of one scoring pass, SURVEY.md §8(d): 8 B packed (locus, ref, alt) entry per scored nonzero
    # + int64 row pointers + 20 B of per-cell outputs + one read of a 12-double per-locus table.  The tiled kernel
    # actually streams a compacted re-encoding of the same entries (2-byte entries in padded groups): those bytes
    # are reported beside it as `moved`.
    n_local, L = data.n_cells, n_loci_used
    algo_bytes = 8 * nnz_used + 8 * (n_local + 1) + 20 * n_local + 96 * L
    moved_bytes = ctx.score_bytes()
    k_s = sc_ms / max(sc_n, 1) * 1e-3
    achieved = algo_bytes / k_s / 1e9 if sc_ms > 0 else 0.0
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "score_kernel_traffic.json")))
        if tj.get("workload") == args.workload and world == 1:  # the capture is of this workload on one GPU
            traffic = tj.get("dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": "per-cell scoring (k_score_tiled + k_score_finish)",
                "algorithmic_bytes_per_launch": int(algo_bytes), "moved_bytes_per_launch": int(moved_bytes),
                "achieved_moved": moved_bytes / k_s / 1e9 if sc_ms > 0 else 0.0,
                "kernel_us": 1e6 * k_s,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)"}

    # end to end through the C ABI with HOST buffers: load (H2D of the COO + GPU CSR/CSC build), EM to convergence,
    # posteriors, and every per-cell / per-locus result copied back — the whole job a user of the CLI runs.
    e2e = None
    if not args.no_e2e:
        n_e2e, ts, iters = max(3, min(args.steps, 5)), [], 0
        # a host parser that keeps ref | alt << 16 per entry hands 12 instead of 16 bytes per entry to the device
        # (clx_load_coo_packed); counts above 65535 need the four-array call
        packed = os.environ.get("CLX_BENCH_E2E_PACKED", "1") == "1" and (
            data.ref.size == 0 or max(int(data.ref.max()), int(data.alt.max())) <= 0xFFFF)
        if packed:
            pin_ra = pinned_copy(data.ref.astype(np.uint32) | (data.alt.astype(np.uint32) << 16))
            e2e_host = [host[0], host[1], pin_ra[1]]
        else:
            e2e_host = host
        for i in range(n_e2e + 1):
            barrier()
            t0 = time.perf_counter()
            if packed:
                ctx.load_cell_data_packed(*e2e_host, data.n_loci, data.n_cells)
            else:
                ctx.load_cell_data(*host, data.n_loci, data.n_cells)
            stats = ctx.run_em(5.0)
            cv, lv, pv = ctx.cell_vectors(), ctx.locus_vectors(), ctx.calculate_posteriors()
            barrier()
            if i > 0:
                ts.append(time.perf_counter() - t0)
            iters = len(stats)
        t = max_over_ranks(statistics.median(ts))
        d2h = sum(v.nbytes for v in cv.values()) + sum(v.nbytes for v in lv.values()) + sum(v.nbytes for v in pv.values())
        e2e = {"value": nnz_total * iters / t, "unit": UNIT, "h2d_bytes_per_step": int(sum(h.nbytes for h in e2e_host)),
               "d2h_bytes_per_step": int(d2h), "seconds_per_job": t, "seconds_per_job_rank0": [round(x, 6) for x in ts], "em_iterations_per_job": iters,
               "entry": "clx_load_coo_packed" if packed else "clx_load_coo",
               "what": "load from pinned host COO + EM to convergence + posteriors + all result vectors to host; value = scored nonzeros x EM iterations / job seconds (median job, max over ranks)"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args, data)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": config_dict(args.workload, cfg, world, {"nnz_scored_total": int(nnz_total), "loci_used": int(n_loci_used)}),
                "wall_ms_per_step_incl_l2_flush": wall_ms, "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": int(launches),
                "roofline": roofline, "cpu_baseline": cpu,
                "phase_us_per_step": {k: 1e3 * v[0] / args.steps for k, v in phases.items()},
                "ms_per_step_with_phase_events": ms_per_step_prof,
                "timing": "value: K iterations timed with CUDA events around each clx_em_iteration (graph replay, no inner events); "
                          "roofline/phase_us: a second K-iteration pass with an event pair around every kernel phase"}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
