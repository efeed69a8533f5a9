#!/usr/bin/env python
"""bench.py — scored cell x locus nonzeros / second per EM iteration (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2]

A step is ONE EM iteration (compute_new_excluded, cellector/src/main.rs:308-347: alpha/beta tables, per-cell
scoring, IQR threshold, exclusion, per-locus column pass, -80 locus filter) over the synthetic workload, driven
through the C ABI of libcellector_cuda.so.  L2 is flushed between timed steps.  One JSON line on rank 0.

  N = 1   the headline is BASELINE.json configs[1] (C2: 10k cells x 100k loci, 2 % foreign cells, host-generated):
          `value`, `e2e` (host buffers through the C ABI), `cpu_baseline` + `parity` (the oracle over ALL of C2, timed
          and compared with the GPU), `roofline`.  C2 is 31 MB of algorithmic bytes and latency-bound, so the line also
          carries `large` / `roofline_large`: configs[2] (C3: 100k x 500k, 5e8 nonzeros) generated on the device and run
          whole on the one GPU, with full-size conservation properties and an oracle comparison on a cell slice.
  N > 1   STRONG scaling: the one fixed matrix C3 (the same generation blocks at every N) with its cells sharded over
          the ranks, per-locus arrays exchanged over NVLink peer memory / NCCL, global threshold.  Before timing, a
          small matrix sharded over the same ranks is compared with the CPU oracle (`sharded_parity`), and rank 0
          alone runs the whole of C3 once (`same_matrix_on_one_gpu`): the denominator of this line's speed-up.
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
    ap.add_argument("--workload", default=None, choices=list(synth.CONFIGS),
                    help="default: C2 (+ the large workload) on one GPU, C3 sharded over the ranks on several")
    ap.add_argument("--large", default="C3", choices=["C3", "C5"], help="large single-GPU workload reported next to the C2 headline")
    ap.add_argument("--no-large", action="store_true")
    ap.add_argument("--no-base", action="store_true", help="N > 1: skip the one-GPU run of the same matrix on rank 0")
    ap.add_argument("--equal-cells", action="store_true", help="N > 1: shards of equal cell counts instead of nnz-balanced")
    ap.add_argument("--slice-cells", type=int, default=1500, help="cells of the large matrix given to the CPU oracle")
    ap.add_argument("--cpu-sample-cells", type=int, default=5000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-soup", action="store_true", help="skip the souporcell EM leg on the headline matrix")
    ap.add_argument("--no-cli-e2e", action="store_true", help="skip the `cellector` CLI run on the workload written as .mtx text")
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
    world = max(world, args.gpus)
    name = args.workload or ("C2" if world == 1 else "C3")
    args.workload = name
    if name in ("C1", "C2", "C3s"):
        cfg = workload_config(name, 1)
        data = synth.generate(cfg)
        cfg_d = config_dict(name, cfg, 1)
    else:
        # the large matrices exist only as generation blocks (synth_device.py); the CPU sample is the first blocks, drawn on
        # the CPU with the same code, i.e. the very cells the GPU arm holds first
        from cellector_b200 import synth_device as sd
        cfg = synth.CONFIGS[name]
        nb = max(1, min(sd.GEN_BLOCKS, -(-args.cpu_sample_cells * sd.GEN_BLOCKS // cfg.n_cells)))
        data = sd.to_host(sd.generate(cfg, 0, nb, device="cpu"))
        cfg_d = large_config_dict(name, cfg, world, True)
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
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak" if world == 1 else "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg_d,
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


class Dist:
    """torch.distributed plumbing of the bench (barrier, max / sum over ranks); no-ops at world == 1."""

    def __init__(self, world, local_rank):
        import torch
        self.torch, self.world = torch, world
        torch.cuda.set_device(local_rank)
        if world > 1:
            import torch.distributed as dist
            self.dist = dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def _red(self, x, op):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, x):
        return self._red(x, self.dist.ReduceOp.MAX if self.world > 1 else None)

    def sum(self, x):
        return self._red(x, self.dist.ReduceOp.SUM if self.world > 1 else None)

    def bcast_obj(self, obj, src=0):
        if self.world == 1:
            return obj
        box = [obj]
        self.dist.broadcast_object_list(box, src=src)
        return box[0]

    def gather_obj(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def measure(ctx, dd, args, local_rank, n_local, L, nnz_used, nnz_total, collective=True):
    """K timed EM iterations (pass 1: graph replay, no inner events -> the metric; pass 2: an event pair around every
    kernel phase -> per-kernel durations for the roofline).  L2 is flushed before every step.  `collective` = the
    timing is shared by all ranks (barriers, max over ranks)."""
    no_flush = os.environ.get("CLX_BENCH_NO_FLUSH") == "1"  # diagnostics only: a reported number always flushes
    bar = dd.barrier if collective else dd.torch.cuda.synchronize
    mx = dd.max if collective else (lambda x: x)

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
        bar()
        ms, w0 = [], time.perf_counter()
        with ClockSampler(local_rank) as clk:
            for _ in range(args.steps):
                ms.append(step().device_ms)
            bar()
        wall = 1e3 * (time.perf_counter() - w0) / args.steps
        return ms, wall, ctx.launch_count() - l0, clk

    dev_ms, wall_ms, launches, clocks = timed_loop()
    ms_per_step = mx(sum(dev_ms) / len(dev_ms))
    ctx.set_profiling(True)
    dev_ms_prof, _, _, _ = timed_loop()
    phases = ctx.phase_ms()
    ctx.set_profiling(False)
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    sc_ms, sc_n = phases["score"]
    # ALGORITHMIC bytes of one scoring pass, SURVEY.md §8(d): 8 B packed (locus, ref, alt) entry per scored nonzero
    # + int64 row pointers + 20 B of per-cell outputs + one read of a 12-double per-locus table.  The run-stream kernel
    # actually moves a compacted re-encoding of the same entries (2-byte words in padded groups) plus the staged table
    # tiles once per cell block: those bytes are reported beside it as `moved`.
    algo_bytes = 8 * nnz_used + 8 * (n_local + 1) + 20 * n_local + 96 * L
    k_s = sc_ms / max(sc_n, 1) * 1e-3
    achieved = algo_bytes / k_s / 1e9 if sc_ms > 0 else 0.0
    moved = ctx.score_bytes()
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": "per-cell scoring (k_score_runs + k_score_finish)",
                "algorithmic_bytes_per_launch": int(algo_bytes), "moved_bytes_per_launch": int(moved),
                "achieved_moved": moved / k_s / 1e9 if sc_ms > 0 else 0.0, "kernel_us": 1e6 * k_s,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)"}
    return {"ms_per_step": ms_per_step, "value": nnz_total / (ms_per_step * 1e-3), "wall_ms": wall_ms, "launches": int(launches),
            "clocks": clocks.summary(), "roofline": roofline,
            "phase_us_per_step": {k: 1e3 * v[0] / args.steps for k, v in phases.items()},
            "ms_per_step_with_phase_events": mx(sum(dev_ms_prof) / len(dev_ms_prof))}


def apply_traffic(roofline, name):
    """roofline.traffic = DRAM bytes of one scoring launch from the committed ncu capture of this workload (profiles/)."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "score_kernel_traffic.json")))["records"].get(name)
    except Exception:
        rec = None
    if rec:
        roofline["traffic"] = rec.get("dram_bytes_per_launch")
        roofline["traffic_source"] = rec.get("source")


def _max_rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    d = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
    d = np.where((a == b) | (np.abs(a - b) <= 1e-12), 0.0, d)
    return float(d.max())


def parity_vs_oracle(gpu, coo, n_loci, n_cells, label, time_budget_s=None):
    """EM to convergence + posteriors on the GPU context and on the CPU oracle for the same COO: integers, excluded
    sets and labels must be identical, fp64 sums within 1e-9 relative (the bar of tests/parity.py).  Also returns the
    oracle's per-iteration seconds (the timed CPU baseline comes from this very run)."""
    from tests.oracle import Oracle
    orc = Oracle()
    g = gpu.load_cell_data(*coo, n_loci, n_cells)
    o = orc.load_cell_data(*coo, n_loci, n_cells)
    rep = {"what": label, "loaded_identical": bool(tuple(g) == tuple(o)), "iterations": 0, "max_rel_ll": 0.0, "max_rel_expected_ll": 0.0,
           "max_rel_locus_ll": 0.0, "max_rel_threshold": 0.0, "integers_identical": True, "excluded_identical": True}
    rep["integers_identical"] &= bool(np.array_equal(gpu.locus_ids(), orc.locus_ids()))
    times = []
    for it in range(50):
        gs = gpu.compute_new_excluded(5.0)
        t0 = time.perf_counter()
        os_ = orc.compute_new_excluded(5.0)
        times.append(time.perf_counter() - t0)
        gc, oc = gpu.cell_vectors(), orc.cell_vectors()
        gl, ol = gpu.locus_vectors(), orc.locus_vectors()
        rep["iterations"] += 1
        rep["excluded_identical"] &= bool(np.array_equal(gc["excluded"], oc["excluded"]))
        ints = all(np.array_equal(gl[k], ol[k]) for k in ("cells_minority", "cells_majority", "ref_minority", "alt_minority",
                                                           "ref_majority", "alt_majority", "loci_used"))
        ints &= bool(np.array_equal(gc["n_loci"], oc["n_loci"]))
        ints &= all(getattr(gs, k) == getattr(os_, k) for k in ("any_change", "n_new_excluded", "n_rescued", "n_excluded",
                                                                 "n_loci_filtered", "n_loci_in_use"))
        rep["integers_identical"] &= bool(ints)
        rep["max_rel_ll"] = max(rep["max_rel_ll"], _max_rel(gc["ll"], oc["ll"]), _max_rel(gc["norm"], oc["norm"]))
        rep["max_rel_expected_ll"] = max(rep["max_rel_expected_ll"], _max_rel(gc["expected_ll"], oc["expected_ll"]))
        rep["max_rel_locus_ll"] = max(rep["max_rel_locus_ll"], _max_rel(gl["ll_minority"], ol["ll_minority"]),
                                      _max_rel(gl["ll_majority"], ol["ll_majority"]))
        rep["max_rel_threshold"] = max(rep["max_rel_threshold"], _max_rel([gs.threshold, gs.median, gs.q1, gs.q3],
                                                                          [os_.threshold, os_.median, os_.q1, os_.q3]))
        if not os_.any_change:
            break
        if time_budget_s is not None and sum(times) > time_budget_s:
            rep["stopped_early"] = "CPU time budget"
            break
    gp, op = gpu.calculate_posteriors(), orc.calculate_posteriors()
    rep["labels_identical"] = bool(all(np.array_equal(gp[k], op[k]) for k in ("label", "anomaly", "qual")))
    rep["max_rel_posterior_ll"] = max(_max_rel(gp["ll_major"], op["ll_major"]), _max_rel(gp["ll_minor"], op["ll_minor"]))
    rep["n_excluded"] = int(gs.n_excluded)
    rep["label_counts"] = np.bincount(gp["label"], minlength=4).tolist()
    rep["ok"] = bool(rep["loaded_identical"] and rep["integers_identical"] and rep["excluded_identical"] and rep["labels_identical"]
                     and max(rep["max_rel_ll"], rep["max_rel_expected_ll"], rep["max_rel_locus_ll"], rep["max_rel_threshold"],
                             rep["max_rel_posterior_ll"]) <= 1e-9)
    orc.close()
    return rep, times, int(o[1])


def cpu_baseline_and_parity(args, gpu, data):
    """All of the per-GPU workload when it is small enough (C2: all 10,000 cells, ~10 s of single-thread CPU), else its
    first cells: the oracle iterations are timed (cpu_baseline) AND compared with the GPU (parity)."""
    n_s = min(args.cpu_sample_cells * 2, data.n_cells)
    keep = data.cell < n_s
    coo = [a[keep] for a in (data.locus, data.cell, data.ref, data.alt)]
    rep, times, nnz_used = parity_vs_oracle(gpu, coo, data.n_loci, n_s,
                                            f"first {n_s} of {data.n_cells} cells x all loci of {args.workload}, EM to convergence + posteriors",
                                            time_budget_s=30.0)
    t = sum(times) / len(times)
    out = {"value": nnz_used / t, "unit": UNIT, "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
           "sample": f"{len(times)} EM iterations (mean; faithful cost: expected-LL per entry and the per-entry records) over the "
                     f"first {n_s} cells x all loci of {args.workload} ({nnz_used} scored nonzeros, {t:.2f} s each, {sum(times):.1f} s timed)"}
    ctx_all = all_cores_context(coo, data.n_loci, n_s)
    if ctx_all:
        out["all_cores_context"] = ctx_all
    return out, rep


def large_config_dict(name, cfg, world, balanced, extra=None):
    d = {"workload": f"{name}: synthetic vartrix ref/alt generated on the device (cellector_b200/synth_device.py), {cfg.n_cells} cells x "
                     f"{cfg.n_loci} loci in total, density {cfg.density}, row-weight sigma {cfg.row_sigma}, {cfg.minority_frac:.0%} minority "
                     f"genotype, min_alt=min_ref=4; ONE fixed matrix, cells sharded over {world} GPU(s)",
         "cells_total": cfg.n_cells, "loci": cfg.n_loci, "density": cfg.density, "seed": cfg.seed,
         "sharding": f"cells x{world}, contiguous ranges " + ("balanced by expected nonzeros" if balanced else "of equal cell counts"),
         "l2": "flushed between timed steps (512 MiB overwrite, then 256 MiB of it read back so the resident lines are clean)"}
    if extra:
        d.update(extra)
    return d


def run_large(args, dd, lib, name, rank, world, local_rank, balanced=True, slice_parity=True):
    """One fixed large matrix (C3 / C4 / C5), generated on the device in generation blocks, cells sharded over `world`
    ranks (world == 1: the whole matrix on this GPU).  Returns the measurement of K EM iterations."""
    import torch
    from cellector_b200 import Cellector, synth_device as sd
    cfg = synth.CONFIGS[name]
    cut = sd.shard_blocks(cfg, world, balanced)
    t0 = time.perf_counter()
    d = sd.generate(cfg, int(cut[rank]), int(cut[rank + 1]), device=f"cuda:{local_rank}")
    torch.cuda.synchronize()
    gen_s = time.perf_counter() - t0
    torch.cuda.empty_cache()  # the generator's temporaries go back to the driver: the library allocates from its own pool
    ctx = Cellector(local_rank, args.flags, lib)
    collective = world > 1
    if collective:
        uid = dd.bcast_obj(Cellector.nccl_unique_id(lib) if rank == 0 else None)
        ctx.comm_init(rank, world, uid)
    t0 = time.perf_counter()
    n_loci_used, nnz_used = ctx.load_cell_data_ptrs(d.locus.data_ptr(), d.cell.data_ptr(), d.ref.data_ptr(), d.alt.data_ptr(),
                                                    d.nnz, d.n_loci, d.n_cells)
    load_s = time.perf_counter() - t0
    sl = sd.to_host(d, 0, min(d.n_cells, args.slice_cells)) if (slice_parity and rank == 0 and args.slice_cells > 0) else None
    n_local, nnz0_local = d.n_cells, d.nnz
    del d
    torch.cuda.empty_cache()
    nnz_total = dd.sum(float(nnz_used)) if collective else float(nnz_used)
    m = measure(ctx, dd, args, local_rank, n_local, n_loci_used, nnz_used, nnz_total, collective=collective)
    # size-independent properties at full size: every scored entry is counted once per cell and once per locus, and
    # the log likelihood summed over cells equals the one summed over loci (the same entries, regrouped)
    ctx.reset()
    ctx.compute_new_excluded(5.0)
    cv, lv = ctx.cell_vectors(), ctx.locus_vectors()
    s_nl = float(cv["n_loci"].astype(np.int64).sum())
    s_ll = float(cv["ll"].sum())
    if collective:
        s_nl, s_ll = dd.sum(s_nl), dd.sum(s_ll)
    l_cnt = float(lv["cells_minority"].astype(np.int64).sum() + lv["cells_majority"].astype(np.int64).sum())
    l_ll = float(lv["ll_minority"].sum() + lv["ll_majority"].sum())
    props = {"entries_counted_per_cell_equal_per_locus": bool(s_nl == l_cnt and s_nl == nnz_total),
             "sum_ll_cells_vs_loci_rel": abs(s_ll - l_ll) / max(abs(l_ll), 1e-300)}
    props["ok"] = bool(props["entries_counted_per_cell_equal_per_locus"] and props["sum_ll_cells_vs_loci_rel"] <= 1e-9)
    imbalance = None
    if collective:
        per_rank = dd.gather_obj({"nnz_used": int(nnz_used), "cells": int(n_local),
                                  "score_us": m["phase_us_per_step"]["score"], "column_us": m["phase_us_per_step"]["column"]})
        mean_nnz = sum(p["nnz_used"] for p in per_rank) / world
        imbalance = {"nnz_max_over_mean": max(p["nnz_used"] for p in per_rank) / mean_nnz,
                     "score_us_max_over_mean": max(p["score_us"] for p in per_rank) / (sum(p["score_us"] for p in per_rank) / world),
                     "per_rank": per_rank}
    ctx.close()
    if world == 1:
        apply_traffic(m["roofline"], name)
    out = dict(m)
    out.update({"name": name, "n_gpus": world, "nnz_scored_total": int(nnz_total), "loci_used": int(n_loci_used), "nnz0_local": int(nnz0_local),
                "generate_s": gen_s, "load_s": load_s, "properties": props, "imbalance": imbalance, "cfg": cfg, "balanced": balanced})
    if sl is not None:  # the oracle on a cell slice of the same matrix against a one-GPU context loaded with that slice
        with Cellector(local_rank, 0, lib) as g1:
            rep, _, _ = parity_vs_oracle(g1, [sl.locus, sl.cell, sl.ref, sl.alt], sl.n_loci, sl.n_cells,
                                         f"first {sl.n_cells} cells x all {sl.n_loci} loci of {name} (its own locus filter), EM to convergence + posteriors",
                                         time_budget_s=20.0)
        out["parity_slice"] = rep
    return out


def sharded_parity(dd, lib, rank, world, local_rank):
    """Small matrix, cells sharded over all ranks, every iteration against the CPU oracle on the whole matrix (the check of
    tests/multi_worker.py, run here because the driver's pytest box has one GPU)."""
    from cellector_b200 import Cellector
    cfg = synth.SynthConfig(n_cells=6000, n_loci=5000, density=0.02, minority_frac=0.06, seed=13, deep_frac=0.01)
    d = synth.generate(cfg, shard=rank, n_shards=world)
    gpu = Cellector(local_rank, 0, lib)
    gpu.comm_init(rank, world, dd.bcast_obj(Cellector.nccl_unique_id(lib) if rank == 0 else None))
    g = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
    shards = dd.gather_obj((d.cell_offset, d.n_cells, d.locus, d.cell, d.ref, d.alt))
    loads = dd.gather_obj(g)
    rep = {"what": f"{cfg.n_cells} cells x {cfg.n_loci} loci over {world} ranks vs the CPU oracle on the whole matrix", "ok": True,
           "max_rel_ll": 0.0, "max_rel_locus_ll": 0.0, "iterations": 0}
    orc = None
    if rank == 0:
        from tests.oracle import Oracle
        locus = np.concatenate([s[2] for s in shards]).astype(np.uint64)
        cell = np.concatenate([s[3].astype(np.uint64) + s[0] for s in shards])
        ref, alt = np.concatenate([s[4] for s in shards]), np.concatenate([s[5] for s in shards])
        order = np.lexsort((cell, locus))
        orc = Oracle()
        o = orc.load_cell_data(locus[order].astype(np.uint32), cell[order].astype(np.uint32), ref[order], alt[order], d.n_loci,
                               sum(s[1] for s in shards))
        rep["ok"] &= bool(loads[0][0] == o[0] and sum(x[1] for x in loads) == o[1] and np.array_equal(gpu.locus_ids(), orc.locus_ids()))
    for it in range(30):
        gs = gpu.compute_new_excluded(5.0)
        cv, lv = gpu.cell_vectors(), gpu.locus_vectors()
        allc = dd.gather_obj(cv)
        stop = None
        if rank == 0:
            os_ = orc.compute_new_excluded(5.0)
            oc, ol = orc.cell_vectors(), orc.locus_vectors()
            cells = {k: np.concatenate([a[k] for a in allc]) for k in cv}
            ok = bool(np.array_equal(cells["excluded"], oc["excluded"]) and np.array_equal(cells["n_loci"], oc["n_loci"]))
            ok &= all(np.array_equal(lv[k], ol[k]) for k in ("cells_minority", "cells_majority", "ref_minority", "alt_minority",
                                                              "ref_majority", "alt_majority", "loci_used"))
            ok &= gs.n_new_excluded == os_.n_new_excluded and gs.n_rescued == os_.n_rescued and gs.any_change == os_.any_change
            rep["max_rel_ll"] = max(rep["max_rel_ll"], _max_rel(cells["ll"], oc["ll"]), _max_rel(cells["expected_ll"], oc["expected_ll"]))
            rep["max_rel_locus_ll"] = max(rep["max_rel_locus_ll"], _max_rel(lv["ll_minority"], ol["ll_minority"]),
                                          _max_rel(lv["ll_majority"], ol["ll_majority"]), _max_rel([gs.threshold], [os_.threshold]))
            rep["ok"] &= bool(ok)
            rep["iterations"] += 1
            stop = not os_.any_change
        stop = dd.bcast_obj(stop)
        if stop:
            break
    pv = gpu.calculate_posteriors()
    allp = dd.gather_obj(pv)
    if rank == 0:
        op = orc.calculate_posteriors()
        gp = {k: np.concatenate([a[k] for a in allp]) for k in pv}
        rep["labels_identical"] = bool(all(np.array_equal(gp[k], op[k]) for k in ("label", "anomaly", "qual")))
        rep["max_rel_posterior_ll"] = max(_max_rel(gp["ll_major"], op["ll_major"]), _max_rel(gp["ll_minor"], op["ll_minor"]))
        rep["ok"] = bool(rep["ok"] and rep["labels_identical"] and max(rep["max_rel_ll"], rep["max_rel_locus_ll"], rep["max_rel_posterior_ll"]) <= 1e-9)
        rep["transport"] = "NCCL" if os.environ.get("CLX_P2P") == "0" else "peer memory (NVLink) with NCCL at load"
        orc.close()
    gpu.close()
    return dd.bcast_obj(rep)


def e2e_job(args, dd, ctx, data, host, nnz_total):
    """End to end through the C ABI with HOST buffers: load (H2D of the COO + GPU CSR/CSC/stream build), EM to convergence,
    posteriors, and every per-cell / per-locus result copied back — the whole job a user of the CLI runs."""
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
        dd.barrier()
        t0 = time.perf_counter()
        if packed:
            ctx.load_cell_data_packed(*e2e_host, data.n_loci, data.n_cells)
        else:
            ctx.load_cell_data(*host, data.n_loci, data.n_cells)
        stats = ctx.run_em(5.0)
        cv, lv, pv = ctx.cell_vectors(), ctx.locus_vectors(), ctx.calculate_posteriors()
        dd.barrier()
        if i > 0:
            ts.append(time.perf_counter() - t0)
        iters = len(stats)
    t = dd.max(statistics.median(ts))
    d2h = sum(v.nbytes for v in cv.values()) + sum(v.nbytes for v in lv.values()) + sum(v.nbytes for v in pv.values())
    return {"value": nnz_total * iters / t, "unit": UNIT, "h2d_bytes_per_step": int(sum(h.nbytes for h in e2e_host)),
            "d2h_bytes_per_step": int(d2h), "seconds_per_job": t, "seconds_per_job_rank0": [round(x, 6) for x in ts],
            "em_iterations_per_job": iters, "entry": "clx_load_coo_packed" if packed else "clx_load_coo",
            "what": "load from pinned host COO + EM to convergence + posteriors + all result vectors to host; value = scored "
                    "nonzeros x EM iterations / job seconds (median job, max over ranks)"}


def cli_e2e(args, data, cpu_em_s_per_iteration):
    """The drop-in surface itself: the workload written as a real alt.mtx / ref.mtx pair, the `cellector` binary run on
    it (process start -> last file closed; median of three runs), next to the CPU oracle
    doing the same stages single-threaded like the reference (text parse with orc_read_mtx_pair, load, EM to convergence
    at the per-iteration time measured above, posteriors; it writes no files, so its number is a lower bound)."""
    import ctypes as C
    import shutil
    import subprocess
    import tempfile
    from tests.oracle import Oracle, load_oracle
    cli = os.path.join(ROOT, "cellector_b200", "bin", "cellector")
    if not os.path.exists(cli):
        return {"unavailable": "cellector_b200/bin/cellector is not built"}
    work = tempfile.mkdtemp(prefix="clx_cli_e2e_")
    try:
        import pandas as pd
        header = "%%MatrixMarket matrix coordinate real general\n% written by sprs\n" + f"{data.n_loci}\t{data.n_cells}\t{data.nnz}\n"
        l1, c1 = data.locus.astype(np.int64) + 1, data.cell.astype(np.int64) + 1
        for name, vals in (("alt.mtx", data.alt), ("ref.mtx", data.ref)):
            path = os.path.join(work, name)
            with open(path, "w") as f:
                f.write(header)
            pd.DataFrame({"l": l1, "c": c1, "v": vals}).to_csv(path, sep="\t", header=False, index=False, mode="a")
        with open(os.path.join(work, "barcodes.tsv"), "w") as f:
            f.write("\n".join(f"CELL{i:09d}-1" for i in range(data.n_cells)) + "\n")
        mtx_bytes = sum(os.path.getsize(os.path.join(work, n)) for n in ("alt.mtx", "ref.mtx"))
        runs = []
        for rep in range(3):
            out = os.path.join(work, f"out{rep}")
            t0 = time.perf_counter()
            r = subprocess.run([cli, "-a", os.path.join(work, "alt.mtx"), "-r", os.path.join(work, "ref.mtx"), "-b",
                                os.path.join(work, "barcodes.tsv"), "--output_directory", out, "--timing_json", out + ".json"],
                               capture_output=True, text=True, timeout=600)
            wall = time.perf_counter() - t0
            if r.returncode != 0:
                return {"unavailable": "the CLI failed: " + r.stderr[-300:]}
            t = json.load(open(out + ".json"))
            t["process_wall_s"] = wall
            t.pop("em_device_ms", None)
            runs.append(t)
        lib = load_oracle()
        nl, nc, nnz = C.c_uint32(0), C.c_uint32(0), C.c_uint64(0)
        coo = [np.empty(data.nnz, np.uint32) for _ in range(4)]
        t0 = time.perf_counter()
        rc = lib.orc_read_mtx_pair(os.path.join(work, "alt.mtx").encode(), os.path.join(work, "ref.mtx").encode(), C.byref(nl),
                                   C.byref(nc), C.byref(nnz), *[a.ctypes.data for a in coo], data.nnz)
        parse_s = time.perf_counter() - t0
        cpu = {"unavailable": "orc_read_mtx_pair failed"}
        if rc == 0 and nnz.value == data.nnz:
            with Oracle(lib) as orc:
                t0 = time.perf_counter()
                orc.load_cell_data(coo[0], coo[1], coo[2], coo[3], nl.value, nc.value)
                load_s = time.perf_counter() - t0
            em_s = cpu_em_s_per_iteration * runs[-1]["iterations"] if cpu_em_s_per_iteration else None  # same count in every run
            cpu = {"parse_s": parse_s, "load_s": load_s, "em_s": em_s, "cores": 1, "kind": "port",
                   "total_s_lower_bound": (parse_s + load_s + em_s) if em_s is not None else None,
                   "what": "oracle port, single-threaded like the reference: text parse + row lists + EM iterations (per-iteration "
                           "time of cpu_baseline x the CLI's iteration count); posteriors and every output file left out"}
        # CUDA start-up inside the child (0.2-0.9 s on a box whose GPU the bench process also holds) decides the wall time:
        # the median of three runs is reported, all three are listed
        order = sorted(range(len(runs)), key=lambda i: runs[i]["process_wall_s"])
        mid = runs[order[len(runs) // 2]]
        res = {"seconds": mid["process_wall_s"], "workload": args.workload, "mtx_bytes": mtx_bytes, "stages": mid,
               "all_runs_seconds": [r["process_wall_s"] for r in runs], "cpu": cpu,
               "what": "cellector_b200/bin/cellector on the workload as alt.mtx/ref.mtx text: process start -> exit, every output "
                       "file of the reference written (iteration_*.tsv, locus contributions, assignments); median of three runs"}
        if cpu.get("total_s_lower_bound"):
            res["speedup_vs_cpu_lower_bound"] = cpu["total_s_lower_bound"] / res["seconds"]
        return res
    except Exception as e:  # the bench line must not die on the optional leg
        return {"unavailable": f"{type(e).__name__}: {e}"}
    finally:
        shutil.rmtree(work, ignore_errors=True)


def soup_leg(ctx, data, k=2, restarts=16):
    """souporcell's k-cluster EM on the matrix the context holds: time per EM iteration for a single
    run and per restart-iteration with the restarts batched, next to the oracle in the reference's f32 arithmetic on one
    core, with the parity of the same run (GPU vs the oracle's f64 statement; assignments vs its f32 one)."""
    try:
        from cellector_b200 import soup
        from tests.soup_oracle import SoupOracle
        n_used = ctx.n_loci_used
        if ctx.n_cells != data.n_cells:
            return {"unavailable": "the context holds a cell slice of the workload"}
        init = np.stack([soup.init_cluster_centers_uniform(n_used, k, 4, r) for r in range(restarts)])
        ctx.soup_em(init[0])  # warm-up
        t0 = time.perf_counter()
        one = ctx.soup_em(init[0])
        t_one = time.perf_counter() - t0
        ctx.soup_restarts(init[:2])
        t0 = time.perf_counter()
        many = ctx.soup_restarts(init)
        t_many = time.perf_counter() - t0
        with SoupOracle() as orc:
            orc.load_cell_data(data.locus, data.cell, data.ref, data.alt, data.n_loci, data.n_cells)
            t0 = time.perf_counter()
            o32 = orc.soup_em(init[0], precise=False)
            t_cpu = time.perf_counter() - t0
            o64 = orc.soup_em(init[0], precise=True)
            entries = orc.n_entries
        it_many = int(many["iterations"].sum())
        rel = float(np.max(np.abs(one["log_probs"] - o64["log_probs"]) / np.abs(o64["log_probs"])))
        return {"k": k, "entries": int(entries), "iterations": one["iterations"],
                "gpu_us_per_iteration": 1e6 * t_one / one["iterations"], "gpu_entries_per_s": entries * one["iterations"] / t_one,
                "batched": {"restarts": restarts, "restart_iterations": it_many, "us_per_restart_iteration": 1e6 * t_many / it_many,
                            "entries_per_s": entries * it_many / t_many, "best": many["best"]},
                "cpu_f32": {"iterations": o32["iterations"], "us_per_iteration": 1e6 * t_cpu / o32["iterations"],
                            "entries_per_s": entries * o32["iterations"] / t_cpu, "cores": 1, "kind": "port"},
                "parity": {"iterations_gpu_f64_f32": [one["iterations"], o64["iterations"], o32["iterations"]],
                           "max_rel_log_prob_vs_f64": rel, "loss_rel_vs_f32": abs(one["loss"] - o32["loss"]) / abs(o32["loss"]),
                           "assignments_identical_to_f32": bool(np.array_equal(np.argmax(one["log_probs"], 1), np.argmax(o32["log_probs"], 1))),
                           "ok": bool(rel <= 1e-9 and one["iterations"] == o64["iterations"])},
                "what": "clx_soup_em / clx_soup_restarts wall time (set-up, iterations, results to the host) on the headline matrix"}
    except Exception as e:  # optional leg
        return {"unavailable": f"{type(e).__name__}: {e}"}


def strip(m):
    return {k: v for k, v in m.items() if k not in ("cfg", "balanced", "clocks")}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return 0
    from cellector_b200 import Cellector, api
    dd = Dist(world, local_rank)
    lib = api.load_library()
    timing = ("value: K iterations timed with CUDA events around each clx_em_iteration (graph replay, no inner events), max over ranks; "
              "roofline/phase_us: a second K-iteration pass with an event pair around every kernel phase")
    rc = 0
    if world == 1 and args.workload in (None, "C1", "C2", "C3s"):
        # ---- the headline: BASELINE.json configs[1] (C2) on one GPU, host-generated, plus (default) the large workload
        name = args.workload or "C2"
        args.workload = name
        cfg = workload_config(name, 1)
        data = synth.generate(cfg)
        pins = [pinned_copy(a) for a in (data.locus, data.cell, data.ref, data.alt)]
        host = [p[1] for p in pins]
        ctx = Cellector(local_rank, args.flags, lib)
        n_loci_used, nnz_used = ctx.load_cell_data(*host, data.n_loci, data.n_cells)
        m = measure(ctx, dd, args, local_rank, data.n_cells, n_loci_used, nnz_used, float(nnz_used))
        apply_traffic(m["roofline"], name)
        e2e = None if args.no_e2e else e2e_job(args, dd, ctx, data, host, float(nnz_used))
        cpu, par = (None, None) if args.no_cpu_baseline else cpu_baseline_and_parity(args, ctx, data)
        soup_res = None if (args.no_soup or name not in ("C1", "C2")) else soup_leg(ctx, data)
        ctx.close()
        cli = None
        if not args.no_cli_e2e and not args.no_e2e and name in ("C1", "C2"):
            em_s = None
            if cpu is not None:
                em_s = float(nnz_used) / cpu["value"]
            cli = cli_e2e(args, data, em_s)
        large = None
        if not args.no_large and name == "C2":
            lg = run_large(args, dd, lib, args.large, 0, 1, local_rank)
            large = strip(lg)
            large["config"] = large_config_dict(args.large, lg["cfg"], 1, True)
        line = {"metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": config_dict(name, cfg, 1, {"nnz_scored_total": int(nnz_used), "loci_used": int(n_loci_used)}),
                "wall_ms_per_step_incl_l2_flush": m["wall_ms"], "clocks": m["clocks"], "e2e": e2e, "gpu_launches": m["launches"],
                "roofline": m["roofline"], "cpu_baseline": cpu, "parity": par, "phase_us_per_step": m["phase_us_per_step"],
                "ms_per_step_with_phase_events": m["ms_per_step_with_phase_events"], "timing": timing}
        if cli is not None:
            line["cli_e2e"] = cli
        if soup_res is not None:
            line["souporcell_em"] = soup_res
        if large is not None:
            line["large"] = large
            line["roofline_large"] = large["roofline"]
            rc |= 0 if large["properties"]["ok"] and large.get("parity_slice", {}).get("ok", True) else 1
        if par is not None and not par["ok"]:
            rc |= 1
        print(json.dumps(line), flush=True)
    else:
        # ---- one fixed large matrix, cells sharded over the ranks: strong scaling
        name = args.workload or "C3"
        sp = sharded_parity(dd, lib, rank, world, local_rank) if world > 1 else None
        base = None
        if world > 1 and not args.no_base and synth.CONFIGS[name].n_cells * synth.CONFIGS[name].n_loci * synth.CONFIGS[name].density < 1.5e9:
            # the same matrix on ONE GPU (rank 0 alone), so that the speed-up of this line has its own denominator
            if rank == 0:
                b = run_large(args, dd, lib, name, 0, 1, local_rank, slice_parity=False)
                base = {"n_gpus": 1, "value": b["value"], "ms_per_step": b["ms_per_step"], "phase_us_per_step": b["phase_us_per_step"],
                        "roofline_frac": b["roofline"]["frac"], "nnz_scored_total": b["nnz_scored_total"]}
            dd.barrier()
        lg = run_large(args, dd, lib, name, rank, world, local_rank, balanced=not args.equal_cells)
        if rank == 0:
            line = {"metric": METRIC, "value": lg["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
                    "warmup": max(args.warmup, 3), "ms_per_step": lg["ms_per_step"], "higher_is_better": True, "scaling": "strong",
                    "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                    "config": large_config_dict(name, lg["cfg"], world, lg["balanced"],
                                                {"nnz_scored_total": lg["nnz_scored_total"], "loci_used": lg["loci_used"]}),
                    "wall_ms_per_step_incl_l2_flush": lg["wall_ms"], "clocks": lg["clocks"], "e2e": None, "gpu_launches": lg["launches"],
                    "roofline": lg["roofline"], "cpu_baseline": None, "phase_us_per_step": lg["phase_us_per_step"],
                    "ms_per_step_with_phase_events": lg["ms_per_step_with_phase_events"], "properties": lg["properties"],
                    "parity_slice": lg.get("parity_slice"), "imbalance": lg["imbalance"], "sharded_parity": sp,
                    "same_matrix_on_one_gpu": base, "generate_s": lg["generate_s"], "load_s": lg["load_s"], "timing": timing}
            if base:
                line["speedup_vs_one_gpu_same_matrix"] = lg["value"] / base["value"]
            print(json.dumps(line), flush=True)
        ok = lg["properties"]["ok"] and (sp is None or sp["ok"]) and lg.get("parity_slice", {"ok": True})["ok"]
        rc |= 0 if ok else 1
    dd.close()
    return rc


if __name__ == "__main__":
    sys.exit(main())
