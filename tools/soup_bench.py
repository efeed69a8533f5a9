"""souporcell EM (clx_soup_em) on a named workload: time per EM iteration on the GPU next to the CPU oracle's f32
statement (the reference's arithmetic, single-threaded like one restart of the reference).  One JSON line.
usage: python tools/soup_bench.py [C1|C2] [k] [restarts]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from cellector_b200 import Cellector, soup, synth
from tests.soup_oracle import SoupOracle

name = sys.argv[1] if len(sys.argv) > 1 else "C2"
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
restarts = int(sys.argv[3]) if len(sys.argv) > 3 else 3
d = synth.generate(synth.CONFIGS[name])
out = {"workload": name, "k": k, "cells": d.n_cells}
with Cellector(0) as gpu:
    n_used, nnz = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
    out.update(loci_used=n_used, entries=nnz)
    gpu.soup_em(soup.init_cluster_centers_uniform(n_used, k, 4, 99))  # warm-up
    runs = []
    for r in range(restarts):
        init = soup.init_cluster_centers_uniform(n_used, k, 4, r)
        t = time.perf_counter()
        res = gpu.soup_em(init)
        dt = time.perf_counter() - t
        runs.append(dict(restart=r, iterations=res["iterations"], loss=res["loss"], seconds=dt,
                         us_per_iteration=1e6 * dt / res["iterations"]))
    out["gpu_runs"] = runs
    it = sum(r["iterations"] for r in runs)
    sec = sum(r["seconds"] for r in runs)
    out["gpu_us_per_iteration"] = 1e6 * sec / it
    out["gpu_entries_per_s"] = nnz * it / sec
    # restarts batched over the same passes of the matrix (clx_soup_restarts)
    n_batch = int(os.environ.get("SOUP_BENCH_RESTARTS", "16"))
    init = np.stack([soup.init_cluster_centers_uniform(n_used, k, 4, r) for r in range(n_batch)])
    gpu.soup_restarts(init[:2])
    t = time.perf_counter()
    rb = gpu.soup_restarts(init)
    dt = time.perf_counter() - t
    itb = int(rb["iterations"].sum())
    out["gpu_batched"] = dict(restarts=n_batch, seconds=dt, restart_iterations=itb, us_per_restart_iteration=1e6 * dt / itb,
                              entries_per_s=nnz * itb / dt, best=rb["best"], best_loss=rb["loss"],
                              iterations=[int(x) for x in rb["iterations"]])
    g0 = gpu.soup_em(soup.init_cluster_centers_uniform(n_used, k, 4, 0))
with SoupOracle() as orc:
    orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
    init = soup.init_cluster_centers_uniform(n_used, k, 4, 0)
    t = time.perf_counter()
    o32 = orc.soup_em(init, precise=False)
    dt = time.perf_counter() - t
    out["cpu_f32"] = dict(iterations=o32["iterations"], seconds=dt, us_per_iteration=1e6 * dt / o32["iterations"],
                          entries_per_s=orc.n_entries * o32["iterations"] / dt, loss=o32["loss"])
    o64 = orc.soup_em(init, precise=True)
out["parity"] = dict(iterations_gpu=g0["iterations"], iterations_f64=o64["iterations"], iterations_f32=o32["iterations"],
                     max_rel_lp_vs_f64=float(np.max(np.abs(g0["log_probs"] - o64["log_probs"]) / np.abs(o64["log_probs"]))),
                     loss_rel_vs_f32=abs(g0["loss"] - o32["loss"]) / abs(o32["loss"]),
                     same_assignment_as_f32=float((np.argmax(g0["log_probs"], 1) == np.argmax(o32["log_probs"], 1)).mean()))
out["speedup_per_iteration"] = out["cpu_f32"]["us_per_iteration"] / out["gpu_us_per_iteration"]
out["speedup_per_restart_iteration_batched"] = out["cpu_f32"]["us_per_iteration"] / out["gpu_batched"]["us_per_restart_iteration"]
print(json.dumps(out))
