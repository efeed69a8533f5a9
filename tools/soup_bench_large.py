"""souporcell EM (clx_soup_em / clx_soup_restarts) on a LARGE named workload generated on the device (C3: 100 k cells x
500 k loci, 5e8 entries — too big for the host generator and for the CPU oracle): time per EM iteration, and the
size-independent checks the domain offers: the loss the device reports equals the sum over cells of logsumexp of the
log-probabilities it returns (numpy on the host: another code path), the centres respect their clamps, a second run
returns the same bits.  One JSON line.  usage: python tools/soup_bench_large.py [C3] [k] [restarts]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from scipy.special import logsumexp

from cellector_b200 import Cellector, soup, synth, synth_device as sd

name = sys.argv[1] if len(sys.argv) > 1 else "C3"
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
restarts = int(sys.argv[3]) if len(sys.argv) > 3 else 4
cfg = synth.CONFIGS[name]
cut = sd.shard_blocks(cfg, 1, True)
d = sd.generate(cfg, int(cut[0]), int(cut[1]), device="cuda:0")
torch.cuda.synchronize()
out = {"workload": name, "k": k, "cells": int(d.n_cells)}
with Cellector(0) as gpu:
    n_used, nnz = gpu.load_cell_data_ptrs(d.locus.data_ptr(), d.cell.data_ptr(), d.ref.data_ptr(), d.alt.data_ptr(), d.nnz,
                                          d.n_loci, d.n_cells)
    del d
    torch.cuda.empty_cache()
    out.update(loci_used=n_used, entries=nnz)
    init = np.stack([soup.init_cluster_centers_uniform(n_used, k, 4, r) for r in range(restarts)])
    gpu.soup_em(init[0])  # warm-up (allocations)
    t = time.perf_counter()
    one = gpu.soup_em(init[0])
    dt = time.perf_counter() - t
    out["single"] = dict(iterations=one["iterations"], seconds=dt, us_per_iteration=1e6 * dt / one["iterations"],
                         entries_per_s=nnz * one["iterations"] / dt, loss=one["loss"])
    again = gpu.soup_em(init[0])
    t = time.perf_counter()
    many = gpu.soup_restarts(init)
    dt = time.perf_counter() - t
    itb = int(many["iterations"].sum())
    out["batched"] = dict(restarts=restarts, restart_iterations=itb, seconds=dt, us_per_restart_iteration=1e6 * dt / itb,
                          entries_per_s=nnz * itb / dt, best=many["best"], iterations=[int(x) for x in many["iterations"]])
host_loss = float(logsumexp(one["log_probs"], axis=1).sum())
out["properties"] = dict(
    loss_equals_sum_of_logsumexp_rel=abs(host_loss - one["loss"]) / abs(one["loss"]),
    centres_within_clamps=bool((one["centers"] >= np.float64(np.float32(0.01))).all() and (one["centers"] <= np.float64(np.float32(0.99))).all()),
    second_run_bit_identical=bool(np.array_equal(one["log_probs"], again["log_probs"]) and one["loss"] == again["loss"]),
    batched_restart0_equals_single=bool(many["losses"][0] == one["loss"] and int(many["iterations"][0]) == one["iterations"]),
    assignment_counts=np.bincount(np.argmax(one["log_probs"], axis=1), minlength=k).tolist())
out["properties"]["ok"] = bool(out["properties"]["loss_equals_sum_of_logsumexp_rel"] <= 1e-12 and out["properties"]["centres_within_clamps"]
                               and out["properties"]["second_run_bit_identical"] and out["properties"]["batched_restart0_equals_single"])
print(json.dumps(out))
