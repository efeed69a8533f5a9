"""Shared parity checks: CUDA path (cellector_b200.Cellector) vs the CPU oracle on the same COO input.

Bars (BASELINE.json north_star): locus filter, per-locus integer counts, excluded sets and final labels
bit-exact; per-cell log likelihoods / posteriors within RTOL = 1e-9 relative (fp64)."""
from __future__ import annotations

import numpy as np

RTOL = 1e-9
# absolute floor for values that are sums of O(1) terms and may cancel to ~0 (e.g. a cell with one zero-total entry)
ATOL = 1e-12


def assert_close(a, b, what, rtol=RTOL, atol=ATOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, what
    both_nan = np.isnan(a) & np.isnan(b)
    err = np.abs(a - b)
    tol = atol + rtol * np.abs(b)
    bad = ~(both_nan | (err <= tol) | ((a == b)))
    if bad.any():
        i = int(np.argmax(np.where(bad, err / np.maximum(np.abs(b), 1e-300), 0)))
        raise AssertionError(f"{what}: {bad.sum()} of {a.size} outside rtol={rtol}; worst idx {i}: got {a.flat[i]!r} want {b.flat[i]!r}")


def assert_equal(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    if not np.array_equal(a, b):
        i = int(np.flatnonzero(a != b)[0])
        raise AssertionError(f"{what}: {np.count_nonzero(a != b)} of {a.size} differ; first idx {i}: got {a.flat[i]} want {b.flat[i]}")


def compare_iteration(gs, os_, gpu, orc, it):
    tag = f"iter {it}"
    for k in ("any_change", "iteration", "n_new_excluded", "n_rescued", "n_excluded", "n_loci_filtered", "n_loci_in_use"):
        assert getattr(gs, k) == getattr(os_, k), f"{tag} stats.{k}: got {getattr(gs, k)} want {getattr(os_, k)}"
    for k in ("median", "q1", "q3", "threshold"):
        assert_close(getattr(gs, k), getattr(os_, k), f"{tag} stats.{k}")
    assert_close(gs.iqr, os_.iqr, f"{tag} stats.iqr", rtol=1e-7)  # difference of two 1e-9-close numbers
    gc, oc = gpu.cell_vectors(), orc.cell_vectors()
    assert_equal(gc["n_loci"], oc["n_loci"], f"{tag} n_loci")
    assert_equal(gc["excluded"], oc["excluded"], f"{tag} excluded set")
    assert_close(gc["ll"], oc["ll"], f"{tag} log_likelihood")
    assert_close(gc["expected_ll"], oc["expected_ll"], f"{tag} expected_log_likelihood")
    assert_close(gc["norm"], oc["norm"], f"{tag} normalized log_likelihood")
    gl, ol = gpu.locus_vectors(), orc.locus_vectors()
    for k in ("cells_minority", "cells_majority", "ref_minority", "alt_minority", "ref_majority", "alt_majority", "loci_used"):
        assert_equal(gl[k], ol[k], f"{tag} locus {k}")
    assert_close(gl["ll_minority"], ol["ll_minority"], f"{tag} locus ll_minority")
    assert_close(gl["ll_majority"], ol["ll_majority"], f"{tag} locus ll_majority")


def run_parity(gpu, orc, data, min_ref=4, min_alt=4, iqr=5.0, max_iter=30, posterior_threshold=0.999, min_loci_used=30):
    """Load the same matrix into both, iterate to convergence comparing everything every iteration."""
    g = gpu.load_cell_data(data.locus, data.cell, data.ref, data.alt, data.n_loci, data.n_cells, min_ref, min_alt)
    o = orc.load_cell_data(data.locus, data.cell, data.ref, data.alt, data.n_loci, data.n_cells, min_ref, min_alt)
    assert g == o, f"(n_loci_used, nnz_used): got {g} want {o}"
    assert_equal(gpu.locus_ids(), orc.locus_ids(), "locus_ids")
    n_it = 0
    for it in range(max_iter):
        gs, os_ = gpu.compute_new_excluded(iqr), orc.compute_new_excluded(iqr)
        compare_iteration(gs, os_, gpu, orc, it)
        n_it += 1
        if not os_.any_change:
            break
    gp, op = gpu.calculate_posteriors(posterior_threshold, min_loci_used), orc.calculate_posteriors(posterior_threshold, min_loci_used)
    for k in ("label", "anomaly", "qual"):
        assert_equal(gp[k], op[k], f"posterior {k}")
    for k in ("ll_major", "ll_minor"):
        assert_close(gp[k], op[k], f"posterior {k}")
    # posteriors are exp() of differences of ~1e3-magnitude log likelihoods: 1e-9 relative on those is an absolute
    # error of ~1e-6 in the exponent, so compare in log space at the same bar and directly where not saturated
    for k in ("posterior", "doublet_posterior"):
        a, b = gp[k], op[k]
        with np.errstate(divide="ignore"):
            la, lb = np.log(a), np.log(b)
        scale = np.maximum(np.abs(op["ll_major"]), np.abs(op["ll_minor"]))
        ok = (a == b) | (np.abs(la - lb) <= 4 * RTOL * scale + 1e-12)
        assert ok.all(), f"posterior {k}: {np.count_nonzero(~ok)} cells outside tolerance"
    return n_it
