"""Pins for the CPU oracle (the reference ships none — "parity unpinned"):
known-answer tests of the statrs numerics against mpmath / scipy, hand-worked quantiles, an independent
numpy/scipy statement of the algorithm on small matrices, and the committed golden vectors."""
import ctypes as C
import json
import os

import mpmath
import numpy as np
import pytest
from scipy import stats as sps

from cellector_b200 import synth
from tests import pyref
from tests.oracle import Oracle

mpmath.mp.dps = 40


def test_ln_gamma_vs_mpmath(oracle_lib):
    xs = [1.0, 1.5, 2.0, 3.0, 4.5, 10.0, 37.25, 100.0, 1234.5, 1e4, 123456.75, 1e6, 3e7]
    for x in xs:
        want = float(mpmath.loggamma(x))
        got = oracle_lib.orc_ln_gamma(x)
        assert abs(got - want) <= 5e-15 * max(1.0, abs(want)), (x, got, want)
    # the Lanczos series is not exact at the integers: statrs gives -4.44e-16 at 1.0 (and so does the reference)
    assert oracle_lib.orc_ln_gamma(1.0) == -4.440892098500626e-16
    # reflection branch (x < 0.5)
    assert abs(oracle_lib.orc_ln_gamma(0.25) - float(mpmath.loggamma(0.25))) < 1e-14


def test_ln_factorial_and_binomial(oracle_lib):
    assert oracle_lib.orc_ln_factorial(0) == 0.0
    assert oracle_lib.orc_ln_factorial(5) == np.log(120.0)
    for n, k in [(1, 0), (3, 1), (10, 3), (100, 50), (170, 85), (171, 3), (5000, 2500)]:
        want = float(mpmath.log(mpmath.binomial(n, k)))
        got = oracle_lib.orc_ln_binomial(n, k)
        assert abs(got - want) <= 2e-12 * max(1.0, abs(want)), (n, k, got, want)
    assert oracle_lib.orc_ln_binomial(3, 5) == -np.inf
    # n <= 170 goes through ln(FCACHE[n]) — exactly ln of the f64 product table
    f = 1.0
    for i in range(1, 171):
        f *= float(i)
    assert oracle_lib.orc_ln_factorial(170) == np.log(f)


def test_log_beta_binomial_pmf_vs_scipy(oracle_lib):
    rng = np.random.default_rng(0)
    for _ in range(200):
        a, r = int(rng.integers(0, 6)), int(rng.integers(0, 6))
        alpha, beta = float(rng.integers(1, 3000)), float(rng.integers(1, 3000))
        lbc = oracle_lib.orc_ln_binomial(a + r, a)
        got = oracle_lib.orc_log_beta_binomial_pmf(float(a), float(r), alpha, beta, lbc)
        want = sps.betabinom.logpmf(a, a + r, alpha, beta)
        assert abs(got - want) <= 1e-9 * max(1.0, abs(want)) + 1e-11, (a, r, alpha, beta, got, want)
    # zero-total entry is exactly 0.0: (0 + den) - den
    assert oracle_lib.orc_log_beta_binomial_pmf(0.0, 0.0, 17.0, 5.0, 0.0) == 0.0


def test_expected_log_likelihood_definition(oracle_lib):
    # stats.rs:8-33: E = log sum_k pmf_k^2 (left-to-right 2-argument logsumexp), var = sum_k pmf_k (ll_k - E)^2
    for total, alpha, beta in [(0, 3.0, 4.0), (1, 10.0, 2.0), (2, 50.0, 70.0), (3, 5.0, 900.0), (12, 33.0, 21.0)]:
        e, v = C.c_double(), C.c_double()
        oracle_lib.orc_expected_log_beta_binomial_pmf(total, alpha, beta, C.byref(e), C.byref(v))
        ks = np.arange(total + 1)
        ll = sps.betabinom.logpmf(ks, total, alpha, beta)
        want_e = np.log(np.sum(np.exp(2 * ll)))
        want_v = np.sum(np.exp(ll) * (ll - want_e) ** 2)
        assert abs(e.value - want_e) < 1e-9 * max(1, abs(want_e)) + 1e-10
        assert abs(v.value - want_v) < 1e-8 * max(1, abs(want_v)) + 1e-10


def test_quantiles_hand_worked(oracle_lib):
    def q(x, tau):
        a = np.asarray(x, np.float64)
        return oracle_lib.orc_quantile(a.ctypes.data_as(C.c_void_p), len(a), tau)

    def med(x):
        a = np.asarray(x, np.float64)
        return oracle_lib.orc_median(a.ctypes.data_as(C.c_void_p), len(a))
    # R-8: h = (N + 1/3) tau + 1/3.  N=4, tau=.25: h=1.41667 -> x1 + .41667 (x2 - x1)
    assert q([4, 1, 3, 2], 0.25) == pytest.approx(1 + (13 / 12 + 1 / 3 - 1) * 1, abs=1e-15)
    assert q([4, 1, 3, 2], 0.75) == pytest.approx(3 + ((4 + 1 / 3) * 0.75 + 1 / 3 - 3), abs=1e-15)
    # N=5, tau=.25: h = 1.6667 ; tau=.75: h = 4.3333
    assert q([10, 20, 30, 40, 50], 0.25) == pytest.approx(10 + (2 / 3) * 10, abs=1e-12)
    assert q([50, 40, 30, 20, 10], 0.75) == pytest.approx(40 + (1 / 3) * 10, abs=1e-12)
    # N=1 and N=2 hit the min/max edge rules
    assert q([7.0], 0.25) == 7.0 and q([7.0], 0.75) == 7.0
    assert q([1.0, 2.0], 0.25) == 1.0 and q([1.0, 2.0], 0.75) == 2.0
    assert med([3, 1, 2]) == 2.0 and med([4, 1, 3, 2]) == 2.5
    rng = np.random.default_rng(1)
    for n in (10, 11, 100, 1001):
        x = rng.normal(size=n)
        for tau in (0.25, 0.75):
            assert q(x, tau) == pytest.approx(pyref.quantile_r8(x, tau), abs=1e-15)
            # scipy's "median_unbiased" is the same R-8 rule
            assert q(x, tau) == pytest.approx(np.quantile(x, tau, method="median_unbiased"), abs=1e-12)
        assert med(x) == np.median(x)


def _check_against_pyref(data, iters=6, **kw):
    orc = Oracle()
    ref = pyref.PyRef(data.locus, data.cell, data.ref, data.alt, data.n_loci, data.n_cells,
                      kw.get("min_ref", 4), kw.get("min_alt", 4))
    n_used, nnz_used = orc.load_cell_data(data.locus, data.cell, data.ref, data.alt, data.n_loci, data.n_cells,
                                          kw.get("min_ref", 4), kw.get("min_alt", 4))
    assert n_used == ref.L and nnz_used == len(ref.l)
    assert np.array_equal(orc.locus_ids(), ref.locus_ids)
    for it in range(iters):
        st = orc.compute_new_excluded(5.0)
        want = ref.iteration(5.0)
        cv, lv = orc.cell_vectors(), orc.locus_vectors()
        assert np.array_equal(cv["n_loci"], want["n_loci"])
        assert np.array_equal(cv["excluded"].astype(bool), want["excluded"])
        np.testing.assert_allclose(cv["ll"], want["ll"], rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(cv["norm"], want["norm"], rtol=1e-9, atol=1e-12)
        for k in ("median", "q1", "q3", "threshold"):
            assert getattr(st, k) == pytest.approx(want[k], rel=1e-9, abs=1e-12)
        assert (st.n_new_excluded, st.n_rescued) == (want["n_new"], want["n_rescued"])
        for k in ("cells_minority", "cells_majority", "ref_minority", "alt_minority", "ref_majority", "alt_majority"):
            assert np.array_equal(lv[k], want[k].astype(lv[k].dtype)), k
        assert np.array_equal(lv["loci_used"].astype(bool), want["loci_used"])
        np.testing.assert_allclose(lv["ll_minority"], want["ll_minority"], rtol=1e-9, atol=1e-8)
        np.testing.assert_allclose(lv["ll_majority"], want["ll_majority"], rtol=1e-9, atol=1e-8)
        if not st.any_change:
            break
    po, pw = orc.calculate_posteriors(0.999, 30), ref.posteriors(0.999, 30)
    np.testing.assert_allclose(po["ll_major"], pw["ll_major"], rtol=1e-9, atol=1e-8)
    np.testing.assert_allclose(po["ll_minor"], pw["ll_minor"], rtol=1e-9, atol=1e-8)
    np.testing.assert_allclose(po["posterior"], pw["posterior"], rtol=1e-5, atol=1e-12)
    assert np.array_equal(po["label"], pw["label"])
    assert np.array_equal(po["anomaly"], pw["anomaly"])
    orc.close()


def test_oracle_vs_independent_python_small():
    _check_against_pyref(synth.generate(synth.SynthConfig(n_cells=400, n_loci=1500, density=0.05, minority_frac=0.1, seed=2)))


def test_oracle_vs_independent_python_deep_and_doublets():
    _check_against_pyref(synth.generate(synth.SynthConfig(n_cells=500, n_loci=1200, density=0.05, minority_frac=0.1, seed=4,
                                                          deep_frac=0.05, deep_mean=50.0, doublet_frac=0.05)))


def test_oracle_edge_matrix():
    from tests.test_gpu_parity import _edge_matrix
    _check_against_pyref(_edge_matrix())


def test_mtx_text_roundtrip(tmp_path, oracle_lib):
    d = synth.generate(synth.SynthConfig(n_cells=50, n_loci=80, density=0.1, seed=6))
    p = synth.write_mtx_pair(d, str(tmp_path))
    nl, nc, nnz = C.c_uint32(), C.c_uint32(), C.c_uint64()
    bufs = [np.zeros(d.nnz, np.uint32) for _ in range(4)]
    rc = oracle_lib.orc_read_mtx_pair(p["alt.mtx"].encode(), p["ref.mtx"].encode(), C.byref(nl), C.byref(nc), C.byref(nnz),
                                      *[b.ctypes.data_as(C.c_void_p) for b in bufs], d.nnz)
    assert rc == 0 and (nl.value, nc.value, nnz.value) == (d.n_loci, d.n_cells, d.nnz)
    for got, want in zip(bufs, (d.locus, d.cell, d.ref, d.alt)):
        assert np.array_equal(got, want)


GOLDEN_DIR = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("fixture", ["oracle_c1_slice.json", "oracle_c1.json", "oracle_c2.json"])
def test_oracle_matches_committed_golden(fixture):
    """The fixtures were written by tests/golden/make_golden.py from this oracle: they freeze the oracle's behaviour (a
    later change of the restatement cannot silently move the target), at a small deep case and at BASELINE.json
    configs[0] / configs[1] in full."""
    g = json.load(open(os.path.join(GOLDEN_DIR, fixture)))
    d = synth.generate(synth.SynthConfig(**g["config"]))
    orc = Oracle()
    assert list(orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)) == g["loaded"]
    for want in g["iterations"]:
        st = orc.compute_new_excluded(5.0).as_dict()
        for k, v in want["stats"].items():
            assert st[k] == v, (k, st[k], v)   # bit-exact: same code, same libm
        cv, lv = orc.cell_vectors(), orc.locus_vectors()
        assert float(cv["ll"].sum()) == want["ll_sum"]
        assert float(cv["expected_ll"].sum()) == want["expected_ll_sum"]
        assert int(cv["n_loci"].sum()) == want["n_loci_sum"]
        assert np.flatnonzero(cv["excluded"]).tolist() == want["excluded"]
        assert cv["ll"][want["cell_idx"]].tolist() == want["cell_ll"]
        assert lv["ll_minority"][want["locus_idx"]].tolist() == want["locus_ll_minority"]
        assert lv["ll_majority"][want["locus_idx"]].tolist() == want["locus_ll_majority"]
    p = orc.calculate_posteriors()
    assert np.bincount(p["label"], minlength=4).tolist() == g["label_counts"]
    assert float(p["ll_minor"].sum()) == g["ll_minor_sum"]
    orc.close()
