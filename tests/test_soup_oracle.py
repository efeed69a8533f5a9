"""CPU checks of the souporcell oracle (oracle/souporcell_oracle.cc): its loader rules against hand-worked cases, its
f64 statement of the EM against an independent dense numpy statement of main.rs:109-280, and its f32 (reference
arithmetic) run against the f64 one."""
import numpy as np
import pytest

from cellector_b200 import soup, synth
from tests.soup_oracle import SoupOracle


def _dense(d, locus_ids):
    """souporcell's loaded matrix as dense [cells][used loci] ref / alt (last line of a coordinate wins)"""
    idx = -np.ones(d.n_loci, np.int64)
    idx[locus_ids] = np.arange(locus_ids.shape[0])
    R = np.zeros((d.n_cells, locus_ids.shape[0]), np.int64)
    A = np.zeros_like(R)
    keep = idx[d.locus] >= 0
    R[d.cell[keep], idx[d.locus[keep]]] = d.ref[keep]   # numpy fancy assignment: the last write wins
    A[d.cell[keep], idx[d.locus[keep]]] = d.alt[keep]
    return R, A


def _numpy_em(R, A, centers_init):
    from scipy.special import gammaln, logsumexp
    N, L = R.shape
    k = centers_init.shape[0]
    T = R + A
    lbc = (gammaln(T + 1.0) - gammaln(A + 1.0) - gammaln(R + 1.0)).astype(np.float32).astype(np.float64)
    lbc_sum = np.where(T > 0, lbc, 0.0).sum(axis=1)
    total_alleles = T.sum(axis=1).astype(np.float32)
    cen = centers_init.astype(np.float64)
    log_prior = np.log(np.float64(np.float32(1.0) / np.float32(k)))
    limit = np.float64(np.float32(0.01) * np.float32(N))
    last, it, trace = -np.inf, 0, []
    for ts in range(9):
        change = 10000.0
        while change > limit and it < 1000:
            lp = log_prior + lbc_sum[:, None] + A @ np.log(cen).T + R @ np.log(1.0 - cen).T
            loss = logsumexp(lp, axis=1).sum()
            temp = np.maximum(total_alleles / (np.float32(20.0) * np.float32(2.0) ** np.float32(ts)), np.float32(1.0))
            if ts == 8:
                temp = np.ones_like(temp)
            nlp = lp / temp.astype(np.float64)[:, None]
            p = np.exp(nlp - logsumexp(nlp, axis=1)[:, None])
            cen = np.clip((1.0 + p.T @ A) / (2.0 + p.T @ T), np.float64(np.float32(0.01)), np.float64(np.float32(0.99)))
            change, last = loss - last, loss
            trace.append(loss)
            it += 1
    return dict(log_probs=lp, centers=cen, loss=loss, iterations=it, loss_trace=np.array(trace))


def test_loader_rules():
    # locus 0: kept (2 ref cells, 2 alt cells); cell 0 appears twice at locus 0: the LAST line (5, 1) wins, but both
    # lines count for the filter (main.rs:418-421); locus 1 fails min_alt; zero-total entry at (0, 2) is skipped
    locus = np.array([0, 0, 0, 0, 1, 1], np.uint32)
    cell = np.array([0, 1, 0, 2, 0, 1], np.uint32)
    ref = np.array([1, 0, 5, 0, 3, 2], np.uint32)
    alt = np.array([0, 2, 1, 0, 0, 1], np.uint32)
    with SoupOracle() as o:
        n_used, n_entries = o.load_cell_data(locus, cell, ref, alt, 2, 3, min_ref=2, min_alt=2)
        assert n_used == 1 and o.locus_ids().tolist() == [0]
        assert n_entries == 2          # (cell 0: 5/1) and (cell 1: 0/2); the zero-total entry of cell 2 is gone
        res = o.soup_em(np.full((1, 1), 0.25, np.float32), precise=True)
        # k = 1: log_probs = ln(1) + ln C(n, a) (as f32) + a ln c + r ln(1 - c) with the centre of the last E step
        assert res["iterations"] >= 9
    with SoupOracle() as o:  # UMI minima (main.rs:440): locus 0 has 6 ref and 3 alt UMIs
        assert o.load_cell_data(locus, cell, ref, alt, 2, 3, 2, 2, min_ref_umis=7)[0] == 0
        assert o.load_cell_data(locus, cell, ref, alt, 2, 3, 2, 2, min_ref_umis=6, min_alt_umis=3)[0] == 1


CASES = {
    "two": (synth.SynthConfig(n_cells=300, n_loci=1500, density=0.05, minority_frac=0.3, seed=41), 2),
    "three_deep": (synth.SynthConfig(n_cells=250, n_loci=1200, density=0.06, minority_frac=0.4, seed=42, deep_frac=0.05,
                                     deep_mean=30.0, doublet_frac=0.05), 3),
}


@pytest.mark.parametrize("name", list(CASES))
def test_precise_oracle_matches_numpy_statement(name):
    cfg, k = CASES[name]
    d = synth.generate(cfg)
    with SoupOracle() as o:
        n_used, _ = o.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        assert 0 < n_used < d.n_loci
        init = soup.init_cluster_centers_uniform(n_used, k, seed=4)
        res = o.soup_em(init, precise=True)
        R, A = _dense(d, o.locus_ids())
    ref = _numpy_em(R, A, init)
    assert res["iterations"] == ref["iterations"] and 9 <= res["iterations"] < 1000
    np.testing.assert_allclose(res["loss_trace"], ref["loss_trace"], rtol=1e-10)
    np.testing.assert_allclose(res["log_probs"], ref["log_probs"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(res["centers"], ref["centers"], rtol=1e-9)
    assert res["loss"] == res["loss_trace"][-1]


@pytest.mark.parametrize("name", list(CASES))
def test_reference_arithmetic_is_within_f32_rounding_of_the_f64_statement(name):
    cfg, k = CASES[name]
    d = synth.generate(cfg)
    with SoupOracle() as o:
        n_used, _ = o.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        init = soup.init_cluster_centers_uniform(n_used, k, seed=4)
        f64 = o.soup_em(init, precise=True)
        f32 = o.soup_em(init, precise=False)
    assert abs(f32["iterations"] - f64["iterations"]) <= 2
    np.testing.assert_allclose(f32["loss"], f64["loss"], rtol=1e-4)
    np.testing.assert_allclose(f32["log_probs"], f64["log_probs"], rtol=2e-3, atol=1e-2)
    clear = np.abs(np.sort(f64["log_probs"], axis=1)[:, -1] - np.sort(f64["log_probs"], axis=1)[:, -2]) > 0.1
    assert clear.mean() > 0.9
    assert (np.argmax(f32["log_probs"], axis=1) == np.argmax(f64["log_probs"], axis=1))[clear].all()


GOLDEN = ["soup_small_k3.json", "soup_c1_k2.json"]


def load_golden(name):
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name)))
    d = synth.generate(synth.SynthConfig(**g["config"]))
    from tests.soup_rng import souporcell_initial_centers
    init = souporcell_initial_centers(g["seed"], 1, 1, g["k"], g["loaded"][0])[0]
    assert float(init.astype(np.float64).sum()) == g["init_sum"]
    return g, d, init


def check_against_golden(g, res, rtol):
    """res = a run in the f64 statement (the oracle: rtol 0 = bit for bit; the CUDA path: 1e-9)"""
    w = g["f64"]
    assert res["iterations"] == w["iterations"]
    close = (lambda a, b: np.testing.assert_array_equal(np.asarray(a), np.asarray(b))) if rtol == 0 else \
        (lambda a, b: np.testing.assert_allclose(np.asarray(a), np.asarray(b), rtol=rtol, atol=1e-12))
    close(res["loss"], w["loss"])
    close(res["log_probs"].sum(axis=0), w["log_probs_sum"])
    close(res["log_probs"][w["cell_idx"]], w["cell_log_probs"])
    close(res["centers"][:, w["locus_idx"]], w["locus_centers"])
    assert np.bincount(np.argmax(res["log_probs"], axis=1), minlength=g["k"]).tolist() == w["assignment_counts"]


@pytest.mark.parametrize("name", GOLDEN)
def test_oracle_reproduces_the_frozen_outputs(name):
    g, d, init = load_golden(name)
    with SoupOracle() as o:
        assert list(o.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)) == g["loaded"]
        res = o.soup_em(init, precise=True)
        np.testing.assert_array_equal(res["loss_trace"], np.asarray(g["f64"]["loss_trace"]))
        check_against_golden(g, res, 0)
        f32 = o.soup_em(init, precise=False)
    assert f32["iterations"] == g["f32"]["iterations"] and f32["loss"] == g["f32"]["loss"]
    assert np.bincount(np.argmax(f32["log_probs"], axis=1), minlength=g["k"]).tolist() == g["f32"]["assignment_counts"]


def test_python_chacha20_block_against_rfc7539():
    """tests/soup_rng.py states the generator of the initial centers a second time (plain Python): its block function
    against the RFC 7539 section 2.3.2 vector, and the thread-major seeding order of souporcell_main (main.rs:62-67)"""
    from tests.soup_rng import StdRng07, chacha20_block, souporcell_initial_centers
    key = bytes(range(32))
    state = [0x61707865, 0x3320646E, 0x79622D32, 0x6B206574] + [int.from_bytes(key[4 * i:4 * i + 4], "little") for i in range(8)] + \
            [1, 0x09000000, 0x4A000000, 0]
    out = chacha20_block(state)
    assert [f"{w:08x}" for w in out[:4]] == ["e4e7f110", "15593bd1", "1fdd0f50", "c47120a3"]
    assert [f"{w:08x}" for w in out[12:]] == ["d19c12b5", "b94e16de", "e883d0cb", "4e3c50a2"]
    # two threads, three restarts -> ceil(3 / 2) = 2 runs per thread; thread 1's generator is seeded from the SECOND
    # 32 u8 draws of the main generator
    c = souporcell_initial_centers(seed=4, threads=2, restarts=3, k=2, n_loci=5)
    assert c.shape == (4, 2, 5) and c.dtype == np.float32 and (c >= np.float32(0.0001)).all() and (c <= np.float32(0.9999)).all()
    main = StdRng07(bytes([4]) * 32)
    main.new_seed()
    t1 = StdRng07(main.new_seed())
    first = min(max(t1.gen_f32(), np.float32(0.0001)), np.float32(0.9999))
    assert c[2, 0, 0] == first
