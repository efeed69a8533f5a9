"""GPU parity tests proper: every call goes through the C ABI of libcellector_cuda.so and is checked against
the CPU oracle on the same seeded input (sizes the oracle finishes in seconds)."""
import os

import numpy as np
import pytest

from cellector_b200 import synth
from tests.parity import assert_equal, run_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def libs(oracle_lib):
    from cellector_b200 import api
    return api.load_library(), oracle_lib


def _pair(libs, flags=0):
    from cellector_b200 import Cellector
    from tests.oracle import Oracle
    return Cellector(0, flags, libs[0]), Oracle(libs[1])


CASES = {
    "small": synth.SynthConfig(n_cells=600, n_loci=3000, density=0.03, minority_frac=0.08, seed=3),
    "c1_slice": synth.SynthConfig(n_cells=2000, n_loci=8000, density=0.02, minority_frac=0.05, seed=1),
    "deep": synth.SynthConfig(n_cells=800, n_loci=2500, density=0.04, minority_frac=0.1, seed=5, deep_frac=0.05, deep_mean=60.0),
    "skewed": synth.SynthConfig(n_cells=3000, n_loci=6000, density=0.01, minority_frac=0.05, seed=7, row_sigma=2.0),
    "doublets": synth.SynthConfig(n_cells=1500, n_loci=5000, density=0.03, minority_frac=0.1, seed=9, doublet_frac=0.05),
    "big_select": synth.SynthConfig(n_cells=40000, n_loci=1500, density=0.004, minority_frac=0.05, seed=11),
    # the oracle rescues 2 cells and the -80 rule switches 11 loci off over 5 iterations (searched with the oracle)
    "rescue_filter": synth.SynthConfig(n_cells=700, n_loci=2000, density=0.05, minority_frac=0.1, seed=2, deep_frac=0.08,
                                       deep_mean=100.0, doublet_frac=0.05),
}


@pytest.mark.parametrize("name", list(CASES))
def test_parity_synthetic(libs, name):
    gpu, orc = _pair(libs)
    with gpu, orc:
        n_it = run_parity(gpu, orc, synth.generate(CASES[name]))
        assert n_it >= 1


@pytest.mark.parametrize("flags", [16, 1, 1 | 16, 2], ids=["no_graph", "profile_graph", "profile_no_graph", "no_expected"])
def test_parity_flag_variants(libs, flags):
    """CLX_FLAG_NO_GRAPH (plain launches), CLX_FLAG_PROFILE (phase events, also as graph nodes), CLX_FLAG_NO_EXPECTED."""
    gpu, orc = _pair(libs, flags)
    with gpu, orc:
        d = synth.generate(CASES["deep"])
        if flags & 2:
            # expected_ll is not computed: compare everything else
            gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
            orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
            for _ in range(20):
                gs, os_ = gpu.compute_new_excluded(), orc.compute_new_excluded()
                gc, oc = gpu.cell_vectors(), orc.cell_vectors()
                assert_equal(gc["excluded"], oc["excluded"], "excluded")
                assert_equal(gc["n_loci"], oc["n_loci"], "n_loci")
                np.testing.assert_allclose(gc["ll"], oc["ll"], rtol=1e-9, atol=1e-12)
                assert not gc["expected_ll"].any()
                if not os_.any_change:
                    break
        else:
            run_parity(gpu, orc, d)
            if flags & 1:
                ph = gpu.phase_ms()
                assert ph["score"][1] >= 1 and ph["score"][0] > 0.0, ph


def test_parity_min_filters_and_iqr(libs):
    gpu, orc = _pair(libs)
    with gpu, orc:
        run_parity(gpu, orc, synth.generate(CASES["small"]), min_ref=2, min_alt=7, iqr=2.0, min_loci_used=10,
                   posterior_threshold=0.9)


def _edge_matrix():
    """Hand-built matrix with the reference's corner cases: a locus failing min_alt, zero-total
    entries, a duplicated coordinate, a cell with no used loci, counts > 100 and > 65535, unsorted file order."""
    rng = np.random.default_rng(0)
    n_cells, n_loci = 64, 40
    rows = []
    for l in range(n_loci - 4):
        for c in rng.choice(n_cells - 2, size=20, replace=False):
            t = 1 + rng.poisson(0.7)
            a = rng.binomial(t, 0.5 if l % 3 else 0.95)
            rows.append((l, int(c), t - a, a))
    rows += [(n_loci - 4, c, 1, 0) for c in range(10)]          # never any alt -> filtered
    rows += [(n_loci - 3, c, 0, 0) for c in range(12)]          # zero-total only -> filtered (no cell has ref>0)
    rows += [(0, 5, 0, 0), (1, 5, 0, 0)]                         # zero-total entries on used loci are kept
    rows += [(2, 7, 1, 1), (2, 7, 1, 1)]                         # duplicated coordinate kept twice
    rows += [(3, 9, 150, 70), (4, 9, 70000, 3), (5, 10, 2, 66000)]  # > 100 (ln_binomial switch) and > u16
    rows += [(n_loci - 4, n_cells - 1, 3, 0)]                    # last cell only touches a filtered locus
    rng.shuffle(rows)                                            # file order is NOT sorted
    arr = np.array(rows, dtype=np.uint32)
    return synth.SynthData(n_cells=n_cells, n_loci=n_loci, locus=arr[:, 0].copy(), cell=arr[:, 1].copy(),
                           ref=arr[:, 2].copy(), alt=arr[:, 3].copy(), is_minority=np.zeros(n_cells, bool))


def test_parity_edge_cases(libs):
    gpu, orc = _pair(libs)
    with gpu, orc:
        run_parity(gpu, orc, _edge_matrix(), min_ref=4, min_alt=4, min_loci_used=5)


def test_empty_and_all_filtered(libs):
    gpu, orc = _pair(libs)
    with gpu, orc:
        d = _edge_matrix()
        # thresholds nobody meets: zero loci used, every cell has norm 0.0, nobody is excluded
        run_parity(gpu, orc, d, min_ref=1000, min_alt=1000)
    gpu, orc = _pair(libs)
    with gpu, orc:
        e = synth.SynthData(n_cells=5, n_loci=7, locus=np.zeros(0, np.uint32), cell=np.zeros(0, np.uint32),
                            ref=np.zeros(0, np.uint32), alt=np.zeros(0, np.uint32), is_minority=np.zeros(5, bool))
        run_parity(gpu, orc, e)


def test_out_of_range_coordinate_is_an_error(libs):
    gpu, _ = _pair(libs)
    with gpu:
        with pytest.raises(RuntimeError, match="out of range"):
            gpu.load_cell_data([0, 9], [0, 0], [1, 1], [1, 1], n_loci0=5, n_cells=3)
        with pytest.raises(RuntimeError, match="out of range"):  # a cell index, at a locus the filter drops anyway
            gpu.load_cell_data([0, 1], [0, 3], [1, 1], [1, 1], n_loci0=5, n_cells=3)
        with pytest.raises(RuntimeError, match="out of range"):  # a cell index among the kept entries
            gpu.load_cell_data([0] * 9, [0, 1, 2, 0, 1, 2, 0, 1, 7], [2] * 9, [2] * 9, n_loci0=2, n_cells=3)
        with pytest.raises(RuntimeError, match="out of range"):
            gpu.load_cell_data([0], [0], [1], [1], n_loci0=5, n_cells=0)
        gpu.load_cell_data([0, 1], [0, 0], [5, 5], [5, 5], n_loci0=5, n_cells=3)  # the context is still usable


def test_final_locus_counts_and_reset(libs):
    from cellector_b200 import api
    gpu, orc = _pair(libs, flags=8)  # CLX_FLAG_KEEP_COO
    with gpu, orc:
        d = synth.generate(CASES["small"])
        run_parity(gpu, orc, d)
        gm, gj = gpu.final_locus_counts()
        om, oj = orc.final_locus_counts()
        assert_equal(gm, om, "final minority counts")
        assert_equal(gj, oj, "final majority counts")
        # reset + rerun reproduces the first iteration bit for bit (deterministic reductions)
        first = None
        for _ in range(2):
            gpu.reset()
            gpu.compute_new_excluded()
            cv = gpu.cell_vectors()
            if first is None:
                first = cv
            else:
                for k in cv:
                    assert np.array_equal(cv[k], first[k]), k


def test_determinism_bitwise(libs):
    gpu, _ = _pair(libs)
    d = synth.generate(CASES["c1_slice"])
    outs = []
    with gpu:
        for _ in range(2):
            gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
            gpu.run_em()
            outs.append((gpu.cell_vectors(), gpu.locus_vectors(), gpu.calculate_posteriors()))
    for a, b in zip(outs[0], outs[1]):
        for k in a:
            assert np.array_equal(a[k], b[k]), k


def test_packed_load_is_the_same_load(libs):
    """clx_load_coo_packed (ref | alt << 16) must build the very state clx_load_coo builds."""
    gpu, _ = _pair(libs)
    d = synth.generate(CASES["c1_slice"])
    outs = []
    with gpu:
        for packed in (False, True):
            if packed:
                ra = d.ref.astype(np.uint32) | (d.alt.astype(np.uint32) << 16)
                used = gpu.load_cell_data_packed(d.locus, d.cell, ra, d.n_loci, d.n_cells)
            else:
                used = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
            gpu.run_em()
            outs.append((used, gpu.locus_ids(), gpu.cell_vectors(), gpu.locus_vectors(), gpu.calculate_posteriors()))
    assert outs[0][0] == outs[1][0]
    assert np.array_equal(outs[0][1], outs[1][1])
    for a, b in zip(outs[0][2:], outs[1][2:]):
        for k in a:
            assert np.array_equal(a[k], b[k]), k


@pytest.mark.parametrize("fixture", ["oracle_c1_slice.json", "oracle_c1.json", "oracle_c2.json"])
def test_gpu_against_committed_golden(libs, fixture):
    """The CUDA path against the committed oracle outputs with no oracle in the loop — a small deep case and BASELINE.json
    configs[0] / configs[1] at full size: integers and the excluded sets identical, fp64 sums, sampled per-cell and
    per-locus values and quantiles within 1e-9 relative."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", fixture)))
    d = synth.generate(synth.SynthConfig(**g["config"]))
    gpu, _ = _pair(libs)
    close = lambda got, want: np.testing.assert_allclose(np.asarray(got, np.float64), np.asarray(want, np.float64), rtol=1e-9, atol=1e-12)
    with gpu:
        assert list(gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)) == g["loaded"]
        for want in g["iterations"]:
            st = gpu.compute_new_excluded(5.0).as_dict()
            for k, v in want["stats"].items():
                if isinstance(v, float):
                    assert st[k] == pytest.approx(v, rel=1e-9, abs=1e-12), k
                else:
                    assert st[k] == v, (k, st[k], v)
            cv = gpu.cell_vectors()
            assert float(cv["ll"].sum()) == pytest.approx(want["ll_sum"], rel=1e-9)
            assert float(cv["expected_ll"].sum()) == pytest.approx(want["expected_ll_sum"], rel=1e-9)
            assert int(cv["n_loci"].sum()) == want["n_loci_sum"]
            assert np.flatnonzero(cv["excluded"]).tolist() == want["excluded"]
            if "cell_idx" in want:
                lv = gpu.locus_vectors()
                close(cv["ll"][want["cell_idx"]], want["cell_ll"])
                assert cv["n_loci"][want["cell_idx"]].astype(int).tolist() == want["cell_n_loci"]
                close(lv["ll_minority"][want["locus_idx"]], want["locus_ll_minority"])
                close(lv["ll_majority"][want["locus_idx"]], want["locus_ll_majority"])
                assert float(lv["ll_minority"].sum()) == pytest.approx(want["locus_ll_minority_sum"], rel=1e-9, abs=1e-9)
                assert float(lv["ll_majority"].sum()) == pytest.approx(want["locus_ll_majority_sum"], rel=1e-9)
                assert int(lv["cells_minority"].astype(np.int64).sum()) == want["locus_cells_minority_sum"]
                assert int(lv["cells_majority"].astype(np.int64).sum()) == want["locus_cells_majority_sum"]
        p = gpu.calculate_posteriors()
        assert np.bincount(p["label"], minlength=4).tolist() == g["label_counts"]
        assert float(p["ll_minor"].sum()) == pytest.approx(g["ll_minor_sum"], rel=1e-9)


def test_more_than_int32_nonzeros_is_refused(libs):
    """Every cub call of the load takes a 32-bit signed item count: 2^31 - 1 nonzeros on one GPU must be refused before
    anything is read (the arrays handed in here are 4 entries long), with the advice to shard."""
    gpu, _ = _pair(libs)
    four = np.zeros(4, np.uint32)
    ptr = four.ctypes.data
    with gpu:
        for nnz0 in (0x7FFFFFFF, 1 << 33):
            with pytest.raises(RuntimeError, match="2\\^31-2 nonzeros per GPU: shard"):
                gpu.load_cell_data_ptrs(ptr, ptr, ptr, ptr, nnz0, 5, 3)
        # the context is still usable
        gpu.load_cell_data([0, 1], [0, 0], [5, 5], [5, 5], n_loci0=5, n_cells=3)


def test_parity_dense_tiles(libs, monkeypatch):
    """Nearly dense matrix in ONE cell block of 3072 cells: runs of ~440 entries per cell and tile, three groups of
    ~30 batches per warp and tile, i.e. ~90 packets per warp and work item — the descriptor-window refills of
    k_score_tiled (windows of 32 packets) and many-batch groups, which the sparse cases never reach."""
    monkeypatch.setenv("CLX_TILED_NB", "3072")
    gpu, orc = _pair(libs)
    cfg = synth.SynthConfig(n_cells=3072, n_loci=1536, density=2.0, minority_frac=0.05, seed=13)
    with gpu, orc:
        run_parity(gpu, orc, synth.generate(cfg))


def test_parity_gather_scoring_forced(libs, monkeypatch):
    """CLX_SCORE_PATH=gather: the per-row gather kernel scores, the class-count column pass stays (the combination very
    sparse matrices take by themselves)."""
    monkeypatch.setenv("CLX_SCORE_PATH", "gather")
    gpu, orc = _pair(libs)
    with gpu, orc:
        run_parity(gpu, orc, synth.generate(CASES["rescue_filter"]))


def test_parity_gather_rows_in_segments(libs, monkeypatch):
    """Rows of ~7000 entries through the gather kernel: two row segments each (ROW_SEG = 4096), partial sums combined
    by k_score_rows_combine — the path of the deep cells of a skewed matrix."""
    monkeypatch.setenv("CLX_SCORE_PATH", "gather")
    gpu, orc = _pair(libs)
    cfg = synth.SynthConfig(n_cells=160, n_loci=12_000, density=0.6, minority_frac=0.1, seed=23, deep_frac=0.01)
    with gpu, orc:
        run_parity(gpu, orc, synth.generate(cfg))


def test_parity_very_sparse_takes_gather_scoring(libs, capfd, monkeypatch):
    """Many cell blocks (forced small: 24 blocks of 256 cells) over ~60 locus tiles: ~100 MB of table staging per pass for
    a few hundred entries per (cell block, tile) unit, so the load picks the gather kernel by itself (checked on the load
    statistics line).  The shape of C5 (200 000 cells x 2 M loci at density 5e-4) in small."""
    monkeypatch.setenv("CLX_TILED_STATS", "1")
    monkeypatch.setenv("CLX_TILED_NB", "256")
    gpu, orc = _pair(libs)
    cfg = synth.SynthConfig(n_cells=6000, n_loci=100_000, density=0.002, minority_frac=0.06, seed=21)
    with gpu, orc:
        run_parity(gpu, orc, synth.generate(cfg))
    assert "gather scoring" in capfd.readouterr().err


def _worker(case, flags=0, env=None):
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "parity_worker.py"), case, str(flags)], capture_output=True, text=True,
                       timeout=900, cwd=root, env={**os.environ, **(env or {})})
    assert r.returncode == 0 and "PARITY_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_parity_rescued_cells_and_filtered_loci(libs):
    """The case must really exercise the rescue counter and the -80 rule (checked on the oracle's own statistics)."""
    gpu, orc = _pair(libs)
    d = synth.generate(CASES["rescue_filter"])
    with gpu, orc:
        run_parity(gpu, orc, d)
        orc.reset()
        rescued = filtered = 0
        for _ in range(30):
            st = orc.compute_new_excluded(5.0)
            rescued += st.n_rescued
            filtered += st.n_loci_filtered
            if not st.any_change:
                break
        assert rescued >= 1 and filtered >= 2, (rescued, filtered)


def test_parity_gather_kernel_flag(libs):
    """CLX_FLAG_KERNEL_V0: the CSR gather scorer (k_score_rows) instead of the run-stream kernel, EM and posterior pass."""
    gpu, orc = _pair(libs, flags=4)
    with gpu, orc:
        run_parity(gpu, orc, synth.generate(CASES["deep"]))


def test_parity_multi_launch_select():
    """CLX_SELECT_MODE=2: the multi-launch radix select (the path of N > 2.4 M gathered cells), forced on a small case."""
    _worker("c1_slice", env={"CLX_SELECT_MODE": "2"})
    _worker("deep", env={"CLX_SELECT_MODE": "2"})


def test_parity_select_four_keys_per_thread():
    _worker("select_r4")


def test_parity_giant_heavy_entries():
    _worker("giant")
