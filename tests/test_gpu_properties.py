"""Size-independent properties at the bench workload's FULL size (C2: 10 k cells x 100 k loci, ~1e7 stored
entries) — the oracle is too slow to run there in a test, so the GPU path is checked against conservation laws
of the algorithm (a checksum of checksums) and against itself under a cell permutation."""
import numpy as np
import pytest

from cellector_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2():
    return synth.generate(synth.CONFIGS["C2"])


def _run(d, perm=None):
    from cellector_b200 import Cellector
    cell = d.cell if perm is None else perm[d.cell].astype(np.uint32)
    with Cellector(0) as g:
        n_loci, nnz = g.load_cell_data(d.locus, cell, d.ref, d.alt, d.n_loci, d.n_cells)
        ids = g.locus_ids()
        stats, first_c, first_l = [], None, None
        for _ in range(50):
            st = g.compute_new_excluded(5.0)
            stats.append(st.as_dict())
            if first_c is None:
                first_c, first_l = g.cell_vectors(), g.locus_vectors()
            if not st.any_change:
                break
        return dict(n_loci=n_loci, nnz=nnz, ids=ids, stats=stats, first_c=first_c, first_l=first_l, last_c=g.cell_vectors(),
                    last_l=g.locus_vectors(), post=g.calculate_posteriors())


def test_conservation_laws_at_full_size(c2):
    d = c2
    r = _run(d)
    # locus filter (load_data.rs:268-273) against numpy on the raw COO
    cr = np.bincount(d.locus[d.ref > 0], minlength=d.n_loci)
    ca = np.bincount(d.locus[d.alt > 0], minlength=d.n_loci)
    used = (cr >= 4) & (ca >= 4)
    assert r["n_loci"] == int(used.sum()) and np.array_equal(r["ids"], np.flatnonzero(used).astype(np.uint32))
    kept = used[d.locus]
    assert r["nnz"] == int(kept.sum())
    # iteration 0 scores every kept entry once: per cell and per locus
    assert np.array_equal(r["first_c"]["n_loci"], np.bincount(d.cell[kept], minlength=d.n_cells).astype(np.uint32))
    fl = r["first_l"]
    per_locus = np.bincount(d.locus[kept], minlength=d.n_loci)[used]
    assert np.array_equal(fl["cells_minority"].astype(np.int64) + fl["cells_majority"], per_locus)
    ref_tot = np.bincount(d.locus[kept], weights=d.ref[kept].astype(np.float64), minlength=d.n_loci)[used].astype(np.uint64)
    alt_tot = np.bincount(d.locus[kept], weights=d.alt[kept].astype(np.float64), minlength=d.n_loci)[used].astype(np.uint64)
    assert np.array_equal(fl["ref_minority"] + fl["ref_majority"], ref_tot)
    assert np.array_equal(fl["alt_minority"] + fl["alt_majority"], alt_tot)
    # the two ways of adding up the log likelihood of iteration 0 agree: over cells and over loci (checksum of checksums)
    total_cells = float(np.sum(r["first_c"]["ll"]))
    total_loci = float(np.sum(fl["ll_minority"]) + np.sum(fl["ll_majority"]))
    assert abs(total_cells - total_loci) <= 1e-9 * abs(total_cells)
    # excluded set bookkeeping across iterations
    n_ex = 0
    for st in r["stats"]:
        n_ex += st["n_new_excluded"] - st["n_rescued"]
        assert st["n_excluded"] == n_ex
        assert st["q1"] <= st["median"] <= st["q3"] and st["threshold"] <= st["q1"]
    assert int(r["last_c"]["excluded"].sum()) == n_ex
    assert np.array_equal(r["last_c"]["excluded"] == 1, r["last_c"]["norm"] < r["stats"][-1]["threshold"])
    # the planted 2 % minority genotype is found
    found = r["post"]["label"] == 0
    assert (found == d.is_minority).mean() > 0.995


def test_cell_permutation_at_full_size(c2):
    d = c2
    rng = np.random.default_rng(7)
    perm = rng.permutation(d.n_cells)
    # file order must stay sorted by (locus, cell) for the summation order inside a row to stay "file order"
    order = np.lexsort((perm[d.cell], d.locus))
    dp = synth.SynthData(n_cells=d.n_cells, n_loci=d.n_loci, locus=d.locus[order], cell=d.cell[order], ref=d.ref[order],
                         alt=d.alt[order], is_minority=d.is_minority)
    a, b = _run(d), _run(dp, perm)
    assert a["n_loci"] == b["n_loci"] and a["nnz"] == b["nnz"] and len(a["stats"]) == len(b["stats"])
    for k in ("n_loci", "excluded"):
        assert np.array_equal(a["last_c"][k], b["last_c"][k][perm])
    np.testing.assert_allclose(b["last_c"]["ll"][perm], a["last_c"]["ll"], rtol=1e-12, atol=1e-12)
    for k in ("label", "anomaly", "qual"):
        assert np.array_equal(a["post"][k], b["post"][k][perm])
    for k in ("cells_minority", "cells_majority", "ref_minority", "alt_minority", "ref_majority", "alt_majority", "loci_used"):
        assert np.array_equal(a["last_l"][k], b["last_l"][k])
    for sa, sb in zip(a["stats"], b["stats"]):
        assert sa["threshold"] == sb["threshold"] or abs(sa["threshold"] - sb["threshold"]) <= 1e-12 * abs(sa["threshold"])
