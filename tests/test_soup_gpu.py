"""souporcell's k-cluster EM on the GPU (clx_soup_em, through the C ABI) against the CPU oracle
(oracle/souporcell_oracle.cc) on the same seeded inputs and the same initial centers.

Tolerances: the CUDA path states the reference's formulas in f64 with fixed-order sums; against the oracle's f64
statement everything must agree to 1e-9 relative (iteration count and loss trace included); against the oracle's f32
statement (the reference's own arithmetic: f32 variables, serial sums) the results must agree within f32 rounding and
every clearly assigned cell must get the same cluster."""
import numpy as np
import pytest

from cellector_b200 import soup, synth
from tests.soup_oracle import SoupOracle

pytestmark = pytest.mark.gpu

CASES = {
    "k2": (synth.SynthConfig(n_cells=1200, n_loci=4000, density=0.04, minority_frac=0.3, seed=51), 2),
    "k3_deep": (synth.SynthConfig(n_cells=900, n_loci=3000, density=0.05, minority_frac=0.4, seed=52, deep_frac=0.05,
                                  deep_mean=40.0, doublet_frac=0.05), 3),
    "k5_skewed": (synth.SynthConfig(n_cells=1500, n_loci=5000, density=0.02, minority_frac=0.2, seed=53, row_sigma=1.5), 5),
    "k8": (synth.SynthConfig(n_cells=500, n_loci=2000, density=0.05, minority_frac=0.5, seed=54), 8),
    "k1": (synth.SynthConfig(n_cells=300, n_loci=1000, density=0.05, minority_frac=0.1, seed=55), 1),
}


def _both(d, k, seed=4):
    from cellector_b200 import Cellector
    with Cellector(0) as gpu, SoupOracle() as orc:
        n_used, _ = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        assert orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)[0] == n_used
        np.testing.assert_array_equal(gpu.locus_ids(), orc.locus_ids())
        init = soup.init_cluster_centers_uniform(n_used, k, seed)
        g = gpu.soup_em(init)
        g2 = gpu.soup_em(init)  # a second run on the same context: same bits (fixed-order sums, no atomics)
        o64 = orc.soup_em(init, precise=True)
        o32 = orc.soup_em(init, precise=False)
    assert g2["iterations"] == g["iterations"] and g2["loss"] == g["loss"]
    np.testing.assert_array_equal(g2["log_probs"], g["log_probs"])
    return g, o64, o32


@pytest.mark.parametrize("name", list(CASES))
def test_soup_em_matches_oracle(name):
    cfg, k = CASES[name]
    g, o64, o32 = _both(synth.generate(cfg), k)
    assert g["iterations"] == o64["iterations"] and 9 <= g["iterations"] < 1000
    np.testing.assert_allclose(g["loss"], o64["loss"], rtol=1e-10)
    np.testing.assert_allclose(g["log_probs"], o64["log_probs"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(g["centers"], o64["centers"], rtol=1e-9)
    # the reference's f32 arithmetic
    assert abs(g["iterations"] - o32["iterations"]) <= 2
    np.testing.assert_allclose(g["loss"], o32["loss"], rtol=1e-4)
    # five clusters on two individuals split arbitrarily: the EM is ill-conditioned there and f32 rounding moves
    # single cells' log-probabilities by a few tenths (oracle f32 against oracle f64: 0.27), not their loss or assignment
    np.testing.assert_allclose(g["log_probs"], o32["log_probs"], rtol=2e-3, atol=0.5 if name == "k5_skewed" else 1e-2)
    if k > 1:
        s = np.sort(g["log_probs"], axis=1)
        clear = (s[:, -1] - s[:, -2]) > 0.1
        assert (np.argmax(g["log_probs"], axis=1) == np.argmax(o32["log_probs"], axis=1))[clear].all()


@pytest.mark.parametrize("name", ["soup_small_k3.json", "soup_c1_k2.json"])
def test_soup_em_against_committed_golden(name):
    """the frozen oracle outputs (tests/golden/make_soup_golden.py), C1 at full size included: 1e-9 on the GPU"""
    from cellector_b200 import Cellector
    from tests.test_soup_oracle import check_against_golden, load_golden
    g, d, init = load_golden(name)
    with Cellector(0) as gpu:
        n_used, _ = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        assert n_used == g["loaded"][0]
        res = gpu.soup_em(init)
    check_against_golden(g, res, 1e-9)
    # and the reference's f32 arithmetic gives the same clustering
    assert np.bincount(np.argmax(res["log_probs"], axis=1), minlength=g["k"]).tolist() == g["f32"]["assignment_counts"]
    assert abs(res["loss"] - g["f32"]["loss"]) <= 1e-5 * abs(g["f32"]["loss"])


def test_soup_restarts_recover_the_two_individuals():
    """souporcell_main (main.rs:62-106): best of a few restarts separates the two simulated individuals"""
    from cellector_b200 import Cellector
    cfg = synth.SynthConfig(n_cells=2000, n_loci=6000, density=0.04, minority_frac=0.25, seed=56)
    d = synth.generate(cfg)
    with Cellector(0) as gpu:
        gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        best = soup.souporcell_main(gpu, k=2, restarts=4, seed=4)
    truth = np.asarray(d.is_minority, dtype=np.int64)
    agree = (best["assignment"] == truth).mean()
    assert max(agree, 1.0 - agree) > 0.98   # cluster ids are arbitrary


@pytest.mark.parametrize("k,n_restarts", [(2, 11), (3, 7), (8, 3)])
def test_soup_restarts_equal_single_runs(k, n_restarts):
    """restarts run R at a time (R = 8, 5, 2 here; the last batch is partly empty): every restart's loss and iteration
    count equal the oracle's f64 run from the same centers, and the best one is the reference's choice"""
    from cellector_b200 import Cellector
    d = synth.generate(synth.SynthConfig(n_cells=700, n_loci=2500, density=0.04, minority_frac=0.35, seed=58, deep_frac=0.02))
    with Cellector(0) as gpu, SoupOracle() as orc:
        n_used, _ = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        init = np.stack([soup.init_cluster_centers_uniform(n_used, k, 4, r) for r in range(n_restarts)])
        g = gpu.soup_restarts(init)
        single = gpu.soup_em(init[g["best"]])
        runs = [orc.soup_em(init[r], precise=True) for r in range(n_restarts)]
    assert [int(x) for x in g["iterations"]] == [r["iterations"] for r in runs]
    np.testing.assert_allclose(g["losses"], [r["loss"] for r in runs], rtol=1e-10)
    want_best = int(np.argmax([r["loss"] for r in runs]))  # first maximum, like the strict > of main.rs:73
    assert g["best"] == want_best
    np.testing.assert_allclose(g["log_probs"], runs[want_best]["log_probs"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(g["centers"], runs[want_best]["centers"], rtol=1e-9)
    # a restart's result does not depend on the slot or the batch it ran in
    np.testing.assert_array_equal(single["log_probs"], g["log_probs"])
    assert single["loss"] == g["loss"]


def test_soup_refuses_duplicated_coordinates_and_clipped_counts():
    from cellector_b200 import Cellector
    d = synth.generate(synth.SynthConfig(n_cells=200, n_loci=600, density=0.08, minority_frac=0.3, seed=57))
    with Cellector(0) as gpu:
        # one line of a locus that passes the filter, stored a second time at the end of the file
        kept = (np.bincount(d.locus[d.ref > 0], minlength=d.n_loci) >= 1) & (np.bincount(d.locus[d.alt > 0], minlength=d.n_loci) >= 1)
        i = int(np.flatnonzero(kept[d.locus])[5])
        locus = np.concatenate([d.locus, d.locus[i:i + 1]])
        cell = np.concatenate([d.cell, d.cell[i:i + 1]])
        ref = np.concatenate([d.ref, d.ref[i:i + 1]])
        alt = np.concatenate([d.alt, d.alt[i:i + 1]])
        n_used, _ = gpu.load_cell_data(locus, cell, ref, alt, d.n_loci, d.n_cells, 1, 1)
        with pytest.raises(RuntimeError, match="stored more than once"):
            gpu.soup_em(soup.init_cluster_centers_uniform(n_used, 2, 4))
        ref2 = d.ref.copy()
        ref2[i] = 70000
        n_used, _ = gpu.load_cell_data(d.locus, d.cell, ref2, d.alt, d.n_loci, d.n_cells, 1, 1)
        with pytest.raises(RuntimeError, match="65534"):
            gpu.soup_em(soup.init_cluster_centers_uniform(n_used, 2, 4))
        with pytest.raises(RuntimeError, match=r"\[1, 8\]"):
            gpu.soup_em(soup.init_cluster_centers_uniform(n_used, 9, 4))


def test_souporcell_cli_matches_oracle(tmp_path):
    """the `souporcell` binary end to end (reference flags in, `barcode<TAB>cluster<TAB>log probabilities` out,
    main.rs:91-105) against the oracle run from the same initial centers (tests/soup_rng.py states the draws separately)"""
    import os
    import subprocess
    from tests.soup_rng import souporcell_initial_centers
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cli = os.path.join(root, "cellector_b200", "bin", "souporcell")
    d = synth.generate(synth.SynthConfig(n_cells=400, n_loci=1200, density=0.06, minority_frac=0.3, seed=59))
    paths = synth.write_mtx_pair(d, str(tmp_path / "in"))
    k, restarts, threads, seed = 2, 5, 2, 7
    r = subprocess.run([cli, "-a", paths["alt.mtx"], "-r", paths["ref.mtx"], "--barcodes", paths["barcodes.tsv"], "-k", str(k),
                        "--restarts", str(restarts), "-t", str(threads), "--seed", str(seed)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    with SoupOracle() as orc:
        n_used, _ = orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        assert f"total loci used {n_used}" in r.stderr
        init = souporcell_initial_centers(seed, threads, restarts, k, n_used)
        assert init.shape[0] == 6  # ceil(5 / 2) runs on each of the 2 threads
        runs = [orc.soup_em(c, precise=True) for c in init]
    best = int(np.argmax([x["loss"] for x in runs]))
    want = runs[best]["log_probs"]
    m = [l for l in r.stderr.split("\n") if l.startswith("best total log probability = ")]
    assert len(m) == 1 and float(m[0].split("= ")[1]) == pytest.approx(runs[best]["loss"], rel=1e-6)
    rows = [l.split("\t") for l in r.stdout.strip().split("\n")]
    barcodes = open(paths["barcodes.tsv"]).read().split()
    assert [x[0] for x in rows] == barcodes and all(len(x) == 2 + k for x in rows)
    got = np.array([[float(v) for v in x[2:]] for x in rows])
    np.testing.assert_allclose(got, want, rtol=2e-7, atol=1e-30)  # printed as f32
    assert [int(x[1]) for x in rows] == np.argmax(got.astype(np.float32), axis=1).tolist()
    truth = np.asarray(d.is_minority, dtype=np.int64)
    agree = (np.array([int(x[1]) for x in rows]) == truth).mean()
    assert max(agree, 1 - agree) > 0.95
    # what the reference asserts on or this build leaves out is refused, not ignored
    # the UMI minima of souporcell's locus filter (main.rs:440), applied by the host before the load
    r2 = subprocess.run([cli, "-a", paths["alt.mtx"], "-r", paths["ref.mtx"], "-b", paths["barcodes.tsv"], "-k", "2", "--restarts", "1",
                         "--min_ref_umis", "12", "--min_alt_umis", "8", "--seed", str(seed)], capture_output=True, text=True, timeout=300)
    assert r2.returncode == 0, r2.stderr[-2000:]
    with SoupOracle() as orc:
        n_umi, _ = orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells, 4, 4, 12, 8)
        assert 0 < n_umi < n_used and f"total loci used {n_umi}" in r2.stderr
        want2 = orc.soup_em(souporcell_initial_centers(seed, 1, 1, k, n_umi)[0], precise=True)
    got2 = np.array([[float(v) for v in l.split("\t")[2:]] for l in r2.stdout.strip().split("\n")])
    np.testing.assert_allclose(got2, want2["log_probs"], rtol=2e-7, atol=1e-30)
    for extra in (["--initialization_strategy", "kmeans++"], ["--known_cell_assignments", "x.tsv"],
                  ["--min_alt_umis", "3", "--min_alt", "0", "--min_ref", "0"]):
        bad = subprocess.run([cli, "-a", paths["alt.mtx"], "-r", paths["ref.mtx"], "-b", paths["barcodes.tsv"], "-k", "2", *extra],
                             capture_output=True, text=True, timeout=60)
        assert bad.returncode != 0 and ("not" in bad.stderr or "need" in bad.stderr), bad.stderr


def test_souporcell_cli_known_genotypes(tmp_path):
    """`souporcell -g genotypes.vcf.gz --known_genotypes_sample_names ...` (init_cluster_centers_known_genotypes,
    main.rs:309-338): centers from the samples' GT fields, one deterministic EM run; against the oracle from the same
    centers built here.  Cluster s is sample s, so the clusters are not arbitrary: 1 = the minority individual."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cli = os.path.join(root, "cellector_b200", "bin", "souporcell")
    cfg = synth.SynthConfig(n_cells=500, n_loci=1500, density=0.05, minority_frac=0.3, seed=60)
    d = synth.generate(cfg)
    paths = synth.write_mtx_pair(d, str(tmp_path / "in"))
    p_major, p_minor = synth.locus_model(cfg)
    gt_of = {0.01: "0/0", 0.5: "0|1", 0.99: "1/1"}
    rng = np.random.Generator(np.random.Philox(7))
    gts = []
    vcf = tmp_path / "in" / "known.vcf"
    with open(vcf, "w") as f:
        f.write("##fileformat=VCFv4.2\n##source=synthetic\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tother\tmin\tmaj\n")
        for i in range(d.n_loci):
            g_maj, g_min = gt_of[float(p_major[i])], gt_of[float(p_minor[i])]
            u = rng.random()
            if u < 0.05:
                g_maj = "./."            # missing: the centre keeps 0.5
            elif u < 0.08:
                g_min = "2/1"            # alleles are capped at 1 (main.rs:325-326)
            gts.append((g_maj, g_min))
            f.write(f"chr1\t{100 + i}\t.\tA\tG\t.\t.\t.\tDP:GT:GQ\t9:0/1:3\t12:{g_min}:40\t7:{g_maj}:30\n")
    subprocess.check_call(["gzip", str(vcf)])
    r = subprocess.run([cli, "-a", paths["alt.mtx"], "-r", paths["ref.mtx"], "-b", paths["barcodes.tsv"], "-k", "2",
                        "-g", str(vcf) + ".gz", "--known_genotypes_sample_names", "maj", "min", "--restarts", "3"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]

    def centre(gt):
        if gt[0] == ".":
            return np.float32(0.5)
        v = np.float32(min(int(gt[0]), 1) + min(int(gt[2]), 1)) / np.float32(2.0)
        return min(max(v, np.float32(0.01)), np.float32(0.99))

    with SoupOracle() as orc:
        n_used, _ = orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        ids = orc.locus_ids()
        init = np.array([[centre(gts[l][s]) for l in ids] for s in (0, 1)], np.float32)
        want = orc.soup_em(init, precise=True)
    rows = [l.split("\t") for l in r.stdout.strip().split("\n")]
    got = np.array([[float(v) for v in x[2:]] for x in rows])
    np.testing.assert_allclose(got, want["log_probs"], rtol=2e-7, atol=1e-30)
    m = [l for l in r.stderr.split("\n") if l.startswith("best total log probability = ")]
    assert float(m[0].split("= ")[1]) == pytest.approx(want["loss"], rel=1e-6)
    assign = np.array([int(x[1]) for x in rows])
    assert (assign == np.asarray(d.is_minority, dtype=np.int64)).mean() > 0.97
    # more sample names than clusters is the reference's index panic
    bad = subprocess.run([cli, "-a", paths["alt.mtx"], "-r", paths["ref.mtx"], "-b", paths["barcodes.tsv"], "-k", "2", "-g",
                          str(vcf) + ".gz", "--known_genotypes_sample_names", "maj", "min", "other"], capture_output=True, text=True, timeout=60)
    assert bad.returncode != 0 and "index out of bounds" in bad.stderr
