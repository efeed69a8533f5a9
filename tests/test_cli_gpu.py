"""The `cellector` CLI end to end on a GPU: reference flags in, reference files out (SURVEY.md §8b outer
boundary: iteration_{i}.tsv, iteration_{i}_threshold.tsv, iteration_{i}_locus_contribution.tsv,
cellector_assignments.tsv, cellector.vcf, stdout), checked against the CPU oracle on the same matrix."""
import os
import re
import subprocess

import numpy as np
import pytest

from cellector_b200 import api, synth
from tests.parity import RTOL, assert_close, assert_equal

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "cellector_b200", "bin", "cellector")


def read_tsv(path):
    with open(path) as f:
        rows = [line.rstrip("\n").split("\t") for line in f]
    return rows[0], rows[1:]


def col(rows, i, typ=str):
    return np.array([typ(r[i]) for r in rows]) if typ is not str else [r[i] for r in rows]


@pytest.mark.parametrize("with_vcf,gpus", [(False, 1), (True, 1), (False, 2), (True, 2)],
                         ids=["plain", "vcf_gz", "plain_gpus2", "vcf_gz_gpus2"])
def test_cli_outputs_match_oracle(tmp_path, oracle_lib, with_vcf, gpus):
    """gpus = 2: `--gpus 2` (one forked worker per GPU, cells sharded, rank 0 writes) must produce the same files."""
    from tests.oracle import Oracle
    if gpus > 1:
        import torch
        if torch.cuda.device_count() < gpus:
            pytest.skip(f"needs {gpus} GPUs")
    cfg = synth.SynthConfig(n_cells=1800, n_loci=5000, density=0.03, minority_frac=0.07, seed=23, deep_frac=0.01, deep_mean=50.0)
    d = synth.generate(cfg)
    paths = synth.write_mtx_pair(d, str(tmp_path / "in"))
    out = tmp_path / "out"
    cmd = [CLI, "-a", paths["alt.mtx"], "-r", paths["ref.mtx"], "-b", paths["barcodes.tsv"], "-g", paths["gt.tsv"],
           "--output_directory", str(out), "--timing_json", str(tmp_path / "timing.json")]
    if with_vcf:
        vcf = tmp_path / "in" / "variants.vcf"
        with open(vcf, "w") as f:
            f.write("##fileformat=VCFv4.2\n##source=synthetic\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
            for i in range(d.n_loci):
                f.write(f"chr{1 + i % 3}\t{100 + i}\t.\tA\tG\t.\t.\t.\n")
        subprocess.check_call(["gzip", "-k", paths["alt.mtx"]])
        cmd[2] = paths["alt.mtx"] + ".gz"
        cmd += ["-v", str(vcf), "--min_alt", "5", "--min_ref=6", "--interquartile_range_multiple", "4"]
    if gpus > 1:
        cmd += ["--gpus", str(gpus)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    min_ref, min_alt, iqr = (6, 5, 4.0) if with_vcf else (4, 4, 5.0)

    barcodes = open(paths["barcodes.tsv"]).read().split()
    gt = ["minority" if m else "majority" for m in d.is_minority]
    with Oracle(oracle_lib) as orc:
        orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells, min_ref, min_alt)
        ids = orc.locus_ids()
        stdout_lines = r.stdout.split("\n")
        detected = [l for l in stdout_lines if l.startswith("detected ")]
        medians = [l for l in stdout_lines if l.startswith("median normalized")]
        filtering = [l for l in stdout_lines if l.startswith("filtering locus")]
        n_filtered = 0
        it = 0
        while True:
            st = orc.compute_new_excluded(iqr)
            cv, lv = orc.cell_vectors(), orc.locus_vectors()
            # stdout (main.rs:338-339)
            assert detected[it] == (f"detected {st.n_new_excluded} new anomylous cells and rescued {st.n_rescued} cells to the "
                                    f"majority in iteration {it + 1}")
            m = re.match(r"median normalized log likelihood (\S+) with interquartile range (\S+), threshold (\S+)$", medians[it])
            assert_close(float(m.group(1)), st.median, "stdout median")
            assert_close(float(m.group(3)), st.threshold, "stdout threshold")
            n_filtered += st.n_loci_filtered
            # iteration_{i}.tsv (main.rs:349-360)
            hdr, rows = read_tsv(out / f"iteration_{it}.tsv")
            assert hdr == ["cell_id", "barcode", "assignment", "log_likelihood", "expected_log_likelihood", "num_loci_used"]
            assert col(rows, 0) == [str(i) for i in range(d.n_cells)] and col(rows, 1) == barcodes and col(rows, 2) == gt
            assert_close(col(rows, 3, float), cv["ll"], f"iteration_{it}.tsv log_likelihood")
            assert_close(col(rows, 4, float), cv["expected_ll"], f"iteration_{it}.tsv expected_log_likelihood")
            assert col(rows, 5) == [str(int(v)) for v in cv["n_loci"]]
            thr = open(out / f"iteration_{it}_threshold.tsv").read()
            assert not thr.endswith("\n")
            assert_close(float(thr), st.threshold, "threshold file")
            # iteration_{i}_locus_contribution.tsv (main.rs:422-498): 17 columns, one row per used locus, stable argsort order
            hdr, rows = read_tsv(out / f"iteration_{it}_locus_contribution.tsv")
            assert len(hdr) == 17 and hdr[0] == "locus_id" and hdr[-1] == "majority_af"
            assert len(rows) == len(ids)
            order = np.argsort(col(rows, 0, int))
            assert_equal(np.sort(col(rows, 0, int)), np.sort(ids.astype(np.int64)), "locus ids of the contribution file")
            by_id = np.argsort(ids)  # oracle index for ascending locus id
            for name, c, exact in (("ll_minority", 3, False), ("ll_majority", 4, False), ("ll_minority", 5, False), ("ll_majority", 6, False),
                                   ("cells_minority", 7, True), ("cells_majority", 8, True), ("alt_minority", 11, True),
                                   ("ref_minority", 12, True), ("alt_majority", 13, True), ("ref_majority", 14, True)):
                got = col(rows, c, float if not exact else int)[order]
                want = lv[name][by_id]
                (assert_equal if exact else assert_close)(got, want.astype(np.int64) if exact else want, f"contribution {hdr[c]}")
            per_cell = col(rows, 9, float)
            assert np.all(np.diff(per_cell) >= -1e-9 * np.abs(per_cell[1:])), "rows are ordered by minority log likelihood per cell"
            chrom = col(rows, 1)
            assert all(c.startswith("chr") for c in chrom) if with_vcf else all(c == "na" for c in chrom)
            it += 1
            if not st.any_change:
                break
        assert len(detected) == it and not os.path.exists(out / f"iteration_{it}.tsv")
        assert len(filtering) == n_filtered
        # cellector_assignments.tsv (main.rs:133-174)
        pv = orc.calculate_posteriors(0.999, 30)
        cv = orc.cell_vectors()
        hdr, rows = read_tsv(out / "cellector_assignments.tsv")
        assert hdr == ["barcode", "posterior_assignment", "anomally_assignment", "log_likelihood_loci_normalized", "loci_used",
                       "posterior_assign_qual", "majority_log_likelihood", "minority_log_likelihood", "ground_truth_assignment"]
        assert col(rows, 0) == barcodes and col(rows, 8) == gt
        assert col(rows, 1) == [api.LABELS[v] for v in pv["label"]]
        assert col(rows, 2) == [str(int(v)) for v in pv["anomaly"]]
        assert_close(col(rows, 3, float), cv["norm"], "assignments normalized ll")
        assert col(rows, 4) == [str(int(v)) for v in cv["n_loci"]]
        assert col(rows, 5) == [str(int(v)) for v in pv["qual"]]
        assert_close(col(rows, 6, float), pv["ll_major"], "assignments majority ll")
        assert_close(col(rows, 7, float), pv["ll_minor"], "assignments minority ll")
        assert "cellector assignment   " in r.stdout and "cell_hashing" in r.stdout and "minority" in r.stdout
        if with_vcf:
            mn, mj = orc.final_locus_counts()
            lines = open(out / "cellector.vcf").read().split("\n")
            assert lines[0] == "##fileformat=VCFv4.2" and lines[2].endswith("\tINFO\tmajority\tminority")
            recs = [l.split("\t") for l in lines[3:] if l]
            assert len(recs) == d.n_loci
            for i in (0, 1, 17, d.n_loci - 1):
                assert recs[i][8] == "GT:GP:AO:RO"
                maj, mino = recs[i][9].split(":"), recs[i][10].split(":")
                assert (int(maj[2]), int(maj[3])) == (int(mj[i, 1]), int(mj[i, 0]))
                assert (int(mino[2]), int(mino[3])) == (int(mn[i, 1]), int(mn[i, 0]))
                assert maj[0] in ("./.", "0/0", "0/1", "1/1") and 1 / 3 - 1e-9 <= float(maj[1]) <= 1.0 + 1e-12
            ao = np.array([int(x[9].split(":")[2]) for x in recs])
            assert_equal(ao, mj[:, 1].astype(np.int64), "vcf majority AO")
            # GT / GP of EVERY record against main.rs:79-118 restated with the oracle's statrs Binomial::pmf
            pmf = oracle_lib.orc_binomial_pmf

            def call(a, r, soup):
                ps = [(1.0 - 0.03) * q + 0.03 * soup for q in (0.99, 0.5, 0.01)]
                like = [pmf(q, int(a + r), int(a)) for q in ps]
                den = 1.0 / 3.0 * like[0] + 1.0 / 3.0 * like[1] + 1.0 / 3.0 * like[2]
                post = [l * 1.0 / 3.0 / den for l in like]
                gt_s = "./."
                for pp, name in zip(post, ("1/1", "0/1", "0/0")):
                    if pp > 0.99:
                        gt_s = name
                        break
                return gt_s, max(post)

            n_called = 0
            for i, rec in enumerate(recs):
                ta, tr = int(mn[i, 1] + mj[i, 1]), int(mn[i, 0] + mj[i, 0])
                soup = ta / (ta + tr) if ta + tr > 0 else 0.5
                for field, cnt in ((rec[9], mj[i]), (rec[10], mn[i])):
                    g, gp, a_s, r_s = field.split(":")
                    want_gt, want_gp = call(int(cnt[1]), int(cnt[0]), soup)
                    assert (int(a_s), int(r_s)) == (int(cnt[1]), int(cnt[0]))
                    assert g == want_gt, (i, field, want_gt)
                    assert float(gp) == pytest.approx(want_gp, rel=1e-9, abs=1e-300), (i, field, want_gp)
                    n_called += g != "./."
            assert n_called > d.n_loci // 4  # the check above is not vacuous: many genotypes are called
        else:
            assert not os.path.exists(out / "cellector.vcf")
    import json
    t = json.load(open(tmp_path / "timing.json"))
    assert t["iterations"] == it and t["nnz0"] == d.nnz and t["gpu_launches"] > 0 and t["gpus"] == gpus
