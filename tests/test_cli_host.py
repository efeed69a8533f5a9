"""Host side of the `cellector` CLI that needs no GPU: the single-pass multi-threaded mtx pair parser
(replaces load_data.rs:190-223, 240-251) and the Rust-Display number formatting of the writers."""
import gzip
import os
import struct
import subprocess
from decimal import Decimal

import numpy as np
import pytest

from cellector_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "cellector_b200", "bin", "cellector")


@pytest.fixture(scope="module", autouse=True)
def _built():
    if not os.path.exists(CLI):
        from cellector_b200 import build
        build.build_all()
    assert os.path.exists(CLI)


def fnv1a(a):
    h = 1469598103934665603
    for v in np.asarray(a, np.uint64).tolist():
        h = ((h ^ v) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def parse_only(alt, ref, threads=4, extra=()):
    r = subprocess.run([CLI, "-a", alt, "-r", ref, "-b", "/dev/null", "--output_directory", "/tmp", "--parse_only", "--threads", str(threads),
                        *extra], capture_output=True, text=True, timeout=120)
    return r


def fields(out):
    t = out.split()
    return {t[i]: t[i + 1] for i in range(1, len(t) - 1, 2)}


@pytest.mark.parametrize("threads", [1, 3, 8])
def test_parser_matches_numpy(tmp_path, threads):
    d = synth.generate(synth.SynthConfig(n_cells=700, n_loci=2500, density=0.03, seed=4, deep_frac=0.02, deep_mean=300.0))
    paths = synth.write_mtx_pair(d, str(tmp_path))
    r = parse_only(paths["alt.mtx"], paths["ref.mtx"], threads)
    assert r.returncode == 0, r.stderr
    f = fields(r.stdout)
    assert int(f["loci"]) == d.n_loci and int(f["cells"]) == d.n_cells and int(f["pairs"]) == d.nnz
    assert int(f["fnv_locus"]) == fnv1a(d.locus) and int(f["fnv_cell"]) == fnv1a(d.cell)
    assert int(f["fnv_ref"]) == fnv1a(d.ref) and int(f["fnv_alt"]) == fnv1a(d.alt)


def test_parser_gz_crlf_and_ragged(tmp_path):
    # gz by extension (load_data.rs:246-247), CRLF / mixed whitespace (split_whitespace), no trailing newline,
    # header nnz ignored, and izip! stopping at the shorter file
    alt = "%%MatrixMarket matrix coordinate real general\r\n% c\r\n5 4 999\r\n1\t1\t2\r\n  3 2   0\r\n5\t4\t7\r\n4 4 1"
    ref = "%%MatrixMarket matrix coordinate real general\n% c\n5\t4\t0\n9 9 1\n9 9 +3\n9 9 0\n"
    a, r = tmp_path / "alt.mtx.gz", tmp_path / "ref.mtx"
    with gzip.open(a, "wt", newline="") as f:
        f.write(alt)
    r.write_text(ref)
    res = parse_only(str(a), str(r), 2)
    assert res.returncode == 0, res.stderr
    f = fields(res.stdout)
    assert (int(f["loci"]), int(f["cells"]), int(f["pairs"])) == (5, 4, 3)
    assert int(f["fnv_locus"]) == fnv1a([0, 2, 4]) and int(f["fnv_cell"]) == fnv1a([0, 1, 3])
    assert int(f["fnv_alt"]) == fnv1a([2, 0, 7]) and int(f["fnv_ref"]) == fnv1a([1, 3, 0])


@pytest.mark.parametrize("bad", ["1 1 2.0\n", "1 1\n", "0 1 1\n", "\n", "1 x 1\n"])
def test_parser_rejects_what_the_reference_panics_on(tmp_path, bad):
    hdr = "%%MatrixMarket matrix coordinate real general\n% c\n5 4 1\n"
    (tmp_path / "alt.mtx").write_text(hdr + bad)
    (tmp_path / "ref.mtx").write_text(hdr + "1 1 1\n")
    res = parse_only(str(tmp_path / "alt.mtx"), str(tmp_path / "ref.mtx"))
    assert res.returncode == 101 and "malformed" in res.stderr


def rust_display(x: float) -> str:
    if x != x:
        return "NaN"
    if x in (float("inf"), float("-inf")):
        return "inf" if x > 0 else "-inf"
    s = format(Decimal(repr(x)), "f")
    if "." in s:
        s = s.rstrip("0").rstrip(".")
    return s


def test_number_formatting_is_rust_display():
    rng = np.random.default_rng(5)
    vals = [0.0, -0.0, 17.0, 1e-7, 1e21, -2.055795012245653, 0.1, 1 / 3, 5e-324, 1.7976931348623157e308, 123456789.125,
            float("nan"), float("inf"), float("-inf"), 0.999, 255.0, -80.0, 1e15, 1e16, 1e17, 123456789012345680.0]
    vals += list(rng.standard_normal(200) * 10.0 ** rng.integers(-12, 12, 200))
    vals += list(-np.abs(rng.standard_normal(100)) * 300)
    inp = "\n".join(format(struct.unpack("<Q", struct.pack("<d", float(v)))[0], "x") for v in vals) + "\n"
    r = subprocess.run([CLI, "--format_selftest"], input=inp, capture_output=True, text=True, timeout=60)
    got = r.stdout.split("\n")[:-1]
    assert len(got) == len(vals)
    for v, g in zip(vals, got):
        assert g == rust_display(float(v)), (v, g, rust_display(float(v)))
        if v == v:
            assert float(g) == float(v)  # round-trips


def test_cli_flag_errors(tmp_path):
    r = subprocess.run([CLI, "-a", "x"], capture_output=True, text=True)
    assert r.returncode == 1 and "required arguments" in r.stderr
    r = subprocess.run([CLI, "--bogus", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "wasn't expected" in r.stderr


SOUP_CLI = os.path.join(ROOT, "cellector_b200", "bin", "souporcell")


def test_souporcell_rng_selftest():
    """ChaCha20 block function of the souporcell host's StdRng against the RFC 7539 section 2.3.2 vector, its first
    draws against the plain-Python statement in tests/soup_rng.py, and the f32 Display formatting"""
    from tests.soup_rng import StdRng07
    r = subprocess.run([SOUP_CLI, "--rng_selftest"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().split("\n")
    assert lines[0] == ("e4e7f110 15593bd1 1fdd0f50 c47120a3 c7f4d1c7 0368c033 9aaa2204 4e6cd4c3 "
                        "466482d2 09aa9f07 05d7c214 a2028bd9 d19c12b5 b94e16de e883d0cb 4e3c50a2")
    rng = StdRng07(bytes([4]) * 32)
    u8 = [rng.gen_u8() for _ in range(3)]
    f32 = [rng.gen_f32() for _ in range(3)]
    t = lines[1].split()
    assert t[0] == "u8" and [int(x) for x in t[1:4]] == u8
    assert t[4] == "f32" and [np.float32(x) for x in t[5:8]] == f32
    assert lines[2].split() == ["f32", "0.1", "-2181.0225", "5", "0.0000001", "340282350000000000000000000000000000000", "-0"]


@pytest.mark.parametrize("gpus", [1, 2, 3, 8])
def test_cli_gpus_partition(tmp_path, gpus):
    """the cell ranges `cellector --gpus N` gives its ranks (host logic, no CUDA): contiguous, covering, every rank at
    least one cell, entry shares balanced although the rows are heavily skewed"""
    d = synth.generate(synth.SynthConfig(n_cells=1500, n_loci=2500, density=0.02, minority_frac=0.05, seed=61, row_sigma=2.0))
    paths = synth.write_mtx_pair(d, str(tmp_path / "in"))
    r = subprocess.run([CLI, "-a", paths["alt.mtx"], "-r", paths["ref.mtx"], "-b", paths["barcodes.tsv"], "--output_directory",
                        str(tmp_path / "out"), "--gpus", str(gpus), "--partition_only"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    rows = [l.split() for l in r.stdout.strip().split("\n") if l.startswith("partition rank")]
    assert [int(x[2]) for x in rows] == list(range(gpus))
    lo, hi, ent = [int(x[4]) for x in rows], [int(x[5]) for x in rows], [int(x[7]) for x in rows]
    assert lo[0] == 0 and hi[-1] == d.n_cells and lo[1:] == hi[:-1] and all(h > l for l, h in zip(lo, hi))
    per_cell = np.bincount(d.cell, minlength=d.n_cells)
    assert ent == [int(per_cell[l:h].sum()) for l, h in zip(lo, hi)] and sum(ent) == d.nnz
    if gpus > 1:
        # a contiguous split cannot beat the deepest single cell; beyond that the shares stay near the mean
        assert max(ent) <= d.nnz / gpus + per_cell.max() + 1


@pytest.mark.parametrize("args,needle", [
    (["-a", "a", "-r", "r", "-b", "b"], "--num_clusters"),                                      # required argument missing
    (["-a", "a", "-r", "r", "-b", "b", "-k", "2", "--bogus", "1"], "wasn't expected"),
    (["-a", "a", "-r", "r", "-b", "b", "-k", "2", "--seed", "256"], "seed is a u8"),
    (["-a", "a", "-r", "r", "-b", "b", "-k", "2", "--initialization_strategy", "kmeans++"], "kmeans++ not yet implemented"),
    (["-a", "a", "-r", "r", "-b", "b", "-k", "2", "--initialization_strategy", "random_cell_assignment"], "not built"),
    (["-a", "a", "-r", "r", "-b", "b", "-k", "2", "--known_cell_assignments", "x"], "known cell assignments not yet implemented"),
    (["-a", "a", "-r", "r", "-b", "b", "-k", "9"], "between 1 and 8"),
    (["-a", "a", "-r", "r", "-b", "b", "-k", "2", "-t", "0"], "divide by zero"),
])
def test_souporcell_cli_argument_errors(args, needle):
    """the souporcell host refuses what the reference refuses (its asserts and unwraps) and what this build leaves out,
    before it touches a file or the GPU"""
    r = subprocess.run([SOUP_CLI, *args], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and needle in r.stderr, r.stderr
