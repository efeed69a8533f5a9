"""Seeded synthetic vartrix-style ref/alt count matrices.

The reference ships no data; its only fixture writer is the `combiner` tool
(/root/reference/combiner/src/main.rs:67-70, 112-115), which needs two real vartrix datasets.  This generator
produces the same *format* (coordinates shared by alt.mtx and ref.mtx, sorted by (locus, cell), 1-based in
text) from a two-individual mixture model with known labels:

  per locus   population allele frequency f ~ Beta(0.5, 0.5) clipped to [0.02, 0.98];
              each individual's genotype ~ HWE(f) -> alt-read probability p in {0.01, 0.5, 0.99}
  per cell    minority individual with probability `minority_frac`
  per entry   present with probability density * row_weight * col_weight,
              total = 1 + Poisson(depth_lambda), alt ~ Binomial(total, p[individual, locus])

Everything is drawn from numpy Philox streams keyed by (seed, stream id, shard), so a cell shard can be
regenerated independently of the others (locus-level draws depend on `seed` only).
"""
from __future__ import annotations

import dataclasses
import os

import numpy as np


@dataclasses.dataclass
class SynthConfig:
    n_cells: int
    n_loci: int
    density: float
    minority_frac: float = 0.05
    depth_lambda: float = 0.4
    seed: int = 42
    row_sigma: float = 0.0      # log-normal sigma of per-cell coverage weights (C5: 2.0)
    deep_frac: float = 0.0      # fraction of entries given a deep total (exercises the heavy path)
    deep_mean: float = 40.0
    doublet_frac: float = 0.0   # fraction of cells that mix both individuals 50/50


# The named workloads of BASELINE.json: configs (densities for C1/C2 are ours).
CONFIGS = {
    "C1": SynthConfig(n_cells=5_000, n_loci=20_000, density=0.02, minority_frac=0.05),
    "C2": SynthConfig(n_cells=10_000, n_loci=100_000, density=0.01, minority_frac=0.02),
    "C3": SynthConfig(n_cells=100_000, n_loci=500_000, density=0.01, minority_frac=0.05),
    # one rank's share of C3 on 8 GPUs (12.5k cells x all 500k loci): the per-GPU shape of the sharded run
    "C3s": SynthConfig(n_cells=12_500, n_loci=500_000, density=0.01, minority_frac=0.05),
    # halves and quarters of C3's cells (diagnostics of the sharded shapes on one GPU)
    "C3h": SynthConfig(n_cells=50_000, n_loci=500_000, density=0.01, minority_frac=0.05),
    "C3q": SynthConfig(n_cells=25_000, n_loci=500_000, density=0.01, minority_frac=0.05),
    "C4": SynthConfig(n_cells=1_000_000, n_loci=1_000_000, density=0.005, minority_frac=0.05),
    "C5": SynthConfig(n_cells=200_000, n_loci=2_000_000, density=0.0005, minority_frac=0.05, row_sigma=2.0),
}


@dataclasses.dataclass
class SynthData:
    n_cells: int            # cells in THIS shard
    n_loci: int
    locus: np.ndarray       # uint32 [nnz], 0-based, file order = sorted by (locus, cell)
    cell: np.ndarray        # uint32 [nnz], 0-based, shard-local
    ref: np.ndarray         # uint32 [nnz]
    alt: np.ndarray         # uint32 [nnz]
    is_minority: np.ndarray  # bool [n_cells] ground truth
    cell_offset: int = 0

    @property
    def nnz(self) -> int:
        return int(self.locus.shape[0])


def _rng(seed: int, stream: int, shard: int = 0) -> np.random.Generator:
    return np.random.Generator(np.random.Philox(key=[(seed << 20) ^ (stream << 12) ^ shard, 0x9E3779B97F4A7C15 & ((1 << 63) - 1)]))


def locus_model(cfg: SynthConfig):
    """Per-locus alt-read probability for the majority and the minority individual."""
    r = _rng(cfg.seed, 1)
    f = np.clip(r.beta(0.5, 0.5, size=cfg.n_loci), 0.02, 0.98)
    levels = np.array([0.01, 0.5, 0.99])
    g_major = (r.random(cfg.n_loci) < f).astype(np.int8) + (r.random(cfg.n_loci) < f).astype(np.int8)
    g_minor = (r.random(cfg.n_loci) < f).astype(np.int8) + (r.random(cfg.n_loci) < f).astype(np.int8)
    return levels[g_major], levels[g_minor]


def generate(cfg: SynthConfig, shard: int = 0, n_shards: int = 1) -> SynthData:
    """Generate the cells [shard * n/n_shards, (shard+1) * n/n_shards) of the configured matrix."""
    p_major, p_minor = locus_model(cfg)
    c0 = (cfg.n_cells * shard) // n_shards
    c1 = (cfg.n_cells * (shard + 1)) // n_shards
    n_local = c1 - c0
    r = _rng(cfg.seed, 2, shard)
    is_minor = r.random(n_local) < cfg.minority_frac
    is_doublet = r.random(n_local) < cfg.doublet_frac if cfg.doublet_frac > 0 else np.zeros(n_local, bool)

    expected = cfg.density * n_local * cfg.n_loci
    n_draw = int(r.poisson(expected))
    if cfg.row_sigma > 0:
        w = r.lognormal(0.0, cfg.row_sigma, size=n_local)
        w /= w.sum()
        cells = r.choice(n_local, size=n_draw, p=w).astype(np.uint64)
    else:
        cells = r.integers(0, n_local, size=n_draw, dtype=np.uint64)
    loci = r.integers(0, cfg.n_loci, size=n_draw, dtype=np.uint64)
    # unique (locus, cell) pairs, sorted by (locus, cell) like vartrix / combiner output
    key = np.unique(loci * np.uint64(n_local) + cells)
    loci = (key // np.uint64(n_local)).astype(np.uint32)
    cells = (key % np.uint64(n_local)).astype(np.uint32)
    nnz = key.shape[0]
    total = 1 + r.poisson(cfg.depth_lambda, size=nnz)
    if cfg.deep_frac > 0:
        deep = r.random(nnz) < cfg.deep_frac
        total = np.where(deep, 1 + r.poisson(cfg.deep_mean, size=nnz), total)
    p = np.where(is_minor[cells], p_minor[loci], p_major[loci])
    if cfg.doublet_frac > 0:
        p = np.where(is_doublet[cells], 0.5 * (p_minor[loci] + p_major[loci]), p)
    alt = r.binomial(total, p).astype(np.uint32)
    ref = (total - alt).astype(np.uint32)
    return SynthData(n_cells=n_local, n_loci=cfg.n_loci, locus=loci, cell=cells, ref=ref, alt=alt,
                     is_minority=is_minor, cell_offset=c0)


def write_mtx_pair(data: SynthData, out_dir: str) -> dict:
    """Write alt.mtx / ref.mtx / barcodes.tsv / gt.tsv in the combiner's format (combiner/src/main.rs:67-70)."""
    os.makedirs(out_dir, exist_ok=True)
    header = "%%MatrixMarket matrix coordinate real general\n% written by sprs\n"
    paths = {k: os.path.join(out_dir, k) for k in ("alt.mtx", "ref.mtx", "barcodes.tsv", "gt.tsv")}
    l1 = data.locus.astype(np.int64) + 1
    c1 = data.cell.astype(np.int64) + 1
    for name, vals in (("alt.mtx", data.alt), ("ref.mtx", data.ref)):
        with open(paths[name], "w") as f:
            f.write(header)
            f.write(f"{data.n_loci}\t{data.n_cells}\t{data.nnz}\n")
            np.savetxt(f, np.stack([l1, c1, vals.astype(np.int64)], axis=1), fmt="%d", delimiter="\t")
    barcodes = [f"CELL{(i + data.cell_offset):09d}-1" for i in range(data.n_cells)]
    with open(paths["barcodes.tsv"], "w") as f:
        f.write("\n".join(barcodes) + "\n")
    with open(paths["gt.tsv"], "w") as f:
        for b, m in zip(barcodes, data.is_minority):
            f.write(f"{b}\t{'minority' if m else 'majority'}\n")
    return paths
