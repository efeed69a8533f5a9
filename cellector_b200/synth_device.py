"""Device-side generator for the large named workloads (BASELINE.json configs 3-5: 2e8 .. 5e9 nonzeros).

Same mixture model as synth.py — per-locus alt-read probabilities from synth.locus_model, minority
cells, presence with probability density x row weight, total = 1 + Poisson(lambda), alt ~ Binomial(total, p) — but
drawn on the GPU with torch generators, because numpy needs minutes and tens of GB of host memory for 5e8 draws.
The matrix is defined by fixed GENERATION BLOCKS of cells (each with its own seed), so the global matrix is identical
for every way of sharding it over ranks: a rank generates the blocks of its cell range.  torch is plumbing here (device
memory + RNG); the product path never imports this module.
"""
from __future__ import annotations

import dataclasses

import numpy as np

from . import synth

MAX_EXTRA_DEPTH = 8  # total = 1 + min(Poisson(lambda), 8)


def _s64(x: int) -> int:
    x &= (1 << 64) - 1
    return x - (1 << 64) if x >= (1 << 63) else x


def _mix64(x):
    """splitmix64-style finaliser on int64 tensors (wrapping multiplies, logical shifts by masking)."""
    m32 = (1 << 32) - 1
    x = x * _s64(0x9E3779B97F4A7C15)
    x = x ^ ((x >> 32) & m32)
    x = x * _s64(0xD6E8FEB86659FD93)
    x = x ^ ((x >> 32) & m32)
    x = x * _s64(0xD6E8FEB86659FD93)
    x = x ^ ((x >> 32) & m32)
    return x


def _unit(h):
    """uniform [0, 1) float32 from the top 24 bits of a hash"""
    return ((h >> 40) & ((1 << 24) - 1)).to(h.dtype).float() * (1.0 / (1 << 24))


GEN_BLOCKS = 1024  # generation granularity (cells / GEN_BLOCKS per block); shard boundaries fall on block boundaries


@dataclasses.dataclass
class DeviceData:
    n_cells: int        # cells of THIS shard
    n_loci: int
    locus: "object"     # torch.int32 [nnz] on the device, file order = sorted by (locus, cell)
    cell: "object"      # torch.int32 [nnz], shard-local
    ref: "object"
    alt: "object"
    is_minority: np.ndarray
    cell_offset: int

    @property
    def nnz(self) -> int:
        return int(self.locus.shape[0])


def block_bounds(cfg: synth.SynthConfig) -> np.ndarray:
    return (np.arange(GEN_BLOCKS + 1, dtype=np.int64) * cfg.n_cells) // GEN_BLOCKS


def row_weights(cfg: synth.SynthConfig) -> np.ndarray:
    """Per-cell coverage weights (mean 1): log-normal for the skewed workload, else all ones.  Host-side and
    deterministic, so every rank can balance the shards by EXPECTED nonzeros without generating anything."""
    if cfg.row_sigma <= 0:
        return np.ones(cfg.n_cells)
    r = np.random.Generator(np.random.Philox(key=[(cfg.seed << 20) ^ (7 << 12), 0x51ED27]))
    w = r.lognormal(0.0, cfg.row_sigma, size=cfg.n_cells)
    return w * (cfg.n_cells / w.sum())


def shard_blocks(cfg: synth.SynthConfig, world: int, balanced: bool = True) -> np.ndarray:
    """Block boundaries b[0..world]: rank r generates blocks [b[r], b[r+1]).  balanced = by expected nonzeros
    (cellector_b200.shard.partition_cells on the per-block weight sums), else equal block counts."""
    if not balanced:
        return (np.arange(world + 1, dtype=np.int64) * GEN_BLOCKS) // world
    from .shard import partition_cells
    bb = block_bounds(cfg)
    w = row_weights(cfg)
    per_block = np.add.reduceat(w, bb[:-1]) if cfg.n_cells >= GEN_BLOCKS else np.ones(GEN_BLOCKS)
    return partition_cells(np.round(per_block * 1e6).astype(np.int64), world)


def generate(cfg: synth.SynthConfig, block0: int = 0, block1: int = GEN_BLOCKS, device="cuda") -> DeviceData:
    """Cells of generation blocks [block0, block1) of the configured matrix, as device tensors."""
    import torch
    p_major, p_minor = synth.locus_model(cfg)
    bb = block_bounds(cfg)
    c0, c1 = int(bb[block0]), int(bb[block1])
    n_local = c1 - c0
    w_all = row_weights(cfg)
    # per-cell labels: one host stream over all cells (cheap), sliced
    r = np.random.Generator(np.random.Philox(key=[(cfg.seed << 20) ^ (9 << 12), 0x51ED27]))
    is_minor_all = r.random(cfg.n_cells) < cfg.minority_frac
    is_minor = is_minor_all[c0:c1]
    dev = torch.device(device)
    pm = torch.from_numpy(np.stack([p_major, p_minor]).astype(np.float32)).to(dev)  # [2, L]
    minor_t = torch.from_numpy(is_minor.astype(np.int64)).to(dev)
    keys = []
    for b in range(block0, block1):
        b0, b1 = int(bb[b]), int(bb[b + 1])
        nb = b1 - b0
        if nb == 0:
            continue
        g = torch.Generator(device=dev)
        g.manual_seed((cfg.seed * 1_000_003 + b) & 0x7FFFFFFFFFFF)
        w = w_all[b0:b1]
        expected = cfg.density * cfg.n_loci * float(w.sum())
        n_draw = int(torch.poisson(torch.tensor([expected], dtype=torch.float64, device=dev), generator=g).item())
        if n_draw == 0:
            continue
        if cfg.row_sigma > 0:
            cdf = torch.from_numpy(np.cumsum(w) / w.sum()).to(dev)
            u = torch.rand(n_draw, dtype=torch.float64, device=dev, generator=g)
            cells = torch.searchsorted(cdf, u).clamp_(max=nb - 1)
        else:
            cells = torch.randint(0, nb, (n_draw,), device=dev, generator=g, dtype=torch.int64)
        loci = torch.randint(0, cfg.n_loci, (n_draw,), device=dev, generator=g, dtype=torch.int64)
        keys.append(loci * n_local + (cells + (b0 - c0)))
    if keys:
        key = torch.unique(torch.cat(keys))  # sorted: file order (locus, cell); duplicate draws collapse
        del keys
    else:
        key = torch.zeros(0, dtype=torch.int64, device=dev)
    loci = torch.div(key, n_local, rounding_mode="floor")
    cells = key - loci * n_local
    del key
    # counts are a pure function of (seed, locus, global cell): identical for every sharding of the matrix
    h = _mix64(loci * cfg.n_cells + (cells + c0) + (cfg.seed << 40))
    u = _unit(h)
    total = torch.ones_like(loci, dtype=torch.int32)
    cdf, term = 0.0, float(np.exp(-cfg.depth_lambda))
    for k in range(MAX_EXTRA_DEPTH):  # total = 1 + Poisson(lambda) by inverse CDF
        cdf += term
        total += (u > cdf).to(torch.int32)
        term *= cfg.depth_lambda / (k + 1)
    del u
    p = pm[minor_t[cells], loci]
    alt = torch.zeros_like(total)
    for t in range(1, MAX_EXTRA_DEPTH + 2):  # alt ~ Binomial(total, p) as a sum of Bernoulli draws
        ut = _unit(_mix64(h + _s64(t * 0x632BE59BD9B4E019)))
        alt += ((ut < p) & (total >= t)).to(torch.int32)
        del ut
    ref = total - alt
    del total, p, h
    return DeviceData(n_cells=n_local, n_loci=cfg.n_loci, locus=loci.to(torch.int32), cell=cells.to(torch.int32), ref=ref, alt=alt,
                      is_minority=is_minor, cell_offset=c0)


def to_host(d: DeviceData, cell_lo: int = 0, cell_hi: int | None = None) -> synth.SynthData:
    """The entries of shard-local cells [cell_lo, cell_hi) as a host SynthData (file order kept, cell ids rebased): what
    the CPU oracle is run on for the sliced parity checks of the large workloads."""
    cell_hi = d.n_cells if cell_hi is None else cell_hi
    keep = (d.cell >= cell_lo) & (d.cell < cell_hi)
    f = lambda t: t[keep].cpu().numpy().astype(np.uint32)
    return synth.SynthData(n_cells=cell_hi - cell_lo, n_loci=d.n_loci, locus=f(d.locus), cell=f(d.cell) - np.uint32(cell_lo),
                           ref=f(d.ref), alt=f(d.alt), is_minority=d.is_minority[cell_lo:cell_hi], cell_offset=d.cell_offset + cell_lo)
