"""Device generator of the large workloads (cellector_b200/synth_device.py) on the CPU: the matrix must not depend on
how it is sharded, the model must look like synth.py's, and the nnz-balanced block partition must balance skew."""
import numpy as np

from cellector_b200 import synth, synth_device as sd


def _global(d):
    return np.stack([d.locus.numpy().astype(np.int64), d.cell.numpy().astype(np.int64) + d.cell_offset,
                     d.ref.numpy().astype(np.int64), d.alt.numpy().astype(np.int64)], axis=1)


def test_sharding_invariance_and_model():
    cfg = synth.SynthConfig(n_cells=4096, n_loci=3000, density=0.01, minority_frac=0.05, seed=5)
    whole = sd.generate(cfg, device="cpu")
    parts = [sd.generate(cfg, b0, b1, device="cpu") for b0, b1 in ((0, 300), (300, 1024))]
    a = _global(whole)
    b = np.concatenate([_global(p) for p in parts])
    a = a[np.lexsort((a[:, 1], a[:, 0]))]
    b = b[np.lexsort((b[:, 1], b[:, 0]))]
    assert np.array_equal(a, b)
    # file order: sorted by (locus, cell), unique coordinates
    k = a[:, 0] * cfg.n_cells + a[:, 1]
    assert np.all(np.diff(k) > 0)
    tot = a[:, 2] + a[:, 3]
    assert tot.min() == 1 and abs(tot.mean() - 1.4) < 0.02          # 1 + Poisson(0.4)
    assert abs(a.shape[0] / (cfg.n_cells * cfg.n_loci) - cfg.density) < 0.001
    # alt fraction of majority cells at loci where the majority genotype is homozygous alt (p = 0.99)
    p_major, _ = synth.locus_model(cfg)
    maj = ~whole.is_minority[a[:, 1]]
    sel = maj & (p_major[a[:, 0]] > 0.9)
    assert abs(a[sel, 3].sum() / tot[sel].sum() - 0.99) < 0.01
    host = sd.to_host(whole, 100, 400)
    assert host.n_cells == 300 and host.cell.max() < 300 and host.nnz == int(((a[:, 1] >= 100) & (a[:, 1] < 400)).sum())


def test_balanced_blocks_under_skew():
    cfg = synth.SynthConfig(n_cells=20000, n_loci=1000, density=0.01, row_sigma=2.0, seed=3)
    w = sd.row_weights(cfg)
    bb = sd.block_bounds(cfg)
    for balanced in (True, False):
        cut = sd.shard_blocks(cfg, 8, balanced)
        assert cut[0] == 0 and cut[-1] == sd.GEN_BLOCKS and np.all(np.diff(cut) >= 0)
        share = np.array([w[bb[cut[r]]:bb[cut[r + 1]]].sum() for r in range(8)])
        imb = share.max() / share.mean()
        if balanced:
            assert imb < 1.15, imb
        else:
            unbalanced = imb
    assert unbalanced > 1.0
