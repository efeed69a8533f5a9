"""Host side of souporcell's clustering above `clx_soup_em` (souporcell/souporcell/src/main.rs:62-106, 354-364):
the initial centers and the best-of-restarts loop.  The reference draws from rand 0.7's StdRng, a third-party
generator that is not part of the repository; the draws here come from numpy's Philox (same distribution: uniform
f32 clipped to [0.0001, 0.9999]), so a run is reproducible from its seed but not draw-for-draw equal to the Rust binary.
"""
from __future__ import annotations

import numpy as np


def init_cluster_centers_uniform(n_loci: int, k: int, seed: int, restart: int = 0) -> np.ndarray:
    """init_cluster_centers_uniform (main.rs:354-364): [k][n_loci] f32."""
    rng = np.random.Generator(np.random.Philox(key=[seed, restart]))
    c = rng.random((k, n_loci), dtype=np.float32)
    return np.clip(c, np.float32(0.0001), np.float32(0.9999))


def souporcell_main(ctx, k: int, restarts: int = 100, seed: int = 4):
    """Best of `restarts` EM runs and the argmax assignment (main.rs:69-106).  `ctx` is a loaded context with a
    `soup_restarts` method (cellector_b200.Cellector): the restarts share the passes over the matrix on the GPU."""
    init = np.stack([init_cluster_centers_uniform(ctx.n_loci_used, k, seed, r) for r in range(restarts)])
    best = ctx.soup_restarts(init)
    # first maximum wins, like the strict > of main.rs:96-99
    best["assignment"] = np.argmax(best["log_probs"], axis=1).astype(np.uint32)
    return best
