"""Regenerates the golden fixtures of souporcell's EM from its CPU oracle (oracle/souporcell_oracle.cc; the reference
cannot run here): soup_small_k3.json (deep entries, doublets, k = 3) and soup_c1_k2.json (BASELINE.json configs[0] at
full size, k = 2 as cellector_pipeline.py:233-240 runs it).  Initial centers: tests/soup_rng.py (ChaCha20 in plain
Python: the same bits on any machine).  Usage: python tests/golden/make_soup_golden.py"""
import dataclasses
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from cellector_b200 import synth  # noqa: E402
from tests.soup_oracle import SoupOracle  # noqa: E402
from tests.soup_rng import souporcell_initial_centers  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
SAMPLES = 24


def make(cfg: dict, k: int, seed: int, out_file: str) -> None:
    d = synth.generate(synth.SynthConfig(**cfg))
    with SoupOracle() as orc:
        n_used, n_entries = orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        init = souporcell_initial_centers(seed, 1, 1, k, n_used)[0]
        f64 = orc.soup_em(init, precise=True)
        f32 = orc.soup_em(init, precise=False)
    ci = np.linspace(0, d.n_cells - 1, SAMPLES).astype(int)
    li = np.linspace(0, n_used - 1, SAMPLES).astype(int)
    out = {"config": cfg, "k": k, "seed": seed, "loaded": [n_used, n_entries],
           "init_sum": float(init.astype(np.float64).sum()),
           "f64": {"iterations": f64["iterations"], "loss": f64["loss"], "loss_trace": f64["loss_trace"].tolist(),
                   "log_probs_sum": f64["log_probs"].sum(axis=0).tolist(), "cell_idx": ci.tolist(),
                   "cell_log_probs": f64["log_probs"][ci].tolist(), "locus_idx": li.tolist(),
                   "locus_centers": f64["centers"][:, li].tolist(),
                   "assignment_counts": np.bincount(np.argmax(f64["log_probs"], axis=1), minlength=k).tolist()},
           "f32": {"iterations": f32["iterations"], "loss": f32["loss"],
                   "assignment_counts": np.bincount(np.argmax(f32["log_probs"], axis=1), minlength=k).tolist()}}
    json.dump(out, open(os.path.join(HERE, out_file), "w"), indent=1)
    print(out_file, out["loaded"], "iterations", f64["iterations"], f32["iterations"], "loss", f64["loss"], f32["loss"],
          out["f64"]["assignment_counts"])


if __name__ == "__main__":
    make(dict(n_cells=900, n_loci=3000, density=0.05, minority_frac=0.4, depth_lambda=0.4, seed=52, row_sigma=0.0,
              deep_frac=0.05, deep_mean=40.0, doublet_frac=0.05), 3, 4, "soup_small_k3.json")
    make(dataclasses.asdict(synth.CONFIGS["C1"]), 2, 4, "soup_c1_k2.json")
