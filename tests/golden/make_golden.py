"""Regenerates tests/golden/oracle_c1_slice.json from the CPU oracle (the reference itself cannot run here:
no Rust toolchain, prebuilt binary absent).  Usage: python tests/golden/make_golden.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from cellector_b200 import synth  # noqa: E402
from tests.oracle import Oracle  # noqa: E402

cfg = dict(n_cells=1200, n_loci=5000, density=0.03, minority_frac=0.06, depth_lambda=0.4, seed=17, row_sigma=0.0,
           deep_frac=0.01, deep_mean=40.0, doublet_frac=0.0)
d = synth.generate(synth.SynthConfig(**cfg))
orc = Oracle()
out = {"config": cfg, "loaded": list(orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)), "iterations": []}
while True:
    st = orc.compute_new_excluded(5.0)
    cv = orc.cell_vectors()
    s = st.as_dict()
    s.pop("device_ms")
    out["iterations"].append({"stats": s, "ll_sum": float(cv["ll"].sum()), "expected_ll_sum": float(cv["expected_ll"].sum()),
                              "n_loci_sum": int(cv["n_loci"].sum()), "excluded": np.flatnonzero(cv["excluded"]).tolist()})
    if not st.any_change:
        break
p = orc.calculate_posteriors()
out["label_counts"] = np.bincount(p["label"], minlength=4).tolist()
out["ll_minor_sum"] = float(p["ll_minor"].sum())
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_c1_slice.json"), "w"), indent=1)
print("iterations:", len(out["iterations"]), "labels:", out["label_counts"])
