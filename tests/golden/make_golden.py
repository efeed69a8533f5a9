"""Regenerates the golden fixtures from the CPU oracle (the reference itself cannot run here: no Rust toolchain,
prebuilt binary absent): oracle_c1_slice.json (small, deep entries), oracle_c1.json and oracle_c2.json (BASELINE.json
configs[0] and configs[1] at full size).  Usage: python tests/golden/make_golden.py"""
import dataclasses
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from cellector_b200 import synth  # noqa: E402
from tests.oracle import Oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
SAMPLES = 24  # per-cell / per-locus values kept verbatim (evenly spaced indices)


def make(cfg: dict, out_file: str) -> None:
    d = synth.generate(synth.SynthConfig(**cfg))
    orc = Oracle()
    out = {"config": cfg, "loaded": list(orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)), "iterations": []}
    while True:
        st = orc.compute_new_excluded(5.0)
        cv, lv = orc.cell_vectors(), orc.locus_vectors()
        s = st.as_dict()
        s.pop("device_ms")
        ci = np.linspace(0, len(cv["ll"]) - 1, SAMPLES).astype(int)
        li = np.linspace(0, len(lv["ll_minority"]) - 1, SAMPLES).astype(int)
        out["iterations"].append({
            "stats": s, "ll_sum": float(cv["ll"].sum()), "expected_ll_sum": float(cv["expected_ll"].sum()),
            "n_loci_sum": int(cv["n_loci"].sum()), "excluded": np.flatnonzero(cv["excluded"]).tolist(),
            "cell_idx": ci.tolist(), "cell_ll": cv["ll"][ci].tolist(), "cell_n_loci": cv["n_loci"][ci].astype(int).tolist(),
            "locus_idx": li.tolist(), "locus_ll_minority": lv["ll_minority"][li].tolist(), "locus_ll_majority": lv["ll_majority"][li].tolist(),
            "locus_ll_minority_sum": float(lv["ll_minority"].sum()), "locus_ll_majority_sum": float(lv["ll_majority"].sum()),
            "locus_cells_minority_sum": int(lv["cells_minority"].astype(np.int64).sum()),
            "locus_cells_majority_sum": int(lv["cells_majority"].astype(np.int64).sum())})
        if not st.any_change:
            break
    p = orc.calculate_posteriors()
    out["label_counts"] = np.bincount(p["label"], minlength=4).tolist()
    out["ll_minor_sum"] = float(p["ll_minor"].sum())
    json.dump(out, open(os.path.join(HERE, out_file), "w"), indent=1)
    print(out_file, "loaded:", out["loaded"], "iterations:", len(out["iterations"]), "labels:", out["label_counts"])


if __name__ == "__main__":
    make(dict(n_cells=1200, n_loci=5000, density=0.03, minority_frac=0.06, depth_lambda=0.4, seed=17, row_sigma=0.0,
              deep_frac=0.01, deep_mean=40.0, doublet_frac=0.0), "oracle_c1_slice.json")
    for name in ("C1", "C2"):
        make(dataclasses.asdict(synth.CONFIGS[name]), f"oracle_{name.lower()}.json")
