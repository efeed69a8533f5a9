"""Runs one named parity case in its own process (so that environment knobs the library reads once per process, like
CLX_SELECT_MODE, can be set per case): python tests/parity_worker.py <case> [clx_create flags]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cellector_b200 import Cellector, synth  # noqa: E402
from tests.oracle import Oracle  # noqa: E402
from tests.parity import run_parity  # noqa: E402

CASES = {
    "c1_slice": synth.SynthConfig(n_cells=2000, n_loci=8000, density=0.02, minority_frac=0.05, seed=1),
    "deep": synth.SynthConfig(n_cells=800, n_loci=2500, density=0.04, minority_frac=0.1, seed=5, deep_frac=0.05, deep_mean=60.0),
    # more than 151,552 cells: the cooperative select holds 4 keys per thread (R = 4)
    "select_r4": synth.SynthConfig(n_cells=160_000, n_loci=600, density=0.02, minority_frac=0.05, seed=19),
    # entries with total > 1000 (k_heavy_giant: one CTA per entry)
    "giant": synth.SynthConfig(n_cells=600, n_loci=2000, density=0.03, minority_frac=0.08, seed=23, deep_frac=0.003, deep_mean=1500.0),
}

if __name__ == "__main__":
    case = sys.argv[1]
    flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    d = synth.generate(CASES[case])
    with Cellector(0, flags) as gpu, Oracle() as orc:
        n_it = run_parity(gpu, orc, d)
    print(f"PARITY_OK case={case} flags={flags} iterations={n_it} CLX_SELECT_MODE={os.environ.get('CLX_SELECT_MODE')}", flush=True)
