#!/usr/bin/env python
"""Stage-by-stage wall-clock of one whole job through the C ABI (diagnostics; not a bench number)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cellector_b200 import Cellector, api, synth

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    cfg = synth.CONFIGS[name]
    data = synth.generate(cfg)
    pins = []
    for a in (data.locus, data.cell, data.ref, data.alt):
        t = torch.empty(a.shape, dtype=torch.int32, pin_memory=True)
        v = t.numpy().view(np.uint32); v[...] = a
        pins.append((t, v))
    host = [p[1] for p in pins]
    ctx = Cellector(0, flags)
    def sync():
        torch.cuda.synchronize()
    for rep in range(4):
        sync(); t0 = time.perf_counter()
        ctx.load_cell_data(*host, data.n_loci, data.n_cells); sync(); t1 = time.perf_counter()
        stats = ctx.run_em(5.0); sync(); t2 = time.perf_counter()
        cv = ctx.cell_vectors(); lv = ctx.locus_vectors(); sync(); t3 = time.perf_counter()
        pv = ctx.calculate_posteriors(); sync(); t4 = time.perf_counter()
        print(f"rep {rep}: load {1e3*(t1-t0):.2f} ms | em x{len(stats)} {1e3*(t2-t1):.2f} ms "
              f"(device {sum(s.device_ms for s in stats):.3f}) | vectors {1e3*(t3-t2):.2f} ms | posteriors {1e3*(t4-t3):.2f} ms "
              f"| total {1e3*(t4-t0):.2f} ms", flush=True)
    ctx.close()

if __name__ == "__main__":
    main()
