#!/usr/bin/env python
"""End-to-end seconds of the `cellector` CLI on a named synthetic workload written as real .mtx text
(BASELINE.json metric, second half: "end-to-end sec"): process start -> last file closed."""
import json, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cellector_b200 import synth

def write_fast(d, out_dir):
    import pandas as pd
    os.makedirs(out_dir, exist_ok=True)
    header = "%%MatrixMarket matrix coordinate real general\n% written by sprs\n" + f"{d.n_loci}\t{d.n_cells}\t{d.nnz}\n"
    for name, vals in (("alt.mtx", d.alt), ("ref.mtx", d.ref)):
        p = os.path.join(out_dir, name)
        with open(p, "w") as f:
            f.write(header)
        pd.DataFrame({"l": d.locus.astype(np.int64) + 1, "c": d.cell.astype(np.int64) + 1, "v": vals}).to_csv(
            p, sep="\t", header=False, index=False, mode="a")
    with open(os.path.join(out_dir, "barcodes.tsv"), "w") as f:
        f.write("\n".join(f"CELL{i:09d}-1" for i in range(d.n_cells)) + "\n")
    with open(os.path.join(out_dir, "gt.tsv"), "w") as f:
        f.write("".join(f"CELL{i:09d}-1\t{'minority' if m else 'majority'}\n" for i, m in enumerate(d.is_minority)))

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    work = sys.argv[2] if len(sys.argv) > 2 else "/tmp/clx_e2e"
    d = synth.generate(synth.CONFIGS[name])
    write_fast(d, work + "/in")
    cli = os.path.join(ROOT, "cellector_b200", "bin", "cellector")
    res = []
    for rep, extra in enumerate(([], [], ["--no_iteration_files"])):
        out = f"{work}/out{rep}"
        t0 = time.perf_counter()
        r = subprocess.run([cli, "-a", work + "/in/alt.mtx", "-r", work + "/in/ref.mtx", "-b", work + "/in/barcodes.tsv", "-g",
                            work + "/in/gt.tsv", "--output_directory", out, "--timing_json", out + ".json", *extra],
                           capture_output=True, text=True)
        wall = time.perf_counter() - t0
        assert r.returncode == 0, r.stderr[-2000:]
        t = json.load(open(out + ".json"))
        t["process_wall_s"] = wall
        t["flags"] = extra
        res.append(t)
    print(json.dumps({"workload": name, "mtx_bytes": os.path.getsize(work + "/in/alt.mtx") + os.path.getsize(work + "/in/ref.mtx"), "runs": res}))
    print(r.stdout[-1500:], file=sys.stderr)

if __name__ == "__main__":
    main()
