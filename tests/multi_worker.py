"""Worker of tests/test_gpu_multi.py: one process per GPU (torchrun), cells sharded across ranks; every
iteration rank 0 compares the gathered results with the CPU oracle run on the whole matrix."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from cellector_b200 import Cellector, api, synth  # noqa: E402
from tests import parity  # noqa: E402


class Gathered:
    """Duck-types the CtxApi read calls over results gathered from all ranks."""

    def __init__(self, cells, loci, post=None):
        self._c, self._l, self._p = cells, loci, post

    def cell_vectors(self):
        return self._c

    def locus_vectors(self):
        return self._l


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    case = sys.argv[1] if len(sys.argv) > 1 else "multi"
    cfgs = {
        "multi": synth.SynthConfig(n_cells=3000, n_loci=5000, density=0.02, minority_frac=0.06, seed=13, deep_frac=0.01),
        "multi_big": synth.SynthConfig(n_cells=36000, n_loci=1500, density=0.004, minority_frac=0.05, seed=17),
    }
    cfg = cfgs[case]
    d = synth.generate(cfg, shard=rank, n_shards=world)
    lib = api.load_library()
    gpu = Cellector(local, 0, lib)
    uid = [Cellector.nccl_unique_id(lib) if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    gpu.comm_init(rank, world, uid[0])
    g = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
    shards = [None] * world
    dist.all_gather_object(shards, (d.cell_offset, d.n_cells, d.locus, d.cell, d.ref, d.alt))
    orc = None
    if rank == 0:
        from tests.oracle import Oracle
        locus = np.concatenate([s[2] for s in shards]).astype(np.uint64)
        cell = np.concatenate([s[3].astype(np.uint64) + s[0] for s in shards])
        ref = np.concatenate([s[4] for s in shards])
        alt = np.concatenate([s[5] for s in shards])
        order = np.lexsort((cell, locus))  # file order: sorted by (locus, cell)
        n_total = sum(s[1] for s in shards)
        whole = (locus[order].astype(np.uint32), cell[order].astype(np.uint32), ref[order], alt[order], d.n_loci, n_total)
        orc = Oracle()
        o = orc.load_cell_data(locus[order].astype(np.uint32), cell[order].astype(np.uint32), ref[order], alt[order],
                               d.n_loci, n_total)
        nnz_all = [None] * world
    else:
        nnz_all = [None] * world
    dist.all_gather_object(nnz_all, g)
    if rank == 0:
        assert nnz_all[0][0] == o[0], f"n_loci_used {nnz_all[0][0]} vs {o[0]}"
        assert sum(x[1] for x in nnz_all) == o[1], f"nnz_used {sum(x[1] for x in nnz_all)} vs {o[1]}"
        parity.assert_equal(gpu.locus_ids(), orc.locus_ids(), "locus_ids")
    n_it = 0
    for it in range(30):
        gs = gpu.compute_new_excluded(5.0)
        cv, lv = gpu.cell_vectors(), gpu.locus_vectors()
        allc = [None] * world
        dist.all_gather_object(allc, cv)
        stop = [False]
        if rank == 0:
            os_ = orc.compute_new_excluded(5.0)
            cells = {k: np.concatenate([a[k] for a in allc]) for k in cv}
            parity.compare_iteration(gs, os_, Gathered(cells, lv), orc, it)
            stop[0] = not os_.any_change
        dist.broadcast_object_list(stop, src=0)
        n_it += 1
        assert stop[0] == (not gs.any_change), "ranks disagree with the oracle about convergence"
        if stop[0]:
            break
    pv = gpu.calculate_posteriors(0.999, 30)
    allp = [None] * world
    dist.all_gather_object(allp, pv)
    if rank == 0:
        op = orc.calculate_posteriors(0.999, 30)
        gp = {k: np.concatenate([a[k] for a in allp]) for k in pv}
        for k in ("label", "anomaly", "qual"):
            parity.assert_equal(gp[k], op[k], f"posterior {k}")
        for k in ("ll_major", "ll_minor"):
            parity.assert_close(gp[k], op[k], f"posterior {k}")
        print(f"MULTI_OK case={case} world={world} iterations={n_it} cells={len(gp['label'])} CLX_P2P={os.environ.get('CLX_P2P')}", flush=True)
    # souporcell's EM on the same shards (soup.cu: partial sums and the loss are all-reduced), single run and batched
    from cellector_b200 import soup
    inits = np.stack([soup.init_cluster_centers_uniform(gpu.n_loci_used, 2, 4, r) for r in range(3)])
    s1 = gpu.soup_em(inits[0])
    s3 = gpu.soup_restarts(inits)
    alls = [None] * world
    dist.all_gather_object(alls, (s1["log_probs"], s3["log_probs"]))
    if rank == 0:
        from tests.soup_oracle import SoupOracle
        with SoupOracle() as so:
            so.load_cell_data(*whole)
            runs = [so.soup_em(inits[r], precise=True) for r in range(3)]
        assert s1["iterations"] == runs[0]["iterations"], (s1["iterations"], runs[0]["iterations"])
        np.testing.assert_allclose(s1["loss"], runs[0]["loss"], rtol=1e-10)
        np.testing.assert_allclose(np.concatenate([a[0] for a in alls]), runs[0]["log_probs"], rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(s1["centers"], runs[0]["centers"], rtol=1e-9)
        assert [int(x) for x in s3["iterations"]] == [r["iterations"] for r in runs]
        np.testing.assert_allclose(s3["losses"], [r["loss"] for r in runs], rtol=1e-10)
        b = int(np.argmax([r["loss"] for r in runs]))
        assert s3["best"] == b
        np.testing.assert_allclose(np.concatenate([a[1] for a in alls]), runs[b]["log_probs"], rtol=1e-9, atol=1e-9)
        print(f"MULTI_SOUP_OK case={case} world={world} iterations={s1['iterations']}", flush=True)
    gpu.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
