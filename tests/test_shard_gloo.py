"""N > 1 host logic on CPU: two gloo ranks shard the cells, all-reduce the locus-filter counts and gather
per-cell vectors; the result must be what the one-process oracle computes on the whole matrix."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from cellector_b200 import shard, synth
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    d = synth.generate(synth.SynthConfig(n_cells=900, n_loci=2500, density=0.02, seed=31, row_sigma=1.5))
    row_nnz = np.bincount(d.cell, minlength=d.n_cells)
    bounds = shard.partition_cells(row_nnz, world)
    locus, cell, ref, alt, n_local = shard.shard_coo(d.locus, d.cell, d.ref, d.alt, bounds, rank)
    # the shard keeps file order and covers exactly its cells
    assert n_local == bounds[rank + 1] - bounds[rank] and (cell < n_local).all()
    assert np.all(np.diff(locus.astype(np.int64)) >= 0)
    counts = shard.allreduce_sum(shard.locus_filter_counts(locus, ref, alt, d.n_loci))
    used = (counts[:, 0] >= 4) & (counts[:, 1] >= 4)
    nnz_used_local = int(used[locus].sum())
    nnz_used = int(shard.allreduce_sum(np.array([nnz_used_local]))[0])
    # a per-cell vector computed per shard comes back in global order
    local_rows = np.bincount(cell[used[locus]], minlength=n_local)
    rows = shard.gather_cells(local_rows, bounds)
    if rank == 0:
        np.savez(os.path.join(out_dir, "r0.npz"), used=used, nnz_used=nnz_used, rows=rows, bounds=bounds, row_nnz=row_nnz)
    dist.destroy_process_group()


def test_two_rank_sharding_matches_oracle(tmp_path, oracle_lib):
    import torch.multiprocessing as mp
    from cellector_b200 import synth
    from tests.oracle import Oracle
    world = 2
    mp.spawn(_worker, args=(world, 29561, str(tmp_path)), nprocs=world, join=True)
    r = np.load(tmp_path / "r0.npz")
    d = synth.generate(synth.SynthConfig(n_cells=900, n_loci=2500, density=0.02, seed=31, row_sigma=1.5))
    with Oracle(oracle_lib) as orc:
        n_loci_used, nnz_used = orc.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells)
        ids = orc.locus_ids()
        orc.compute_new_excluded(5.0)
        n_loci_per_cell = orc.cell_vectors()["n_loci"]
    assert int(r["used"].sum()) == n_loci_used and int(r["nnz_used"]) == nnz_used
    assert np.array_equal(np.flatnonzero(r["used"]).astype(np.uint32), ids)
    # iteration 0 uses every kept locus, so loci_used_per_cell is the kept-entry count of the row
    assert np.array_equal(r["rows"].astype(np.uint32), n_loci_per_cell)
    b = r["bounds"]
    assert b[0] == 0 and b[-1] == 900 and np.all(np.diff(b) >= 0)
    halves = [int(r["row_nnz"][b[i]:b[i + 1]].sum()) for i in range(world)]
    assert abs(halves[0] - halves[1]) <= int(r["row_nnz"].max()), "nnz-balanced contiguous ranges"


def test_partition_edge_cases():
    from cellector_b200 import shard
    assert list(shard.partition_cells(np.zeros(5, int), 2)) == [0, 0, 5]
    assert list(shard.partition_cells(np.array([10]), 4)) == [0, 0, 0, 1, 1] or list(shard.partition_cells(np.array([10]), 4))[-1] == 1
    b = shard.partition_cells(np.ones(100, int), 8)
    assert b[0] == 0 and b[-1] == 100 and np.all(np.diff(b) >= 12) and np.all(np.diff(b) <= 13)
    assert list(shard.partition_cells(np.array([], int), 3)) == [0, 0, 0, 0]
