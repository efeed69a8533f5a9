#!/usr/bin/env python
"""Host-side model of the tiled scorer's layout (csrc/tiled.cu): for a synthetic config and a cell-block size NB it
counts runs, groups, padded lane-iterations and stream bytes, i.e. how much of the lookup loop is padding.
No GPU needed.  Usage: python tools/tiled_model.py C2 [NB ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cellector_b200 import synth  # noqa: E402

TILE_L, WARPS, BATCH_ITERS, STAGE = 512, 32, 15, 1024


def filtered(d, min_ref=4, min_alt=4):
    """The locus filter of load_data.rs:254-280 and the dense renumbering."""
    nref = np.bincount(d.locus[d.ref > 0], minlength=d.n_loci)
    nalt = np.bincount(d.locus[d.alt > 0], minlength=d.n_loci)
    used = (nref >= min_ref) & (nalt >= min_alt)
    new_id = np.cumsum(used) - 1
    keep = used[d.locus]
    return new_id[d.locus[keep]].astype(np.int64), d.cell[keep].astype(np.int64), (d.ref[keep] + d.alt[keep]).astype(np.int64), int(used.sum())


def model(locus, cell, total, n_cells, L, NB):
    light = (total >= 1) & (total <= 3)
    locus, cell = locus[light], cell[light]
    nT = (L + TILE_L - 1) // TILE_L
    cb, cl, t = cell // NB, cell % NB, locus // TILE_L
    w = cl % WARPS
    # run = (cell block, warp, tile, cell); sub-unit = (cell block, warp, tile)
    su = (cb * WARPS + w) * nT + t
    run_key = su * NB + cl
    uk, cnt = np.unique(run_key, return_counts=True)
    su_of_run = uk // NB
    # inside a sub-unit: longest first, 32 runs per group; a group costs max(len) iterations of 32 lanes
    order = np.lexsort((-cnt, su_of_run))
    su_s, cnt_s = su_of_run[order], cnt[order]
    first = np.r_[True, su_s[1:] != su_s[:-1]]
    start_idx = np.maximum.accumulate(np.where(first, np.arange(len(su_s)), 0))
    pos = np.arange(len(su_s)) - start_idx
    leader = pos % 32 == 0
    group_iters = cnt_s[leader]                   # sorted descending: the leader is the longest run of its group
    n_groups = int(leader.sum())
    lane_iters = int(group_iters.sum()) * 32
    batches = int(np.ceil(group_iters / BATCH_ITERS).sum())
    stream_bytes = 64 * int(group_iters.sum()) + 32 * n_groups + 16 * batches
    n_sub_nonempty = int(first.sum())
    # what run packing could reach: a sub-unit's runs spread over 32 lane slots, several short runs per slot, one extra
    # slot entry per additional run (the in-line cell index): iterations = max(longest run, ceil(slots needed / 32))
    su_id = np.cumsum(first) - 1
    su_entries = np.bincount(su_id, weights=cnt_s).astype(np.int64)
    su_runs = np.bincount(su_id).astype(np.int64)
    su_longest = cnt_s[first]
    packed_iters = np.maximum(su_longest, -(-(su_entries + np.maximum(su_runs - 32, 0)) // 32))
    packed_lane_iters = int(packed_iters.sum()) * 32
    # a packing the device builder can do by formula: sorted runs dealt to the lanes boustrophedon (0..31, 31..0, ...);
    # a lane's load = its entries + one slot per extra run; the sub-unit runs for the heaviest lane
    rnd, idx = pos // 32, pos % 32
    lane = np.where(rnd % 2 == 0, idx, 31 - idx)
    slot = su_id * 32 + lane
    load = np.bincount(slot, weights=cnt_s + (rnd > 0), minlength=int(su_id.max() + 1) * 32).reshape(-1, 32)
    snake_lane_iters = int(load.max(axis=1).sum()) * 32
    return {"NB": NB, "cell_blocks": int((n_cells + NB - 1) // NB), "tiles": int(nT), "light_entries": int(light.sum()),
            "runs": int(len(uk)), "mean_run": float(cnt.mean()), "groups": n_groups,
            "lane_iterations": lane_iters, "padding_frac": 1.0 - light.sum() / lane_iters,
            "stream_MB": stream_bytes / 1e6, "bytes_per_entry": stream_bytes / light.sum(),
            "table_MB_read": (n_cells + NB - 1) // NB * nT * TILE_L * 104 / 1e6, "nonempty_sub_units": n_sub_nonempty,
            "packed_lane_iterations": packed_lane_iters, "packed_padding_frac": 1.0 - light.sum() / packed_lane_iters,
            "snake_lane_iterations": snake_lane_iters, "snake_padding_frac": 1.0 - light.sum() / snake_lane_iters}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    nbs = [int(x) for x in sys.argv[2:]] or [1024, 2048, 3072]
    d = synth.generate(synth.CONFIGS[name])
    locus, cell, total, L = filtered(d)
    print(f"{name}: {d.n_cells} cells, {L} loci used, {len(locus)} scored nonzeros")
    for NB in nbs:
        m = model(locus, cell, total, d.n_cells, L, NB)
        print("  ".join(f"{k}={v:.3f}" if isinstance(v, float) else f"{k}={v}" for k, v in m.items()))


if __name__ == "__main__":
    main()
