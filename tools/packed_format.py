#!/usr/bin/env python
"""Blueprint of the PACKED run stream for the tiled scorer — a host model, not product code.

Today a (cell block, tile, warp) sub-unit is stored as groups of 32 runs, one run per lane, padded to the group's
longest run (51 % null entries at C2, 37 % at a C3 shard).  Packed: the sub-unit is ONE group of M iterations; a lane
slot holds several runs back to back.

16-bit slot values
    bits 0-3   slot of T(alt | total) in the 13-double record (0..8); total follows from the slot (0-1: 1, 2-4: 2, 5-8: 3)
    bits 4-13  locus index inside the tile (0..511); 512 = the all-zero record, i.e. NULL = 0x2000 (padding)
    bit  14    the run ends with this entry: add the lane's sums to the cell's accumulator and clear them
    bit  15    not an entry: the low 8 bits are the cell index (q = cell_local / 32) of the lane's NEXT run
The first run of every lane takes its cell index from the 32-byte table in the first batch header (0xFF = empty lane),
as today.  Packing: runs longest first, each into the least loaded lane (ties: lowest lane); a run that is not the
first of its lane costs one extra slot.  M = the heaviest lane.

`pack_sub_unit` / `consume` are the builder and the inner loop in plain Python; tests/test_packed_format_model.py
checks consume(pack(x)) against direct sums and the padding the model predicts."""
from __future__ import annotations

import numpy as np

NULL = 0x2000
END = 0x4000
CELL = 0x8000
TOTAL_OF_SLOT = (1, 1, 2, 2, 2, 3, 3, 3, 3)


def entry(ll: int, slot: int) -> int:
    assert 0 <= ll < 512 and 0 <= slot < 9
    return (ll << 4) | slot


def pack_sub_unit(runs):
    """runs: list of (q, [entry, ...]) with non-empty entry lists, any order.  Returns (first_cell[32], lanes[32][M])."""
    order = sorted(range(len(runs)), key=lambda i: (-len(runs[i][1]), i))  # longest first, stable
    load = [0] * 32
    lanes = [[] for _ in range(32)]
    first_cell = [0xFF] * 32
    for i in order:
        q, ent = runs[i]
        assert 0 <= q < 0xFF and len(ent) > 0
        lane = min(range(32), key=lambda l: (load[l], l))
        if lanes[lane]:
            lanes[lane].append(CELL | q)
        else:
            first_cell[lane] = q
        lanes[lane].extend(ent[:-1])
        lanes[lane].append(ent[-1] | END)
        load[lane] = len(lanes[lane])
    M = max(load) if runs else 0
    for l in range(32):
        lanes[l].extend([NULL] * (M - len(lanes[l])))
    return first_cell, lanes


def consume(first_cell, lanes, tab13):
    """The inner loop of the kernel, lane by lane.  tab13: [513][13] (record 512 all zero).  Returns {q: (sumT, sumE)}."""
    acc = {}
    for l in range(32):
        cell = first_cell[l]
        sT = sE = 0.0
        for v in lanes[l]:
            if v & CELL:
                cell = v & 0xFF
                continue
            rec = tab13[(v >> 4) & 0x3FF]
            slot = v & 15
            sT += rec[slot]
            sE += rec[9 + TOTAL_OF_SLOT[slot] - 1] if ((v >> 4) & 0x3FF) < 512 else 0.0
            if v & END:
                assert cell != 0xFF
                a = acc.get(cell, (0.0, 0.0))
                acc[cell] = (a[0] + sT, a[1] + sE)
                sT = sE = 0.0
        assert sT == 0.0 and sE == 0.0  # nothing is left behind an END-less tail: only NULLs follow the last run
    return acc


def padding_stats(run_lengths):
    """(lane-iterations today, lane-iterations packed) for one sub-unit given its run lengths."""
    ls = sorted(run_lengths, reverse=True)
    today = sum(ls[g] for g in range(0, len(ls), 32)) * 32
    load = [0] * 32
    for n in ls:
        lane = min(range(32), key=lambda l: (load[l], l))
        load[lane] += n + (1 if load[lane] else 0)
    return today, max(load) * 32 if ls else 0


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for mean, runs in ((5.1, 32), (5.1, 64), (5.1, 79), (5.1, 96)):
        t = p = e = 0
        for _ in range(2000):
            ls = [int(x) for x in rng.poisson(mean, size=runs) if x > 0]
            a, b = padding_stats(ls)
            t, p, e = t + a, p + b, e + sum(ls)
        print(f"runs/sub-unit {runs}: padding today {1 - e / t:.3f}, packed {1 - e / p:.3f} (lane-iterations {p / t:.2f}x)")
