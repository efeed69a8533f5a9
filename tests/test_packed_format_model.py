"""Host model of the packed run stream (tools/packed_format.py): the inner loop applied to what the
builder emits gives every cell exactly its own entries, and the stream is no longer than the model says."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import packed_format as pf  # noqa: E402


def _random_sub_unit(rng, n_runs, mean):
    qs = rng.choice(96, size=n_runs, replace=False)
    runs = []
    for q in qs:
        n = 1 + rng.poisson(mean)
        runs.append((int(q), [pf.entry(int(rng.integers(512)), int(rng.integers(9))) for _ in range(n)]))
    return runs


@pytest.mark.parametrize("n_runs,mean", [(1, 3.0), (7, 5.0), (32, 5.0), (33, 5.0), (79, 5.0), (96, 1.0), (96, 40.0)])
def test_consume_of_pack_is_the_direct_sum(n_runs, mean):
    rng = np.random.default_rng(n_runs * 1000 + int(mean))
    tab = np.vstack([rng.normal(size=(512, 13)), np.zeros((1, 13))])
    for _ in range(20):
        runs = _random_sub_unit(rng, n_runs, mean)
        first_cell, lanes = pf.pack_sub_unit(runs)
        got = pf.consume(first_cell, lanes, tab)
        assert set(got) == {q for q, _ in runs}
        for q, ent in runs:
            sT = sE = 0.0
            for v in ent:  # file order inside the run, like the kernel
                rec = tab[v >> 4]
                sT += rec[v & 15]
                sE += rec[9 + pf.TOTAL_OF_SLOT[v & 15] - 1]
            assert got[q] == (sT, sE)  # same order of additions inside a run: bit-identical
        M = len(lanes[0])
        assert all(len(x) == M for x in lanes)
        n_entries = sum(len(e) for _, e in runs)
        assert M >= max(len(e) for _, e in runs) and 32 * M >= n_entries + max(n_runs - 32, 0)
        today, packed = pf.padding_stats([len(e) for _, e in runs])
        assert packed == 32 * M and (packed <= today or n_runs > 32)


def test_empty_sub_unit():
    first_cell, lanes = pf.pack_sub_unit([])
    assert first_cell == [0xFF] * 32 and all(x == [] for x in lanes)
    assert pf.consume(first_cell, lanes, np.zeros((513, 13))) == {}
