"""clx_locus_filter_static — the entry point that allocates nothing (its scratch is a static __device__ array inside
the library; include/cellector_cuda.h) — against the existing entry point clx_load_coo + clx_get_locus_ids, through the
C ABI, on two calls in a row.  Integer work: the comparison is bit-exact, like every other locus-filter check here.

The file sorts last on purpose: a failure here hides no other test.
The second call uses FEWER loci than the first and other counts: stale counters or a stale kept-count left in the
shared static array by the first call would show up in its output."""

import numpy as np
import pytest

from cellector_b200 import synth

pytestmark = pytest.mark.gpu

CASES = [
    synth.SynthConfig(n_cells=900, n_loci=7000, density=0.02, minority_frac=0.08, seed=31),
    synth.SynthConfig(n_cells=400, n_loci=2500, density=0.03, minority_frac=0.1, seed=32, deep_frac=0.05),
]


def _existing(lib, d, min_ref, min_alt):
    """used-locus ids from the existing path: clx_load_coo's locus filter + clx_get_locus_ids"""
    from cellector_b200 import Cellector
    with Cellector(0, 0, lib) as gpu:
        n_used, _ = gpu.load_cell_data(d.locus, d.cell, d.ref, d.alt, d.n_loci, d.n_cells, min_ref, min_alt)
        ids = gpu.locus_ids()
    assert ids.shape[0] == n_used
    return ids


def _static(lib, bufs, d, min_ref, min_alt):
    """the same filter through clx_locus_filter_static; the caller (this test) owns every device buffer"""
    import torch
    nnz = d.locus.shape[0]
    for k, a in (("locus", d.locus), ("ref", d.ref), ("alt", d.alt)):
        bufs[k][:nnz].copy_(torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32)))
    bufs["used"].fill_(7)      # whatever the kernel does not write stays visible
    bufs["n_used"].fill_(-1)
    torch.cuda.synchronize()
    rc = lib.clx_locus_filter_static(0, nnz, d.n_loci, bufs["locus"].data_ptr(), bufs["ref"].data_ptr(),
                                     bufs["alt"].data_ptr(), min_ref, min_alt, bufs["used"].data_ptr(),
                                     bufs["n_used"].data_ptr(), None)
    assert rc == 0, lib.clx_last_error(None)
    used = bufs["used"].cpu().numpy()
    assert set(np.unique(used[:d.n_loci])) <= {0, 1}
    assert (used[d.n_loci:] == 7).all(), "wrote past n_loci0"
    return np.flatnonzero(used[:d.n_loci]).astype(np.uint32), int(bufs["n_used"].cpu().numpy()[0])


def test_static_locus_filter_matches_load_coo_twice():
    import torch
    from cellector_b200 import api
    lib = api.load_library()
    data = [synth.generate(c) for c in CASES]
    cap_nnz = max(d.locus.shape[0] for d in data)
    cap_loci = max(d.n_loci for d in data)
    dev = torch.device("cuda:0")
    bufs = {k: torch.zeros(cap_nnz, dtype=torch.int32, device=dev) for k in ("locus", "ref", "alt")}
    bufs["used"] = torch.zeros(cap_loci + 64, dtype=torch.uint8, device=dev)
    bufs["n_used"] = torch.zeros(1, dtype=torch.int32, device=dev)

    # two calls in a row (larger then smaller), then the first input again with other minima
    for d, (min_ref, min_alt) in ((data[0], (4, 4)), (data[1], (4, 4)), (data[0], (2, 6))):
        want = _existing(lib, d, min_ref, min_alt)
        got, n_used = _static(lib, bufs, d, min_ref, min_alt)
        assert n_used == want.shape[0]
        np.testing.assert_array_equal(got, want)
        assert 0 < want.shape[0] < d.n_loci  # the filter both keeps and drops loci in these cases


def test_static_locus_filter_rejects_what_it_cannot_hold():
    import torch
    from cellector_b200 import api
    lib = api.load_library()
    one = torch.zeros(4, dtype=torch.int32, device="cuda:0")
    used = torch.zeros(8, dtype=torch.uint8, device="cuda:0")
    max_loci = 1 << 21  # CLX_STATIC_MAX_LOCI
    rc = lib.clx_locus_filter_static(0, 0, max_loci + 1, None, None, None, 4, 4, used.data_ptr(), None, None)
    assert rc == 2 and b"CLX_STATIC_MAX_LOCI" in lib.clx_last_error(None)
    # a locus index outside [0, n_loci0) is an argument error, as in clx_load_coo
    one[:] = torch.tensor([0, 1, 9, 2], dtype=torch.int32)
    rc = lib.clx_locus_filter_static(0, 4, 4, one.data_ptr(), one.data_ptr(), one.data_ptr(), 1, 1, used.data_ptr(),
                                     None, None)
    assert rc == 2 and b"out of range" in lib.clx_last_error(None)
    # and the next call is not poisoned by that error flag (the tail words are cleared per call)
    one[:] = torch.tensor([0, 1, 1, 3], dtype=torch.int32)
    n_used = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    rc = lib.clx_locus_filter_static(0, 4, 4, one.data_ptr(), one.data_ptr(), one.data_ptr(), 1, 1, used.data_ptr(),
                                     n_used.data_ptr(), None)
    assert rc == 0, lib.clx_last_error(None)
    assert used[:4].cpu().tolist() == [0, 1, 0, 1] and int(n_used.item()) == 2
