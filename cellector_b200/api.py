"""ctypes binding of the C ABI in include/cellector_cuda.h.

This is glue for tests and bench.py; the product is libcellector_cuda.so (and the `cellector` CLI built on it).
The method names mirror the reference functions they stand for (cellector/src/main.rs, load_data.rs):
load_cell_data -> clx_load_coo, compute_new_excluded -> clx_em_iteration, calculate_posteriors -> clx_posteriors.

There is no CPU fallback: `Cellector()` raises if the CUDA library is missing or no device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcellector_cuda.so")

FLAG_PROFILE = 1
FLAG_NO_EXPECTED = 2
FLAG_KERNEL_V0 = 4
FLAG_KEEP_COO = 8
FLAG_NO_GRAPH = 16

PHASES = ("tables", "score", "heavy", "select", "column", "comm")
LABELS = ("0", "1", "doublet", "unassigned")


class IterStats(C.Structure):
    _fields_ = [
        ("any_change", C.c_int32), ("iteration", C.c_uint32),
        ("n_new_excluded", C.c_uint64), ("n_rescued", C.c_uint64), ("n_excluded", C.c_uint64),
        ("median", C.c_double), ("q1", C.c_double), ("q3", C.c_double), ("iqr", C.c_double),
        ("threshold", C.c_double),
        ("n_loci_filtered", C.c_uint64), ("n_loci_in_use", C.c_uint64),
        ("device_ms", C.c_float), ("reserved", C.c_float),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class LocusArrays(C.Structure):
    _fields_ = [
        ("ll_minority", C.c_void_p), ("ll_majority", C.c_void_p),
        ("cells_minority", C.c_void_p), ("cells_majority", C.c_void_p),
        ("ref_minority", C.c_void_p), ("alt_minority", C.c_void_p),
        ("ref_majority", C.c_void_p), ("alt_majority", C.c_void_p),
        ("loci_used", C.c_void_p),
    ]


# symbol -> (restype, argtypes); the exported surface of include/cellector_cuda.h
ABI = {
    "create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_uint]),
    "destroy": (None, [C.c_void_p]),
    "last_error": (C.c_char_p, [C.c_void_p]),
    "nccl_unique_id": (C.c_int, [C.c_void_p]),
    "comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "load_coo": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                           C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]),
    "load_coo_packed": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]),
    "get_locus_ids": (C.c_int, [C.c_void_p, C.c_void_p]),
    "reset": (C.c_int, [C.c_void_p]),
    "em_iteration": (C.c_int, [C.c_void_p, C.c_double, C.POINTER(IterStats)]),
    "get_cell_vectors": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "get_locus_vectors": (C.c_int, [C.c_void_p, C.POINTER(LocusArrays)]),
    "posteriors": (C.c_int, [C.c_void_p, C.c_double, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                             C.c_void_p, C.c_void_p, C.c_void_p]),
    "final_locus_counts": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "get_phase_ms": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
    "reset_phase_ms": (C.c_int, [C.c_void_p]),
    "set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "get_stream": (C.c_void_p, [C.c_void_p]),
    "score_bytes": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "launch_count": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "flush_l2": (C.c_int, [C.c_void_p]),
    "soup_em": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double),
                          C.POINTER(C.c_uint32)]),
    "soup_restarts": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.POINTER(C.c_uint32)]),
    "locus_filter_static": (C.c_int, [C.c_int, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                      C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "alloc_pinned": (C.c_void_p, [C.c_size_t]),
    "free_pinned": (None, [C.c_void_p]),
}


def bind(lib: C.CDLL, prefix: str, names=None):
    for name in (names or ABI):
        fn = getattr(lib, prefix + name)
        fn.restype, fn.argtypes = ABI[name]
    return lib


def load_library(path: str = "") -> C.CDLL:
    path = path or os.environ.get("CLX_LIB_PATH") or LIB_PATH  # CLX_LIB_PATH: A/B runs of two builds of the library
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    return bind(C.CDLL(path, mode=C.RTLD_GLOBAL), "clx_")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class CtxApi:
    """Thin object wrapper over one context of a library exporting `<prefix>*` (see ABI)."""

    def __init__(self, lib: C.CDLL, prefix: str, handle):
        self._lib, self._p, self._h = lib, prefix, handle
        self.n_cells = 0
        self.n_loci0 = 0
        self.n_loci_used = 0
        self.nnz_used = 0

    def _f(self, name):
        return getattr(self._lib, self._p + name)

    def _check(self, rc, what):
        if rc != 0:
            msg = self._f("last_error")(self._h)
            raise RuntimeError(f"{self._p}{what} failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if self._h:
            self._f("destroy")(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # load_data::load_cell_data (load_data.rs:134-181)
    def load_cell_data(self, locus, cell, ref, alt, n_loci0, n_cells, min_ref=4, min_alt=4):
        arrs = [np.ascontiguousarray(x, dtype=np.uint32) for x in (locus, cell, ref, alt)]
        nnz0 = arrs[0].shape[0]
        assert all(a.shape[0] == nnz0 for a in arrs)
        nl, nz = C.c_uint32(0), C.c_uint64(0)
        self._check(self._f("load_coo")(self._h, nnz0, n_loci0, n_cells, *[_ptr(a) for a in arrs], min_ref, min_alt,
                                        C.byref(nl), C.byref(nz)), "load_coo")
        self.n_cells, self.n_loci0 = int(n_cells), int(n_loci0)
        self.n_loci_used, self.nnz_used = int(nl.value), int(nz.value)
        return self.n_loci_used, self.nnz_used

    def load_cell_data_ptrs(self, locus_ptr, cell_ptr, ref_ptr, alt_ptr, nnz0, n_loci0, n_cells, min_ref=4, min_alt=4):
        """load_cell_data from raw addresses of four uint32 arrays — host or DEVICE memory (clx_load_coo copies with
        cudaMemcpyDefault), e.g. torch tensors' data_ptr() for matrices generated on the GPU."""
        nl, nz = C.c_uint32(0), C.c_uint64(0)
        self._check(self._f("load_coo")(self._h, int(nnz0), n_loci0, n_cells, C.c_void_p(locus_ptr), C.c_void_p(cell_ptr),
                                        C.c_void_p(ref_ptr), C.c_void_p(alt_ptr), min_ref, min_alt, C.byref(nl), C.byref(nz)),
                    "load_coo")
        self.n_cells, self.n_loci0 = int(n_cells), int(n_loci0)
        self.n_loci_used, self.nnz_used = int(nl.value), int(nz.value)
        return self.n_loci_used, self.nnz_used

    def load_cell_data_packed(self, locus, cell, ref_alt, n_loci0, n_cells, min_ref=4, min_alt=4):
        """load_cell_data with both counts of an entry in one word (ref | alt << 16)."""
        arrs = [np.ascontiguousarray(x, dtype=np.uint32) for x in (locus, cell, ref_alt)]
        nnz0 = arrs[0].shape[0]
        nl, nz = C.c_uint32(0), C.c_uint64(0)
        self._check(self._f("load_coo_packed")(self._h, nnz0, n_loci0, n_cells, *[_ptr(a) for a in arrs], min_ref, min_alt,
                                               C.byref(nl), C.byref(nz)), "load_coo_packed")
        self.n_cells, self.n_loci0 = int(n_cells), int(n_loci0)
        self.n_loci_used, self.nnz_used = int(nl.value), int(nz.value)
        return self.n_loci_used, self.nnz_used

    def locus_ids(self):
        out = np.empty(self.n_loci_used, np.uint32)
        self._check(self._f("get_locus_ids")(self._h, _ptr(out)), "get_locus_ids")
        return out

    def reset(self):
        self._check(self._f("reset")(self._h), "reset")

    # compute_new_excluded (main.rs:308-347)
    def compute_new_excluded(self, iqr_multiple=5.0) -> IterStats:
        st = IterStats()
        self._check(self._f("em_iteration")(self._h, float(iqr_multiple), C.byref(st)), "em_iteration")
        return st

    def cell_vectors(self):
        n = self.n_cells
        out = dict(ll=np.empty(n), expected_ll=np.empty(n), n_loci=np.empty(n, np.uint32), norm=np.empty(n),
                   excluded=np.empty(n, np.uint8))
        self._check(self._f("get_cell_vectors")(self._h, *[_ptr(out[k]) for k in
                                                           ("ll", "expected_ll", "n_loci", "norm", "excluded")]),
                    "get_cell_vectors")
        return out

    def locus_vectors(self):
        n = self.n_loci_used
        out = dict(ll_minority=np.empty(n), ll_majority=np.empty(n),
                   cells_minority=np.empty(n, np.uint32), cells_majority=np.empty(n, np.uint32),
                   ref_minority=np.empty(n, np.uint64), alt_minority=np.empty(n, np.uint64),
                   ref_majority=np.empty(n, np.uint64), alt_majority=np.empty(n, np.uint64),
                   loci_used=np.empty(n, np.uint8))
        la = LocusArrays(**{k: out[k].ctypes.data for k in out})
        self._check(self._f("get_locus_vectors")(self._h, C.byref(la)), "get_locus_vectors")
        return out

    # calculate_posteriors (main.rs:228-280) + assignment rule (main.rs:141-170)
    def calculate_posteriors(self, posterior_threshold=0.999, min_loci_used=30):
        n = self.n_cells
        out = dict(posterior=np.empty(n), doublet_posterior=np.empty(n), ll_major=np.empty(n), ll_minor=np.empty(n),
                   label=np.empty(n, np.uint8), anomaly=np.empty(n, np.uint8), qual=np.empty(n, np.uint8))
        self._check(self._f("posteriors")(self._h, float(posterior_threshold), int(min_loci_used),
                                          *[_ptr(out[k]) for k in ("posterior", "doublet_posterior", "ll_major",
                                                                   "ll_minor", "label", "anomaly", "qual")]),
                    "posteriors")
        return out

    def final_locus_counts(self):
        mn = np.empty((self.n_loci0, 2), np.uint64)
        mj = np.empty((self.n_loci0, 2), np.uint64)
        self._check(self._f("final_locus_counts")(self._h, _ptr(mn), _ptr(mj)), "final_locus_counts")
        return mn, mj

    # souporcell's EM (souporcell/souporcell/src/main.rs:109-186) from the caller's initial centers [k][n_loci_used]
    def soup_em(self, centers_init):
        cen = np.ascontiguousarray(centers_init, dtype=np.float32)
        k = cen.shape[0]
        assert cen.shape == (k, self.n_loci_used)
        lp = np.empty((self.n_cells, k), np.float64)
        out = np.empty((k, self.n_loci_used), np.float64)
        loss, it = C.c_double(0), C.c_uint32(0)
        self._check(self._f("soup_em")(self._h, k, _ptr(cen), _ptr(lp), _ptr(out), C.byref(loss), C.byref(it)), "soup_em")
        return dict(log_probs=lp, centers=out, loss=loss.value, iterations=int(it.value))

    # best of several restarts (souporcell_main, main.rs:69-88); centers_init [n_restarts][k][n_loci_used]
    def soup_restarts(self, centers_init):
        cen = np.ascontiguousarray(centers_init, dtype=np.float32)
        n, k = cen.shape[0], cen.shape[1]
        assert cen.shape == (n, k, self.n_loci_used)
        lp = np.empty((self.n_cells, k), np.float64)
        out = np.empty((k, self.n_loci_used), np.float64)
        losses, its, best = np.empty(n, np.float64), np.empty(n, np.uint32), C.c_uint32(0)
        self._check(self._f("soup_restarts")(self._h, k, n, _ptr(cen), _ptr(lp), _ptr(out), _ptr(losses), _ptr(its),
                                             C.byref(best)), "soup_restarts")
        return dict(log_probs=lp, centers=out, losses=losses, iterations=its, best=int(best.value),
                    loss=float(losses[best.value]))

    # cellector() driver loop (main.rs:36-46)
    def run_em(self, iqr_multiple=5.0, max_iterations=1000):
        stats = []
        for _ in range(max_iterations):
            st = self.compute_new_excluded(iqr_multiple)
            stats.append(st)
            if not st.any_change:
                break
        return stats


class Cellector(CtxApi):
    """One GPU context of libcellector_cuda.so."""

    def __init__(self, device: int = 0, flags: int = 0, lib: C.CDLL | None = None):
        lib = lib or load_library()
        h = C.c_void_p()
        rc = lib.clx_create(C.byref(h), device, flags)
        if rc != 0:
            msg = lib.clx_last_error(None)
            raise RuntimeError(f"clx_create failed ({rc}): {msg.decode() if msg else ''} — no CPU fallback exists")
        super().__init__(lib, "clx_", h)

    def comm_init(self, rank: int, world: int, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        self._check(self._lib.clx_comm_init(self._h, rank, world, buf), "comm_init")

    @staticmethod
    def nccl_unique_id(lib: C.CDLL | None = None) -> bytes:
        lib = lib or load_library()
        buf = C.create_string_buffer(128)
        rc = lib.clx_nccl_unique_id(buf)
        if rc != 0:
            raise RuntimeError(f"clx_nccl_unique_id failed ({rc})")
        return buf.raw

    def phase_ms(self):
        out = {}
        for i, name in enumerate(PHASES):
            ms, n = C.c_double(0), C.c_uint64(0)
            self._check(self._lib.clx_get_phase_ms(self._h, i, C.byref(ms), C.byref(n)), "get_phase_ms")
            out[name] = (ms.value, int(n.value))
        return out

    def set_profiling(self, on: bool):
        self._check(self._lib.clx_set_profiling(self._h, 1 if on else 0), "set_profiling")

    def reset_phase_ms(self):
        self._check(self._lib.clx_reset_phase_ms(self._h), "reset_phase_ms")

    def stream(self) -> int:
        return int(self._lib.clx_get_stream(self._h) or 0)

    def score_bytes(self) -> int:
        b = C.c_uint64(0)
        self._check(self._lib.clx_score_bytes(self._h, C.byref(b)), "score_bytes")
        return int(b.value)

    def launch_count(self) -> int:
        b = C.c_uint64(0)
        self._check(self._lib.clx_launch_count(self._h, C.byref(b)), "launch_count")
        return int(b.value)

    def flush_l2(self):
        self._check(self._lib.clx_flush_l2(self._h), "flush_l2")
