"""ctypes binding of the CPU oracle of souporcell's k-cluster EM (oracle/souporcell_oracle.cc).  TEST INFRASTRUCTURE
ONLY: imported by tests/ (and bench.py's optional soup line as its checker), never by the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
SOUP_SO = os.path.join(ORACLE_DIR, "_build", "libsouporcell_oracle.so")


def load_soup_oracle() -> C.CDLL:
    src = [os.path.join(ORACLE_DIR, f) for f in ("souporcell_oracle.cc", "statrs_compat.h")]
    if not os.path.exists(SOUP_SO) or any(os.path.getmtime(s) > os.path.getmtime(SOUP_SO) for s in src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    lib = C.CDLL(SOUP_SO)
    u32, u64, vp = C.c_uint32, C.c_uint64, C.c_void_p
    lib.sorc_create.restype, lib.sorc_create.argtypes = vp, []
    lib.sorc_destroy.restype, lib.sorc_destroy.argtypes = None, [vp]
    lib.sorc_last_error.restype, lib.sorc_last_error.argtypes = C.c_char_p, [vp]
    lib.sorc_load_coo.restype = C.c_int
    lib.sorc_load_coo.argtypes = [vp, u64, u32, u32, vp, vp, vp, vp, u32, u32, u32, u32, C.POINTER(u32), C.POINTER(u64)]
    lib.sorc_get_locus_ids.restype, lib.sorc_get_locus_ids.argtypes = C.c_int, [vp, vp]
    lib.sorc_em.restype = C.c_int
    lib.sorc_em.argtypes = [vp, u32, vp, C.c_int, vp, vp, C.POINTER(C.c_double), C.POINTER(u32), vp, u32]
    return lib


class SoupOracle:
    """souporcell's load_cell_data + EM on the CPU (main.rs:390-468, 109-186)."""

    def __init__(self, lib: C.CDLL | None = None):
        self._lib = lib or load_soup_oracle()
        self._h = C.c_void_p(self._lib.sorc_create())
        self.n_cells = self.n_loci_used = self.n_entries = 0

    def close(self):
        if self._h:
            self._lib.sorc_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def load_cell_data(self, locus, cell, ref, alt, n_loci0, n_cells, min_ref=4, min_alt=4, min_ref_umis=0, min_alt_umis=0):
        arrs = [np.ascontiguousarray(x, dtype=np.uint32) for x in (locus, cell, ref, alt)]
        nl, ne = C.c_uint32(0), C.c_uint64(0)
        rc = self._lib.sorc_load_coo(self._h, arrs[0].shape[0], n_loci0, n_cells, *[a.ctypes.data for a in arrs], min_ref,
                                     min_alt, min_ref_umis, min_alt_umis, C.byref(nl), C.byref(ne))
        if rc != 0:
            raise RuntimeError(self._lib.sorc_last_error(self._h).decode())
        self.n_cells, self.n_loci_used, self.n_entries = int(n_cells), int(nl.value), int(ne.value)
        return self.n_loci_used, self.n_entries

    def locus_ids(self):
        out = np.empty(self.n_loci_used, np.uint32)
        self._lib.sorc_get_locus_ids(self._h, out.ctypes.data)
        return out

    def soup_em(self, centers_init, precise=True, trace=1000):
        cen = np.ascontiguousarray(centers_init, dtype=np.float32)
        k = cen.shape[0]
        assert cen.shape == (k, self.n_loci_used)
        lp = np.empty((self.n_cells, k), np.float64)
        out = np.empty((k, self.n_loci_used), np.float64)
        tr = np.full(trace, np.nan)
        loss, it = C.c_double(0), C.c_uint32(0)
        rc = self._lib.sorc_em(self._h, k, cen.ctypes.data, 1 if precise else 0, lp.ctypes.data, out.ctypes.data,
                               C.byref(loss), C.byref(it), tr.ctypes.data, trace)
        if rc != 0:
            raise RuntimeError(self._lib.sorc_last_error(self._h).decode())
        return dict(log_probs=lp, centers=out, loss=loss.value, iterations=int(it.value), loss_trace=tr[:min(trace, it.value)])
