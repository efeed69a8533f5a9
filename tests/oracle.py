"""ctypes binding of the CPU oracle (oracle/cellector_oracle.cc).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by
the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

from cellector_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "_build", "libcellector_oracle.so")

_CTX_NAMES = ["destroy", "last_error", "load_coo", "get_locus_ids", "reset", "em_iteration", "get_cell_vectors",
              "get_locus_vectors", "posteriors", "final_locus_counts"]


def build_oracle(force=False):
    src = [os.path.join(ORACLE_DIR, f) for f in ("cellector_oracle.cc", "statrs_compat.h")]
    if force or not os.path.exists(ORACLE_SO) or any(os.path.getmtime(s) > os.path.getmtime(ORACLE_SO) for s in src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    return ORACLE_SO


def load_oracle() -> C.CDLL:
    lib = C.CDLL(build_oracle())
    api.bind(lib, "orc_", _CTX_NAMES)
    lib.orc_create.restype, lib.orc_create.argtypes = C.c_void_p, []
    d, u = C.c_double, C.c_uint64
    for name, res, args in [
        ("ln_gamma", d, [d]), ("ln_factorial", d, [u]), ("ln_binomial", d, [u, u]), ("logsumexp", d, [d, d]),
        ("log_beta_calc", d, [d, d]), ("log_beta_binomial_pmf", d, [d, d, d, d, d]),
        ("expected_log_beta_binomial_pmf", None, [u, d, d, C.POINTER(d), C.POINTER(d)]),
        ("quantile", d, [C.c_void_p, u, d]), ("median", d, [C.c_void_p, u]), ("binomial_pmf", d, [d, u, u]),
        ("em_iteration_ex", C.c_int, [C.c_void_p, d, C.POINTER(api.IterStats), C.c_int]),
        ("last_timings", None, [C.c_void_p, C.POINTER(d * 3)]),
        ("set_faithful_posteriors", None, [C.c_void_p, C.c_int]),
        ("read_mtx_pair", C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                    C.POINTER(u), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, u]),
    ]:
        fn = getattr(lib, "orc_" + name)
        fn.restype, fn.argtypes = res, args
    return lib


class Oracle(api.CtxApi):
    """CPU reference context with the same methods as cellector_b200.Cellector."""

    def __init__(self, lib: C.CDLL | None = None):
        lib = lib or load_oracle()
        super().__init__(lib, "orc_", C.c_void_p(lib.orc_create()))

    def compute_new_excluded(self, iqr_multiple=5.0, faithful=True):
        st = api.IterStats()
        self._check(self._lib.orc_em_iteration_ex(self._h, float(iqr_multiple), C.byref(st), 1 if faithful else 0),
                    "em_iteration_ex")
        return st

    def last_timings(self):
        t = (C.c_double * 3)()
        self._lib.orc_last_timings(self._h, C.byref(t))
        return dict(score_s=t[0], locus_s=t[1], total_s=t[2])
