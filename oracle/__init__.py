"""oracle -- TEST INFRASTRUCTURE (the checker), never the product path.

ctypes front-end for the plain-C restatement in oracle/oracle.c and
oracle/oracle_drivers.c, plus access to the compiled unmodified reference in
oracle/_ref when it is present.  Only tests/, __graft_entry__.smoke() and
bench.py's CPU-baseline legs may import this package.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
_SRC = [os.path.join(HERE, f) for f in ("oracle.c", "oracle_drivers.c")]
_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement with gcc (seconds)."""
    deps = _SRC + [os.path.join(HERE, "gammainc.h")]
    if (not force and os.path.exists(_LIB_PATH)
            and all(os.path.getmtime(s) <= os.path.getmtime(_LIB_PATH) for s in deps)):
        return _LIB_PATH
    os.makedirs(os.path.dirname(_LIB_PATH), exist_ok=True)
    cmd = ["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fvisibility=hidden",
           "-fno-fast-math", "-ffp-contract=off", "-o", _LIB_PATH] + _SRC + ["-lm"]
    subprocess.check_call(cmd)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        c = ctypes
        u8p, u32p, i64p, u64p, f64p = (c.POINTER(c.c_uint8), c.POINTER(c.c_uint32),
                                       c.POINTER(c.c_int64), c.POINTER(c.c_uint64),
                                       c.POINTER(c.c_double))
        L.obo_next_combmask.argtypes = [u64p, c.c_int]
        L.obo_count_table.argtypes = [u8p, c.c_size_t, c.c_size_t, u32p, u32p, c.c_int, i64p]
        L.obo_bdeu_from_counts.argtypes = [i64p, u32p, c.c_int, c.c_int, c.c_double]
        L.obo_bdeu_from_counts.restype = c.c_double
        L.obo_citest_from_counts.argtypes = [i64p, u32p, c.c_int, c.c_int, c.c_int, c.c_int,
                                             f64p, u64p, f64p]
        L.obo_pchisq_export.argtypes = [c.c_double, c.c_uint64]
        L.obo_pchisq_export.restype = c.c_double
        L.obo_exhaustive_search.argtypes = [u8p, c.c_size_t, c.c_size_t, u32p, c.c_int, c.c_double,
                                            c.c_int, u64p, f64p, u8p, f64p]
        L.obo_pc.argtypes = [u8p, c.c_size_t, c.c_size_t, u32p, c.c_int, c.c_int, c.c_double,
                             c.c_int, c.c_int, u8p, u8p, u64p]
        _lib = L
    return _lib


def _p(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct))


class Dataset:
    """Column-major uint8 codes + cardinalities (what DataframeWrapper holds,
    include/base/dataframe_wrapper.h:33-45)."""

    def __init__(self, codes_rows_by_vars: np.ndarray, cards=None):
        codes = np.asarray(codes_rows_by_vars)
        assert codes.ndim == 2
        self.n_rows, self.n_vars = codes.shape
        self.colmajor = np.ascontiguousarray(codes.T.astype(np.uint8))   # [n_vars, n_rows]
        if cards is None:
            cards = [int(self.colmajor[v].max()) + 1 if self.n_rows else 1 for v in range(self.n_vars)]
        self.cards = np.asarray(cards, dtype=np.uint32)


def next_combmask(comb: int, n: int):
    c = ctypes.c_uint64(comb)
    more = lib().obo_next_combmask(ctypes.byref(c), n)
    return c.value, bool(more)


def count_table(ds: Dataset, var_ids) -> np.ndarray:
    """Dense int64 counts in reference key order (first var most significant)."""
    var_ids = np.asarray(var_ids, dtype=np.uint32)
    cells = int(np.prod(ds.cards[var_ids].astype(np.int64))) if len(var_ids) else 1
    out = np.zeros(cells, dtype=np.int64)
    rc = lib().obo_count_table(_p(ds.colmajor, ctypes.c_uint8), ds.n_rows, ds.n_rows,
                               _p(ds.cards, ctypes.c_uint32), _p(var_ids, ctypes.c_uint32),
                               len(var_ids), _p(out, ctypes.c_int64))
    if rc != 0:
        raise ValueError("var_ids must be sorted")
    return out


def bdeu_local_score(ds: Dataset, child: int, parents, ess: float = 1.0) -> float:
    var_ids = sorted(list(parents) + [child])
    counts = count_table(ds, var_ids)
    cards = np.ascontiguousarray(ds.cards[np.asarray(var_ids, dtype=np.int64)])
    return float(lib().obo_bdeu_from_counts(_p(counts, ctypes.c_int64), _p(cards, ctypes.c_uint32),
                                            len(var_ids), var_ids.index(child), ess))


def citest(ds: Dataset, x: int, y: int, z, kind: int):
    """Returns (stat, dof, p). kind 0 = chi-square, 1 = G-square."""
    if x == y:
        raise ValueError("x and y must differ")
    var_ids = sorted(list(z) + [x, y])
    counts = count_table(ds, var_ids)
    cards = np.ascontiguousarray(ds.cards[np.asarray(var_ids, dtype=np.int64)])
    st, p, dof = ctypes.c_double(), ctypes.c_double(), ctypes.c_uint64()
    lib().obo_citest_from_counts(_p(counts, ctypes.c_int64), _p(cards, ctypes.c_uint32),
                                 len(var_ids), var_ids.index(x), var_ids.index(y), kind,
                                 ctypes.byref(st), ctypes.byref(dof), ctypes.byref(p))
    return st.value, dof.value, p.value


def pchisq(x: float, dof: int) -> float:
    return float(lib().obo_pchisq_export(x, dof))


def exhaustive_search(ds: Dataset, ess: float, max_parents: int, want_tables: bool = False):
    n = ds.n_vars
    masks = np.zeros(n, dtype=np.uint64)
    gs = ch = ls = None
    if want_tables:
        gs = np.zeros(1 << n, dtype=np.float64)
        ch = np.zeros(1 << n, dtype=np.uint8)
        ls = np.zeros(n << n, dtype=np.float64)
    rc = lib().obo_exhaustive_search(
        _p(ds.colmajor, ctypes.c_uint8), ds.n_rows, ds.n_rows, _p(ds.cards, ctypes.c_uint32), n,
        ess, max_parents, _p(masks, ctypes.c_uint64),
        _p(gs, ctypes.c_double) if want_tables else None,
        _p(ch, ctypes.c_uint8) if want_tables else None,
        _p(ls, ctypes.c_double) if want_tables else None)
    if rc != 0:
        raise ValueError("max_parents must be in [0, num_of_vars-1]")
    adj = np.zeros((n, n), dtype=np.uint8)
    for v in range(n):
        for p_ in range(n):
            if (int(masks[v]) >> p_) & 1:
                adj[p_, v] = 1
    if want_tables:
        return adj, gs, ch, ls.reshape(n, 1 << n)
    return adj


def pc(ds: Dataset, kind: int, level: float, max_cond_vars: int, stable: bool = True,
       want_details: bool = False):
    n = ds.n_vars
    adj = np.zeros((n, n), dtype=np.uint8)
    sep = np.zeros((n, n, n), dtype=np.uint8)
    nt = np.zeros(max_cond_vars + 1, dtype=np.uint64)
    lib().obo_pc(_p(ds.colmajor, ctypes.c_uint8), ds.n_rows, ds.n_rows, _p(ds.cards, ctypes.c_uint32),
                 n, kind, level, max_cond_vars, int(stable), _p(adj, ctypes.c_uint8),
                 _p(sep, ctypes.c_uint8), _p(nt, ctypes.c_uint64))
    if want_details:
        return adj, sep, nt
    return adj


# ---------------------------------------------------------------------------
# the compiled unmodified reference (oracle/_ref), when present
# ---------------------------------------------------------------------------
def reference_modules():
    """Returns (openbnsllib, obn_refharness) from oracle/_ref or (None, None)."""
    ref_dir = os.path.join(HERE, "_ref")
    if not os.path.isdir(ref_dir):
        return None, None
    if ref_dir not in sys.path:
        sys.path.insert(0, ref_dir)
    try:
        import openbnsllib  # noqa
        import obn_refharness  # noqa
    except ImportError:
        return None, None
    if not hasattr(openbnsllib, "structure_learning"):
        return None, None
    return openbnsllib, obn_refharness
