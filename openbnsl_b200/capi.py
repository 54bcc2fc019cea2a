"""ctypes binding of the C ABI in include/openbnsl_b200.h.

This is the thinnest possible caller of the boundary: numpy arrays in, numpy
arrays out, every call going through the extern "C" entry points of
libopenbnsl_b200.so.  It raises the exception classes the reference raises
for the same conditions (std::invalid_argument -> ValueError,
std::out_of_range -> IndexError, everything else RuntimeError).
There is no fallback: if the library is missing or no GPU is usable, it raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libopenbnsl_b200.so")

OBN_OK, OBN_ERR_INVALID, OBN_ERR_CUDA, OBN_ERR_OOM, OBN_ERR_UNSUPPORTED, OBN_ERR_RANGE, OBN_ERR_COMM = range(7)
COMM_ID_BYTES = 128
STATIC_MAX_CELLS = 65536
CI_CHI2, CI_G2 = 0, 1

EXPORTS = [
    "obn_last_error", "obn_version", "obn_ctx_create", "obn_ctx_destroy", "obn_ctx_stream", "obn_ctx_set_option",
    "obn_comm_unique_id", "obn_comm_init", "obn_comm_destroy", "obn_comm_rank", "obn_comm_world",
    "obn_comm_allgather_inplace", "obn_comm_allgather_host", "obn_comm_allreduce_min_i32",
    "obn_dataset_upload", "obn_dataset_destroy", "obn_host_pin", "obn_host_unpin", "obn_dataset_n_vars", "obn_dataset_n_rows",
    "obn_count_tables", "obn_count_table_static", "obn_bdeu_scores", "obn_bdeu_subset_scores", "obn_ci_tests",
    "obn_ci_tests_submit", "obn_ci_tests_collect", "obn_pc_round", "obn_pc_round_submit", "obn_pc_round_collect",
    "obn_ci_group_plan",
    "obn_es_begin", "obn_es_level_begin", "obn_es_level_scores", "obn_es_level_buffer", "obn_es_level_dp", "obn_es_level_dp_part", "obn_es_level_dp_commit", "obn_es_level_dp_sharded", "obn_lattice_bottom_program",
    "obn_es_lattice", "obn_es_f_buffer", "obn_es_finish", "obn_es_end",
    "obn_profile_enable", "obn_profile_read",
]


class Profile(C.Structure):
    _fields_ = [("count_kernel_ms", C.c_double), ("count_launches", C.c_uint64),
                ("total_launches", C.c_uint64), ("algorithmic_bytes", C.c_double),
                ("tuples", C.c_uint64), ("h2d_bytes", C.c_double), ("d2h_bytes", C.c_double),
                ("host_prep_ms", C.c_double), ("host_wait_ms", C.c_double),
                ("scanned_bytes", C.c_double), ("increments", C.c_double), ("tables", C.c_uint64),
                ("collectives", C.c_uint64), ("collective_ms", C.c_double), ("exact_fallbacks", C.c_uint64)]


_lib = None


def load():
    """dlopen the library and declare the prototypes (does not touch the GPU)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libopenbnsl_b200.so is not built (run python -m openbnsl_b200.build); "
                           "there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u8p, u32p, u64p, i32p, f64p = (C.c_void_p, C.POINTER(C.c_uint8), C.POINTER(C.c_uint32),
                                       C.POINTER(C.c_uint64), C.POINTER(C.c_int32), C.POINTER(C.c_double))
    L.obn_last_error.restype = C.c_char_p
    L.obn_version.restype = C.c_char_p
    L.obn_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.obn_ctx_destroy.argtypes = [vp]
    L.obn_ctx_destroy.restype = None
    L.obn_ctx_stream.argtypes = [vp]
    L.obn_ctx_stream.restype = vp
    L.obn_dataset_upload.argtypes = [vp, u8p, C.c_size_t, C.c_size_t, C.c_size_t, u32p, C.POINTER(vp)]
    L.obn_host_pin.argtypes = [vp, vp, C.c_size_t]
    L.obn_host_unpin.argtypes = [vp, vp]
    L.obn_dataset_destroy.argtypes = [vp]
    L.obn_dataset_destroy.restype = None
    L.obn_dataset_n_vars.argtypes = [vp]
    L.obn_dataset_n_vars.restype = C.c_size_t
    L.obn_dataset_n_rows.argtypes = [vp]
    L.obn_dataset_n_rows.restype = C.c_size_t
    L.obn_count_tables.argtypes = [vp, vp, u32p, u64p, C.c_size_t, i32p, u64p]
    L.obn_count_table_static.argtypes = [C.c_int, vp, vp, C.c_size_t, C.c_size_t, u32p, u32p, C.c_size_t, vp, C.c_size_t]
    L.obn_bdeu_scores.argtypes = [vp, vp, u32p, u64p, C.c_size_t, u32p, C.c_double, C.c_int, f64p]
    L.obn_bdeu_subset_scores.argtypes = [vp, vp, u32p, u64p, C.c_size_t, C.c_double, C.c_int, f64p]
    L.obn_ci_tests.argtypes = [vp, vp, u32p, u64p, C.c_size_t, u32p, u32p, C.c_int, f64p, u64p, f64p]
    L.obn_ci_tests_submit.argtypes = [vp, vp, u32p, u64p, C.c_size_t, u32p, u32p, C.c_int, C.c_int]
    L.obn_ci_tests_collect.argtypes = [vp, f64p, u64p, f64p]
    L.obn_pc_round.argtypes = [vp, vp, u32p, u64p, C.c_size_t, u32p, u32p, u32p, C.c_size_t, C.c_int, C.c_double,
                               C.c_int, i32p]
    L.obn_pc_round_submit.argtypes = [vp, vp, u32p, u64p, C.c_size_t, u32p, u32p, u32p, C.c_size_t, C.c_int,
                                      C.c_double, C.c_int, C.c_int]
    L.obn_pc_round_collect.argtypes = [vp, i32p]
    L.obn_ci_group_plan.argtypes = [u32p, C.c_size_t, u32p, u64p, C.c_size_t, u32p, u32p, C.c_int, C.c_int, u32p, u64p, u32p,
                                    u32p, C.POINTER(C.c_size_t)]
    L.obn_ctx_set_option.argtypes = [vp, C.c_char_p, C.c_double]
    L.obn_comm_unique_id.argtypes = [u8p]
    L.obn_comm_init.argtypes = [vp, u8p, C.c_int, C.c_int]
    L.obn_comm_destroy.argtypes = [vp]
    L.obn_comm_destroy.restype = None
    L.obn_comm_rank.argtypes = [vp]
    L.obn_comm_world.argtypes = [vp]
    L.obn_comm_allgather_inplace.argtypes = [vp, vp, C.c_size_t]
    L.obn_comm_allgather_host.argtypes = [vp, vp, vp, C.c_size_t]
    L.obn_comm_allreduce_min_i32.argtypes = [vp, i32p, C.c_size_t]
    L.obn_profile_enable.argtypes = [vp, C.c_int]
    L.obn_profile_read.argtypes = [vp, C.POINTER(Profile), C.c_int]
    _lib = L
    return L


def _check(rc):
    if rc == OBN_OK:
        return
    msg = load().obn_last_error().decode()
    if rc == OBN_ERR_INVALID:
        raise ValueError(msg)
    if rc == OBN_ERR_RANGE:
        raise IndexError(msg)
    if rc == OBN_ERR_OOM:
        raise MemoryError(msg)
    raise RuntimeError(msg)


def _ptr(a, ct):
    return a.ctypes.data_as(C.POINTER(ct)) if a is not None else None


def _csr(tuples):
    """list of int sequences -> (var_ids uint32, offsets uint64)."""
    offs = np.zeros(len(tuples) + 1, dtype=np.uint64)
    if len(tuples):
        offs[1:] = np.cumsum([len(t) for t in tuples])
    flat = np.fromiter((v for t in tuples for v in t), dtype=np.uint32, count=int(offs[-1]))
    return flat, offs


def ci_group_plan(cards, tuples, x, y, rank=0, world=1):
    """The table sharing obn_ci_tests / obn_pc_round use (host only).  Returns (groups, group_of_test,
    digits_of_test): groups = list of ascending variable lists, digits = (px, py) positions of x < y.
    rank / world: the share of a sharded round (group_of_test is 0xffffffff for the other ranks' tests)."""
    flat, offs = _csr(tuples)
    cards = np.ascontiguousarray(np.asarray(cards, dtype=np.uint32))
    xa = np.ascontiguousarray(np.asarray(x, dtype=np.uint32))
    ya = np.ascontiguousarray(np.asarray(y, dtype=np.uint32))
    n = len(tuples)
    gv = np.zeros(max(1, len(flat)), dtype=np.uint32)
    go = np.zeros(n + 1, dtype=np.uint64)
    group_of, digits = np.zeros(max(1, n), dtype=np.uint32), np.zeros(max(1, n), dtype=np.uint32)
    n_groups = C.c_size_t(0)
    _check(load().obn_ci_group_plan(_ptr(cards, C.c_uint32), len(cards), _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64),
                                    n, _ptr(xa, C.c_uint32), _ptr(ya, C.c_uint32), int(rank), int(world),
                                    _ptr(gv, C.c_uint32), _ptr(go, C.c_uint64), _ptr(group_of, C.c_uint32), _ptr(digits, C.c_uint32),
                                    C.byref(n_groups)))
    groups = [gv[int(go[q]):int(go[q + 1])].tolist() for q in range(n_groups.value)]
    return groups, group_of[:n], [(int(d) & 0xff, (int(d) >> 8) & 0xff) for d in digits[:n]]


def count_table_static(device, stream, data_dev, n_rows, row_stride, cols, cards, counts_out_dev):
    """obn_count_table_static: one table into the caller's DEVICE buffer, no allocation inside the library
    (global scratch = a static __device__ array; at most STATIC_MAX_CELLS cells).  data_dev / counts_out_dev / stream
    are integers (device addresses, a cudaStream_t); asynchronous on `stream`."""
    cols = np.ascontiguousarray(np.asarray(cols, dtype=np.uint32))
    cards = np.ascontiguousarray(np.asarray(cards, dtype=np.uint32))
    n_cells = 1
    for c in cards.tolist():
        n_cells *= int(c)
    _check(load().obn_count_table_static(int(device), C.c_void_p(stream or None), C.c_void_p(data_dev), int(n_rows),
                                         int(row_stride), _ptr(cols, C.c_uint32), _ptr(cards, C.c_uint32), len(cols),
                                         C.c_void_p(counts_out_dev), n_cells))
    return n_cells


def comm_unique_id() -> bytes:
    buf = np.zeros(COMM_ID_BYTES, dtype=np.uint8)
    _check(load().obn_comm_unique_id(_ptr(buf, C.c_uint8)))
    return buf.tobytes()


class Context:
    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        _check(load().obn_ctx_create(device, C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            load().obn_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self) -> int:
        return load().obn_ctx_stream(self._h) or 0

    def upload(self, codes_rows_by_vars: np.ndarray, cards) -> "Dataset":
        return Dataset(self, codes_rows_by_vars, cards)

    def set_option(self, name: str, value: float):
        _check(load().obn_ctx_set_option(self._h, name.encode(), float(value)))

    def comm_init(self, id_bytes: bytes, rank: int, world: int):
        buf = np.frombuffer(id_bytes, dtype=np.uint8).copy()
        _check(load().obn_comm_init(self._h, _ptr(buf, C.c_uint8), int(rank), int(world)))

    def profile_enable(self, on=True):
        _check(load().obn_profile_enable(self._h, int(on)))

    def profile_read(self, reset=True) -> Profile:
        p = Profile()
        _check(load().obn_profile_read(self._h, C.byref(p), int(reset)))
        return p


class Dataset:
    """Device-resident column-major uint8 sample matrix."""

    def __init__(self, ctx: Context, codes_rows_by_vars: np.ndarray, cards, colmajor=None):
        if colmajor is None:
            codes = np.asarray(codes_rows_by_vars)
            colmajor = np.ascontiguousarray(codes.T.astype(np.uint8, copy=False))
        self.ctx = ctx
        self.n_vars, self.n_rows = colmajor.shape
        self.cards = np.ascontiguousarray(np.asarray(cards, dtype=np.uint32))
        self._h = C.c_void_p()
        _check(load().obn_dataset_upload(ctx._h, _ptr(colmajor, C.c_uint8), self.n_vars, self.n_rows,
                                         self.n_rows, _ptr(self.cards, C.c_uint32), C.byref(self._h)))

    def close(self):
        if self._h:
            if self.ctx._h:             # a ctx that is already gone took its device memory with it
                load().obn_dataset_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- batched hot path -------------------------------------------------
    def count_tables(self, tuples):
        flat, offs = _csr(tuples)
        sizes = [int(np.prod(self.cards[np.asarray(t, dtype=np.int64)].astype(np.int64))) if len(t) else 1
                 for t in tuples]
        out_off = np.zeros(len(tuples) + 1, dtype=np.uint64)
        out_off[1:] = np.cumsum(sizes)
        out = np.zeros(int(out_off[-1]), dtype=np.int32)
        _check(load().obn_count_tables(self.ctx._h, self._h, _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64),
                                       len(tuples), _ptr(out, C.c_int32), _ptr(out_off, C.c_uint64)))
        return [out[int(out_off[i]):int(out_off[i + 1])] for i in range(len(tuples))]

    def bdeu_scores(self, tuples, child, ess=1.0, exact_order=False):
        flat, offs = _csr(tuples)
        ch = np.ascontiguousarray(np.asarray(child, dtype=np.uint32))
        out = np.zeros(len(tuples), dtype=np.float64)
        _check(load().obn_bdeu_scores(self.ctx._h, self._h, _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64),
                                      len(tuples), _ptr(ch, C.c_uint32), float(ess), int(exact_order),
                                      _ptr(out, C.c_double)))
        return out

    def bdeu_subset_scores(self, tuples, ess=1.0, exact_order=False):
        flat, offs = _csr(tuples)
        out = np.zeros(int(offs[-1]), dtype=np.float64)
        _check(load().obn_bdeu_subset_scores(self.ctx._h, self._h, _ptr(flat, C.c_uint32),
                                             _ptr(offs, C.c_uint64), len(tuples), float(ess),
                                             int(exact_order), _ptr(out, C.c_double)))
        return out, offs

    def ci_tests_csr(self, flat, offs, x, y, kind):
        n = len(offs) - 1
        stat = np.zeros(n, dtype=np.float64)
        dof = np.zeros(n, dtype=np.uint64)
        p = np.zeros(n, dtype=np.float64)
        _check(load().obn_ci_tests(self.ctx._h, self._h, _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64), n,
                                   _ptr(x, C.c_uint32), _ptr(y, C.c_uint32), int(kind),
                                   _ptr(stat, C.c_double), _ptr(dof, C.c_uint64), _ptr(p, C.c_double)))
        return stat, dof, p

    def pc_round(self, flat, offs, x, y, slot_of_test, n_slots, kind, level, shard=False, split=False):
        """obn_pc_round (or its submit / collect halves): first passing test index per edge slot."""
        n = len(offs) - 1
        slot = np.ascontiguousarray(np.asarray(slot_of_test, dtype=np.uint32))
        first = np.full(max(1, n_slots), 2**31 - 1, dtype=np.int32)    # collect leaves it untouched after an empty submit
        if split:
            _check(load().obn_pc_round_submit(self.ctx._h, self._h, _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64), n,
                                              _ptr(x, C.c_uint32), _ptr(y, C.c_uint32), _ptr(slot, C.c_uint32),
                                              int(n_slots), int(kind), float(level), int(shard), 0))
            _check(load().obn_pc_round_collect(self.ctx._h, _ptr(first, C.c_int32)))
        else:
            _check(load().obn_pc_round(self.ctx._h, self._h, _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64), n,
                                       _ptr(x, C.c_uint32), _ptr(y, C.c_uint32), _ptr(slot, C.c_uint32), int(n_slots),
                                       int(kind), float(level), int(shard), _ptr(first, C.c_int32)))
        return first[:n_slots]

    def pc_round_submit(self, flat, offs, x, y, slot_of_test, n_slots, kind, level, shard=False, chain=False):
        """obn_pc_round_submit alone (up to two rounds in flight; chain: skip the edges the previous round decided)."""
        slot = np.ascontiguousarray(np.asarray(slot_of_test, dtype=np.uint32))
        _check(load().obn_pc_round_submit(self.ctx._h, self._h, _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64), len(offs) - 1,
                                          _ptr(x, C.c_uint32), _ptr(y, C.c_uint32), _ptr(slot, C.c_uint32),
                                          int(n_slots), int(kind), float(level), int(shard), int(chain)))

    def pc_round_collect(self, n_slots):
        first = np.full(max(1, n_slots), 2**31 - 1, dtype=np.int32)
        _check(load().obn_pc_round_collect(self.ctx._h, _ptr(first, C.c_int32)))
        return first[:n_slots]

    def ci_tests_submit_collect(self, flat, offs, x, y, kind, between=None):
        """The same batch through obn_ci_tests_submit / obn_ci_tests_collect; `between` runs while it is in flight."""
        n = len(offs) - 1
        _check(load().obn_ci_tests_submit(self.ctx._h, self._h, _ptr(flat, C.c_uint32), _ptr(offs, C.c_uint64), n,
                                          _ptr(x, C.c_uint32), _ptr(y, C.c_uint32), int(kind), 1))
        if between is not None:
            between()
        stat = np.zeros(n, dtype=np.float64)
        dof = np.zeros(n, dtype=np.uint64)
        p = np.zeros(n, dtype=np.float64)
        _check(load().obn_ci_tests_collect(self.ctx._h, _ptr(stat, C.c_double), _ptr(dof, C.c_uint64), _ptr(p, C.c_double)))
        return stat, dof, p

    def ci_tests(self, x, y, zs, kind):
        """x[t], y[t], zs[t] (sequence of conditioning ids) -> (stat, dof, p)."""
        tuples = [sorted(list(z) + [int(a), int(b)]) for a, b, z in zip(x, y, zs)]
        flat, offs = _csr(tuples)
        xa = np.ascontiguousarray(np.asarray(x, dtype=np.uint32))
        ya = np.ascontiguousarray(np.asarray(y, dtype=np.uint32))
        return self.ci_tests_csr(flat, offs, xa, ya, kind)
