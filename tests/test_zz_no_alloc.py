"""obn_count_table_static (include/openbnsl_b200.h): the counting kernel with its global scratch table in a static
__device__ array, so that the call allocates nothing.  Checked through the C ABI against the existing entry point
(obn_count_tables) -- integer counts, so the tolerance is the suite's tolerance for counts: exact -- on calls in a row
that share the scratch.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from openbnsl_b200 import capi
from openbnsl_b200 import networks as nw

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup(gpu_ctx):
    import torch
    net = nw.alarm()
    codes = net.sample(20011, 3)                          # not a multiple of 4 or 16: exercises the tail rows
    cards = np.array([int(codes[:, v].max()) + 1 for v in range(net.n)], dtype=np.uint32)
    ds = gpu_ctx.upload(codes, cards)
    n_rows, n_vars = codes.shape
    stride = (n_rows + 127) // 128 * 128
    col_major = np.zeros((n_vars, stride), dtype=np.uint8)
    col_major[:, :n_rows] = codes.T
    dev = torch.from_numpy(col_major).to("cuda:0")        # the CALLER's device memory (torch: plumbing only)
    return ds, cards, dev, n_rows, stride


def _static(dev, n_rows, stride, cards, cols, out, stream=0, offset=0):
    cells = capi.count_table_static(0, stream, dev.data_ptr() + offset, n_rows, stride, cols,
                                    cards[np.asarray(cols, dtype=np.int64)], out.data_ptr())
    return cells


def test_static_matches_count_tables_on_calls_in_a_row(setup):
    import torch
    ds, cards, dev, n_rows, stride = setup
    rng = np.random.default_rng(11)
    tuples = [[], [5], [0, 1], [3, 7, 20], [9, 22, 26, 33, 34, 35, 36], [1, 9, 22, 26, 33, 34, 35, 36]]   # .. 16384, 49152 cells
    for k in (2, 3, 4, 5, 6, 7, 8):
        for _ in range(3):
            t = sorted(rng.choice(37, size=k, replace=False).tolist())
            if int(np.prod(cards[t].astype(np.int64))) <= capi.STATIC_MAX_CELLS:
                tuples.append(t)
    tuples += [tuples[5], tuples[5], tuples[2]]           # the same table twice in a row, then a smaller one
    want = ds.count_tables(tuples)                        # the existing entry point
    sizes = set()
    out = torch.empty(capi.STATIC_MAX_CELLS, dtype=torch.int32, device="cuda:0")
    for t, w in zip(tuples, want):
        out.fill_(-7)                                     # stale output must be overwritten, cells beyond left alone
        cells = _static(dev, n_rows, stride, cards, t, out)
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        assert cells == len(w)
        sizes.add(cells)
        assert np.array_equal(got[:cells], w), t
        assert got[:cells].sum() == n_rows
        assert (got[cells:] == -7).all()
    assert min(sizes) == 1 and max(sizes) > 8192          # shared-memory and global-table variants both ran


def test_static_back_to_back_without_sync_and_unaligned(setup):
    """Two calls enqueued on one stream with no host synchronisation between them (stream order is what separates
    them), and a matrix whose columns are not 4-byte aligned (byte-load variant)."""
    import torch
    ds, cards, dev, n_rows, stride = setup
    a, b = [1, 4, 9, 30], [2, 8]
    wa, wb = ds.count_tables([a, b])
    s = torch.cuda.Stream(device="cuda:0")
    oa = torch.zeros(len(wa), dtype=torch.int32, device="cuda:0")
    ob = torch.zeros(len(wb), dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    _static(dev, n_rows, stride, cards, a, oa, stream=s.cuda_stream)
    _static(dev, n_rows, stride, cards, b, ob, stream=s.cuda_stream)
    s.synchronize()
    assert np.array_equal(oa.cpu().numpy(), wa)
    assert np.array_equal(ob.cpu().numpy(), wb)
    # rows 1 .. n_rows-1 of every column: the pointers are odd, the table is the one of the shifted sample
    shifted = dev[:, 1:n_rows].cpu().numpy()
    key = shifted[2].astype(np.int64) * int(cards[8]) + shifted[8]
    want = np.bincount(key, minlength=len(wb))
    _static(dev, n_rows - 1, stride, cards, b, ob, offset=1)
    torch.cuda.synchronize()
    assert np.array_equal(ob.cpu().numpy(), want)


def test_static_allocates_nothing_and_rejects_oversize(setup):
    import torch
    ds, cards, dev, n_rows, stride = setup
    out = torch.empty(capi.STATIC_MAX_CELLS, dtype=torch.int32, device="cuda:0")
    t = [0, 1, 2, 3]
    _static(dev, n_rows, stride, cards, t, out)           # warm: module load, symbol address
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(20):
        _static(dev, n_rows, stride, cards, t, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 >= free0                                 # no device memory taken by the calls
    # beyond the static scratch: a clear error, nothing launched
    big = sorted(range(37), key=lambda v: -int(cards[v]))[:12]
    big = sorted(big)
    assert int(np.prod(cards[big].astype(np.int64))) > capi.STATIC_MAX_CELLS
    with pytest.raises(RuntimeError, match="OBN_STATIC_MAX_CELLS"):
        _static(dev, n_rows, stride, cards, big, out)
    with pytest.raises(RuntimeError, match="at most 16"):
        _static(dev, n_rows, stride, np.ones(37, dtype=np.uint32), list(range(17)), out)
    with pytest.raises(ValueError, match="sorted"):
        _static(dev, n_rows, stride, cards, [3, 1], out)
    L = capi.load()
    one = np.array([2], dtype=np.uint32)
    rc = L.obn_count_table_static(0, None, C.c_void_p(dev.data_ptr()), n_rows, stride, capi._ptr(one, C.c_uint32),
                                  capi._ptr(one, C.c_uint32), 1, C.c_void_p(out.data_ptr()), 3)
    assert rc == capi.OBN_ERR_INVALID and b"n_cells" in L.obn_last_error()
    # and the scratch is still usable afterwards
    w = ds.count_tables([t])[0]
    _static(dev, n_rows, stride, cards, t, out)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy()[:len(w)], w)
