"""GPU parity of the CI path's table sharing (obn_ci_groups.inl + the member epilogue of count_kernel)
and of obn_pc_round (decision on the device) against the checker: frontier-shaped batches where every
variable set is asked by several (x, y) assignments, all table tiers, duplicated ids, both halves of
the submit / collect split."""
import itertools

import numpy as np
import pytest

import oracle
from helpers import close, ranked
from openbnsl_b200 import capi
from openbnsl_b200 import networks as nw
from test_ci_groups import star_frontier

pytestmark = pytest.mark.gpu
INT32_MAX = 2**31 - 1


def check_against_oracle(ds, ods, xs, ys, tuples, kind):
    flat, offs = capi._csr(tuples)
    xa, ya = np.asarray(xs, dtype=np.uint32), np.asarray(ys, dtype=np.uint32)
    stat, dof, p = ds.ci_tests_csr(flat, offs, xa, ya, kind)
    for t in range(len(tuples)):
        z = list(tuples[t])
        z.remove(xs[t]); z.remove(ys[t])            # first occurrences are x and y (citest.h:63-64)
        ws, wd, wp = oracle.citest(ods, xs[t], ys[t], z, kind)
        assert int(dof[t]) == wd, (t, tuples[t], xs[t], ys[t])
        assert close(stat[t], ws, 1e-9, 1e-9), (t, tuples[t], stat[t], ws)
        assert abs(p[t] - wp) <= 1e-9
    return stat, dof, p


@pytest.mark.parametrize("netf,n_rows,depth,degree", [
    (nw.asia, 10000, 1, 4), (nw.alarm, 20011, 1, 5), (nw.alarm, 20011, 2, 5), (nw.alarm, 100000, 3, 5),
    (nw.munin1_like, 30001, 1, 4), (nw.munin1_like, 30001, 2, 4), (nw.hepar2_like, 50000, 4, 6)])
@pytest.mark.parametrize("kind", [0, 1])
def test_grouped_ci_equals_checker(gpu_ctx, netf, n_rows, depth, degree, kind):
    net = netf()
    codes, cards = ranked(net.sample(n_rows, 5))
    ods = oracle.Dataset(codes, cards)
    ds = gpu_ctx.upload(codes, cards)
    rng = np.random.default_rng(7 * depth + degree)
    xs, ys, tuples = star_frontier(rng, net.n, 3, depth, degree)
    groups, _, _ = capi.ci_group_plan(cards, tuples, xs, ys)
    assert depth == 0 or len(groups) * (depth + 1) <= len(tuples)        # the batch really shares tables
    gpu_ctx.profile_read(True)
    stat, dof, p = check_against_oracle(ds, ods, xs, ys, tuples, kind)
    prof = gpu_ctx.profile_read(True)
    assert prof.tables == len(groups) and prof.tuples == len(tuples)
    assert prof.scanned_bytes <= prof.algorithmic_bytes / (depth + 1) + 1
    # the same tests one per batch (no sharing possible): identical tables, so identical dof and statistics
    flat, offs = capi._csr(tuples)
    for t in rng.choice(len(tuples), size=min(12, len(tuples)), replace=False):
        f1, o1 = capi._csr([tuples[t]])
        s1, d1, p1 = ds.ci_tests_csr(f1, o1, np.asarray([xs[t]], dtype=np.uint32), np.asarray([ys[t]], dtype=np.uint32), kind)
        assert int(d1[0]) == int(dof[t]) and close(s1[0], stat[t], 1e-12, 1e-12)
    ds.close()


def test_every_role_assignment_every_tier_and_duplicates(gpu_ctx):
    """All C(k, 2) assignments over one set (x / y not adjacent digits, not the last digits), sets whose tables
    fall in every size class up to the global tier, x > y on input, and repeated ids."""
    net = nw.munin1_like()
    codes, cards = ranked(net.sample(40000, 2))
    ods = oracle.Dataset(codes, cards)
    ds = gpu_ctx.upload(codes, cards)
    order = np.argsort(cards)
    small, big = [int(v) for v in order[:8]], [int(v) for v in order[-6:]]
    sets = [sorted(small[:2]), sorted(small[:4]), sorted(small[:6]), sorted(small[2:5] + big[:1]),
            sorted(big[:2] + small[:2]), sorted(big[:3]), sorted(big[:3] + small[:2]), sorted(big[:4]), sorted(big[:5])]
    xs, ys, tuples = [], [], []
    for vs in sets:
        cells = int(np.prod([cards[v] for v in vs]))
        if cells >= 2**25:
            continue
        for a, b in itertools.combinations(vs, 2):
            xs.append(b); ys.append(a); tuples.append(vs)           # x > y: swapped by the library (citest.h:48)
    v0, v1 = small[0], small[1]
    lo, hi = min(v0, v1), max(v0, v1)
    tuples += [[lo, lo, hi], [lo, hi, hi]]; xs += [lo, lo]; ys += [hi, hi]
    for kind in (0, 1):
        check_against_oracle(ds, ods, xs, ys, tuples, kind)
    ds.close()


@pytest.mark.parametrize("split", [False, True])
def test_pc_round_takes_the_decision_on_the_device(gpu_ctx, split):
    net = nw.alarm()
    codes, cards = ranked(net.sample(30000, 9))
    ds = gpu_ctx.upload(codes, cards)
    rng = np.random.default_rng(3)
    xs, ys, tuples = star_frontier(rng, net.n, 8, 1, 5)
    flat, offs = capi._csr(tuples)
    xa, ya = np.asarray(xs, dtype=np.uint32), np.asarray(ys, dtype=np.uint32)
    pairs = {}
    slot = np.array([pairs.setdefault((min(a, b), max(a, b)), len(pairs)) for a, b in zip(xs, ys)], dtype=np.uint32)
    n_slots = len(pairs) + 2                                         # two edges without any test in this round
    for kind, level in ((1, 0.05), (0, 0.5), (1, 1e-12), (0, 1.1)):
        _, _, p = ds.ci_tests_csr(flat, offs, xa, ya, kind)
        want = np.full(n_slots, INT32_MAX, dtype=np.int64)
        for t in range(len(tuples)):
            if p[t] >= level:
                want[slot[t]] = min(want[slot[t]], t)
        got = ds.pc_round(flat, offs, xa, ya, slot, n_slots, kind, level, split=split)
        assert got.tolist() == want.tolist()
    # an empty round, and a batch call while a round is in flight
    empty = ds.pc_round(np.zeros(0, np.uint32), np.zeros(1, np.uint64), xa[:0], ya[:0], slot[:0], 3, 1, 0.05, split=split)
    assert empty.tolist() == [INT32_MAX] * 3
    with pytest.raises(IndexError):
        ds.pc_round(flat, offs, xa, ya, slot + 1000, n_slots, 1, 0.05)
    ds.close()


def test_two_rounds_in_flight_and_chained_rounds(gpu_ctx):
    """Two PC rounds in flight per ctx; the second one chained to the first: the tests of the edges the first round
    decides are skipped on the device (never counted, never reported), so first[] and the profile counters equal
    those of a host that had waited for round one and filtered round two itself."""
    net = nw.hepar2_like()
    codes, cards = ranked(net.sample(40000, 4))
    ds = gpu_ctx.upload(codes, cards)
    rng = np.random.default_rng(8)
    xs, ys, tuples = star_frontier(rng, net.n, 10, 2, 6)
    n = len(tuples)
    xa, ya = np.asarray(xs, dtype=np.uint32), np.asarray(ys, dtype=np.uint32)
    pairs = {}
    slot = np.array([pairs.setdefault((min(a, b), max(a, b)), len(pairs)) for a, b in zip(xs, ys)], dtype=np.uint32)
    n_slots = len(pairs)
    order = rng.permutation(n)
    ia, ib = np.sort(order[: n // 2]), np.sort(order[n // 2:])          # round one / round two, interleaved edges

    def sub(idx):
        flat, offs = capi._csr([tuples[i] for i in idx])
        return flat, offs, xa[idx], ya[idx], slot[idx]

    for kind, level in ((1, 0.05), (0, 0.3), (1, 1e-9)):
        A, B = sub(ia), sub(ib)
        first_a = ds.pc_round(*A, n_slots, kind, level)
        keep = np.array([first_a[s] == INT32_MAX for s in B[4]])          # what a waiting host would still ask
        assert 0 < keep.sum() < len(ib) or level == 1e-9
        Bf = sub(ib[keep])
        gpu_ctx.profile_enable(True)
        gpu_ctx.profile_read(True)
        want_b = ds.pc_round(*Bf, n_slots, kind, level)
        prof_want = gpu_ctx.profile_read(True)
        back = np.flatnonzero(keep)                                      # index in the filtered round -> index in the full one
        want_b = np.array([INT32_MAX if v == INT32_MAX else back[v] for v in want_b.tolist()])
        # pipelined: both rounds enqueued before anything is collected
        ds.pc_round_submit(*A, n_slots, kind, level)
        ds.pc_round_submit(*B, n_slots, kind, level, chain=True)
        with pytest.raises(ValueError, match="two submitted batches"):
            ds.pc_round_submit(*B, n_slots, kind, level, chain=True)
        with pytest.raises(ValueError, match="waiting to be collected"):
            ds.count_tables([[0, 1]])
        got_a = ds.pc_round_collect(n_slots)
        prof_a = gpu_ctx.profile_read(True)
        got_b = ds.pc_round_collect(n_slots)
        prof_b = gpu_ctx.profile_read(True)
        assert got_a.tolist() == first_a.tolist()
        assert got_b.tolist() == want_b.tolist()
        assert prof_a.tuples == len(ia)
        for f in ("tuples", "tables", "algorithmic_bytes", "scanned_bytes", "increments"):
            assert getattr(prof_b, f) == getattr(prof_want, f), f
        # a third round chained to the second (the first one's staging is reused)
        ds.pc_round_submit(*A, n_slots, kind, level, chain=True)
        got_c = ds.pc_round_collect(n_slots)
        keep_c = np.array([want_b[s] == INT32_MAX for s in A[4]])
        if keep_c.any():
            Af = sub(ia[keep_c])
            back_c = np.flatnonzero(keep_c)
            want_c = [INT32_MAX if v == INT32_MAX else back_c[v] for v in ds.pc_round(*Af, n_slots, kind, level).tolist()]
            assert got_c.tolist() == want_c
        else:
            assert got_c.tolist() == [INT32_MAX] * n_slots
        gpu_ctx.profile_enable(False)
    # chain without a round to continue, or over other slots
    ds.count_tables([[0, 1]])
    A = sub(ia)
    ds.pc_round_submit(*A, n_slots, 1, 0.05)
    with pytest.raises(ValueError, match="chain"):
        ds.pc_round_submit(*A, n_slots + 1, 1, 0.05, chain=True)
    ds.pc_round_collect(n_slots)
    # two plain CI batches in flight, collected in submission order
    L = capi.load()
    import ctypes as C
    flat_a, offs_a = capi._csr([tuples[i] for i in ia])
    flat_b, offs_b = capi._csr([tuples[i] for i in ib])
    for flat, offs, idx in ((flat_a, offs_a, ia), (flat_b, offs_b, ib)):
        capi._check(L.obn_ci_tests_submit(gpu_ctx._h, ds._h, capi._ptr(flat, C.c_uint32), capi._ptr(offs, C.c_uint64), len(idx),
                                          capi._ptr(np.ascontiguousarray(xa[idx]), C.c_uint32),
                                          capi._ptr(np.ascontiguousarray(ya[idx]), C.c_uint32), 1, 1))
    got = []
    for flat, offs, idx in ((flat_a, offs_a, ia), (flat_b, offs_b, ib)):
        stat = np.zeros(len(idx)); dof = np.zeros(len(idx), dtype=np.uint64); p = np.zeros(len(idx))
        capi._check(L.obn_ci_tests_collect(gpu_ctx._h, capi._ptr(stat, C.c_double), capi._ptr(dof, C.c_uint64), capi._ptr(p, C.c_double)))
        got.append((stat, dof, p))
    for (flat, offs, idx), (stat, dof, p) in zip(((flat_a, offs_a, ia), (flat_b, offs_b, ib)), got):
        ws, wd, wp = ds.ci_tests_csr(flat, offs, np.ascontiguousarray(xa[idx]), np.ascontiguousarray(ya[idx]), 1)
        assert np.array_equal(stat, ws) and np.array_equal(dof, wd) and np.array_equal(p, wp)
    ds.close()
