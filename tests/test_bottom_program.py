"""The bottom tier of the marginalisation lattice (openbnsl_b200/csrc/obn_lattice.cuh) runs a fixed
per-thread PROGRAM built on the host.  This CPU test fetches it through the C ABI
(obn_lattice_bottom_program needs no device), replays both phases of the kernel in numpy on random
256-cell chunks and checks every one of the 256 bottom patterns against directly computed
marginals -- so a change of the program builder is caught without a GPU."""
import ctypes as C

import numpy as np
import pytest

from openbnsl_b200 import capi


def _program():
    lib = capi.load()
    lib.obn_lattice_bottom_program.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_int)]
    geo = (C.c_int * 6)()
    assert lib.obn_lattice_bottom_program(None, 0, None, 0, geo) == 0
    bot, threads, quads, tab_quads, upper, slots = list(geo)
    words = np.zeros(upper * threads, dtype=np.uint32)
    pats = np.zeros(slots * threads, dtype=np.uint16)
    assert lib.obn_lattice_bottom_program(words.ctypes.data, words.size, pats.ctypes.data, pats.size, geo) == 0
    # too-small buffers are refused, not overrun
    assert lib.obn_lattice_bottom_program(words.ctypes.data, 3, pats.ctypes.data, pats.size, geo) != 0
    return bot, threads, quads, tab_quads, upper, slots, words.reshape(upper, threads), pats


def test_geometry():
    bot, threads, quads, tab_quads, upper, slots, words, pats = _program()
    assert (bot, upper, quads) == (8, 6, 3 ** 6)
    assert slots * threads >= quads and tab_quads >= quads + 3
    # every quad has a pattern, quads of one pattern are an aligned power-of-two run (the flush relies on it)
    p = pats[:quads].astype(int)
    assert (p != 0xFFFF).all() and (pats[quads:] == 0xFFFF).all()
    start = 0
    while start < quads:
        U = p[start]
        run = 1 << bin(U).count("1")
        assert (p[start:start + run] == U).all() and start % min(run, 32) == 0
        assert start + run == quads or p[start + run] != U
        start += run
    assert sorted(set(p.tolist())) == list(range(64))


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_replay_matches_direct_marginals(seed):
    bot, threads, quads, tab_quads, upper, slots, words, pats = _program()
    rng = np.random.default_rng(seed)
    chunk = rng.integers(0, 1000, size=1 << bot).astype(np.int64)
    tab = np.zeros((tab_quads, 4), dtype=np.int64)            # quads 729.. stay zero / dump
    tab[:64] = chunk.reshape(64, 4)                            # level 6: the chunk itself
    items = [192, 240, 160, 60, 12, 1]                         # C(6,j) 2^j for j = 5 .. 0
    for lvl in range(upper):
        new = {}
        for t in range(threads):
            w = int(words[lvl, t])
            s0, sh, dst = w & 0x3FF, (w >> 10) & 7, w >> 13
            if t < items[lvl]:
                assert dst < quads and s0 + (1 << sh) < quads
                new[dst] = tab[s0] + tab[s0 + (1 << sh)]
            else:                                              # idle thread: zero sources, dump destination
                assert s0 >= quads and dst >= quads
        for d, v in new.items():                               # all loads of a level precede its stores
            tab[d] = v
    # phase 2: nine cells per quad, summed per (upper pattern, which of the two fastest digits stay)
    got = {}
    term = lambda x: int(x) * int(x) + 3 * int(x)              # any function of the count with term(0) = 0
    for q in range(quads):
        U = int(pats[q])
        x0, x1, x2, x3 = (int(v) for v in tab[q])
        m0, m1, n0, n1 = x0 + x1, x2 + x3, x0 + x2, x1 + x3
        for l, cells in ((3, (x0, x1, x2, x3)), (2, (m0, m1)), (1, (n0, n1)), (0, (m0 + m1,))):
            got[(U << 2) | l] = got.get((U << 2) | l, 0) + sum(term(c) for c in cells)
    cube = chunk.reshape([2] * bot)                            # axis 0 = bit 7 ... axis 7 = bit 0
    for p in range(1 << bot):
        axes = tuple(bot - 1 - j for j in range(bot) if not (p >> j) & 1)
        m = cube.sum(axis=axes) if axes else cube
        assert got[p] == sum(term(c) for c in np.asarray(m).reshape(-1)), p
