"""Model check (CPU) of the mbarrier protocol of scan_rows_tma (obn_count_kernel.cuh): the same
sequence of init / arrive.expect_tx / bulk copy / try_wait.parity / arrive per thread, run under
random interleavings with late-completing copies.  It cannot prove the PTX right, but it does pin
the part that hangs a GPU when it is wrong: slots, phases and parities of the `full` / `empty`
rings, the one-tile-late refill, partial last tiles and empty row ranges."""
import random

import pytest

STAGES = 3          # kTmaStages


class Barrier:
    """mbarrier: `count` arrivals (+ outstanding transaction bytes) complete a phase."""

    def __init__(self, count):
        self.count, self.pending, self.tx, self.phase = count, count, 0, 0

    def _maybe_complete(self):
        if self.pending == 0 and self.tx == 0:
            self.phase += 1
            self.pending = self.count

    def arrive(self, expect_tx=0):
        assert self.pending > 0, "more arrivals than the barrier was initialised for"
        self.tx += expect_tx
        self.pending -= 1
        self._maybe_complete()

    def complete_tx(self, nbytes):
        self.tx -= nbytes
        self._maybe_complete()

    def try_wait(self, parity):          # true once the phase with this parity has completed
        return (self.phase & 1) != parity


def run(nthr, n_vec, k, rng):
    tv = nthr
    n_tiles = (n_vec + tv - 1) // tv if n_vec else 0
    full = [Barrier(1) for _ in range(STAGES)]
    empty = [Barrier(nthr) for _ in range(STAGES)]
    slot_tile = [None] * STAGES          # which tile's bytes a slot holds (None while a copy is landing)
    unread = [0] * STAGES                # threads that have not read the slot's current tile yet
    copies = []                          # in-flight bulk copies: (slot, tile, bytes)
    consumed = [[] for _ in range(nthr)]

    def issue(tile):
        s = tile % STAGES
        assert unread[s] == 0, "slot refilled while some thread has not read it"
        nbytes = min(tv, n_vec - tile * tv) * 16
        full[s].arrive(expect_tx=k * nbytes)
        slot_tile[s] = None
        for _ in range(k):
            copies.append((s, tile, nbytes))

    def thread(tid):
        if tid == 0:
            for t in range(min(n_tiles, STAGES)):
                issue(t)
        yield
        for tile in range(n_tiles):
            s = tile % STAGES
            while not full[s].try_wait((tile // STAGES) & 1):
                yield
            assert slot_tile[s] == tile, "read a slot that holds another tile"
            if tile * tv + tid < n_vec:
                consumed[tid].append(tile)
            unread[s] -= 1
            empty[s].arrive()
            yield                           # the increments
            if tid == 0 and tile >= 1:
                pt, nt = tile - 1, tile - 1 + STAGES
                if nt < n_tiles:
                    while not empty[pt % STAGES].try_wait((pt // STAGES) & 1):
                        yield
                    issue(nt)
            yield

    threads = {t: thread(t) for t in range(nthr)}
    steps = 0
    while threads:
        steps += 1
        assert steps < 200000, "deadlock"
        if copies and rng.random() < 0.3:                      # a bulk copy lands
            s, tile, nbytes = copies.pop(rng.randrange(len(copies)))
            full[s].complete_tx(nbytes)
            if not any(c[0] == s for c in copies):
                slot_tile[s] = tile
                unread[s] = nthr
        tid = rng.choice(list(threads))
        try:
            next(threads[tid])
        except StopIteration:
            del threads[tid]
    assert not copies
    for tid in range(nthr):
        assert consumed[tid] == [t for t in range(n_tiles) if t * tv + tid < n_vec]


@pytest.mark.parametrize("seed", range(30))
def test_ring_never_deadlocks_and_never_reads_a_wrong_tile(seed):
    rng = random.Random(seed)
    nthr = rng.choice([1, 2, 4, 8])
    n_vec = rng.choice([0, 1, nthr - 1, nthr, nthr + 1, 3 * nthr, 3 * nthr + 1, 7 * nthr + 2, 20 * nthr])
    run(nthr, max(0, n_vec), rng.randint(1, 10), rng)
