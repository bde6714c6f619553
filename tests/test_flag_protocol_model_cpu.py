"""An executable model of the cross-GPU flag protocol (``csrc/common/symm.cuh``): threads stand for GPUs, numpy arrays
for the symmetric flag pages and data windows.  It checks the DESIGN -- monotonic epochs that are never reset, one slot
per op family, device-side epoch counters that make launches replayable, the open / close barrier pair of the
reduction kernels -- against adversarial schedules (random delays, a rank that runs far ahead).  Memory-model questions
(release / acquire at system scope) are outside what a Python model can show; this pins down the logic:

* nobody leaves barrier ``e`` before everybody has entered it, for thousands of back-to-back barriers without a reset;
* a fast rank cannot lap a slow one (it needs the slow rank's arrival for every epoch);
* 32-bit wrap-around of the epoch comparison ``int32(flag - epoch) >= 0`` is safe;
* two slots used concurrently by different "streams" do not disturb each other;
* the two-shot all-reduce schedule (open barrier, reduce own slice from all peers, store to all peers, close barrier)
  never reads a peer's slice before it is written nor overwrites one that is still being read -- iteration after
  iteration on the SAME buffer.
"""
import random
import threading
import time

import numpy as np
import pytest

MAX_WORLD, MAX_SLOTS = 16, 8
EPOCH_BASE = MAX_SLOTS * MAX_WORLD + 2 * MAX_SLOTS           # as in symm.cuh: arrivals, go[], done[], epoch[]


class Fabric:
    """Flag pages of ``world`` ranks; every rank can write every page (peer-mapped memory)."""

    def __init__(self, world: int, start_epoch: int = 1):
        self.world = world
        self.pages = [np.zeros(EPOCH_BASE + MAX_SLOTS, dtype=np.uint32) for _ in range(world)]
        for p in self.pages:
            p[EPOCH_BASE:EPOCH_BASE + MAX_SLOTS] = start_epoch      # device-side epoch counters start at the same value
            p[:MAX_SLOTS * MAX_WORLD] = np.uint32(start_epoch - 1)  # arrival words: "everything before is done"

    def barrier(self, rank: int, slot: int, epoch: int, jitter: float = 0.0) -> None:
        epoch &= 0xFFFFFFFF
        for t in range(self.world):                               # st.release.sys peer[t].slot[rank] = epoch
            if jitter:
                time.sleep(random.random() * jitter)
            self.pages[t][slot * MAX_WORLD + rank] = epoch
        for t in range(self.world):                               # spin on my own page
            while True:
                diff = (int(self.pages[rank][slot * MAX_WORLD + t]) - epoch) & 0xFFFFFFFF
                if diff < 0x80000000:                             # int32(flag - epoch) >= 0
                    break
                time.sleep(0)

    def take_epoch(self, rank: int, slot: int, n: int) -> int:
        """Device-side counter: read, hand out ``n`` epochs, store the next value (one kernel per slot at a time)."""
        e = int(self.pages[rank][EPOCH_BASE + slot])
        self.pages[rank][EPOCH_BASE + slot] = np.uint32((e + n) & 0xFFFFFFFF)
        return e


def _run(world, fn):
    errors = []

    def wrap(r):
        try:
            fn(r)
        except BaseException as e:  # noqa: BLE001 - reported to the main thread
            errors.append((r, repr(e)))

    threads = [threading.Thread(target=wrap, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert not any(t.is_alive() for t in threads), "deadlock"
    assert not errors, errors


@pytest.mark.parametrize("world", [2, 3, 8])
def test_barrier_separates_phases_without_any_reset(world):
    fab = Fabric(world)
    phase = np.zeros(world, dtype=np.int64)                       # what each rank has published
    iters = 400

    def rank_fn(r):
        rng = random.Random(r)
        for i in range(1, iters + 1):
            if r == 0 and i % 50 == 0:
                time.sleep(0.002)                                 # a slow rank now and then
            phase[r] = i                                          # "write my data for iteration i"
            fab.barrier(r, slot=0, epoch=fab.take_epoch(r, 0, 1), jitter=1e-5 if rng.random() < 0.05 else 0.0)
            seen = phase.copy()
            # everybody has published i; nobody can be further than i + 1 (they need MY arrival for the next barrier)
            assert (seen >= i).all() and (seen <= i + 1).all(), (r, i, seen)

    _run(world, rank_fn)
    assert all(int(p[EPOCH_BASE]) == 1 + iters for p in fab.pages)   # the device counters advanced in lock step


def test_epoch_comparison_survives_wrap_around():
    world = 4
    fab = Fabric(world, start_epoch=0xFFFFFFFF - 5)               # 12 barriers straddle the 32-bit wrap
    phase = np.zeros(world, dtype=np.int64)

    def rank_fn(r):
        for i in range(1, 13):
            phase[r] = i
            fab.barrier(r, 0, fab.take_epoch(r, 0, 1))
            assert (phase >= i).all()

    _run(world, rank_fn)
    assert int(fab.pages[0][EPOCH_BASE]) == (0xFFFFFFFF - 5 + 12) & 0xFFFFFFFF


def test_slots_are_independent():
    """Two op families (say DP buckets on the comm stream, TP collectives on the compute stream) interleave freely."""
    world = 3
    fab = Fabric(world)
    a = np.zeros(world, dtype=np.int64)
    b = np.zeros(world, dtype=np.int64)

    def stream(r, slot, arr, iters, delay):
        for i in range(1, iters + 1):
            arr[r] = i
            if delay and (i + r) % 7 == 0:
                time.sleep(delay)
            fab.barrier(r, slot, fab.take_epoch(r, slot, 1))
            assert (arr >= i).all() and (arr <= i + 1).all()

    def rank_fn(r):
        t1 = threading.Thread(target=stream, args=(r, 3, a, 300, 0.0))
        t2 = threading.Thread(target=stream, args=(r, 4, b, 120, 0.001))
        t1.start(), t2.start()
        t1.join(60), t2.join(60)
        assert not t1.is_alive() and not t2.is_alive()

    _run(world, rank_fn)
    assert int(fab.pages[0][EPOCH_BASE + 3]) == 301 and int(fab.pages[0][EPOCH_BASE + 4]) == 121


@pytest.mark.parametrize("world", [2, 4])
def test_two_shot_allreduce_schedule_on_a_reused_buffer(world):
    """open barrier -> reduce my slice from every peer's window -> store the sum into every peer's window -> close
    barrier; the next iteration overwrites the same window right away (as the gradient buckets do every step)."""
    fab = Fabric(world)
    n = 8 * world
    bufs = [np.zeros(n, dtype=np.float64) for _ in range(world)]
    iters = 150

    def rank_fn(r):
        rng = random.Random(100 + r)
        lo, hi = r * n // world, (r + 1) * n // world
        for i in range(iters):
            bufs[r][:] = (r + 1) * (i + 1)                         # this step's local gradient
            e = fab.take_epoch(r, 4, 2)
            fab.barrier(r, 4, e)                                  # everybody's gradient is in place
            acc = sum(bufs[p][lo:hi].copy() for p in range(world))
            if rng.random() < 0.1:
                time.sleep(1e-4)
            for p in range(world):                                # two-shot: store my reduced slice to every peer
                bufs[p][lo:hi] = acc
            fab.barrier(r, 4, e + 1)                              # results have landed everywhere; buffers may be reused
            want = (i + 1) * world * (world + 1) / 2
            assert np.all(bufs[r] == want), (r, i, bufs[r], want)

    _run(world, rank_fn)


def test_missing_close_barrier_is_caught_by_the_model():
    """Negative control, made deterministic with events: without the closing barrier nothing stops rank 0 from writing
    its NEXT gradient into its window while rank 1 has not reduced the current one yet."""
    world, n, iters = 2, 4, 5
    fab = Fabric(world)
    bufs = [np.zeros(n, dtype=np.float64) for _ in range(world)]
    overwritten = [threading.Event() for _ in range(iters)]
    bad = []

    def rank_fn(r):
        lo, hi = r * n // world, (r + 1) * n // world
        bufs[r][:] = (r + 1) * 1
        for i in range(iters):
            fab.barrier(r, 0, fab.take_epoch(r, 0, 1))            # opening barrier only
            if r == 1:
                overwritten[i].wait(10)                           # rank 0 is already one step ahead (see below)
            acc = sum(bufs[p][lo:hi].copy() for p in range(world))
            if not np.all(acc == (i + 1) * 3):
                bad.append((r, i, acc.tolist()))
            if r == 0:
                bufs[0][:] = 1 * (i + 2)                          # next step's gradient: legal without a closing barrier
                overwritten[i].set()
            else:
                bufs[1][:] = 2 * (i + 2)

    _run(world, rank_fn)
    assert any(r == 1 for r, _, _ in bad), "the model failed to show the race the closing barrier exists for"
