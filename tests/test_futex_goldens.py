"""The reference's own futex tests (modules/fimo_tasks/src/Futex.zig:760-927, "taken from the rust parking lot
implementation": SingleLatchTest + the twelve `futex wake ...` cases), restated over the two `fimo-tasks` symbols the
system-group step imports — `futex_wait` / `futex_wake` (pkgs/fimo_tasks_meta/src/symbols.zig:132-160) — and run
against the CPU oracle's parking lot and the product's (fimo_b200/csrc/futex.cpp).  Fences and the timeline
semaphore of the path (Job.zig:20-173) sit on exactly these two calls.

A latch is a semaphore word: `down` decrements and parks while the old value was <= 0, `up` increments and wakes
exactly one parked thread.  Every thread that was released bumps `num_awake`.  The cases vary repeats, latches,
the delay before each single wake, threads per latch and how many are woken one by one before `wake(all)`.

Python has no atomics, so the read-modify-write of the semaphore word is done under a lock; what is under test —
wait validating `expect` against memory, wake counts, nobody lost, nobody woken twice — happens inside the library
(ctypes releases the GIL around both calls)."""
import ctypes as C
import threading
import time

import pytest

from fimo_b200 import abi, library

USIZE_MAX = (1 << 64) - 1
ALL = abi.FutexFilter(0, 0, 0)  # FIMO_TASKS_META_FUTEX_FILTER_ALL, futex.h:119


@pytest.fixture(scope="module", params=["oracle", "product"])
def lib(request):
    if request.param == "oracle":
        from oracle import loader

        return loader.load()
    return library.load()


class SingleLatch:
    """Futex.zig:764-822."""

    def __init__(self, dll, num_threads: int):
        self.dll, self.num_threads = dll, num_threads
        self.semaphore = C.c_int64(0)
        self.addr = C.addressof(self.semaphore)
        self.rmw = threading.Lock()
        self.num_awake = 0
        self.woken_twice = False
        self.threads = [threading.Thread(target=self.run) for _ in range(num_threads)]
        for t in self.threads:
            t.start()

    def run(self):
        self.down()
        with self.rmw:
            self.num_awake += 1

    def down(self):  # :796-806
        with self.rmw:
            old = self.semaphore.value
            self.semaphore.value = old - 1
        if old > 0:
            return
        while True:
            current = self.semaphore.value
            status = self.dll.fimo_tasks_futex_wait(self.addr, 8, current & USIZE_MAX, 0, None)
            if status == 0:  # FIMO_TASKS_META_FUTEX_STATUS_OK: woken by `up` / `finish`
                return
            assert status == 1, status  # Invalid: the word changed before we parked -> retry (":802 catch continue")

    def up(self):  # :808-820
        with self.rmw:
            old = self.semaphore.value
            self.semaphore.value = old + 1
        if old < 0:
            while True:
                woken = self.dll.fimo_tasks_futex_wake(self.addr, 1, ALL)
                if woken == 1:
                    return
                assert woken == 0, f"should not wake up {woken} threads"

    def wake_one(self):  # :771-777
        before = self.num_awake
        self.up()
        deadline = time.time() + 30
        while self.num_awake != before + 1:
            assert time.time() < deadline, "a single wake released nobody"
            time.sleep(0)

    def finish(self, num_single_unparks: int):  # :779-790
        left = self.num_threads - num_single_unparks
        deadline = time.time() + 60
        while left > 0:
            before = self.num_awake
            woken = self.dll.fimo_tasks_futex_wake(self.addr, USIZE_MAX, ALL)
            assert woken <= left
            while self.num_awake != before + woken:
                assert time.time() < deadline, "wake(all) reported threads that never ran"
                time.sleep(0)
            left -= woken
            assert time.time() < deadline, "threads never parked"
        for t in self.threads:
            t.join(30)
            assert not t.is_alive()
        assert self.num_awake == self.num_threads


def run_case(dll, repeats, num_latches, delay_us, num_threads, num_single_unparks):  # testRun / testWrapper, :824-881
    for _ in range(repeats):
        latches = [SingleLatch(dll, num_threads) for _ in range(num_latches)]
        for _i in range(num_single_unparks):
            time.sleep(delay_us * 1e-6)
            for latch in latches:
                latch.wake_one()
        for latch in latches:
            latch.finish(num_single_unparks)


# (name, repeats, latches, delay in us, threads, single unparks) — Futex.zig:883-929.  Repeats of the four largest
# cases are scaled down (Python threads) so that the whole CPU suite stays within minutes; the shapes are the reference's.
CASES = [
    ("wake all one fast", 200, 1, 0, 1, 0),                      # reference: 1000 repeats
    ("wake all hundred fast", 10, 1, 0, 100, 0),                 # 100
    ("wake one one fast", 200, 1, 0, 1, 1),                      # 1000
    ("wake one hundred fast", 3, 1, 0, 100, 100),                # 20
    ("wake one fifty then fifty all fast", 5, 1, 0, 100, 50),    # 50
    ("wake all one", 20, 1, 10000, 1, 0),                        # 100
    ("wake all hundred", 5, 1, 10000, 100, 0),                   # 100
    ("wake one one", 10, 1, 10000, 1, 1),
    ("wake one fifty", 1, 1, 10000, 50, 50),
    ("wake one fifty then fifty all", 2, 1, 10000, 100, 50),
    ("wake hundred all one fast", 20, 100, 0, 1, 0),             # 100
    ("wake hundred all one", 1, 100, 10000, 1, 0),
]


@pytest.mark.timeout(300)
@pytest.mark.parametrize("case", CASES, ids=[c[0].replace(" ", "-") for c in CASES])
def test_reference_futex_latch_cases(lib, case):
    _name, repeats, latches, delay_us, threads, singles = case
    run_case(lib.dll, repeats, latches, delay_us, threads, singles)


def test_wait_rejects_a_stale_expect_and_wake_counts_nobody(lib):
    """futex.wait semantics the fence relies on (Futex.zig:451-520): a mismatching `expect` returns Invalid without
    parking; wake on a key nobody waits on returns 0."""
    word = C.c_uint64(7)
    assert lib.dll.fimo_tasks_futex_wait(C.addressof(word), 8, 8, 0, None) == 1  # Invalid
    assert lib.dll.fimo_tasks_futex_wake(C.addressof(word), USIZE_MAX, ALL) == 0
    byte = C.c_uint8(1)
    assert lib.dll.fimo_tasks_futex_wait(C.addressof(byte), 1, 0, 0, None) == 1  # key_size 1: the fence's state byte
