// oracle_internal.h — shared declarations of the CPU oracle.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: it is
// built into oracle/_build/libfimo_worlds_oracle.so and may be loaded only by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
//
// Parity status: SCHEDULER semantics are pinned against the reference's own golden
// tests (pkgs/fimo_worlds_meta/src/systems.zig:214-253,427-577,1263-1289;
// resources.zig:122-313; worlds.zig:201-261 — restated in tests/test_oracle_reference_goldens.py).
// ARITHMETIC of the synthetic systems is "parity unpinned": the reference contains
// no transform/cull code at all; oracle/systems_ref.c is the
// definition the CUDA path is compared against.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <mutex>

#include "fimo_worlds_b200.h"

namespace oracle {

[[noreturn]] inline void panic(const char *msg) {
    // The reference @panic()s on contract violations (World.zig:73-74,90,120; SystemGroup.zig:383-386).
    std::fprintf(stderr, "fimo_worlds oracle panic: %s\n", msg);
    std::fflush(stderr);
    std::abort();
}

// ---- futex (oracle_futex.cpp) -------------------------------------------------
// Restates modules/fimo_tasks/src/Futex.zig:451-760 (address-hashed parking lot with
// token filters).  Besides blocking waiters it supports callback waiters, which is how
// the oracle stands in for parked fibers: a parked fiber is "a continuation that the
// waker re-enqueues" (Pool.zig:474-532, 392-401).
FimoTasksMeta_FutexStatus futex_wait(const void *key, size_t key_size, uint64_t expect, size_t token,
                                     const FimoInstant *timeout);
size_t futex_wake(const void *key, size_t max_waiters, FimoTasksMeta_FutexFilter filter);
// Registers `fn` to run (on the waking thread) when `key` is woken with a matching filter.
// Returns false without registering if *key != expect.
bool futex_wait_async(const void *key, size_t key_size, uint64_t expect, size_t token, std::function<void()> fn);

inline void fence_wait(FimoWorldsMeta_Fence *f) { FimoWorldsMeta_fence_wait(f, &futex_wait); }
inline void fence_signal(FimoWorldsMeta_Fence *f) { FimoWorldsMeta_fence_signal(f, &futex_wake); }
// Runs `fn` once the fence is signaled (immediately if it already is).
void fence_then(FimoWorldsMeta_Fence *f, std::function<void()> fn);

// ---- pool (oracle_pool.cpp) ---------------------------------------------------
struct Pool;
Pool *pool_create(const char *label, size_t label_len, size_t workers);
FimoTasksMeta_Pool pool_handle(Pool *p); // {data, vtable}, no reference taken
Pool *pool_current();                    // pool owning the calling worker thread, or null
// A worker that is about to block for real brackets the wait with these so the pool
// keeps `worker_count` runnable threads (the stand-in for "the fiber parks, the OS
// thread keeps working", Worker.zig:304-409).
void pool_blocking_enter();
void pool_blocking_exit();

// ---- simple reader/writer lock usable from any thread --------------------------
// Stand-in for fimo_tasks_meta.sync.RwLock (pkgs/fimo_tasks_meta/src/sync/RwLock.zig):
// writer-preferring, blocking.
class RwLock {
public:
    void lock_read() {
        std::unique_lock<std::mutex> g(m_);
        if (writer_ || writers_waiting_ > 0) {
            pool_blocking_enter();
            cv_.wait(g, [&] { return !writer_ && writers_waiting_ == 0; });
            pool_blocking_exit();
        }
        ++readers_;
    }
    void unlock_read() {
        std::lock_guard<std::mutex> g(m_);
        if (--readers_ == 0) cv_.notify_all();
    }
    void lock_write() {
        std::unique_lock<std::mutex> g(m_);
        if (writer_ || readers_ > 0) {
            ++writers_waiting_;
            pool_blocking_enter();
            cv_.wait(g, [&] { return !writer_ && readers_ == 0; });
            pool_blocking_exit();
            --writers_waiting_;
        }
        writer_ = true;
    }
    void unlock_write() {
        std::lock_guard<std::mutex> g(m_);
        writer_ = false;
        cv_.notify_all();
    }

private:
    std::mutex m_;
    std::condition_variable cv_;
    size_t readers_ = 0, writers_waiting_ = 0;
    bool writer_ = false;
};

// thread-local error name behind the last FIMO_STATUS_ERROR (ctx.setResult stand-in)
void set_last_error(const char *name);

} // namespace oracle
