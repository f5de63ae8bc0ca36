// rt.h — internal declarations of libfimo_worlds_b200 (the product).
//
// Host side of the B200-native `fimo-worlds` system-group step:
//   futex.cpp   parking lot behind the `fimo-tasks` futex symbols   (modules/fimo_tasks/src/Futex.zig:451-760)
//   pool.cpp    worker pool + command-buffer state machine          (modules/fimo_tasks/src/{Pool,CommandBuffer,Worker}.zig)
//   worlds.cpp  universe / world / system group + the 32 exports    (modules/fimo_worlds/src/*.zig)
//   device.cu   HBM resource store, streams, completion thread, DAG -> CUDA graph executor
//   kernels.cu  sm_100a kernels of the built-in systems and the fused world-step
// Nothing in here links, includes or calls anything under oracle/.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#include "fimo_worlds_b200.h"

namespace fb {

// Contract violations @panic in the reference (World.zig:73-74,90,120; SystemGroup.zig:383-386,409,425).
[[noreturn]] inline void panic(const char *msg) {
    std::fprintf(stderr, "fimo_worlds_b200 panic: %s\n", msg);
    std::fflush(stderr);
    std::abort();
}

void set_last_error(const char *name);
const char *last_error();

// ---- futex.cpp --------------------------------------------------------------------------------
FimoTasksMeta_FutexStatus futex_wait(const void *key, size_t key_size, uint64_t expect, size_t token,
                                     const FimoInstant *timeout);
size_t futex_wake(const void *key, size_t max_waiters, FimoTasksMeta_FutexFilter filter);
// The futex symbols in use: the built-in parking lot, or whatever `bind_tasks` injected.
extern FimoTasksMeta_futex_wait g_futex_wait;
extern FimoTasksMeta_futex_wake g_futex_wake;
inline void fence_wait(FimoWorldsMeta_Fence *f) { FimoWorldsMeta_fence_wait(f, g_futex_wait); }
inline void fence_signal(FimoWorldsMeta_Fence *f) { FimoWorldsMeta_fence_signal(f, g_futex_wake); }

// ---- pool.cpp ---------------------------------------------------------------------------------
struct Pool;
Pool *pool_create(const char *label, size_t label_len, size_t workers);
FimoTasksMeta_Pool pool_handle(Pool *p);
// A pool thread that is about to block for real tells its pool, so the pool can keep
// `worker_count` threads runnable — the OS-thread equivalent of "the fiber parks, the worker
// keeps fetching tasks" (Worker.zig:304-409).
struct BlockingScope {
    BlockingScope();
    ~BlockingScope();
    BlockingScope(const BlockingScope &) = delete;
    BlockingScope &operator=(const BlockingScope &) = delete;
};

// ---- tracing.cpp ------------------------------------------------------------------------------
// fimo tracing subscribers attached with fimo_worlds_b200_trace_subscribe (tracing.zig:387-565).
bool trace_bridge_active();
void trace_bridge_generation(const void *group, const FimoB200_TraceEvent &event);
void trace_bridge_group_destroyed(const void *group);

// Reader/writer lock with try-lock; may be released by a different thread than the one that
// acquired it (a device generation locks on the scheduling thread and unlocks on the
// completion thread).  Writer-preferring like sync/RwLock.zig.
class RwLock {
public:
    void lock_read();
    bool try_lock_read();
    void unlock_read();
    void lock_write();
    bool try_lock_write();
    void unlock_write();

private:
    std::mutex m_;
    std::condition_variable cv_;
    size_t readers_ = 0, writers_waiting_ = 0;
    bool writer_ = false;
};

} // namespace fb
