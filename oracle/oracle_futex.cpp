// oracle_futex.cpp — user-space futex of the CPU oracle (TEST INFRASTRUCTURE ONLY).
//
// Restates the contract of modules/fimo_tasks/src/Futex.zig:
//   wait      :451-559  compare *key with `expect` under the bucket lock, park with a token
//   wakeFilter:561-623  wake up to `max_waiters` waiters of `key` whose token passes `filter`
// and the filter semantics of pkgs/fimo_tasks_meta/src/sync/Futex.zig:37-114
// (`checkToken`: mask, optional deref, typed compare).
#include <chrono>
#include <cstring>
#include <deque>
#include <memory>
#include <vector>

#include "oracle_internal.h"

namespace oracle {
namespace {

struct Waiter {
    const void *key;
    size_t token;
    bool woken = false;
    std::function<void()> callback; // empty => blocking waiter
    std::condition_variable cv;
};

struct Bucket {
    std::mutex m;
    std::deque<std::shared_ptr<Waiter>> waiters; // FIFO (Futex.zig keeps an intrusive FIFO per bucket)
};

constexpr size_t kBuckets = 256;
Bucket g_buckets[kBuckets];

Bucket &bucket_for(const void *key) {
    // Futex.zig:207-230 hashes the address with a Fibonacci multiplier.
    uint64_t h = (uint64_t)(uintptr_t)key * 0x9E3779B97F4A7C15ull;
    return g_buckets[h >> 56];
}

bool key_equals(const void *key, size_t key_size, uint64_t expect) {
    // Futex.zig:419-449
    switch (key_size) {
        case 1: return __atomic_load_n((const uint8_t *)key, __ATOMIC_ACQUIRE) == (uint8_t)expect;
        case 2: return __atomic_load_n((const uint16_t *)key, __ATOMIC_ACQUIRE) == (uint16_t)expect;
        case 4: return __atomic_load_n((const uint32_t *)key, __ATOMIC_ACQUIRE) == (uint32_t)expect;
        case 8: return __atomic_load_n((const uint64_t *)key, __ATOMIC_ACQUIRE) == expect;
        default: return false;
    }
}

template <typename T> bool cmp(unsigned op, T a, T b) {
    switch (op) {
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_EQ: return a == b;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_NE: return a != b;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LT: return a < b;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LE: return a <= b;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_GT: return a > b;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_GE: return a >= b;
        default: return false;
    }
}

template <typename T> bool check_typed(const FimoTasksMeta_FutexFilter &f, size_t masked) {
    const bool tok_deref = (f.op & 1u) != 0;
    const unsigned cmp_op = (unsigned)((f.op >> 3) & 7u);
    const bool arg_deref = ((f.op >> 6) & 1u) != 0;
    const T val = tok_deref ? *(const T *)(uintptr_t)masked : (T)masked;
    const T arg = arg_deref ? *(const T *)(uintptr_t)f.cmp_arg : (T)f.cmp_arg;
    return cmp<T>(cmp_op, val, arg);
}

// sync/Futex.zig:60-114 `checkToken`
bool check_token(const FimoTasksMeta_FutexFilter &f, size_t token) {
    const size_t masked = token & f.token_mask;
    switch ((f.op >> 1) & 3u) {
        case FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U8: return check_typed<uint8_t>(f, masked);
        case FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U16: return check_typed<uint16_t>(f, masked);
        case FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U32: return check_typed<uint32_t>(f, masked);
        default: return check_typed<uint64_t>(f, masked);
    }
}

} // namespace

FimoTasksMeta_FutexStatus futex_wait(const void *key, size_t key_size, uint64_t expect, size_t token,
                                     const FimoInstant *timeout) {
    if (key_size != 1 && key_size != 2 && key_size != 4 && key_size != 8) return FIMO_TASKS_META_FUTEX_STATUS_KEY_ERROR;
    Bucket &b = bucket_for(key);
    auto w = std::make_shared<Waiter>();
    w->key = key;
    w->token = token;
    std::unique_lock<std::mutex> g(b.m);
    if (!key_equals(key, key_size, expect)) return FIMO_TASKS_META_FUTEX_STATUS_INVALID;
    b.waiters.push_back(w);
    pool_blocking_enter();
    bool timed_out = false;
    if (timeout) {
        // FimoInstant counts from the steady clock's epoch (pkgs/fimo_std/src/time.zig).
        auto deadline = std::chrono::steady_clock::time_point(std::chrono::seconds(timeout->secs) +
                                                              std::chrono::nanoseconds(timeout->nanos));
        timed_out = !w->cv.wait_until(g, deadline, [&] { return w->woken; });
    } else {
        w->cv.wait(g, [&] { return w->woken; });
    }
    if (timed_out) {
        for (auto it = b.waiters.begin(); it != b.waiters.end(); ++it)
            if (it->get() == w.get()) {
                b.waiters.erase(it);
                break;
            }
    }
    g.unlock();
    pool_blocking_exit();
    return timed_out ? FIMO_TASKS_META_FUTEX_STATUS_TIMEOUT : FIMO_TASKS_META_FUTEX_STATUS_OK;
}

bool futex_wait_async(const void *key, size_t key_size, uint64_t expect, size_t token, std::function<void()> fn) {
    Bucket &b = bucket_for(key);
    auto w = std::make_shared<Waiter>();
    w->key = key;
    w->token = token;
    w->callback = std::move(fn);
    std::lock_guard<std::mutex> g(b.m);
    if (!key_equals(key, key_size, expect)) return false;
    b.waiters.push_back(std::move(w));
    return true;
}

size_t futex_wake(const void *key, size_t max_waiters, FimoTasksMeta_FutexFilter filter) {
    Bucket &b = bucket_for(key);
    std::vector<std::shared_ptr<Waiter>> callbacks;
    size_t woken = 0;
    {
        std::lock_guard<std::mutex> g(b.m);
        for (auto it = b.waiters.begin(); it != b.waiters.end() && woken < max_waiters;) {
            Waiter &w = **it;
            if (w.key == key && check_token(filter, w.token)) {
                w.woken = true;
                if (w.callback) callbacks.push_back(*it);
                else w.cv.notify_one();
                it = b.waiters.erase(it);
                ++woken;
            } else {
                ++it;
            }
        }
    }
    for (auto &w : callbacks) w->callback(); // outside the bucket lock: callbacks may re-enter the futex
    return woken;
}

void fence_then(FimoWorldsMeta_Fence *f, std::function<void()> fn) {
    // Continuation form of Fence.wait (Job.zig:33-50): mark contended, park, re-check on wake.
    for (;;) {
        uint8_t state = __atomic_load_n(&f->state, __ATOMIC_ACQUIRE);
        if (state & FIMO_WORLDS_META_FENCE_SIGNALED) {
            fn();
            return;
        }
        if (!(state & FIMO_WORLDS_META_FENCE_CONTENDED)) {
            uint8_t expected = FIMO_WORLDS_META_FENCE_UNSIGNALED;
            if (!__atomic_compare_exchange_n(&f->state, &expected, FIMO_WORLDS_META_FENCE_CONTENDED, true,
                                             __ATOMIC_RELAXED, __ATOMIC_RELAXED))
                continue;
        }
        auto again = [f, fn]() { fence_then(f, fn); };
        if (futex_wait_async(&f->state, 1, FIMO_WORLDS_META_FENCE_CONTENDED, 0, again)) return;
    }
}

namespace {
thread_local const char *t_last_error = "";
}
void set_last_error(const char *name) { t_last_error = name; }

} // namespace oracle

extern "C" {
FimoTasksMeta_FutexStatus fimo_tasks_futex_wait(const void *key, FimoUSize key_size, FimoU64 expect, FimoUSize token,
                                                const FimoInstant *timeout) {
    return oracle::futex_wait(key, key_size, expect, token, timeout);
}
FimoUSize fimo_tasks_futex_wake(const void *key, FimoUSize max_waiters, FimoTasksMeta_FutexFilter filter) {
    return oracle::futex_wake(key, max_waiters, filter);
}
void fimo_worlds_b200_bind_tasks(FimoTasksMeta_futex_wait, FimoTasksMeta_futex_wake) {
    // The oracle always uses its own futex.
}
void fimo_worlds_b200_fence_wait(FimoWorldsMeta_Fence *fence) { oracle::fence_wait(fence); }
void fimo_worlds_b200_fence_signal(FimoWorldsMeta_Fence *fence) { oracle::fence_signal(fence); }
const char *fimo_worlds_b200_last_error(void) { return oracle::t_last_error; }
}
