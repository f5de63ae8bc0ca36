// futex.cpp — the `fimo-tasks` futex symbols the system-group step imports
// (pkgs/fimo_tasks_meta/src/symbols.zig:132-182), as a standalone address-keyed parking lot.
//
// Contract followed: modules/fimo_tasks/src/Futex.zig
//   wait       :451-559  compare *key against `expect` under the bucket lock, then park with a token
//   wakeFilter :561-623  wake up to `max_waiters` parked on `key` whose token passes the filter
// Filter evaluation: pkgs/fimo_tasks_meta/src/sync/Futex.zig:37-114 (mask, optional deref, typed compare).
// A host application that runs the real fimo_tasks module replaces both entry points with
// `fimo_worlds_b200_bind_tasks` so that its fibers park through their own pool.
#include <chrono>

#include "rt.h"

namespace fb {
namespace {

// Waiters live on the stack of the parked thread and are linked into their bucket's list.
struct Parked {
    const void *key;
    size_t token;
    Parked *next = nullptr;
    bool released = false;
    std::condition_variable cv;
};

struct Slot {
    std::mutex lock;
    Parked *head = nullptr;
    Parked *tail = nullptr;

    void append(Parked *p) {
        p->next = nullptr;
        if (tail) tail->next = p;
        else head = p;
        tail = p;
    }
    void unlink(Parked *p) {
        Parked *prev = nullptr;
        for (Parked *it = head; it; prev = it, it = it->next) {
            if (it != p) continue;
            if (prev) prev->next = it->next;
            else head = it->next;
            if (tail == it) tail = prev;
            return;
        }
    }
};

constexpr unsigned kSlotBits = 9;
Slot g_slots[1u << kSlotBits];

Slot &slot_of(const void *key) {
    const uint64_t mixed = ((uint64_t)(uintptr_t)key >> 0) * 0x9E3779B97F4A7C15ull;
    return g_slots[mixed >> (64 - kSlotBits)];
}

bool key_matches(const void *key, size_t key_size, uint64_t expect) {
    switch (key_size) {
        case 1: return __atomic_load_n((const uint8_t *)key, __ATOMIC_ACQUIRE) == (uint8_t)expect;
        case 2: return __atomic_load_n((const uint16_t *)key, __ATOMIC_ACQUIRE) == (uint16_t)expect;
        case 4: return __atomic_load_n((const uint32_t *)key, __ATOMIC_ACQUIRE) == (uint32_t)expect;
        default: return __atomic_load_n((const uint64_t *)key, __ATOMIC_ACQUIRE) == expect;
    }
}

template <typename T> bool typed_pass(const FimoTasksMeta_FutexFilter &f, size_t masked) {
    const T lhs = (f.op & 1u) ? *(const T *)(uintptr_t)masked : (T)masked;
    const T rhs = ((f.op >> 6) & 1u) ? *(const T *)(uintptr_t)f.cmp_arg : (T)f.cmp_arg;
    switch ((f.op >> 3) & 7u) {
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_EQ: return lhs == rhs;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_NE: return lhs != rhs;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LT: return lhs < rhs;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LE: return lhs <= rhs;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_GT: return lhs > rhs;
        case FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_GE: return lhs >= rhs;
        default: return false;
    }
}

bool token_passes(const FimoTasksMeta_FutexFilter &f, size_t token) {
    const size_t masked = token & f.token_mask;
    switch ((f.op >> 1) & 3u) {
        case FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U8: return typed_pass<uint8_t>(f, masked);
        case FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U16: return typed_pass<uint16_t>(f, masked);
        case FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U32: return typed_pass<uint32_t>(f, masked);
        default: return typed_pass<uint64_t>(f, masked);
    }
}

thread_local const char *t_error = "";

} // namespace

void set_last_error(const char *name) { t_error = name; }
const char *last_error() { return t_error; }

FimoTasksMeta_FutexStatus futex_wait(const void *key, size_t key_size, uint64_t expect, size_t token,
                                     const FimoInstant *timeout) {
    if (key_size != 1 && key_size != 2 && key_size != 4 && key_size != 8) return FIMO_TASKS_META_FUTEX_STATUS_KEY_ERROR;
    Slot &slot = slot_of(key);
    Parked self{key, token};
    std::unique_lock<std::mutex> guard(slot.lock);
    if (!key_matches(key, key_size, expect)) return FIMO_TASKS_META_FUTEX_STATUS_INVALID;
    slot.append(&self);
    BlockingScope blocking; // the pool of a worker thread keeps its thread budget while we sleep
    if (!timeout) {
        self.cv.wait(guard, [&] { return self.released; });
        return FIMO_TASKS_META_FUTEX_STATUS_OK;
    }
    // FimoInstant is measured on the monotonic clock (pkgs/fimo_std/src/time.zig).
    const auto deadline = std::chrono::steady_clock::time_point(std::chrono::seconds(timeout->secs) +
                                                                std::chrono::nanoseconds(timeout->nanos));
    if (self.cv.wait_until(guard, deadline, [&] { return self.released; })) return FIMO_TASKS_META_FUTEX_STATUS_OK;
    slot.unlink(&self);
    return FIMO_TASKS_META_FUTEX_STATUS_TIMEOUT;
}

size_t futex_wake(const void *key, size_t max_waiters, FimoTasksMeta_FutexFilter filter) {
    Slot &slot = slot_of(key);
    size_t released = 0;
    std::lock_guard<std::mutex> guard(slot.lock);
    Parked *prev = nullptr;
    for (Parked *it = slot.head; it && released < max_waiters;) {
        Parked *next = it->next;
        if (it->key == key && token_passes(filter, it->token)) {
            if (prev) prev->next = next;
            else slot.head = next;
            if (slot.tail == it) slot.tail = prev;
            it->released = true;
            it->cv.notify_one(); // `it` stays valid: its owner needs slot.lock to return
            released++;
        } else {
            prev = it;
        }
        it = next;
    }
    return released;
}

FimoTasksMeta_futex_wait g_futex_wait = nullptr;
FimoTasksMeta_futex_wake g_futex_wake = nullptr;

// ---- RwLock -------------------------------------------------------------------------------------
void RwLock::lock_read() {
    std::unique_lock<std::mutex> g(m_);
    if (writer_ || writers_waiting_ > 0) {
        BlockingScope blocking;
        cv_.wait(g, [&] { return !writer_ && writers_waiting_ == 0; });
    }
    ++readers_;
}
bool RwLock::try_lock_read() {
    std::lock_guard<std::mutex> g(m_);
    if (writer_ || writers_waiting_ > 0) return false;
    ++readers_;
    return true;
}
void RwLock::unlock_read() {
    std::lock_guard<std::mutex> g(m_);
    if (--readers_ == 0) cv_.notify_all();
}
void RwLock::lock_write() {
    std::unique_lock<std::mutex> g(m_);
    if (writer_ || readers_ > 0) {
        ++writers_waiting_;
        {
            BlockingScope blocking;
            cv_.wait(g, [&] { return !writer_ && readers_ == 0; });
        }
        --writers_waiting_;
    }
    writer_ = true;
}
bool RwLock::try_lock_write() {
    std::lock_guard<std::mutex> g(m_);
    if (writer_ || readers_ > 0) return false;
    writer_ = true;
    return true;
}
void RwLock::unlock_write() {
    std::lock_guard<std::mutex> g(m_);
    writer_ = false;
    cv_.notify_all();
}

} // namespace fb

extern "C" {

FimoTasksMeta_FutexStatus fimo_tasks_futex_wait(const void *key, FimoUSize key_size, FimoU64 expect, FimoUSize token,
                                                const FimoInstant *timeout) {
    return fb::futex_wait(key, key_size, expect, token, timeout);
}
FimoUSize fimo_tasks_futex_wake(const void *key, FimoUSize max_waiters, FimoTasksMeta_FutexFilter filter) {
    return fb::futex_wake(key, max_waiters, filter);
}
void fimo_worlds_b200_bind_tasks(FimoTasksMeta_futex_wait futex_wait, FimoTasksMeta_futex_wake futex_wake) {
    fb::g_futex_wait = futex_wait ? futex_wait : &fimo_tasks_futex_wait;
    fb::g_futex_wake = futex_wake ? futex_wake : &fimo_tasks_futex_wake;
}
void fimo_worlds_b200_fence_wait(FimoWorldsMeta_Fence *fence) { fb::fence_wait(fence); }
void fimo_worlds_b200_fence_signal(FimoWorldsMeta_Fence *fence) { fb::fence_signal(fence); }
const char *fimo_worlds_b200_last_error(void) { return fb::last_error(); }

__attribute__((constructor)) static void fimo_b200_bind_default_futex(void) {
    fb::g_futex_wait = &fimo_tasks_futex_wait;
    fb::g_futex_wake = &fimo_tasks_futex_wake;
}
}
