// pool.cpp — worker pool behind the `fimo-tasks` Pool vtable (pkgs/fimo_tasks_meta/include/fimo_tasks_meta/pool.h:27-62)
// and the command-buffer entry set (command_buffer.h:44-86).
//
// What the reference does (modules/fimo_tasks/src): every entry, completion and wake-up of a pool is
// funnelled through ONE event-loop thread (Pool.zig:660-765) which interprets command buffers
// (CommandBuffer.zig:150-201) and hands tasks to worker threads that run them as stackful fibers
// (Worker.zig:250-411).  What this file does instead: there is no event-loop thread.  A command
// buffer is a small lock-protected state machine (`Cmd`) that is advanced by whichever thread makes
// progress possible — the enqueuing thread, or the worker that just finished one of its tasks — so a
// system hand-off costs one queue hop instead of two plus a serialisation point.  Entry semantics
// follow the documented meaning in command_buffer.h:56-86:
//   in-order issue; a blocked wait entry stalls every later entry (CommandBuffer.zig:156-188)
//   wait_on_command_indirect(k): until entry (index - k) has completed  (:390-411, not quirk 1 of SURVEY.md §5)
//   wait_on_barrier: until all earlier entries completed                 (:339-348)
//   enqueue_command_buffer: child buffer, completes the entry when done  (:315-337)
//   wait_on_command_buffer: until another buffer's handle completed      (:350-388)
// Tasks run on OS threads (boost.context is an un-vendored network dependency of the reference,
// modules/fimo_tasks/build.zig.zon:8-11); a task that parks on the futex keeps the pool's thread
// budget through BlockingScope.
#include <algorithm>
#include <deque>
#include <thread>

#include "rt.h"

namespace fb {

struct Cmd;

struct WorkItem {
    Cmd *cmd;
    size_t entry;
    FimoTasksMeta_Task *task;
};

struct Pool {
    size_t id = 0;
    std::string label;
    size_t worker_count = 0;
    std::atomic<size_t> refs{1};
    std::atomic<bool> closed{false};

    std::mutex m;
    std::condition_variable cv;
    std::deque<WorkItem> queue;
    size_t runnable = 0;     // threads that are not parked inside a task
    size_t idle = 0;         // runnable threads waiting for work
    size_t live_threads = 0; // regular + spare threads still alive
    size_t open_cmds = 0;    // command buffers accepted and not yet finished
    bool stopping = false;
};

namespace {

std::atomic<size_t> g_pool_ids{1};
thread_local Pool *t_pool = nullptr;

enum : uint8_t { kPending = 0, kRunning = 1, kDone = 2 };

std::mutex g_link_lock; // guards Cmd::blocked_on_me across command buffers

} // namespace

struct Cmd {
    Pool *pool = nullptr;
    FimoTasksMeta_CommandBuffer *cb = nullptr;
    std::mutex m;
    std::condition_variable cv;
    std::vector<uint8_t> state;
    size_t issue = 0;    // next entry to issue
    size_t finished = 0; // entries completed
    bool abort_on_error = false;
    bool failed = false;
    bool finalized = false;
    int outcome = 0; // 0 running, 1 completed, 2 aborted (under m)
    std::atomic<int> outcome_published{0};
    std::atomic<size_t> refs{1};
    Cmd *parent = nullptr;
    size_t parent_entry = 0;
    std::vector<Cmd *> blocked_on_me; // under g_link_lock
};

namespace {

void cmd_unref(Cmd *c) {
    if (c->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) delete c;
}

void worker_main(Pool *p, bool spare);

// Called with p->m held.
void spawn_spare_locked(Pool *p) {
    p->runnable++;
    p->live_threads++;
    std::thread(worker_main, p, true).detach();
}

void push_task(Pool *p, WorkItem w) {
    std::lock_guard<std::mutex> g(p->m);
    p->queue.push_back(w);
    if (p->idle > 0) p->cv.notify_one();
    else if (p->runnable < p->worker_count) spawn_spare_locked(p); // every regular thread is busy or parked
}

void resume(Cmd *c, ptrdiff_t done_entry, bool child_aborted);

Cmd *cmd_create(Pool *p, FimoTasksMeta_CommandBuffer *cb) {
    Cmd *c = new Cmd();
    c->pool = p;
    c->cb = cb;
    c->state.assign(cb->entries_len, kPending);
    std::lock_guard<std::mutex> g(p->m);
    p->open_cmds++;
    return c;
}

// Issues entries until one blocks.  c->m is held.  Returns true when the buffer is complete.
bool advance_locked(Cmd *c) {
    const FimoTasksMeta_CommandBuffer *cb = c->cb;
    while (c->issue < cb->entries_len) {
        const size_t idx = c->issue;
        if (c->failed) { // aborted buffers skip what was not issued yet (CommandBuffer.zig:459-513)
            c->state[idx] = kDone;
            c->finished++;
            c->issue++;
            continue;
        }
        const FimoTasksMeta_CommandBufferEntry &e = cb->entries[idx];
        bool done_now = true;
        switch (e.type) {
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR:
                c->abort_on_error = e.payload.abort_on_error;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SET_MIN_STACK_SIZE:
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_WORKER:
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_ANY_WORKER:
                break; // fiber stack sizes and worker pinning do not apply to thread-run tasks
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK:
                c->state[idx] = kRunning;
                push_task(c->pool, WorkItem{c, idx, e.payload.enqueue_task});
                done_now = false;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_COMMAND_BUFFER: {
                Cmd *child = cmd_create(c->pool, e.payload.enqueue_command_buffer);
                child->parent = c;
                child->parent_entry = idx;
                c->refs.fetch_add(1, std::memory_order_relaxed);
                c->state[idx] = kRunning;
                done_now = false;
                // The child takes its first step on a worker (null task), never under our lock.
                push_task(c->pool, WorkItem{child, (size_t)-1, nullptr});
                break;
            }
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_BARRIER:
                if (c->finished != idx) return false;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_BUFFER: {
                Cmd *other = (Cmd *)e.payload.wait_on_command_buffer.data;
                std::lock_guard<std::mutex> link(g_link_lock);
                if (other->outcome_published.load(std::memory_order_acquire) == 0) {
                    if (std::find(other->blocked_on_me.begin(), other->blocked_on_me.end(), c) == other->blocked_on_me.end()) {
                        c->refs.fetch_add(1, std::memory_order_relaxed);
                        other->blocked_on_me.push_back(c);
                    }
                    return false;
                }
                break;
            }
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_INDIRECT: {
                const size_t k = e.payload.wait_on_command_indirect;
                if (k == 0 || k > idx) panic("wait_on_command_indirect out of range");
                if (c->state[idx - k] != kDone) return false;
                break;
            }
            default: panic("unknown command buffer entry");
        }
        if (done_now) {
            c->state[idx] = kDone;
            c->finished++;
        }
        c->issue++;
    }
    if (c->finished != cb->entries_len || c->finalized) return false;
    c->finalized = true;
    return true;
}

void pool_cmd_closed(Pool *p);

void finalize(Cmd *c) {
    // CommandBuffer.zig:196-200: complete/abort hook, publish the outcome, deinit hook, wake dependants.
    FimoTasksMeta_CommandBuffer *cb = c->cb;
    const bool failed = c->failed;
    if (failed) {
        if (cb->on_abort) cb->on_abort(cb);
    } else if (cb->on_complete) {
        cb->on_complete(cb);
    }
    void (*on_deinit)(FimoTasksMeta_CommandBuffer *) = cb->on_deinit;
    std::vector<Cmd *> dependants;
    {
        std::lock_guard<std::mutex> g(c->m);
        c->outcome = failed ? 2 : 1;
        std::lock_guard<std::mutex> link(g_link_lock);
        c->outcome_published.store(c->outcome, std::memory_order_release);
        dependants.swap(c->blocked_on_me);
    }
    c->cv.notify_all();
    if (on_deinit) on_deinit(cb); // `cb` may be gone after this line
    for (Cmd *d : dependants) {
        resume(d, -1, false);
        cmd_unref(d);
    }
    Pool *p = c->pool;
    if (Cmd *parent = c->parent) {
        resume(parent, (ptrdiff_t)c->parent_entry, failed);
        cmd_unref(parent);
    }
    cmd_unref(c);
    pool_cmd_closed(p);
}

void resume(Cmd *c, ptrdiff_t done_entry, bool child_aborted) {
    bool complete;
    {
        std::lock_guard<std::mutex> g(c->m);
        if (done_entry >= 0) {
            c->state[(size_t)done_entry] = kDone;
            c->finished++;
            if (child_aborted && c->abort_on_error) c->failed = true;
        }
        complete = advance_locked(c);
    }
    if (complete) finalize(c);
}

void pool_destroy_if_last_thread_locked(Pool *p, std::unique_lock<std::mutex> &g) {
    const bool last = --p->live_threads == 0;
    g.unlock();
    if (last) delete p;
}

void worker_main(Pool *p, bool spare) {
    t_pool = p;
    std::unique_lock<std::mutex> g(p->m);
    for (;;) {
        if (spare && p->runnable > p->worker_count) break; // parked threads came back: retire
        if (p->queue.empty()) {
            if (p->stopping || spare) break;
            p->idle++;
            p->cv.wait(g);
            p->idle--;
            continue;
        }
        WorkItem w = p->queue.front();
        p->queue.pop_front();
        g.unlock();
        if (w.task) {
            // Task.zig:56-67, Worker.zig:225-248,306-321: body, completion hooks, completion report.
            w.task->on_start(w.task);
            if (w.task->on_complete) w.task->on_complete(w.task);
            if (w.task->on_deinit) w.task->on_deinit(w.task);
            resume(w.cmd, (ptrdiff_t)w.entry, false);
        } else {
            resume(w.cmd, -1, false); // first advance of a child command buffer
        }
        g.lock();
    }
    p->runnable--;
    pool_destroy_if_last_thread_locked(p, g);
}

void maybe_stop_locked(Pool *p) {
    if (p->refs.load(std::memory_order_acquire) == 0 && p->open_cmds == 0 && !p->stopping) {
        p->stopping = true;
        p->cv.notify_all();
    }
}

void pool_cmd_closed(Pool *p) {
    std::lock_guard<std::mutex> g(p->m);
    p->open_cmds--;
    maybe_stop_locked(p);
}

// ---- vtables ------------------------------------------------------------------------------------
FimoTasksMeta_PoolId vt_id(void *d) { return ((Pool *)d)->id; }
void vt_acquire(void *d) { ((Pool *)d)->refs.fetch_add(1, std::memory_order_relaxed); }
void vt_release(void *d) {
    Pool *p = (Pool *)d;
    std::lock_guard<std::mutex> g(p->m); // the workers cannot delete `p` while we hold its lock
    if (p->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) {
        p->closed.store(true);
        maybe_stop_locked(p);
    }
}
void vt_request_close(void *d) { ((Pool *)d)->closed.store(true); }
bool vt_accepts_requests(void *d) { return !((Pool *)d)->closed.load(); }
bool vt_owns_current_thread(void *d) { return t_pool == (Pool *)d; }
const char *vt_label(void *d, FimoUSize *len) {
    Pool *p = (Pool *)d;
    *len = p->label.size();
    return p->label.data();
}
FimoUSize vt_workers(void *d, FimoTasksMeta_PoolWorker *ptr, FimoUSize len) {
    Pool *p = (Pool *)d;
    for (size_t i = 0; ptr && i < len && i < p->worker_count; i++) ptr[i] = i;
    return p->worker_count;
}
FimoUSize vt_stack_sizes(void *, FimoTasksMeta_PoolStackSize *ptr, FimoUSize len) {
    if (ptr && len > 0) ptr[0] = (size_t)8 << 20; // FimoTasks.zig:32 default_stack_size
    return 1;
}

void h_acquire(void *d) { ((Cmd *)d)->refs.fetch_add(1, std::memory_order_relaxed); }
void h_release(void *d) { cmd_unref((Cmd *)d); }
FimoTasksMeta_Pool h_owner_pool(void *d) {
    Pool *p = ((Cmd *)d)->pool;
    vt_acquire(p);
    return pool_handle(p);
}
FimoTasksMeta_CommandBufferHandleCompletionStatus h_wait_on(void *d) {
    // CommandBuffer.zig:603-631
    Cmd *c = (Cmd *)d;
    std::unique_lock<std::mutex> g(c->m);
    if (c->outcome == 0) {
        BlockingScope blocking;
        c->cv.wait(g, [&] { return c->outcome != 0; });
    }
    return c->outcome == 1 ? FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_COMPLETED
                           : FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_ABORTED;
}
const FimoTasksMeta_CommandBufferHandleVTable g_handle_vtable = {h_acquire, h_release, h_owner_pool, h_wait_on};

FimoStatus vt_enqueue_buffer(void *d, FimoTasksMeta_CommandBuffer *buffer, FimoTasksMeta_CommandBufferHandle *handle) {
    // Pool.zig:820-842
    Pool *p = (Pool *)d;
    if (p->closed.load()) {
        set_last_error("Closed");
        return FIMO_STATUS_ERROR;
    }
    Cmd *c = cmd_create(p, buffer);
    if (handle) {
        c->refs.fetch_add(1, std::memory_order_relaxed);
        handle->data = c;
        handle->vtable = &g_handle_vtable;
    }
    resume(c, -1, false);
    return FIMO_STATUS_OK;
}

const FimoTasksMeta_PoolVTable g_pool_vtable = {vt_id,    vt_acquire, vt_release,     vt_request_close, vt_accepts_requests,
                                                vt_owns_current_thread, vt_label, vt_workers, vt_stack_sizes,
                                                vt_enqueue_buffer};

} // namespace

FimoTasksMeta_Pool pool_handle(Pool *p) { return FimoTasksMeta_Pool{p, &g_pool_vtable}; }

Pool *pool_create(const char *label, size_t label_len, size_t workers) {
    Pool *p = new Pool();
    p->id = g_pool_ids.fetch_add(1);
    p->label = label && label_len ? std::string(label, label_len) : std::string("<unlabelled>");
    if (workers == 0) workers = std::thread::hardware_concurrency(); // FimoTasks.zig:127-132
    if (workers == 0) workers = 1;
    p->worker_count = workers;
    p->runnable = workers;
    p->live_threads = workers;
    for (size_t i = 0; i < workers; i++) std::thread(worker_main, p, false).detach();
    return p;
}

BlockingScope::BlockingScope() {
    Pool *p = t_pool;
    if (!p) return;
    std::lock_guard<std::mutex> g(p->m);
    p->runnable--;
    if (!p->queue.empty() && p->idle == 0 && p->runnable < p->worker_count) spawn_spare_locked(p);
}
BlockingScope::~BlockingScope() {
    Pool *p = t_pool;
    if (!p) return;
    std::lock_guard<std::mutex> g(p->m);
    p->runnable++;
}

} // namespace fb

extern "C" FimoStatus fimo_tasks_create_worker_pool(const FimoTasksMeta_PoolConfig *config, FimoTasksMeta_Pool *pool) {
    // FimoTasks.zig:201-358: a non-null `next` is rejected (:208-212); label and worker_count are honoured.
    if (config->next != nullptr) {
        fb::set_last_error("InvalidConfig");
        return FIMO_STATUS_ERROR;
    }
    *pool = fb::pool_handle(fb::pool_create(config->label, config->label_len, config->worker_count));
    return FIMO_STATUS_OK;
}
