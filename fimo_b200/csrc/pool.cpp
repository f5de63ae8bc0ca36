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
//   select_worker / select_any_worker: the following tasks run on that worker's thread only (:236-268; an index
//       outside the pool aborts the entry); set_min_stack_size is accepted and has no effect (tasks own a whole thread)
//   abort_on_error + errors (:459-513): a task errors by calling the `abort` symbol (Worker.zig:115-122,
//       symbols.zig:97); with the setting on, every entry not yet issued is skipped — its task / child buffer gets its
//       on_abort and on_deinit hooks — up to an abort_on_error=false entry, and the buffer ends aborted; an aborted
//       child buffer or awaited handle propagates the same way; malformed entries (offset out of range, unknown worker
//       or type) abort instead of panicking; wait_on_command_indirect(0) is a no-op (:397).
// Tasks run on OS threads (boost.context is an un-vendored network dependency of the reference,
// modules/fimo_tasks/build.zig.zon:8-11); a task that parks on the futex keeps the pool's thread
// budget through BlockingScope, and while it is parked a stand-in thread serves the tasks pinned to its worker.
#include <algorithm>
#include <csetjmp>
#include <deque>
#include <thread>

#include "rt.h"

namespace fb {

struct Cmd;

struct WorkItem {
    Cmd *cmd;
    size_t entry;
    FimoTasksMeta_Task *task;
    ptrdiff_t worker = -1; // >= 0: only that worker thread may run it (select_worker)
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
    std::vector<std::deque<WorkItem>> pinned; // per regular worker: tasks enqueued under select_worker
    // Threads currently serving worker index i and not parked inside a task: 1 while the regular thread runs or idles,
    // 0 while it is parked (BlockingScope) — then a stand-in spare adopts the index (worker_main) so that tasks pinned
    // to i keep running, as they do in the reference where only the fiber parks (Worker.zig:58-76,262-280).
    std::vector<size_t> active;
    size_t runnable = 0;     // threads that are not parked inside a task
    size_t idle = 0;         // runnable threads waiting for work
    size_t live_threads = 0; // regular + spare threads still alive
    size_t open_cmds = 0;    // command buffers accepted and not yet finished
    bool stopping = false;
};

namespace {

std::atomic<size_t> g_pool_ids{1};
thread_local Pool *t_pool = nullptr;
thread_local ptrdiff_t t_worker = -1;       // index of this regular worker thread in its pool (-1: spare / foreign thread)
thread_local std::jmp_buf *t_abort = nullptr; // where `abort` unwinds to while a task body runs

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
    bool failed = false; // has_error in the reference
    ptrdiff_t worker = -1; // select_worker state
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

void worker_main(Pool *p, bool spare, ptrdiff_t index);

// Called with p->m held.
void spawn_spare_locked(Pool *p) {
    p->runnable++;
    p->live_threads++;
    std::thread(worker_main, p, true, (ptrdiff_t)-1).detach();
}

// Called with p->m held: worker `index` has nobody running (its thread is parked inside a task) and has pinned work.
void spawn_stand_in_locked(Pool *p, size_t index) {
    p->active[index]++; // counted from here, so that nobody spawns a second one before this one has started
    p->runnable++;
    p->live_threads++;
    std::thread(worker_main, p, true, (ptrdiff_t)index).detach();
}

void push_task(Pool *p, WorkItem w) {
    std::lock_guard<std::mutex> g(p->m);
    if (w.worker >= 0) { // pinned: only a thread serving that worker index takes it
        p->pinned[(size_t)w.worker].push_back(w);
        if (p->active[(size_t)w.worker] == 0) spawn_stand_in_locked(p, (size_t)w.worker);
        p->cv.notify_all();
        return;
    }
    p->queue.push_back(w);
    if (p->idle > 0) p->cv.notify_one();
    else if (p->runnable < p->worker_count) spawn_spare_locked(p); // every regular thread is busy or parked
}

void resume(Cmd *c, ptrdiff_t done_entry, bool child_aborted);
extern const FimoTasksMeta_CommandBufferHandleVTable g_handle_vtable;

Cmd *cmd_create(Pool *p, FimoTasksMeta_CommandBuffer *cb) {
    Cmd *c = new Cmd();
    c->pool = p;
    c->cb = cb;
    c->state.assign(cb->entries_len, kPending);
    std::lock_guard<std::mutex> g(p->m);
    p->open_cmds++;
    return c;
}

// Entry.abort (pkgs/fimo_tasks_meta/src/command_buffer.zig:56-70) + Entry.deinit: an entry that will never be issued.
void abort_entry(const FimoTasksMeta_CommandBufferEntry &e) {
    if (e.type == FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK) {
        FimoTasksMeta_Task *t = e.payload.enqueue_task;
        if (t->on_abort) t->on_abort(t);
        if (t->on_deinit) t->on_deinit(t);
    } else if (e.type == FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_COMMAND_BUFFER) {
        FimoTasksMeta_CommandBuffer *b = e.payload.enqueue_command_buffer;
        if (b->on_abort) b->on_abort(b);
        if (b->on_deinit) b->on_deinit(b);
    }
}
void mark_processed(Cmd *c) {
    c->state[c->issue] = kDone;
    c->finished++;
    c->issue++;
}
// CommandBuffer.zig:490-513.  c->m held.
void abort_forward(Cmd *c) {
    if (!c->abort_on_error) return;
    const FimoTasksMeta_CommandBuffer *cb = c->cb;
    while (c->issue < cb->entries_len) {
        const FimoTasksMeta_CommandBufferEntry &e = cb->entries[c->issue];
        if (e.type == FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR) {
            c->abort_on_error = e.payload.abort_on_error;
            mark_processed(c);
            if (!c->abort_on_error) break;
        } else {
            abort_entry(e);
            mark_processed(c);
        }
    }
    c->failed = c->abort_on_error;
}
// :459-466: the entry being issued is malformed.
void abort_current_and_forward(Cmd *c) {
    abort_entry(c->cb->entries[c->issue]);
    mark_processed(c);
    abort_forward(c);
}
// :468-488: entry `index` (already issued) ended in error.
void abort_from_index(Cmd *c, size_t index) {
    if (!c->abort_on_error) return;
    const FimoTasksMeta_CommandBuffer *cb = c->cb;
    if (c->issue == cb->entries_len) {
        c->failed = true;
        return;
    }
    for (size_t i = index; i < c->issue; i++) // did the setting change between that entry and now?
        if (cb->entries[i].type == FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR && !cb->entries[i].payload.abort_on_error) return;
    abort_forward(c);
}

// Issues entries until one blocks.  c->m is held.  Returns true when the buffer is complete.
bool advance_locked(Cmd *c) {
    const FimoTasksMeta_CommandBuffer *cb = c->cb;
    while (c->issue < cb->entries_len) {
        const size_t idx = c->issue;
        const FimoTasksMeta_CommandBufferEntry &e = cb->entries[idx];
        bool done_now = true;
        switch (e.type) {
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR:
                c->abort_on_error = e.payload.abort_on_error;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SET_MIN_STACK_SIZE:
                break; // a task owns a whole OS thread here: every size is honoured
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_WORKER:
                if (e.payload.select_worker >= c->pool->worker_count) { // CommandBuffer.zig:243-254
                    abort_current_and_forward(c);
                    continue;
                }
                c->worker = (ptrdiff_t)e.payload.select_worker;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_ANY_WORKER:
                c->worker = -1;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK:
                c->state[idx] = kRunning;
                push_task(c->pool, WorkItem{c, idx, e.payload.enqueue_task, c->worker});
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
                if (e.payload.wait_on_command_buffer.vtable != &g_handle_vtable || other->pool != c->pool) { // :357-381
                    abort_current_and_forward(c);
                    continue;
                }
                int outcome;
                {
                    std::lock_guard<std::mutex> link(g_link_lock);
                    outcome = other->outcome_published.load(std::memory_order_acquire);
                    if (outcome == 0) {
                        if (std::find(other->blocked_on_me.begin(), other->blocked_on_me.end(), c) == other->blocked_on_me.end()) {
                            c->refs.fetch_add(1, std::memory_order_relaxed);
                            other->blocked_on_me.push_back(c);
                        }
                        return false;
                    }
                }
                mark_processed(c);
                if (outcome == 2) abort_forward(c); // :385-387
                continue;
            }
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_INDIRECT: {
                const size_t k = e.payload.wait_on_command_indirect;
                if (k == 0) break; // :397
                if (k > idx) {     // :398-408
                    abort_current_and_forward(c);
                    continue;
                }
                if (c->state[idx - k] != kDone) return false;
                break;
            }
            default: // :413-421
                abort_current_and_forward(c);
                continue;
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
        if (done_entry >= 0) { // completeTask (CommandBuffer.zig:127-148): a task or child buffer has ended, possibly in error
            c->state[(size_t)done_entry] = kDone;
            c->finished++;
            if (child_aborted) abort_from_index(c, (size_t)done_entry);
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

void worker_main(Pool *p, bool spare, ptrdiff_t index) {
    t_pool = p;
    t_worker = index;
    std::unique_lock<std::mutex> g(p->m);
    // A spare with an index is a stand-in: it serves ONLY the private queue of a worker whose thread is parked inside a
    // task (fimo_tasks_worker_id reports that index), and leaves as soon as a parked thread of that index is back or
    // the queue is empty.  The returning thread does not wait for it: for the rest of the stand-in's current task two
    // tasks of one worker index run side by side, which fibers on one worker never do.
    const bool stand_in = spare && index >= 0;
    for (;;) {
        if (stand_in) {
            if (p->active[(size_t)index] > 1 || p->pinned[(size_t)index].empty()) break;
        } else if (spare && p->runnable > p->worker_count) {
            break; // parked threads came back: retire
        }
        std::deque<WorkItem> *from = index >= 0 && !p->pinned[(size_t)index].empty() ? &p->pinned[(size_t)index] : &p->queue;
        if (from->empty()) {
            if (p->stopping || spare) break;
            p->idle++;
            p->cv.wait(g);
            p->idle--;
            continue;
        }
        WorkItem w = from->front();
        from->pop_front();
        g.unlock();
        if (w.task) {
            // Task.zig:56-67, Worker.zig:225-248,306-335: body, then the completion OR the abort hook (the task called
            // the `abort` symbol: fimo_tasks_abort unwinds to here), deinit, completion report with is_error.
            std::jmp_buf unwind;
            t_abort = &unwind;
            volatile bool is_error = false;
            if (setjmp(unwind) == 0) w.task->on_start(w.task);
            else is_error = true;
            t_abort = nullptr;
            if (is_error) {
                if (w.task->on_abort) w.task->on_abort(w.task);
            } else if (w.task->on_complete) {
                w.task->on_complete(w.task);
            }
            if (w.task->on_deinit) w.task->on_deinit(w.task);
            resume(w.cmd, (ptrdiff_t)w.entry, is_error);
        } else {
            resume(w.cmd, -1, false); // first advance of a child command buffer
        }
        g.lock();
    }
    if (stand_in) p->active[(size_t)index]--;
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
extern const FimoTasksMeta_CommandBufferHandleVTable g_handle_vtable = {h_acquire, h_release, h_owner_pool, h_wait_on};

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
    p->pinned.resize(workers);
    p->active.assign(workers, 1);
    for (size_t i = 0; i < workers; i++) std::thread(worker_main, p, false, (ptrdiff_t)i).detach();
    return p;
}

BlockingScope::BlockingScope() {
    Pool *p = t_pool;
    if (!p) return;
    std::lock_guard<std::mutex> g(p->m);
    p->runnable--;
    if (t_worker >= 0) { // this worker index is about to have nobody running: its pinned tasks go to a stand-in
        const size_t i = (size_t)t_worker;
        if (--p->active[i] == 0 && !p->pinned[i].empty()) spawn_stand_in_locked(p, i);
    }
    if (!p->queue.empty() && p->idle == 0 && p->runnable < p->worker_count) spawn_spare_locked(p);
}
BlockingScope::~BlockingScope() {
    Pool *p = t_pool;
    if (!p) return;
    std::lock_guard<std::mutex> g(p->m);
    p->runnable++;
    if (t_worker >= 0) p->active[(size_t)t_worker]++; // a stand-in, if any, leaves after its current task
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

// `abort` of fimo-tasks (symbols.zig:94-99; Worker.zig:115-122): ends the calling task in error.  Never returns.
extern "C" void fimo_tasks_abort(void) {
    if (!fb::t_abort) fb::panic("fimo_tasks_abort called outside a task");
    std::longjmp(*fb::t_abort, 1);
}
// `worker_id` (symbols.zig:60-65): the index of the calling worker thread in its pool.
extern "C" bool fimo_tasks_worker_id(FimoTasksMeta_PoolWorker *id) {
    if (fb::t_worker < 0) return false;
    *id = (FimoTasksMeta_PoolWorker)fb::t_worker;
    return true;
}
