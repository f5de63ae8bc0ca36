// oracle_pool.cpp — CPU restatement of the fimo_tasks dispatch path (TEST INFRASTRUCTURE ONLY).
//
// Follows the structure of modules/fimo_tasks/src:
//   Pool.zig:275-294,660-765   one event-loop thread + N worker threads per pool; every command
//                              buffer entry, completion and wake-up is processed on the event loop
//   CommandBuffer.zig:150-201  in-order issue of entries; a blocked entry stalls all later entries
//   CommandBuffer.zig:272-313  enqueue_task     :339-348 wait_on_barrier
//   CommandBuffer.zig:390-411  wait_on_command_indirect (documented meaning, command_buffer.h:83-85;
//                              NOT the double markStatus of SURVEY.md §5 quirk 1)
//   CommandBuffer.zig:127-148,445-457  completeTask / progressCompleted
//   CommandBuffer.zig:236-268  select_worker / select_any_worker (private queue per worker, Worker.zig:58-76)
//   CommandBuffer.zig:459-513  abortCurrentAndForward / abortFromIndex / abortForward;
//                              command_buffer.zig:56-70 Entry.abort; Worker.zig:115-122 abortTask (`abort` symbol)
//   Worker.zig:250-411         worker run loop (task fetch, run, report completion)
//   Worker.zig:225-248,284-321 taskEntry / switch to the task's context and back; Task.zig:56-67 ensureReady;
//   Pool.zig:100-170           StackAllocator (cached stacks); context.zig:67-100,477-535 make/jump context, mmap +
//                              guard page
// Every task runs as a STACKFUL FIBER, as in the reference: a stack from the pool's cache (8 MiB + one guard page,
// mmap'd and mprotect'ed once, FimoTasks.zig:32), a fresh context made on it, one switch in and one switch out
// (boost.context's fcontext is an un-vendored network dependency, modules/fimo_tasks/build.zig.zon:8-11; the two
// routines below restate its x86-64 SysV switch: callee-saved registers + MXCSR / x87 control word + stack pointer).
// `abort` switches away from the fiber for good (Worker.zig:115-122).  Difference that remains, stated once: a task
// that BLOCKS (futex wait) blocks its OS thread instead of parking the fiber; the pool compensates by keeping
// `worker_count` runnable threads (pool_blocking_enter/exit), which is what a parked fiber achieves in the reference.
#include <sys/mman.h>
#include <unistd.h>

#include <algorithm>
#include <deque>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "oracle_internal.h"

// ---- stackful fibers (context.zig:67-100: make_fcontext / jump_fcontext) ---------------------------------------
// fiber_switch(&save_sp, load_sp): pushes the callee-saved state of the running context, stores its stack pointer,
// adopts `load_sp` and pops that context's state.  A new fiber's stack is primed so that the first switch "returns"
// into fiber_trampoline with the Fiber* in r12.
extern "C" void oracle_fiber_switch(void **save_sp, void *load_sp);
extern "C" void oracle_fiber_trampoline();
asm(R"(
    .text
    .globl oracle_fiber_switch
    .type oracle_fiber_switch, @function
oracle_fiber_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    subq $8, %rsp
    stmxcsr (%rsp)
    fnstcw 4(%rsp)
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    ldmxcsr (%rsp)
    fldcw 4(%rsp)
    addq $8, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
    .size oracle_fiber_switch, .-oracle_fiber_switch
    .globl oracle_fiber_trampoline
    .type oracle_fiber_trampoline, @function
oracle_fiber_trampoline:
    movq %r12, %rdi
    call oracle_fiber_main
    ud2
    .size oracle_fiber_trampoline, .-oracle_fiber_trampoline
)");

namespace oracle {
struct Stack { // context.zig:477-535: [guard page | usable]; grows down from memory + size
    void *memory = nullptr;
    size_t size = 0;
};
constexpr size_t kStackSize = 8u << 20; // FimoTasks.zig:32 default_stack_size

struct Fiber {
    FimoTasksMeta_Task *task = nullptr;
    void *sp = nullptr;        // the fiber's saved stack pointer while it is switched out
    void *worker_sp = nullptr; // the worker's, while the fiber runs
    bool aborted = false;
};

thread_local Fiber *t_fiber = nullptr; // the fiber running on this thread
} // namespace oracle

extern "C" void oracle_fiber_main(oracle::Fiber *f) { // Worker.taskEntry (:225-248)
    f->task->on_start(f->task);
    void *dead;
    oracle_fiber_switch(&dead, f->worker_sp); // completeTask: back to the worker for good
    __builtin_unreachable();
}

namespace oracle {

namespace {
std::atomic<size_t> g_next_pool_id{1};
thread_local Pool *t_pool = nullptr;      // pool owning this thread (worker or event loop)
thread_local bool t_is_worker = false;    // true on worker threads only
thread_local bool t_is_spare = false;     // compensation thread
thread_local ptrdiff_t t_worker = -1;     // index of a regular worker thread

} // namespace

struct RtBuffer;

enum class EntryStatus : uint8_t { not_processed, running_task, running_buffer, processed };

struct RtBuffer {
    Pool *owner = nullptr;
    FimoTasksMeta_CommandBuffer *buffer = nullptr;
    std::vector<EntryStatus> status;
    size_t entry_index = 0;     // next entry to issue
    size_t completed_index = 0; // all entries below are processed
    bool abort_on_error = false;
    bool has_error = false;
    ptrdiff_t worker = -1; // select_worker state (CommandBuffer.zig:34)
    bool in_process_list = false;
    RtBuffer *parent = nullptr; // set for enqueue_command_buffer children
    size_t parent_entry = 0;
    std::vector<RtBuffer *> buffer_waiters; // buffers blocked in wait_on_command_buffer on us

    std::atomic<size_t> refs{1};
    std::mutex m; // guards `done` for Handle.waitOn
    std::condition_variable cv;
    int done = 0; // 0 running, 1 completed, 2 aborted
};

struct Msg {
    enum Kind { enqueue, task_done, wake } kind;
    RtBuffer *rb;
    size_t entry;
    bool is_error = false;
};

struct WorkItem {
    RtBuffer *rb;
    size_t entry;
    FimoTasksMeta_Task *task;
};

struct Pool {
    size_t id;
    std::string label;
    size_t worker_count;
    std::atomic<size_t> refs{1};
    std::atomic<bool> closed{false};

    // event loop mailbox (Pool.zig:386-401 IntrusiveMpsc + futex wake)
    std::mutex mbox_m;
    std::condition_variable mbox_cv;
    std::deque<Msg> mbox;
    size_t active_buffers = 0; // event-loop thread only
    std::deque<RtBuffer *> process_list;

    // global task queue (Pool.zig:417-430 SPMC)
    std::mutex q_m;
    std::condition_variable q_cv;
    std::deque<WorkItem> queue;
    std::vector<std::deque<WorkItem>> private_queue; // Worker.zig:58-76, one per regular worker
    std::vector<Stack> stack_cache; // Pool.zig:100-170 StackAllocator free list (guarded by q_m)
    bool stop_workers = false;
    size_t runnable = 0;       // threads currently able to run tasks
    size_t threads_total = 0;  // live worker + spare threads
    std::vector<std::thread> threads;

    std::thread event_thread;
};

static void post(Pool *p, Msg m) {
    {
        std::lock_guard<std::mutex> g(p->mbox_m);
        p->mbox.push_back(m);
    }
    p->mbox_cv.notify_one();
}

static void rb_unref(RtBuffer *rb) {
    if (rb->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) delete rb;
}

// ---- worker threads (Worker.zig:250-411) ---------------------------------------
static void worker_main(Pool *p, bool spare, ptrdiff_t index) {
    t_pool = p;
    t_is_worker = true;
    t_is_spare = spare;
    t_worker = index;
    std::unique_lock<std::mutex> g(p->q_m);
    for (;;) {
        // A compensation thread retires as soon as the pool is back at its thread budget.
        if (spare && p->runnable > p->worker_count) break;
        // Worker.zig:262-280: the private queue first, then the pool's global queue.
        std::deque<WorkItem> *from = &p->queue;
        if (index >= 0 && !p->private_queue[(size_t)index].empty()) from = &p->private_queue[(size_t)index];
        if (from->empty()) {
            if (p->stop_workers) break;
            p->q_cv.wait(g);
            continue;
        }
        WorkItem w = from->front();
        from->pop_front();
        // StackAllocator.allocate (Pool.zig:112-121): a cached stack, else a fresh mapping with its guard page
        Stack stack;
        if (!p->stack_cache.empty()) stack = p->stack_cache.back(), p->stack_cache.pop_back();
        g.unlock();
        if (!stack.memory) {
            const size_t page = (size_t)sysconf(_SC_PAGESIZE);
            stack.size = kStackSize + page;
            stack.memory = mmap(nullptr, stack.size, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_STACK, -1, 0);
            if (stack.memory == MAP_FAILED || mprotect(stack.memory, page, PROT_NONE) != 0) panic("fiber stack allocation failed");
        }
        // Task.ensureReady (Task.zig:56-67): a fresh context on the stack; Worker.run (:284-300): switch to it.  The task
        // body runs on the fiber; it comes back by finishing (taskEntry -> completeTask) or through `abort`.
        Fiber fiber;
        fiber.task = w.task;
        {
            uintptr_t top = ((uintptr_t)stack.memory + stack.size) & ~(uintptr_t)15;
            void **sp = (void **)top;
            *--sp = (void *)&oracle_fiber_trampoline;      // "return address" of the first switch: rsp is 16-byte aligned at the trampoline's `call`
            *--sp = nullptr;                               // rbp
            *--sp = nullptr;                               // rbx
            *--sp = &fiber;                                // r12 = Fiber*
            *--sp = nullptr, *--sp = nullptr, *--sp = nullptr; // r13, r14, r15
            uint32_t *csr = (uint32_t *)(--sp);            // MXCSR + x87 control word: the defaults
            csr[0] = 0x1F80u, csr[1] = 0x037Fu;
            fiber.sp = sp;
        }
        t_fiber = &fiber;
        oracle_fiber_switch(&fiber.worker_sp, fiber.sp);
        t_fiber = nullptr;
        const bool is_error = fiber.aborted; // Worker.zig:115-122 abortTask
        if (is_error) {
            if (w.task->on_abort) w.task->on_abort(w.task);
        } else if (w.task->on_complete) {
            w.task->on_complete(w.task);
        }
        if (w.task->on_deinit) w.task->on_deinit(w.task);
        post(p, Msg{Msg::task_done, w.rb, w.entry, is_error}); // Worker.zig:306-335 "complete" message
        g.lock();
        p->stack_cache.push_back(stack); // StackAllocator.deallocate (:123-130): back to the cache
    }
    p->runnable--;
    p->threads_total--;
}

void pool_blocking_enter() {
    Pool *p = t_pool;
    if (!p || !t_is_worker) return;
    std::lock_guard<std::mutex> g(p->q_m);
    p->runnable--;
    if (p->runnable < p->worker_count && !p->stop_workers) {
        p->runnable++;
        p->threads_total++;
        std::thread(worker_main, p, true, (ptrdiff_t)-1).detach();
    }
}

void pool_blocking_exit() {
    Pool *p = t_pool;
    if (!p || !t_is_worker) return;
    std::lock_guard<std::mutex> g(p->q_m);
    p->runnable++;
    p->q_cv.notify_all(); // lets a spare notice it is over budget and retire
}

// ---- command buffer interpreter (event-loop thread only) ------------------------
static void enqueue_process(Pool *p, RtBuffer *rb) {
    if (rb->in_process_list || rb->done) return;
    rb->in_process_list = true;
    p->process_list.push_back(rb);
}

static void progress_completed(RtBuffer *rb) {
    // CommandBuffer.zig:445-457
    while (rb->completed_index < rb->entry_index && rb->status[rb->completed_index] == EntryStatus::processed)
        rb->completed_index++;
}

static void finish_buffer(Pool *p, RtBuffer *rb);
extern const FimoTasksMeta_CommandBufferHandleVTable g_handle_vtable;

// command_buffer.zig:56-70 Entry.abort, then Entry.deinit: the entry will never be issued.
static void abort_entry(const FimoTasksMeta_CommandBufferEntry &e) {
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

// CommandBuffer.zig:490-513
static void abort_forward(RtBuffer *rb) {
    if (!rb->abort_on_error) return;
    const FimoTasksMeta_CommandBuffer *cb = rb->buffer;
    while (rb->entry_index < cb->entries_len) {
        const FimoTasksMeta_CommandBufferEntry &e = cb->entries[rb->entry_index];
        rb->status[rb->entry_index] = EntryStatus::processed;
        rb->entry_index++;
        if (e.type == FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR) {
            rb->abort_on_error = e.payload.abort_on_error;
            if (!rb->abort_on_error) break;
        } else {
            abort_entry(e);
        }
    }
    rb->has_error = rb->abort_on_error;
}

// CommandBuffer.zig:459-466: the entry at entry_index is malformed.
static void abort_current_and_forward(RtBuffer *rb) {
    abort_entry(rb->buffer->entries[rb->entry_index]);
    rb->status[rb->entry_index] = EntryStatus::processed;
    rb->entry_index++;
    abort_forward(rb);
}

// CommandBuffer.zig:468-488: the already-issued entry `index` ended in error.
static void abort_from_index(RtBuffer *rb, size_t index) {
    if (!rb->abort_on_error) return;
    const FimoTasksMeta_CommandBuffer *cb = rb->buffer;
    if (rb->entry_index == cb->entries_len) {
        rb->has_error = true;
        return;
    }
    for (size_t i = index; i < rb->entry_index; i++) {
        const FimoTasksMeta_CommandBufferEntry &e = cb->entries[i];
        if (e.type == FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR && !e.payload.abort_on_error) return;
    }
    abort_forward(rb);
}

// CommandBuffer.zig:127-148 completeTask
static void complete_entry(Pool *p, RtBuffer *rb, size_t entry, bool is_error) {
    rb->status[entry] = EntryStatus::processed;
    if (is_error) abort_from_index(rb, entry);
    progress_completed(rb);
    enqueue_process(p, rb);
}

// Returns false if the buffer is blocked (error.Block in the reference).
static bool process_buffer(Pool *p, RtBuffer *rb) {
    const FimoTasksMeta_CommandBuffer *cb = rb->buffer;
    while (rb->entry_index < cb->entries_len) {
        const size_t idx = rb->entry_index;
        const FimoTasksMeta_CommandBufferEntry &e = cb->entries[idx];
        switch (e.type) {
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR:
                rb->abort_on_error = e.payload.abort_on_error;
                rb->status[idx] = EntryStatus::processed;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SET_MIN_STACK_SIZE:
                // Stack sizes have no meaning for OS-thread tasks: every request is satisfiable.
                rb->status[idx] = EntryStatus::processed;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_WORKER:
                if (e.payload.select_worker >= p->worker_count) { // :243-254
                    abort_current_and_forward(rb);
                    continue;
                }
                rb->worker = (ptrdiff_t)e.payload.select_worker;
                rb->status[idx] = EntryStatus::processed;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_ANY_WORKER:
                rb->worker = -1;
                rb->status[idx] = EntryStatus::processed;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK: {
                rb->status[idx] = EntryStatus::running_task;
                {
                    std::lock_guard<std::mutex> g(p->q_m);
                    const WorkItem w{rb, idx, e.payload.enqueue_task};
                    if (rb->worker >= 0) p->private_queue[(size_t)rb->worker].push_back(w);
                    else p->queue.push_back(w);
                }
                // a pinned task must reach ITS worker (one shared condition variable: wake all); any worker can
                // take a task of the global queue (wake one — waking all 16 per task cost the CPU arm 2-7x)
                if (rb->worker >= 0) p->q_cv.notify_all();
                else p->q_cv.notify_one();
                break;
            }
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_COMMAND_BUFFER: {
                RtBuffer *child = new RtBuffer();
                child->owner = p;
                child->buffer = e.payload.enqueue_command_buffer;
                child->status.assign(child->buffer->entries_len, EntryStatus::not_processed);
                child->parent = rb;
                child->parent_entry = idx;
                rb->status[idx] = EntryStatus::running_buffer;
                p->active_buffers++;
                enqueue_process(p, child);
                break;
            }
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_BARRIER:
                progress_completed(rb);
                if (rb->completed_index != idx) return false;
                rb->status[idx] = EntryStatus::processed;
                break;
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_BUFFER: {
                RtBuffer *other = (RtBuffer *)e.payload.wait_on_command_buffer.data;
                if (e.payload.wait_on_command_buffer.vtable != &g_handle_vtable || other->owner != p) { // :357-381
                    abort_current_and_forward(rb);
                    continue;
                }
                int other_done;
                {
                    std::lock_guard<std::mutex> g(other->m);
                    other_done = other->done;
                }
                if (!other_done) {
                    auto &w = other->buffer_waiters;
                    if (std::find(w.begin(), w.end(), rb) == w.end()) w.push_back(rb);
                    return false;
                }
                rb->status[idx] = EntryStatus::processed;
                rb->entry_index++;
                if (other_done == 2) abort_forward(rb); // :385-387
                continue;
            }
            case FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_INDIRECT: {
                const size_t k = e.payload.wait_on_command_indirect;
                if (k == 0) { // :397
                    rb->status[idx] = EntryStatus::processed;
                    break;
                }
                if (k > idx) { // :398-408
                    abort_current_and_forward(rb);
                    continue;
                }
                if (rb->status[idx - k] != EntryStatus::processed) return false; // error.Block
                rb->status[idx] = EntryStatus::processed;
                break;
            }
            default: // :413-421
                abort_current_and_forward(rb);
                continue;
        }
        rb->entry_index++;
    }
    progress_completed(rb);
    if (rb->completed_index != rb->entry_index) return false; // CommandBuffer.zig:190-194
    finish_buffer(p, rb);
    return true;
}

static void finish_buffer(Pool *p, RtBuffer *rb) {
    // CommandBuffer.zig:196-200: complete/abort, broadcast to waiters, drop the pool's reference.
    FimoTasksMeta_CommandBuffer *cb = rb->buffer;
    RtBuffer *parent = rb->parent;
    const size_t parent_entry = rb->parent_entry;
    std::vector<RtBuffer *> waiters;
    waiters.swap(rb->buffer_waiters);
    const bool err = rb->has_error;
    if (err) {
        if (cb->on_abort) cb->on_abort(cb);
    } else {
        if (cb->on_complete) cb->on_complete(cb);
    }
    auto on_deinit = cb->on_deinit;
    {
        std::lock_guard<std::mutex> g(rb->m);
        rb->done = err ? 2 : 1;
    }
    rb->cv.notify_all();
    if (on_deinit) on_deinit(cb);
    for (RtBuffer *w : waiters) enqueue_process(p, w);
    if (parent) complete_entry(p, parent, parent_entry, err);
    p->active_buffers--;
    rb_unref(rb);
}

// ---- event loop (Pool.zig:660-765) ----------------------------------------------
static void event_loop(Pool *p) {
    t_pool = p;
    for (;;) {
        std::deque<Msg> batch;
        {
            std::unique_lock<std::mutex> g(p->mbox_m);
            p->mbox_cv.wait(g, [&] {
                return !p->mbox.empty() || (p->closed.load() && p->refs.load() == 0 && p->active_buffers == 0);
            });
            batch.swap(p->mbox);
        }
        if (batch.empty()) break; // closed, unreferenced and drained
        for (const Msg &m : batch) {
            switch (m.kind) {
                case Msg::enqueue:
                    p->active_buffers++;
                    enqueue_process(p, m.rb);
                    break;
                case Msg::task_done: complete_entry(p, m.rb, m.entry, m.is_error); break;
                case Msg::wake: break;
            }
        }
        while (!p->process_list.empty()) {
            RtBuffer *rb = p->process_list.front();
            p->process_list.pop_front();
            rb->in_process_list = false;
            process_buffer(p, rb);
        }
    }
    // shut down workers
    {
        std::lock_guard<std::mutex> g(p->q_m);
        p->stop_workers = true;
    }
    p->q_cv.notify_all();
    for (auto &t : p->threads) t.join();
    // wait for detached compensation threads to retire
    for (;;) {
        {
            std::lock_guard<std::mutex> g(p->q_m);
            if (p->threads_total == 0) break;
        }
        p->q_cv.notify_all();
        std::this_thread::yield();
    }
    for (const Stack &st : p->stack_cache) munmap(st.memory, st.size); // StackAllocator.deinit: every cached stack goes back
    delete p;
}

// ---- vtables ----------------------------------------------------------------------
static FimoTasksMeta_PoolId vt_id(void *d) { return ((Pool *)d)->id; }
static void vt_acquire(void *d) { ((Pool *)d)->refs.fetch_add(1, std::memory_order_relaxed); }
static void vt_release(void *d) {
    Pool *p = (Pool *)d;
    if (p->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) {
        p->closed.store(true);
        post(p, Msg{Msg::wake, nullptr, 0});
    }
}
static void vt_request_close(void *d) { ((Pool *)d)->closed.store(true); }
static bool vt_accepts_requests(void *d) { return !((Pool *)d)->closed.load(); }
static bool vt_owns_current_thread(void *d) { return t_pool == (Pool *)d; }
static const char *vt_label(void *d, FimoUSize *len) {
    Pool *p = (Pool *)d;
    *len = p->label.size();
    return p->label.data();
}
static FimoUSize vt_workers(void *d, FimoTasksMeta_PoolWorker *ptr, FimoUSize len) {
    Pool *p = (Pool *)d;
    for (size_t i = 0; ptr && i < len && i < p->worker_count; i++) ptr[i] = i;
    return p->worker_count;
}
static FimoUSize vt_stack_sizes(void *, FimoTasksMeta_PoolStackSize *ptr, FimoUSize len) {
    if (ptr && len > 0) ptr[0] = 8u << 20; // FimoTasks.zig:32 default_stack_size = 8 MiB
    return 1;
}

static void h_acquire(void *d) { ((RtBuffer *)d)->refs.fetch_add(1, std::memory_order_relaxed); }
static void h_release(void *d) { rb_unref((RtBuffer *)d); }
static FimoTasksMeta_Pool h_owner_pool(void *d) {
    Pool *p = ((RtBuffer *)d)->owner;
    vt_acquire(p);
    return pool_handle(p);
}
static FimoTasksMeta_CommandBufferHandleCompletionStatus h_wait_on(void *d) {
    // CommandBuffer.zig:603-631
    RtBuffer *rb = (RtBuffer *)d;
    std::unique_lock<std::mutex> g(rb->m);
    if (!rb->done) {
        pool_blocking_enter();
        rb->cv.wait(g, [&] { return rb->done != 0; });
        g.unlock();
        pool_blocking_exit();
        g.lock();
    }
    return rb->done == 1 ? FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_COMPLETED
                         : FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_ABORTED;
}
extern const FimoTasksMeta_CommandBufferHandleVTable g_handle_vtable = {h_acquire, h_release, h_owner_pool, h_wait_on};

static FimoStatus vt_enqueue_buffer(void *d, FimoTasksMeta_CommandBuffer *buffer,
                                    FimoTasksMeta_CommandBufferHandle *handle) {
    // Pool.zig:820-842: wrap in runtime state, hand to the event loop, optionally return a handle.
    Pool *p = (Pool *)d;
    if (p->closed.load()) {
        set_last_error("Closed");
        return FIMO_STATUS_ERROR;
    }
    RtBuffer *rb = new RtBuffer();
    rb->owner = p;
    rb->buffer = buffer;
    rb->status.assign(buffer->entries_len, EntryStatus::not_processed);
    if (handle) {
        rb->refs.fetch_add(1, std::memory_order_relaxed);
        handle->data = rb;
        handle->vtable = &g_handle_vtable;
    }
    post(p, Msg{Msg::enqueue, rb, 0});
    return FIMO_STATUS_OK;
}

static const FimoTasksMeta_PoolVTable g_pool_vtable = {
        vt_id,    vt_acquire, vt_release,     vt_request_close, vt_accepts_requests, vt_owns_current_thread,
        vt_label, vt_workers, vt_stack_sizes, vt_enqueue_buffer};

FimoTasksMeta_Pool pool_handle(Pool *p) { return FimoTasksMeta_Pool{p, &g_pool_vtable}; }
Pool *pool_current() { return t_pool; }

Pool *pool_create(const char *label, size_t label_len, size_t workers) {
    Pool *p = new Pool();
    p->id = g_next_pool_id.fetch_add(1);
    p->label = label && label_len ? std::string(label, label_len) : std::string("<unlabelled>");
    if (workers == 0) workers = std::thread::hardware_concurrency(); // FimoTasks.zig:127-132
    if (workers == 0) workers = 1;
    p->worker_count = workers;
    p->runnable = workers;
    p->threads_total = workers;
    p->private_queue.resize(workers);
    for (size_t i = 0; i < workers; i++) p->threads.emplace_back(worker_main, p, false, (ptrdiff_t)i);
    p->event_thread = std::thread(event_loop, p);
    p->event_thread.detach();
    return p;
}

} // namespace oracle

extern "C" FimoStatus fimo_tasks_create_worker_pool(const FimoTasksMeta_PoolConfig *config, FimoTasksMeta_Pool *pool) {
    // FimoTasks.zig:201-358 validates the config; the oracle honours label and worker_count.
    if (config->next != nullptr) {
        oracle::set_last_error("InvalidConfig"); // FimoTasks.zig:208-212 rejects a non-null `next`
        return FIMO_STATUS_ERROR;
    }
    *pool = oracle::pool_handle(oracle::pool_create(config->label, config->label_len, config->worker_count));
    return FIMO_STATUS_OK;
}

// symbols.zig:94-99 `abort`; Worker.zig:115-122
extern "C" void fimo_tasks_abort(void) {
    oracle::Fiber *f = oracle::t_fiber;
    if (!f) oracle::panic("fimo_tasks_abort called outside a task");
    f->aborted = true;
    void *dead;
    oracle_fiber_switch(&dead, f->worker_sp); // the fiber is never resumed; its stack goes back to the cache
    __builtin_unreachable();
}
// symbols.zig:60-65 `worker_id`
extern "C" bool fimo_tasks_worker_id(FimoTasksMeta_PoolWorker *id) {
    if (oracle::t_worker < 0) return false;
    *id = (FimoTasksMeta_PoolWorker)oracle::t_worker;
    return true;
}
