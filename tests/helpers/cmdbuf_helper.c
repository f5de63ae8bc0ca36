/* Test helper: tasks and command buffers whose hooks append to a shared event log.  A task body may end itself through
 * the `abort` symbol of the library under test (a longjmp / fiber switch that must not cross Python frames, hence C). */
#include <sched.h>
#include <stdatomic.h>
#include <stddef.h>
#include <stdint.h>

#include "fimo_worlds_b200.h"

enum { EV_TASK_START = 1, EV_TASK_COMPLETE, EV_TASK_ABORT, EV_TASK_DEINIT, EV_BUF_COMPLETE, EV_BUF_ABORT, EV_BUF_DEINIT };

typedef struct {
    int32_t id, kind;
    int64_t worker; /* worker_id() of the reporting thread, -1 when it has none */
} HelperEvent;

typedef struct {
    atomic_size_t n;
    atomic_int gate; /* tasks with wait_gate spin until this is non-zero */
    size_t cap;
    HelperEvent *events;
} HelperLog;

typedef struct {
    FimoTasksMeta_Task task; /* first: hooks get a pointer to this member */
    HelperLog *log;
    void (*abort_fn)(void);
    bool (*worker_id_fn)(FimoTasksMeta_PoolWorker *);
    int32_t id, should_abort, wait_gate;
    /* set (before the gate opens) by helper_task_wait_on: the body then waits for that command buffer INSIDE the task */
    void *wait_data;
    const FimoTasksMeta_CommandBufferHandleVTable *wait_vtable;
    _Atomic int32_t wait_status; /* -1 until the wait has returned, then the completion status */
} HelperTask;

typedef struct {
    FimoTasksMeta_CommandBuffer buffer;
    HelperLog *log;
    int32_t id;
} HelperBuffer;

static void record(HelperLog *log, int32_t id, int32_t kind, bool (*worker_id_fn)(FimoTasksMeta_PoolWorker *)) {
    const size_t i = atomic_fetch_add(&log->n, 1);
    if (i >= log->cap) return;
    FimoTasksMeta_PoolWorker w = 0;
    log->events[i].id = id;
    log->events[i].kind = kind;
    log->events[i].worker = worker_id_fn && worker_id_fn(&w) ? (int64_t)w : -1;
}

static void task_start(FimoTasksMeta_Task *t) {
    HelperTask *h = (HelperTask *)t;
    record(h->log, h->id, EV_TASK_START, h->worker_id_fn);
    while (h->wait_gate && !atomic_load(&h->log->gate)) sched_yield();
    if (h->wait_vtable) atomic_store(&h->wait_status, (int32_t)h->wait_vtable->wait_on(h->wait_data));
    if (h->should_abort) h->abort_fn();
}
static void task_complete(FimoTasksMeta_Task *t) { record(((HelperTask *)t)->log, ((HelperTask *)t)->id, EV_TASK_COMPLETE, NULL); }
static void task_abort(FimoTasksMeta_Task *t) { record(((HelperTask *)t)->log, ((HelperTask *)t)->id, EV_TASK_ABORT, NULL); }
static void task_deinit(FimoTasksMeta_Task *t) { record(((HelperTask *)t)->log, ((HelperTask *)t)->id, EV_TASK_DEINIT, NULL); }
static void buf_complete(FimoTasksMeta_CommandBuffer *b) { record(((HelperBuffer *)b)->log, ((HelperBuffer *)b)->id, EV_BUF_COMPLETE, NULL); }
static void buf_abort(FimoTasksMeta_CommandBuffer *b) { record(((HelperBuffer *)b)->log, ((HelperBuffer *)b)->id, EV_BUF_ABORT, NULL); }
static void buf_deinit(FimoTasksMeta_CommandBuffer *b) { record(((HelperBuffer *)b)->log, ((HelperBuffer *)b)->id, EV_BUF_DEINIT, NULL); }

size_t helper_sizeof_task(void) { return sizeof(HelperTask); }
size_t helper_sizeof_buffer(void) { return sizeof(HelperBuffer); }

void helper_task_init(HelperTask *h, HelperLog *log, int32_t id, int32_t should_abort, int32_t wait_gate, void (*abort_fn)(void),
                      bool (*worker_id_fn)(FimoTasksMeta_PoolWorker *)) {
    h->task.label = "helper";
    h->task.label_len = 6;
    h->task.on_start = task_start;
    h->task.on_complete = task_complete;
    h->task.on_abort = task_abort;
    h->task.on_deinit = task_deinit;
    h->log = log;
    h->abort_fn = abort_fn;
    h->worker_id_fn = worker_id_fn;
    h->id = id;
    h->should_abort = should_abort;
    h->wait_gate = wait_gate;
    h->wait_data = NULL;
    h->wait_vtable = NULL;
    atomic_store(&h->wait_status, -1);
}

void helper_task_wait_on(HelperTask *h, void *data, const FimoTasksMeta_CommandBufferHandleVTable *vtable) {
    h->wait_data = data;
    h->wait_vtable = vtable;
}
int32_t helper_task_wait_status(HelperTask *h) { return atomic_load(&h->wait_status); }

void helper_buffer_init(HelperBuffer *b, HelperLog *log, int32_t id, FimoTasksMeta_CommandBufferEntry *entries, size_t n) {
    b->buffer.label = "helper";
    b->buffer.label_len = 6;
    b->buffer.entries = entries;
    b->buffer.entries_len = n;
    b->buffer.on_complete = buf_complete;
    b->buffer.on_abort = buf_abort;
    b->buffer.on_deinit = buf_deinit;
    b->log = log;
    b->id = id;
}

void helper_open_gate(HelperLog *log) { atomic_store(&log->gate, 1); }
size_t helper_event_count(HelperLog *log) { return atomic_load(&log->n); }
