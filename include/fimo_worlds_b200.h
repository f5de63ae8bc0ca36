/* fimo_worlds_b200.h — C ABI of the B200-native `fimo-worlds` system-group step.
 *
 * This header is the drop-in boundary.  It is a self-contained,
 * C17 / C++17 / nvcc-clean mirror of the reference's interface packages:
 *
 *   pkgs/fimo_worlds_meta/include/fimo_worlds_meta/{jobs,resources,systems,worlds}.h
 *   pkgs/fimo_tasks_meta/include/fimo_tasks_meta/{command_buffer,futex,pool,task}.h
 *   pkgs/fimo_std/include/fimo_std/error.h (FimoStatus only)
 *
 * The reference's own headers do not compile (jobs.h:42,47; modules.h:284 — see
 * SURVEY.md preamble), and where header and implementation disagree the Zig
 * `symbols.zig` files are authoritative (pkgs/fimo_worlds_meta/src/symbols.zig:60-365,
 * pkgs/fimo_tasks_meta/src/pool.zig:88-104).  Struct layouts are byte-identical to
 * the reference on x86-64 and are pinned by the static asserts at the bottom.
 *
 * The reference exposes its 32 `fimo-worlds` symbols through the fimo module
 * loader (name + namespace + version), not as C-linkage names.  This library
 * exports each of them as `fimo_worlds_<symbol name>` and additionally as a
 * table (`fimo_worlds_b200_exports`) in the order of `all_symbols`
 * (symbols.zig:22-58), which is what a `fimo_exports` shim binds.
 *
 * The same header is implemented twice:
 *   - libfimo_worlds_b200.so  the product: CUDA sm_100a device executor.
 *   - oracle/_build/libfimo_worlds_oracle.so  the CPU restatement (test infrastructure only).
 */
#ifndef FIMO_WORLDS_B200_H
#define FIMO_WORLDS_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define FIMO_B200_EXPORT __attribute__((visibility("default")))
#else
#define FIMO_B200_EXPORT
#endif

/* ------------------------------------------------------------------------- */
/* fimo_std mirrors                                                           */
/* ------------------------------------------------------------------------- */

typedef size_t FimoUSize;
typedef uint8_t FimoU8;
typedef uint64_t FimoU64;
typedef int32_t FimoI32;

/* fimo_std/error.h:175-186 — status of a fallible call. */
typedef FimoI32 FimoStatus;
#define FIMO_STATUS_OK ((FimoStatus)0)
/* error, details parked in the thread-local result (here: fimo_worlds_b200_last_error) */
#define FIMO_STATUS_ERROR ((FimoStatus)-1)
#define FIMO_STATUS_ERROR_NO_REPORT ((FimoStatus)-2)

/* fimo_std/time.h — 16-byte instant / duration. */
typedef struct FimoDuration {
    FimoU64 secs;
    uint32_t nanos;
} FimoDuration;
typedef struct FimoInstant {
    FimoU64 secs;
    uint32_t nanos;
} FimoInstant;

/* ------------------------------------------------------------------------- */
/* fimo_tasks_meta mirrors (types on the boundary of the hot path)            */
/* ------------------------------------------------------------------------- */

typedef FimoUSize FimoTasksMeta_TaskId;
typedef FimoUSize FimoTasksMeta_PoolId;
typedef FimoUSize FimoTasksMeta_PoolWorker;
typedef FimoUSize FimoTasksMeta_PoolStackSize;

/* task.h:17-38 */
typedef struct FimoTasksMeta_Task {
    const char *label;
    FimoUSize label_len;
    void (*on_start)(struct FimoTasksMeta_Task *task);
    void (*on_complete)(struct FimoTasksMeta_Task *task);
    void (*on_abort)(struct FimoTasksMeta_Task *task);
    void (*on_deinit)(struct FimoTasksMeta_Task *task);
} FimoTasksMeta_Task;

typedef struct FimoTasksMeta_CommandBuffer FimoTasksMeta_CommandBuffer;
typedef struct FimoTasksMeta_CommandBufferHandle FimoTasksMeta_CommandBufferHandle;

/* pool.h:27-62; enqueue_buffer returns Status as in pkgs/fimo_tasks_meta/src/pool.zig:99-103 */
typedef struct FimoTasksMeta_PoolVTable {
    FimoTasksMeta_PoolId (*id)(void *pool);
    void (*acquire)(void *pool);
    void (*release)(void *pool);
    void (*request_close)(void *pool);
    bool (*accepts_requests)(void *pool);
    bool (*owns_current_thread)(void *pool);
    const char *(*label)(void *pool, FimoUSize *len);
    FimoUSize (*workers)(void *pool, FimoTasksMeta_PoolWorker *ptr, FimoUSize len);
    FimoUSize (*stack_sizes)(void *pool, FimoTasksMeta_PoolStackSize *ptr, FimoUSize len);
    FimoStatus (*enqueue_buffer)(void *pool, FimoTasksMeta_CommandBuffer *buffer,
                                 FimoTasksMeta_CommandBufferHandle *handle /* nullable: detached */);
} FimoTasksMeta_PoolVTable;

/* pool.h:65-68 */
typedef struct FimoTasksMeta_Pool {
    void *data;
    const FimoTasksMeta_PoolVTable *vtable;
} FimoTasksMeta_Pool;

/* command_buffer.h:17-41 */
typedef enum FimoTasksMeta_CommandBufferHandleCompletionStatus {
    FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_COMPLETED,
    FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_ABORTED,
} FimoTasksMeta_CommandBufferHandleCompletionStatus;

typedef struct FimoTasksMeta_CommandBufferHandleVTable {
    void (*acquire)(void *handle);
    void (*release)(void *handle);
    FimoTasksMeta_Pool (*owner_pool)(void *handle);
    FimoTasksMeta_CommandBufferHandleCompletionStatus (*wait_on)(void *handle);
} FimoTasksMeta_CommandBufferHandleVTable;

struct FimoTasksMeta_CommandBufferHandle {
    void *data;
    const FimoTasksMeta_CommandBufferHandleVTable *vtable;
};

/* command_buffer.h:44-54 */
typedef FimoI32 FimoTasksMeta_CommandBufferEntryType;
enum {
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR = 0,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SET_MIN_STACK_SIZE = 1,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_WORKER = 2,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_ANY_WORKER = 3,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK = 4,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_COMMAND_BUFFER = 5,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_BARRIER = 6,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_BUFFER = 7,
    FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_INDIRECT = 8,
};

/* command_buffer.h:57-86 */
typedef union FimoTasksMeta_CommandBufferEntryPayload {
    bool abort_on_error;
    FimoTasksMeta_PoolStackSize set_min_stack_size;
    FimoTasksMeta_PoolWorker select_worker;
    FimoU8 select_any_worker;
    FimoTasksMeta_Task *enqueue_task;
    FimoTasksMeta_CommandBuffer *enqueue_command_buffer;
    FimoU8 wait_on_barrier;
    FimoTasksMeta_CommandBufferHandle wait_on_command_buffer;
    /* waits for the command at index `this_command - value` */
    FimoUSize wait_on_command_indirect;
} FimoTasksMeta_CommandBufferEntryPayload;

/* command_buffer.h:92-96 */
typedef struct FimoTasksMeta_CommandBufferEntry {
    FimoTasksMeta_CommandBufferEntryType type;
    FimoTasksMeta_CommandBufferEntryPayload payload;
} FimoTasksMeta_CommandBufferEntry;

/* command_buffer.h:99-122 */
struct FimoTasksMeta_CommandBuffer {
    const char *label;
    FimoUSize label_len;
    const FimoTasksMeta_CommandBufferEntry *entries;
    FimoUSize entries_len;
    void (*on_complete)(FimoTasksMeta_CommandBuffer *buffer);
    void (*on_abort)(FimoTasksMeta_CommandBuffer *buffer);
    void (*on_deinit)(FimoTasksMeta_CommandBuffer *buffer);
};

/* pool.h:71-106 */
typedef struct FimoTasksMeta_PoolConfigStackConfig {
    const void *next;
    FimoTasksMeta_PoolStackSize size;
    FimoUSize preallocated_count;
    FimoUSize cold_count;
    FimoUSize hot_count;
    FimoUSize max_allocated;
} FimoTasksMeta_PoolConfigStackConfig;

typedef struct FimoTasksMeta_PoolConfig {
    const void *next;
    const char *label;
    FimoUSize label_len;
    const FimoTasksMeta_PoolConfigStackConfig *stacks;
    FimoUSize stacks_len;
    FimoUSize default_stack_index;
    FimoUSize worker_count; /* 0 = number of CPUs (FimoTasks.zig:127-132) */
    bool is_queryable;
} FimoTasksMeta_PoolConfig;

/* futex.h:14-19 */
typedef FimoI32 FimoTasksMeta_FutexStatus;
enum {
    FIMO_TASKS_META_FUTEX_STATUS_OK = 0,
    FIMO_TASKS_META_FUTEX_STATUS_INVALID = 1,
    FIMO_TASKS_META_FUTEX_STATUS_TIMEOUT = 2,
    FIMO_TASKS_META_FUTEX_STATUS_KEY_ERROR = 3,
};

/* futex.h:22-27 */
typedef struct FimoTasksMeta_FutexKeyExpect {
    const void *key;
    FimoUSize key_size;
    FimoU64 expect;
    FimoUSize token;
} FimoTasksMeta_FutexKeyExpect;

/* futex.h:46-119: op = token_op | token_type<<1 | cmp_op<<3 | cmp_arg_op<<6 */
typedef struct FimoTasksMeta_FutexFilter {
    FimoUSize op;
    FimoUSize token_mask;
    FimoUSize cmp_arg;
} FimoTasksMeta_FutexFilter;

enum { FIMO_TASKS_META_FUTEX_FILTER_TOKEN_OP_NOOP = 0, FIMO_TASKS_META_FUTEX_FILTER_TOKEN_OP_DEREF = 1 };
enum {
    FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U8 = 0,
    FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U16 = 1,
    FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U32 = 2,
    FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U64 = 3,
};
enum {
    FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_EQ = 0,
    FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_NE = 1,
    FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LT = 2,
    FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LE = 3,
    FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_GT = 4,
    FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_GE = 5,
};
enum { FIMO_TASKS_META_FUTEX_FILTER_CMP_ARG_OP_NOOP = 0, FIMO_TASKS_META_FUTEX_FILTER_CMP_ARG_OP_DEREF = 1 };

static inline FimoUSize FimoTasksMeta_futex_filter_op_init(FimoUSize token_op, FimoUSize token_type, FimoUSize cmp_op,
                                                           FimoUSize cmp_arg_op) {
    return (token_op & 1u) | ((token_type & 3u) << 1) | ((cmp_op & 7u) << 3) | ((cmp_arg_op & 1u) << 6);
}

/* The two `fimo-tasks` futex symbols the path imports (pkgs/fimo_tasks_meta/src/symbols.zig:132-160). */
typedef FimoTasksMeta_FutexStatus (*FimoTasksMeta_futex_wait)(const void *key, FimoUSize key_size, FimoU64 expect,
                                                              FimoUSize token, const FimoInstant *timeout);
typedef FimoUSize (*FimoTasksMeta_futex_wake)(const void *key, FimoUSize max_waiters, FimoTasksMeta_FutexFilter filter);

/* ------------------------------------------------------------------------- */
/* fimo_worlds_meta mirrors                                                   */
/* ------------------------------------------------------------------------- */

typedef struct FimoWorldsMeta_World FimoWorldsMeta_World;
typedef struct FimoWorldsMeta_Resource *FimoWorldsMeta_ResourceHandle;
typedef struct FimoWorldsMeta_System *FimoWorldsMeta_SystemHandle;
typedef struct FimoWorldsMeta_SystemGroup FimoWorldsMeta_SystemGroup;
typedef struct FimoWorldsMeta_SystemContext FimoWorldsMeta_SystemContext;

/* jobs.h:15-21; Zig: pkgs/fimo_worlds_meta/src/Job.zig:20-26.
 * One atomic byte: bit0 = signaled, bit1 = contended (a waiter is parked). */
typedef struct FimoWorldsMeta_Fence {
    FimoU8 state; /* accessed with __atomic builtins only */
} FimoWorldsMeta_Fence;
#define FIMO_WORLDS_META_FENCE_UNSIGNALED ((FimoU8)0)
#define FIMO_WORLDS_META_FENCE_SIGNALED ((FimoU8)1)
#define FIMO_WORLDS_META_FENCE_CONTENDED ((FimoU8)2)

/* jobs.h:74-76; Job.zig:66-70 — monotonically increasing u64 counter. */
typedef struct FimoWorldsMeta_TimelineSemaphore {
    FimoU64 state; /* accessed with __atomic builtins only */
} FimoWorldsMeta_TimelineSemaphore;

/* jobs.h:24-71 restated with GCC atomics (the reference's C inline versions do not
 * compile: jobs.h:42 passes &fence->state as a Fence*, jobs.h:47 passes a value as
 * the `expected` pointer).  Semantics follow Job.zig:28-63. */
static inline void FimoWorldsMeta_fence_init(FimoWorldsMeta_Fence *fence) {
    __atomic_store_n(&fence->state, FIMO_WORLDS_META_FENCE_UNSIGNALED, __ATOMIC_RELAXED);
}
static inline bool FimoWorldsMeta_fence_is_signaled(const FimoWorldsMeta_Fence *fence) {
    return (__atomic_load_n(&fence->state, __ATOMIC_ACQUIRE) & FIMO_WORLDS_META_FENCE_SIGNALED) != 0;
}
static inline void FimoWorldsMeta_fence_wait(FimoWorldsMeta_Fence *fence, FimoTasksMeta_futex_wait futex_wait) {
    for (;;) {
        FimoU8 state = __atomic_load_n(&fence->state, __ATOMIC_RELAXED);
        if ((state & FIMO_WORLDS_META_FENCE_SIGNALED) != 0) {
            if (FimoWorldsMeta_fence_is_signaled(fence)) return;
        }
        if ((state & FIMO_WORLDS_META_FENCE_CONTENDED) == 0) {
            FimoU8 expected = FIMO_WORLDS_META_FENCE_UNSIGNALED;
            if (!__atomic_compare_exchange_n(&fence->state, &expected, FIMO_WORLDS_META_FENCE_CONTENDED, true,
                                             __ATOMIC_RELAXED, __ATOMIC_RELAXED))
                continue;
        }
        futex_wait(&fence->state, sizeof(FimoU8), FIMO_WORLDS_META_FENCE_CONTENDED, 0, NULL);
    }
}
static inline void FimoWorldsMeta_fence_signal(FimoWorldsMeta_Fence *fence, FimoTasksMeta_futex_wake futex_wake) {
    const FimoU8 state = __atomic_exchange_n(&fence->state, FIMO_WORLDS_META_FENCE_SIGNALED, __ATOMIC_RELEASE);
    if ((state & FIMO_WORLDS_META_FENCE_CONTENDED) != 0) {
        FimoTasksMeta_FutexFilter all = {0, 0, 0}; /* FIMO_TASKS_META_FUTEX_FILTER_ALL */
        futex_wake(&fence->state, ~(FimoUSize)0, all);
    }
}
static inline void FimoWorldsMeta_fence_reset(FimoWorldsMeta_Fence *fence) {
    (void)__atomic_fetch_and(&fence->state, (FimoU8)~FIMO_WORLDS_META_FENCE_SIGNALED, __ATOMIC_RELEASE);
}

/* jobs.h:79-144; Job.zig:72-173 */
static inline void FimoWorldsMeta_timeline_semaphore_init(FimoWorldsMeta_TimelineSemaphore *s, FimoU64 value) {
    __atomic_store_n(&s->state, value, __ATOMIC_RELAXED);
}
static inline FimoU64 FimoWorldsMeta_timeline_semaphore_counter(const FimoWorldsMeta_TimelineSemaphore *s) {
    return __atomic_load_n(&s->state, __ATOMIC_ACQUIRE);
}
static inline bool FimoWorldsMeta_timeline_semaphore_is_signaled(const FimoWorldsMeta_TimelineSemaphore *s,
                                                                 FimoU64 value) {
    return __atomic_load_n(&s->state, __ATOMIC_ACQUIRE) >= value;
}
static inline void FimoWorldsMeta_timeline_semaphore_wait(const FimoWorldsMeta_TimelineSemaphore *s, FimoU64 value,
                                                          FimoTasksMeta_futex_wait futex_wait) {
    if (FimoWorldsMeta_timeline_semaphore_is_signaled(s, value)) return;
    for (;;) {
        const FimoU64 state = __atomic_load_n(&s->state, __ATOMIC_RELAXED);
        if (state >= value) {
            (void)__atomic_load_n(&s->state, __ATOMIC_ACQUIRE);
            return;
        }
        futex_wait(&s->state, sizeof(FimoU64), state, (FimoUSize)value, NULL);
    }
}
static inline void FimoWorldsMeta_timeline_semaphore_signal(FimoWorldsMeta_TimelineSemaphore *s, FimoU64 value,
                                                            FimoTasksMeta_futex_wake futex_wake) {
    __atomic_store_n(&s->state, value, __ATOMIC_RELEASE);
    FimoTasksMeta_FutexFilter filter;
    filter.token_mask = ~(FimoUSize)0;
    filter.op = FimoTasksMeta_futex_filter_op_init(
            FIMO_TASKS_META_FUTEX_FILTER_TOKEN_OP_NOOP, FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U64,
            FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LE, FIMO_TASKS_META_FUTEX_FILTER_CMP_ARG_OP_NOOP);
    filter.cmp_arg = (FimoUSize)value;
    futex_wake(&s->state, ~(FimoUSize)0, filter);
}

/* resources.h:16-27 */
typedef struct FimoWorldsMeta_ResourceDescriptor {
    const void *next; /* reference: reserved, ignored (FimoWorlds.zig:108-122). Here: FimoB200_ResourceDeviceExt chain. */
    const char *label;
    FimoUSize label_len;
    FimoUSize size;
    FimoUSize alignment;
} FimoWorldsMeta_ResourceDescriptor;

/* systems.h:28-48 */
typedef FimoUSize FimoWorldsMeta_SystemDependencyFlags;
#define FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_WEAK ((FimoUSize)1)
#define FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_IGNORE_DEFERRED ((FimoUSize)2)
typedef struct FimoWorldsMeta_SystemDependency {
    FimoWorldsMeta_SystemHandle system;
    FimoWorldsMeta_SystemDependencyFlags flags;
} FimoWorldsMeta_SystemDependency;

/* systems.h:51-114.  NOTE the naming (Universe.zig:195-219, SystemGroup.zig:225-232):
 * `before` = systems that run BEFORE this one (this system depends on them);
 * `after`  = systems that run AFTER this one. */
typedef struct FimoWorldsMeta_SystemDescriptor {
    const void *next; /* reference: reserved. Here: FimoB200_SystemDeviceExt chain. */
    const char *label;
    FimoUSize label_len;
    const FimoWorldsMeta_ResourceHandle *exclusive_handles;
    FimoUSize exclusive_handles_len;
    const FimoWorldsMeta_ResourceHandle *shared_handles;
    FimoUSize shared_handles_len;
    const FimoWorldsMeta_SystemDependency *before;
    FimoUSize before_len;
    const FimoWorldsMeta_SystemDependency *after;
    FimoUSize after_len;
    const void *factory;
    FimoUSize factory_size;
    FimoUSize factory_alignment;
    void (*factory_destroy)(const void *factory);
    FimoUSize system_size;
    FimoUSize system_alignment;
    bool (*system_create)(const void *factory, FimoWorldsMeta_SystemContext *context, void *system);
    void (*system_destroy)(void *system);
    void (*system_run)(void *system, void *const *exclusive_resources, void *const *shared_resources,
                       FimoWorldsMeta_Fence *fence);
} FimoWorldsMeta_SystemDescriptor;

/* systems.h:117-131 */
typedef struct FimoWorldsMeta_SystemGroupDescriptor {
    const void *next;
    const char *label;
    FimoUSize label_len;
    const FimoTasksMeta_Pool *pool;
    FimoWorldsMeta_World *world;
} FimoWorldsMeta_SystemGroupDescriptor;

/* systems.h:134-155 */
typedef FimoI32 FimoWorldsMeta_SystemAllocatorStrategy;
enum {
    FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_TRANSIENT = 0,
    FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SINGLE_GENERATION = 1,
    FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_MULTI_GENERATION = 2,
    FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT = 3,
};

/* worlds.h:20-32 */
typedef struct FimoWorldsMeta_WorldDescriptor {
    const void *next; /* reference: reserved. Here: FimoB200_WorldBatchExt chain. */
    const char *label;
    FimoUSize label_len;
    const FimoTasksMeta_Pool *pool;
} FimoWorldsMeta_WorldDescriptor;

/* ------------------------------------------------------------------------- */
/* B200 device extension structs (new; ABI-preserving)                        */
/*                                                                            */
/* The reference reserves a Vulkan-style `next` pointer on every descriptor   */
/* and never reads it (FimoWorlds.zig:108-168), so the same descriptors stay  */
/* valid on the reference's CPU path and on the oracle.  This library walks   */
/* the chain and picks out the structs below by `s_type`.                     */
/* ------------------------------------------------------------------------- */

#define FIMO_B200_STYPE_WORLD_BATCH 0xB2000001u
#define FIMO_B200_STYPE_RESOURCE_DEVICE 0xB2000002u
#define FIMO_B200_STYPE_SYSTEM_DEVICE 0xB2000003u
#define FIMO_B200_STYPE_GROUP_DEVICE 0xB2000004u

typedef struct FimoB200_ExtHeader {
    uint32_t s_type;
    uint32_t reserved;
    const void *next;
} FimoB200_ExtHeader;

/* WorldDescriptor.next: one `World` handle stands for `n_worlds` independent
 * worlds of `entities_per_world` entities each, resident on CUDA device `device`.
 * Without this struct a world is a batch of one world with zero entities. */
typedef struct FimoB200_WorldBatchExt {
    FimoB200_ExtHeader header;
    uint64_t n_worlds;
    uint64_t entities_per_world;
    int32_t device; /* CUDA ordinal; -1 = current device */
    uint32_t flags;
} FimoB200_WorldBatchExt;

/* How many elements of a device resource one world owns. */
enum {
    FIMO_B200_SCOPE_ENTITY = 0,      /* count = entities_per_world */
    FIMO_B200_SCOPE_WORLD = 1,       /* count = 1 (per-world constants: frustum, dt, counters) */
    FIMO_B200_SCOPE_ENTITY_BITS = 2, /* count = ceil(entities_per_world / 32) u32 words (bit masks) */
    FIMO_B200_SCOPE_BATCH = 3,       /* count = 1 and a single instance for the whole batch (statistics) */
};

/* ResourceDescriptor.next: the resource is an SoA HBM buffer.  The registered
 * `size` must be sizeof(FimoB200_ResourceView): the host-visible blob of a device
 * resource is a fixed-size view, the data lives in HBM. */
typedef struct FimoB200_ResourceDeviceExt {
    FimoB200_ExtHeader header;
    uint32_t scope;      /* FIMO_B200_SCOPE_* */
    uint32_t components; /* 4-byte words per element: pos 3, rot 4, world 12, frustum 24 ... */
    uint32_t elem_bytes; /* must be 4 */
    uint32_t flags;
} FimoB200_ResourceDeviceExt;

/* The blob of a device resource.  Layout of the data, on host and device alike:
 * dense 4-byte words [n_worlds][components][count] (world-major, then component
 * plane, then element), i.e. comp_stride = 4*count, world_stride = 4*components*count.
 *
 * world_add_resource(world, handle, &view):   in: host_ptr = initial contents or NULL (zero-fill);
 *                                             device_ptr = NULL (library allocates) or external HBM to adopt.
 * world_lock_resources(...) -> FimoB200_ResourceView*: all fields valid, device idle for this resource.
 * world_remove_resource(world, handle, &view): in: view.host_ptr = destination for a final D2H copy or NULL;
 *                                             out: the view (device_ptr is dangling unless external).
 * On the CPU path (reference / oracle) `host_ptr` IS the data and n_worlds == 1. */
typedef struct FimoB200_ResourceView {
    void *device_ptr;
    void *host_ptr;
    uint64_t n_worlds;
    uint64_t count;
    uint32_t components;
    uint32_t elem_bytes;
    uint64_t comp_stride;
    uint64_t world_stride;
    uint32_t scope;
    uint32_t flags; /* bit0: device_ptr is external (not owned) */
} FimoB200_ResourceView;
#define FIMO_B200_VIEW_FLAG_EXTERNAL 1u

typedef struct FimoB200_LaunchInfo {
    void *stream; /* cudaStream_t to launch on */
    uint64_t n_worlds;
    uint64_t entities_per_world;
    int32_t device;
    uint32_t flags; /* bit0: stream is being captured into a CUDA graph (no syncs, no allocations) */
    uint64_t reserved;
} FimoB200_LaunchInfo;
#define FIMO_B200_LAUNCH_FLAG_CAPTURING 1u

/* Device entry point of a system: enqueue the system's kernels for ALL worlds of the
 * batch on `info->stream`.  Same argument meaning and order as `system_run`
 * (systems.h:103-113), with resource views instead of blob pointers and the
 * stream taking the place of the deferred fence (stream order == completion). */
typedef FimoStatus (*FimoB200_device_entry)(void *system, const FimoB200_ResourceView *const *exclusive_resources,
                                            const FimoB200_ResourceView *const *shared_resources,
                                            const FimoB200_LaunchInfo *info);

/* Built-in synthetic systems.  A non-zero id lets the DAG
 * compiler fuse neighbouring built-ins into one world-step kernel. */
enum {
    FIMO_B200_BUILTIN_NONE = 0,
    FIMO_B200_BUILTIN_INTEGRATE = 1, /* X: pos            S: vel, dt */
    FIMO_B200_BUILTIN_PROPAGATE = 2, /* X: world          S: pos, rot, scale, parent, level_start */
    FIMO_B200_BUILTIN_CULL = 3,      /* X: vis_mask, vis_list, vis_count, stats   S: world, aabb_c, aabb_h, frustum */
    /* the 8-system fan-out / fan-in DAG (BASELINE.json configs[2]) adds: */
    FIMO_B200_BUILTIN_FORCES = 4,        /* X: vel            S: gravity (world,3), dt          vel += gravity*dt */
    FIMO_B200_BUILTIN_INTEGRATE_ANG = 5, /* X: rot            S: angvel (entity,3), dt          q += dt/2 * (w (x) q), renormalised */
    FIMO_B200_BUILTIN_WORLD_AABB = 6,    /* X: waabb (entity,6: centre, half extent)   S: world, aabb_c, aabb_h */
    FIMO_B200_BUILTIN_CULL_WAABB = 7,    /* X: vis_mask, vis_list, vis_count, stats   S: waabb, frustum */
    FIMO_B200_BUILTIN_BOUNDS_REDUCE = 8, /* X: wbounds (world,6: min xyz, max xyz)    S: waabb */
    FIMO_B200_BUILTIN_GROUP_STATS = 9,   /* X: gstats (batch,16 words)               S: vis_count, wbounds */
    /* the 16-system DAG with mixed read/write conflicts (configs[4]) adds: */
    FIMO_B200_BUILTIN_PREV_COPY = 10,    /* X: prev_world     S: world                         last step's transforms */
    FIMO_B200_BUILTIN_LOD = 11,          /* X: lod (entity,1 u32)   S: waabb, campos (world,3), lod_dist2 (world,3) */
    FIMO_B200_BUILTIN_VEL_CLAMP = 12,    /* X: vel            S: vmax (world,1) */
    FIMO_B200_BUILTIN_SLEEP_FLAGS = 13,  /* X: sleep (entity bits)  S: vel, sleep_eps (world,1) */
    FIMO_B200_BUILTIN_ANGVEL_DAMP = 14,  /* X: angvel         S: damp (world,1) */
    FIMO_B200_BUILTIN_VIS_BOUNDS = 15,   /* X: vbounds (world,6)    S: waabb, vis_mask           bounds of the visible entities */
    FIMO_B200_BUILTIN_COUNT_
};

/* SystemDescriptor.next: the system runs on the device. `system_run` must still be
 * non-null (ABI); this library never calls it for a device system. */
typedef struct FimoB200_SystemDeviceExt {
    FimoB200_ExtHeader header;
    FimoB200_device_entry device_entry;
    uint32_t builtin_id; /* FIMO_B200_BUILTIN_* */
    uint32_t flags;
} FimoB200_SystemDeviceExt;

/* SystemGroupDescriptor.next: executor options. */
typedef void (*FimoB200_post_step)(void *user, void *stream, uint64_t generation);
typedef struct FimoB200_GroupDeviceExt {
    FimoB200_ExtHeader header;
    uint32_t flags;
    uint32_t reserved;
    /* Called on the launcher thread right after generation `generation` was enqueued on
     * `stream` (e.g. to enqueue the end-of-step statistics allreduce). May be NULL. */
    FimoB200_post_step post_step;
    void *post_step_user;
    /* Read only with FIMO_B200_GROUP_FLAG_STATS_RING: device memory [stats_ring_slots][FIMO_B200_STAT_WORDS] u64.
     * The kernel that publishes the statistics of the group's first cull system also stores them in slot
     * `generation % stats_ring_slots` (generation counted from the group's first one) — a per-step snapshot
     * without an extra launch or copy, e.g. the send buffer of the end-of-step statistics all-reduce. */
    void *stats_ring;
    uint32_t stats_ring_slots;
    uint32_t reserved2;
} FimoB200_GroupDeviceExt;
#define FIMO_B200_GROUP_FLAG_NO_FUSION 1u /* one kernel per system (CUDA graph of the DAG) */
#define FIMO_B200_GROUP_FLAG_NO_GRAPH 2u  /* plain stream launches, no CUDA graph */
#define FIMO_B200_GROUP_FLAG_STATS_RING 4u /* stats_ring / stats_ring_slots are valid */

/* Statistics words written by BUILTIN_CULL into the `stats` resource (u64 each, 8 per batch). */
enum {
    FIMO_B200_STAT_VISIBLE = 0,  /* sum over worlds of vis_count, last step */
    FIMO_B200_STAT_ENTITIES = 1, /* entities tested, last step */
    FIMO_B200_STAT_MASK_SUM = 2, /* sum of all vis_mask words (as u64), last step */
    FIMO_B200_STAT_WORLDS = 3,   /* worlds stepped, last step */
    FIMO_B200_STAT_STEPS = 4,    /* world-steps accumulated since the resource was added (+1 per world per step) */
    FIMO_B200_STAT_WORDS = 8
};
/* Words written by BUILTIN_GROUP_STATS into `gstats` (16 x u32 per batch). */
enum {
    FIMO_B200_GSTAT_VISIBLE_LO = 0, /* u64: sum over worlds of vis_count */
    FIMO_B200_GSTAT_WORLDS_LO = 2,  /* u64: worlds reduced */
    FIMO_B200_GSTAT_BOUNDS = 4,     /* 6 x f32: min xyz, max xyz over all worlds' wbounds */
    FIMO_B200_GSTAT_WORDS = 16
};
#define FIMO_B200_BOUNDS_EMPTY_MIN 3.402823466e+38f   /* bounds of nothing: min = +FLT_MAX, max = -FLT_MAX */
#define FIMO_B200_PARENT_ROOT 0xFFFFFFFFu

/* ------------------------------------------------------------------------- */
/* The 32 `fimo-worlds` entry points                                          */
/* (pkgs/fimo_worlds_meta/src/symbols.zig:22-58 for the list, :60-365 for     */
/*  the signatures; implementation shims modules/fimo_worlds/src/FimoWorlds.zig:108-447) */
/* ------------------------------------------------------------------------- */

/* symbols.zig:60-76   resources.h:32-39 */
FIMO_B200_EXPORT FimoStatus fimo_worlds_resource_register(const FimoWorldsMeta_ResourceDescriptor *resource,
                                                          FimoWorldsMeta_ResourceHandle *handle);
FIMO_B200_EXPORT void fimo_worlds_resource_unregister(FimoWorldsMeta_ResourceHandle handle);

/* symbols.zig:78-94   systems.h:160-167 */
FIMO_B200_EXPORT FimoStatus fimo_worlds_system_register(const FimoWorldsMeta_SystemDescriptor *system,
                                                        FimoWorldsMeta_SystemHandle *handle);
FIMO_B200_EXPORT void fimo_worlds_system_unregister(FimoWorldsMeta_SystemHandle handle);

/* symbols.zig:96-166   systems.h:170-223 */
FIMO_B200_EXPORT FimoStatus fimo_worlds_system_group_create(const FimoWorldsMeta_SystemGroupDescriptor *descriptor,
                                                            FimoWorldsMeta_SystemGroup **group);
FIMO_B200_EXPORT void fimo_worlds_system_group_destroy(FimoWorldsMeta_SystemGroup *group);
FIMO_B200_EXPORT FimoWorldsMeta_World *fimo_worlds_system_group_get_world(FimoWorldsMeta_SystemGroup *group);
FIMO_B200_EXPORT const char *fimo_worlds_system_group_get_label(FimoWorldsMeta_SystemGroup *group, FimoUSize *len);
FIMO_B200_EXPORT FimoTasksMeta_Pool fimo_worlds_system_group_get_pool(FimoWorldsMeta_SystemGroup *group);
FIMO_B200_EXPORT FimoStatus fimo_worlds_system_group_add_systems(FimoWorldsMeta_SystemGroup *group,
                                                                 const FimoWorldsMeta_SystemHandle *systems,
                                                                 FimoUSize systems_len);
FIMO_B200_EXPORT void fimo_worlds_system_group_remove_system(FimoWorldsMeta_SystemGroup *group,
                                                             FimoWorldsMeta_SystemHandle handle,
                                                             FimoWorldsMeta_Fence *signal);
FIMO_B200_EXPORT FimoStatus fimo_worlds_system_group_schedule(FimoWorldsMeta_SystemGroup *group,
                                                              FimoWorldsMeta_Fence *const *wait_on,
                                                              FimoUSize wait_on_len, FimoWorldsMeta_Fence *signal);

/* symbols.zig:168-235   systems.h:226-272 */
FIMO_B200_EXPORT FimoWorldsMeta_SystemGroup *fimo_worlds_system_context_get_group(FimoWorldsMeta_SystemContext *context);
FIMO_B200_EXPORT FimoUSize fimo_worlds_system_context_get_generation(FimoWorldsMeta_SystemContext *context);
FIMO_B200_EXPORT void *fimo_worlds_system_context_allocator_alloc(FimoWorldsMeta_SystemContext *context,
                                                                  FimoWorldsMeta_SystemAllocatorStrategy strategy,
                                                                  FimoUSize size, FimoUSize alignment,
                                                                  FimoUSize ret_addr);
FIMO_B200_EXPORT bool fimo_worlds_system_context_allocator_resize(FimoWorldsMeta_SystemContext *context,
                                                                  FimoWorldsMeta_SystemAllocatorStrategy strategy,
                                                                  void *ptr, FimoUSize size, FimoUSize alignment,
                                                                  FimoUSize new_size, FimoUSize ret_addr);
FIMO_B200_EXPORT void *fimo_worlds_system_context_allocator_remap(FimoWorldsMeta_SystemContext *context,
                                                                  FimoWorldsMeta_SystemAllocatorStrategy strategy,
                                                                  void *ptr, FimoUSize size, FimoUSize alignment,
                                                                  FimoUSize new_size, FimoUSize ret_addr);
FIMO_B200_EXPORT void fimo_worlds_system_context_allocator_free(FimoWorldsMeta_SystemContext *context,
                                                                FimoWorldsMeta_SystemAllocatorStrategy strategy,
                                                                void *ptr, FimoUSize size, FimoUSize alignment,
                                                                FimoUSize ret_addr);

/* symbols.zig:237-365   worlds.h:35-118 */
FIMO_B200_EXPORT FimoStatus fimo_worlds_world_create(const FimoWorldsMeta_WorldDescriptor *descriptor,
                                                     FimoWorldsMeta_World **world);
FIMO_B200_EXPORT void fimo_worlds_world_destroy(FimoWorldsMeta_World *world);
FIMO_B200_EXPORT const char *fimo_worlds_world_get_label(FimoWorldsMeta_World *world, FimoUSize *len);
FIMO_B200_EXPORT FimoTasksMeta_Pool fimo_worlds_world_get_pool(FimoWorldsMeta_World *world);
FIMO_B200_EXPORT bool fimo_worlds_world_has_resource(FimoWorldsMeta_World *world, FimoWorldsMeta_ResourceHandle handle);
FIMO_B200_EXPORT FimoStatus fimo_worlds_world_add_resource(FimoWorldsMeta_World *world,
                                                           FimoWorldsMeta_ResourceHandle handle, const void *value);
FIMO_B200_EXPORT FimoStatus fimo_worlds_world_remove_resource(FimoWorldsMeta_World *world,
                                                              FimoWorldsMeta_ResourceHandle handle, void *value);
FIMO_B200_EXPORT void fimo_worlds_world_lock_resources(FimoWorldsMeta_World *world,
                                                       const FimoWorldsMeta_ResourceHandle *exclusive_handles,
                                                       FimoUSize exclusive_handles_len,
                                                       const FimoWorldsMeta_ResourceHandle *shared_handles,
                                                       FimoUSize shared_handles_len, void **out_resources);
FIMO_B200_EXPORT void fimo_worlds_world_unlock_resource_exclusive(FimoWorldsMeta_World *world,
                                                                  FimoWorldsMeta_ResourceHandle handle);
FIMO_B200_EXPORT void fimo_worlds_world_unlock_resource_shared(FimoWorldsMeta_World *world,
                                                               FimoWorldsMeta_ResourceHandle handle);
FIMO_B200_EXPORT void *fimo_worlds_world_allocator_alloc(FimoWorldsMeta_World *world, FimoUSize size,
                                                         FimoUSize alignment, FimoUSize ret_addr);
FIMO_B200_EXPORT bool fimo_worlds_world_allocator_resize(FimoWorldsMeta_World *world, void *ptr, FimoUSize size,
                                                         FimoUSize alignment, FimoUSize new_size, FimoUSize ret_addr);
FIMO_B200_EXPORT void *fimo_worlds_world_allocator_remap(FimoWorldsMeta_World *world, void *ptr, FimoUSize size,
                                                         FimoUSize alignment, FimoUSize new_size, FimoUSize ret_addr);
FIMO_B200_EXPORT void fimo_worlds_world_allocator_free(FimoWorldsMeta_World *world, void *ptr, FimoUSize size,
                                                       FimoUSize alignment, FimoUSize ret_addr);

/* ------------------------------------------------------------------------- */
/* Module plumbing that the fimo loader provides in the reference             */
/* ------------------------------------------------------------------------- */

/* One row of the export table: what `fimo_exports` lists in FimoWorlds.zig:48-81. */
typedef struct FimoB200_Export {
    const char *name;       /* e.g. "system_group_schedule" */
    const char *name_space; /* "fimo-worlds" (package.h:12) */
    void (*function)(void);
} FimoB200_Export;
FIMO_B200_EXPORT const FimoB200_Export *fimo_worlds_b200_exports(FimoUSize *count);

/* The module record the fimo loader looks for.  The reference's loader `dlopen`s `module.fimo_module` and calls
 * `fimo_impl_module_export_iterator` (pkgs/fimo_std/include/fimo_std/impl/module.h:40,61-62; Zig side
 * pkgs/fimo_std/src/modules/exports.zig:2016-2044), which walks the pointers stored in the ELF section
 * `fimo_module` (`__start_fimo_module` .. `__stop_fimo_module`, NULL entries skipped) and hands each
 * `FimoModulesExport` (modules.h:942-1002) to the inspector.  This library embeds ONE such record: module
 * "fimo_worlds", namespace import "fimo-tasks", symbol imports `futex_wait` / `futex_wake`
 * (pkgs/fimo_tasks_meta/src/symbols.zig:132-160), and the 32 "fimo-worlds" symbols as static exports
 * (the function pointers of the table above), versioned with the context version the build was given
 * (-DFIMO_B200_CONTEXT_VERSION_{MAJOR,MINOR,PATCH}; default 0.2.0 = build.zig.zon of the reference).
 * The structs below mirror modules.h / version.h byte for byte. */
typedef struct FimoB200_Version { /* FimoVersion, version.h:14-22 */
    FimoUSize major, minor, patch;
    const char *pre;
    FimoUSize pre_len;
    const char *build;
    FimoUSize build_len;
} FimoB200_Version;
typedef struct FimoB200_ModulesNamespaceImport { /* modules.h: FimoModulesNamespaceImport */
    const char *name;
} FimoB200_ModulesNamespaceImport;
typedef struct FimoB200_ModulesSymbolImport { /* FimoModulesSymbolImport */
    FimoB200_Version version;
    const char *name;
    const char *ns;
} FimoB200_ModulesSymbolImport;
typedef struct FimoB200_ModulesSymbolExport { /* FimoModulesSymbolExport */
    const void *symbol;
    int32_t linkage; /* FIMO_MODULES_SYMBOL_LINKAGE_GLOBAL = 0 */
    FimoB200_Version version;
    const char *name;
    const char *ns;
} FimoB200_ModulesSymbolExport;
typedef struct FimoB200_ModulesExport { /* struct FimoModulesExport, modules.h:942-1002 */
    const void *next;
    FimoB200_Version version;
    const char *name, *description, *author, *license;
    const void *parameters;
    FimoUSize parameters_count;
    const void *resources;
    FimoUSize resources_count;
    const FimoB200_ModulesNamespaceImport *namespace_imports;
    FimoUSize namespace_imports_count;
    const FimoB200_ModulesSymbolImport *symbol_imports;
    FimoUSize symbol_imports_count;
    const FimoB200_ModulesSymbolExport *symbol_exports;
    FimoUSize symbol_exports_count;
    const void *dynamic_symbol_exports;
    FimoUSize dynamic_symbol_exports_count;
    const void *modifiers;
    FimoUSize modifiers_count;
} FimoB200_ModulesExport;
typedef bool (*FimoB200_ModuleInspector)(const FimoB200_ModulesExport *module, void *data);
FIMO_B200_EXPORT void fimo_impl_module_export_iterator(FimoB200_ModuleInspector inspector, void *data);

/* Name of the error behind the last FIMO_STATUS_ERROR on this thread
 * ("NotFound", "Deadlock", "Duplicate", "InUse", "InitError", "OutOfMemory", ...).
 * Stand-in for the thread-local context result (pkgs/fimo_std/src/ctx.zig:184-186). */
FIMO_B200_EXPORT const char *fimo_worlds_b200_last_error(void);

/* `fimo-tasks` futex symbols the path imports (symbols.zig:132-160).  The library
 * ships a standalone parking-lot implementation; a host application that runs the
 * real fimo_tasks module injects its own with `bind_tasks` so that fibers park
 * through their pool. */
FIMO_B200_EXPORT FimoTasksMeta_FutexStatus fimo_tasks_futex_wait(const void *key, FimoUSize key_size, FimoU64 expect,
                                                                 FimoUSize token, const FimoInstant *timeout);
FIMO_B200_EXPORT FimoUSize fimo_tasks_futex_wake(const void *key, FimoUSize max_waiters,
                                                 FimoTasksMeta_FutexFilter filter);
FIMO_B200_EXPORT void fimo_worlds_b200_bind_tasks(FimoTasksMeta_futex_wait futex_wait,
                                                  FimoTasksMeta_futex_wake futex_wake);

/* Fence helpers bound to the library's futex symbols (what Job.zig:33-62 does
 * through `symbols.futex_*`). */
FIMO_B200_EXPORT void fimo_worlds_b200_fence_wait(FimoWorldsMeta_Fence *fence);
FIMO_B200_EXPORT void fimo_worlds_b200_fence_signal(FimoWorldsMeta_Fence *fence);

/* `create_worker_pool` of fimo-tasks (symbols.zig:35-52, pool.h:145-148), Status-returning
 * as in the Zig implementation.  Needed by worlds/groups that are given an explicit pool. */
FIMO_B200_EXPORT FimoStatus fimo_tasks_create_worker_pool(const FimoTasksMeta_PoolConfig *config,
                                                          FimoTasksMeta_Pool *pool);

/* `abort` of fimo-tasks (symbols.zig:94-99; modules/fimo_tasks/src/Worker.zig:115-122): ends the calling
 * task in error and never returns.  The task's on_abort hook runs instead of on_complete, and its command
 * buffer applies abort_on_error (modules/fimo_tasks/src/CommandBuffer.zig:127-148,468-513).  Calling it
 * outside a task of this library's pools is a contract violation. */
FIMO_B200_EXPORT void fimo_tasks_abort(void);
/* `worker_id` of fimo-tasks (symbols.zig:60-65): index of the calling worker inside its pool; false on any
 * other thread.  Tasks enqueued after a select_worker entry observe exactly that index. */
FIMO_B200_EXPORT bool fimo_tasks_worker_id(FimoTasksMeta_PoolWorker *id);

/* ------------------------------------------------------------------------- */
/* Device-side helpers (product library only; absent from the CPU oracle)     */
/* ------------------------------------------------------------------------- */

/* Fills `desc` (system_size/alignment, create/destroy/run) and `ext` (device entry,
 * builtin id) for a built-in synthetic system and chains `ext` on `desc->next`.
 * The caller still fills label, resource handles and dependencies.  `system_run`
 * is a stub that aborts: built-in systems exist on the device only (no CPU fallback). */
FIMO_B200_EXPORT FimoStatus fimo_b200_builtin_system(uint32_t builtin_id, FimoWorldsMeta_SystemDescriptor *desc,
                                                     FimoB200_SystemDeviceExt *ext);

/* Synchronous copies between a locked device resource and dense host memory
 * laid out as the view describes. */
FIMO_B200_EXPORT FimoStatus fimo_b200_resource_upload(const FimoB200_ResourceView *view, const void *host_src);
FIMO_B200_EXPORT FimoStatus fimo_b200_resource_download(const FimoB200_ResourceView *view, void *host_dst);

/* Stream-ordered transfer of one resource used by the group's systems between HBM and host memory, queued
 * like a generation (system_group_schedule, symbols.zig:150-166): it runs after every generation / transfer
 * queued before it and before everything queued after it, on the group's stream, without the host taking
 * the resource lock (the group holds its resources while work is in flight, SystemGroup.zig:340-351).  It
 * does not consume a generation number.  `host` holds the whole resource ([n_worlds][components][count]
 * words) and must stay valid until `signal` (nullable) is signaled; use pinned memory
 * (fimo_b200_host_alloc, or any cudaHostRegister'ed / torch-pinned buffer) for a truly asynchronous copy. */
enum { FIMO_B200_TRANSFER_HOST_TO_DEVICE = 0, FIMO_B200_TRANSFER_DEVICE_TO_HOST = 1 };
FIMO_B200_EXPORT FimoStatus fimo_b200_group_transfer(FimoWorldsMeta_SystemGroup *group,
                                                     FimoWorldsMeta_ResourceHandle handle, void *host,
                                                     uint32_t direction, FimoWorldsMeta_Fence *signal);
/* One generation with its own host inputs and host results — system_group_schedule (symbols.zig:150-166,
 * SystemGroup.zig:430-452) bracketed by stream-ordered copies: `uploads` are copied host -> HBM in front of the
 * generation, `downloads` HBM -> host behind it, all on the group's stream in this order and ordered with every
 * other generation / transfer of the group.  It consumes one generation number; `signal` (nullable) fires once
 * the generation has completed AND the downloads have landed.  Buffers follow fimo_b200_group_transfer's rules
 * (whole resource, valid until `signal`; pinned memory for asynchronous copies).  Several such steps may be in
 * flight; each needs its own host buffers, and the caller must not write an upload buffer between the call (once
 * `wait_on` is satisfied) and `signal`: pinned uploads are read early, into HBM staging slots beside the generation
 * still running, and reach the resources in stream order in front of their own generation.  The one writer that IS
 * honoured is the group itself: an upload that reads host memory which a download or transfer queued earlier has
 * yet to write (a result fed back as the next step's input) stays in stream order behind that copy. */
typedef struct FimoB200_StepTransfer {
    FimoWorldsMeta_ResourceHandle resource;
    void *host;
} FimoB200_StepTransfer;
FIMO_B200_EXPORT FimoStatus fimo_b200_group_schedule_io(FimoWorldsMeta_SystemGroup *group,
                                                        const FimoB200_StepTransfer *uploads, FimoUSize n_uploads,
                                                        const FimoB200_StepTransfer *downloads, FimoUSize n_downloads,
                                                        FimoWorldsMeta_Fence *const *wait_on, FimoUSize n_wait,
                                                        FimoWorldsMeta_Fence *signal);
/* One generation whose CULL OUTPUTS come back to the host: what `cull` produces for a frame is the per-world visible
 * count, the visibility mask and the ordered list of visible entities — of which only the first `count` entries of
 * each world's `vis_list` are meaningful.  Behind the generation one small pack (two kernels, in stream order) writes
 *     u32 header[8] = {total visible, n_worlds, mask_words, list capacity (entries), truncated (0/1), generation lo, hi, entities}
 *     u32 counts[n_worlds]; u32 offsets[n_worlds] (exclusive prefix of counts)
 *     u32 mask[n_worlds][mask_words]
 *     u32 list[capacity]            the worlds' ordered visible lists back to back (world w at offsets[w])
 * into an HBM staging slot, and ONE copy-engine transfer brings `host_bytes` of it to `host` (page-locked) on a side
 * stream while the next generation already runs.  `capacity` = what fits in host_bytes (at most n_worlds * entities);
 * if more entities are visible the list is cut there and `truncated` is set (counts, offsets and masks are always
 * complete).  Same ordering, generation-number and `signal` rules as fimo_b200_group_schedule_io. */
typedef struct FimoB200_CullReadback {
    FimoWorldsMeta_ResourceHandle vis_mask, vis_list, vis_count;
    void *host;
    uint64_t host_bytes;
} FimoB200_CullReadback;
#define FIMO_B200_READBACK_HEADER_WORDS 8
FIMO_B200_EXPORT FimoStatus fimo_b200_group_schedule_readback(FimoWorldsMeta_SystemGroup *group,
                                                              const FimoB200_StepTransfer *uploads, FimoUSize n_uploads,
                                                              const FimoB200_CullReadback *readback,
                                                              FimoWorldsMeta_Fence *const *wait_on, FimoUSize n_wait,
                                                              FimoWorldsMeta_Fence *signal);
FIMO_B200_EXPORT void *fimo_b200_host_alloc(FimoUSize bytes); /* page-locked host memory; NULL on failure */
FIMO_B200_EXPORT void fimo_b200_host_free(void *ptr);

/* The packed cull read-back for callers that must not allocate device memory.  No counterpart in the reference
 * (it has neither a device nor a read-back; this is a B200-side convenience next to fimo_b200_group_schedule_readback).
 *
 * Who needs it: fimo_b200_group_schedule_readback packs into HBM staging slots that the library cudaMalloc's on
 * first use and again whenever a larger buffer is asked for.  Some hosts cannot tolerate that: a frame loop in
 * which cudaMalloc / cudaFree would stall every stream of the device behind their implicit synchronisation; an
 * engine that claimed the whole HBM budget for its own pool at start-up, so that the driver has nothing left to
 * hand out; a plug-in host that forbids driver allocations once it is past initialisation.  Such a caller owns the
 * cull outputs already (its own device buffers, or the `device_ptr` of locked resource views) and only wants them
 * packed.  This call uses no cudaMalloc, cudaFree or any other allocating routine, on any call.
 *
 * What it does, all queued on `stream` (NULL = the default stream) of the device that is CURRENT on the calling
 * thread, to which every device pointer must belong:
 *   1. clears the list part of the staging area (see "shared" below),
 *   2. one kernel packs `vis_count[n_worlds]`, `vis_mask[n_worlds][mask_words]` and the first vis_count[w] entries
 *      of every `vis_list[n_worlds][entities]` row into the layout documented at FimoB200_CullReadback
 *      (header[8], counts, offsets, masks, lists back to back; `generation` is copied into the header),
 *   3. one copy brings header + counts + offsets + masks + `capacity` list entries to `out`, which may be host
 *      memory (page-locked if the copy is to be asynchronous) or device memory.  capacity = what fits in
 *      `out_bytes`, at most n_worlds * entities; a longer list is cut there and header[4] (`truncated`) is set.
 * It returns once the work is queued; `out` is valid when `stream` has reached that point.
 *
 * Limits: the staging area is a fixed array inside the library's device image — FIMO_B200_PACK_STATIC_MAX_BYTES
 * (8 MiB) of HBM on every GPU that loads the library, held until the process exits.  A call whose packed size
 * (4 * (8 + 2 n_worlds + n_worlds * mask_words + capacity) bytes) exceeds it, or with more than
 * FIMO_B200_PACK_STATIC_MAX_WORLDS worlds, fails with FIMO_STATUS_ERROR and fimo_worlds_b200_last_error() ==
 * "TooLarge" (the numbers go to stderr); nothing is queued.  Hand in a smaller `out_bytes` (a truncated list) or
 * pack the batch in slices of worlds.
 *
 * Shared: every call on a device uses the same array, and it keeps what the previous call left there.  Step 1
 * zeroes the list entries this call may copy out, so entries past the packed total read as 0 and never as another
 * call's indices; everything else that is copied is rewritten by step 2.  Two calls that overlap on the GPU (two
 * streams, two threads, two groups) race on the array: order them on one stream or with events.
 *
 * First call on a device: looks up the array's device address (cudaGetSymbolAddress, which also makes the driver
 * load the kernel image and with it the array).  Make that first call during start-up, before any stream capture is
 * begun on the thread. */
#define FIMO_B200_PACK_STATIC_MAX_BYTES (8u << 20)
#define FIMO_B200_PACK_STATIC_MAX_WORLDS 8192u
FIMO_B200_EXPORT FimoStatus fimo_b200_cull_pack_static(const uint32_t *vis_count, const uint32_t *vis_mask,
                                                       const uint32_t *vis_list, uint64_t n_worlds, uint64_t entities,
                                                       uint64_t mask_words, uint64_t generation, void *out,
                                                       uint64_t out_bytes, void *stream);

/* Tracing bridge.  The reference records spans and events through the context's tracing
 * subsystem (pkgs/fimo_std/src/tracing.zig:387-554: enter_span / exit_span / log_message); device work has no
 * call stack there.  With a sink installed the completion thread reports every finished device generation —
 * group label, generation number, device time between the CUDA events that bracket it, host timestamps
 * (CLOCK_MONOTONIC ns) of the enqueue and of the observed completion, kernels launched — and the module shim
 * forwards it as a span / event of its own subscriber.  The sink runs outside any CUDA callback. */
typedef struct FimoB200_TraceEvent {
    const char *group_label;
    FimoUSize group_label_len;
    uint64_t generation;
    double device_ms;
    int64_t host_enqueue_ns;
    int64_t host_complete_ns;
    uint32_t kernels;
    uint32_t fused_world_step;
} FimoB200_TraceEvent;
typedef void (*FimoB200_trace_sink)(void *user, const FimoB200_TraceEvent *event);
FIMO_B200_EXPORT void fimo_worlds_b200_set_trace_sink(FimoB200_trace_sink sink, void *user);

/* The same information as fimo tracing events.  A fimo tracing *subscriber* (pkgs/fimo_std/include/fimo_std/
 * tracing.h:199-204; tracing.zig:562-565: `{data, on_event(data, *const Event)}`) receives extern structs that
 * start with the `Event` tag (tracing.zig:387-554, tracing.h:72-192).  The structs below are byte-identical
 * mirrors of the ones the bridge emits (tests/golden/abi_layout.json pins them against the reference header), so
 * the module shim can hand the context's own subscribers — the `.ftrdb` writer included — to
 * fimo_worlds_b200_trace_subscribe and device generations show up as spans:
 *
 *   first generation of a group : declare_event_info(span id), create_call_stack(stack = one token per group)
 *   every generation            : enter_span(time = device start, stack, span id, message
 *                                 "<label> generation <n>: <k> kernels, <t> ms on device"),
 *                                 [log_message per kernel-level record when the executor collected one],
 *                                 exit_span(time = device end, stack, is_unwinding = false)
 *   unsubscribe / group destroy : destroy_call_stack(stack)
 *
 * Times are FimoInstant on the host's monotonic clock (the clock fimo_instant_now reads, time.zig:363-420):
 * exit = the completion thread's observation of the end event, enter = exit - device time between the two CUDA
 * events that bracket the generation.  Events of one group are delivered in generation order from the completion
 * thread, never from inside a CUDA callback. */
typedef FimoI32 FimoTracingLevel; /* tracing.h:18-25 */
enum {
    FIMO_TRACING_LEVEL_OFF = 0,
    FIMO_TRACING_LEVEL_ERROR = 1,
    FIMO_TRACING_LEVEL_WARN = 2,
    FIMO_TRACING_LEVEL_INFO = 3,
    FIMO_TRACING_LEVEL_DEBUG = 4,
    FIMO_TRACING_LEVEL_TRACE = 5,
};
typedef struct FimoTracingEventInfo { /* tracing.h:35-56 */
    const char *name;
    const char *target;
    const char *scope;
    const char *file_name;
    FimoI32 line_number;
    FimoTracingLevel level;
} FimoTracingEventInfo;
typedef uint32_t FimoTracingEvent; /* tracing.h:72-94, tracing.zig:389-411 */
enum {
    FIMO_TRACING_EVENT_START = 0,
    FIMO_TRACING_EVENT_FINISH = 1,
    FIMO_TRACING_EVENT_REGISTER_THREAD = 2,
    FIMO_TRACING_EVENT_UNREGISTER_THREAD = 3,
    FIMO_TRACING_EVENT_CREATE_CALL_STACK = 4,
    FIMO_TRACING_EVENT_DESTROY_CALL_STACK = 5,
    FIMO_TRACING_EVENT_UNBLOCK_CALL_STACK = 6,
    FIMO_TRACING_EVENT_SUSPEND_CALL_STACK = 7,
    FIMO_TRACING_EVENT_RESUME_CALL_STACK = 8,
    FIMO_TRACING_EVENT_ENTER_SPAN = 9,
    FIMO_TRACING_EVENT_EXIT_SPAN = 10,
    FIMO_TRACING_EVENT_LOG_MESSAGE = 11,
    FIMO_TRACING_EVENT_DECLARE_EVENT_INFO = 12,
};
typedef struct FimoTracingEventCreateCallStack { /* tracing.h:137-141 */
    FimoTracingEvent event;
    FimoInstant time;
    void *stack;
} FimoTracingEventCreateCallStack;
typedef struct FimoTracingEventDestroyCallStack { /* tracing.h:143-147 */
    FimoTracingEvent event;
    FimoInstant time;
    void *stack;
} FimoTracingEventDestroyCallStack;
typedef struct FimoTracingEventEnterSpan { /* tracing.h:169-176 */
    FimoTracingEvent event;
    FimoInstant time;
    void *stack;
    const FimoTracingEventInfo *span;
    const char *message;
    FimoUSize message_length;
} FimoTracingEventEnterSpan;
typedef struct FimoTracingEventExitSpan { /* tracing.h:178-183 */
    FimoTracingEvent event;
    FimoInstant time;
    void *stack;
    bool is_unwinding;
} FimoTracingEventExitSpan;
typedef struct FimoTracingEventLogMessage { /* tracing.h:185-192 */
    FimoTracingEvent event;
    FimoInstant time;
    void *stack;
    const FimoTracingEventInfo *info;
    const char *message;
    FimoUSize message_length;
} FimoTracingEventLogMessage;
typedef struct FimoTracingEventDeclareEventInfo { /* tracing.zig:505-508 */
    FimoTracingEvent event;
    const FimoTracingEventInfo *info;
} FimoTracingEventDeclareEventInfo;
typedef struct FimoTracingSubscriber { /* tracing.h:199-204 */
    void *ptr;
    void (*on_event)(void *data, const FimoTracingEvent *event);
} FimoTracingSubscriber;
/* Attaches up to 8 subscribers (copied); NULL / 0 detaches all of them (each live group's call stack is destroyed
 * first).  Independent of fimo_worlds_b200_set_trace_sink. */
FIMO_B200_EXPORT FimoStatus fimo_worlds_b200_trace_subscribe(const FimoTracingSubscriber *subscribers, FimoUSize count);

/* What the DAG compiler produced for the group's current system set. */
typedef struct FimoB200_ExecutorInfo {
    uint32_t systems;          /* member systems */
    uint32_t device_systems;   /* of which run on the device */
    uint32_t nodes;            /* executor nodes after fusion */
    uint32_t fused_world_step; /* 1 if integrate->propagate->cull was fused into one kernel */
    uint32_t uses_cuda_graph;
    uint32_t kernel_launches_per_generation;
    uint64_t edges;                      /* dependency edges between nodes */
    uint64_t algorithmic_bytes_per_step; /* live-in read once + live-out written once, 0 if unknown */
    /* > 1: the executor is ONE fused world-step node able to step several generations per launch; plain generations
     * scheduled while two launches are already on the stream then leave together, up to this many per launch
     * (FIMO_B200_GENS caps it).  Results, generation numbers and fences are unchanged; only launch boundaries go. */
    uint64_t max_generations_per_launch;
    uint64_t reserved[3];
} FimoB200_ExecutorInfo;
FIMO_B200_EXPORT FimoStatus fimo_b200_group_executor_info(FimoWorldsMeta_SystemGroup *group,
                                                          FimoB200_ExecutorInfo *info);

/* The same in words — which kernels, with which launch geometry (the text a bench line or a trace quotes).  Call
 * fimo_b200_group_executor_info first (it compiles a dirty group).  Returns the length of the full text; at most
 * buffer_len - 1 characters + NUL are written. */
FIMO_B200_EXPORT FimoUSize fimo_b200_group_executor_describe(FimoWorldsMeta_SystemGroup *group, char *buffer, FimoUSize buffer_len);

/* Kernel launches of all generations of this group enqueued so far (copies and host nodes not counted). */
FIMO_B200_EXPORT uint64_t fimo_b200_group_kernels_launched(FimoWorldsMeta_SystemGroup *group);

/* The CUDA stream (cudaStream_t) generations of this group are enqueued on. */
FIMO_B200_EXPORT void *fimo_b200_group_stream(FimoWorldsMeta_SystemGroup *group);

/* Blocks until every scheduled generation of the group has finished on the device. */
FIMO_B200_EXPORT void fimo_b200_group_wait_idle(FimoWorldsMeta_SystemGroup *group);

/* Average device time per generation in milliseconds over the `*generations` generations completed since
 * the previous call, measured with CUDA events on the group's stream around each generation (0 if none). */
FIMO_B200_EXPORT double fimo_b200_group_device_ms(FimoWorldsMeta_SystemGroup *group, uint64_t *generations);

/* Device time in milliseconds (CUDA events on the group's stream) from the start of the first to the end
 * of the last generation completed since the previous call — i.e. including any gap between generations.
 * Resets the span; call fimo_b200_group_device_ms first if both are wanted for the same window. */
FIMO_B200_EXPORT double fimo_b200_group_span_ms(FimoWorldsMeta_SystemGroup *group);

/* ------------------------------------------------------------------------- */
/* Multi-GPU: worlds sharded by index, one process per GPU    */
/*                                                                            */
/* Worlds are independent units (modules/fimo_worlds/src/World.zig:26,45-52): */
/* rank r owns a contiguous block of global world indices and steps it with   */
/* no data-path collective.  The one exchange is the end-of-step statistics   */
/* (the FIMO_B200_STAT_* words of the group's first cull system), summed over  */
/* ranks for EVERY generation, off the group's stream.  Two transports:       */
/*                                                                            */
/* PEERS — the kernel that publishes a generation's statistics (its last CTA) */
/*   also stores them straight into every peer's inbox in HBM over NVLink     */
/*   (cudaIpc-mapped peer memory): 16 eight-byte {value half, generation tag} */
/*   words per rank and generation, self-validating, so no fence, no extra    */
/*   launch and no collective kernel competing with the persistent grid for   */
/*   SM slots.  Every rank holds every rank's words; the sum is taken on read. */
/* NCCL  — the generation's snapshot goes to a ring in HBM and blocks of      */
/*   `bucket` consecutive generations are all-reduced in place               */
/*   (ncclAllReduce, ncclUint64, ncclSum) on a library side stream, enqueued  */
/*   from the launcher thread; the group's stream waits for a block's         */
/*   collective only before the kernel overwrites that block.  libnccl.so.2   */
/*   is dlopen()ed when this mode is bound (no link-time dependency).         */
/*                                                                            */
/* Both are bound after system_group_create and before the first schedule;    */
/* `generation` below counts the generations the group's first cull system    */
/* has published (== the group's generation number when that system has been  */
/* a member from the start).  All ranks must have drained (a host barrier)    */
/* before any rank destroys a group bound to peers.                           */
/* ------------------------------------------------------------------------- */
#define FIMO_B200_STATS_PEER_HANDLE_BYTES 64   /* cudaIpcMemHandle_t */
#define FIMO_B200_STATS_NCCL_ID_BYTES 128      /* ncclUniqueId */
#define FIMO_B200_STATS_MAX_RANKS 16
enum { FIMO_B200_STATS_MODE_PEERS = 1, FIMO_B200_STATS_MODE_NCCL = 2 };
/* Allocates this rank's inbox ([slots][FIMO_B200_STATS_MAX_RANKS][16] u64, zeroed) and returns the handle the
 * other ranks open; the caller exchanges the handles out of band (e.g. an all-gather of 64 bytes). */
FIMO_B200_EXPORT FimoStatus fimo_b200_group_stats_peers_open(FimoWorldsMeta_SystemGroup *group, uint32_t slots, void *handle_out);
/* `handles`: [nranks][FIMO_B200_STATS_PEER_HANDLE_BYTES], rank order (own entry ignored). */
FIMO_B200_EXPORT FimoStatus fimo_b200_group_stats_peers_bind(FimoWorldsMeta_SystemGroup *group, const void *handles,
                                                             uint32_t rank, uint32_t nranks);
FIMO_B200_EXPORT FimoStatus fimo_b200_stats_nccl_unique_id(void *id_out); /* ncclGetUniqueId; rank 0 calls, broadcasts */
/* The ring holds 4 * bucket generations.  Once bound, queued generations share a launch only up to 3 * bucket at a time
 * (a launch writes its rows before its all-reduces can be enqueued, so it must not wrap into rows still waiting for
 * theirs), and the last 3 * bucket generations stay readable. */
FIMO_B200_EXPORT FimoStatus fimo_b200_group_stats_nccl_bind(FimoWorldsMeta_SystemGroup *group, const void *unique_id,
                                                            uint32_t rank, uint32_t nranks, uint32_t bucket);
/* The statistics of `generation` summed over all ranks (8 u64 words).  PEERS: waits (bounded) until every rank's
 * words of that generation have arrived; valid for the last `slots` generations.  NCCL: collective — every rank
 * calls it for the same generation, with nothing scheduled after it; flushes the partial block. */
FIMO_B200_EXPORT FimoStatus fimo_b200_group_global_stats(FimoWorldsMeta_SystemGroup *group, uint64_t generation,
                                                         uint32_t mode, uint64_t *out_words);

/* ------------------------------------------------------------------------- */
/* Layout pins (sizes measured with gcc on x86-64)       */
/* ------------------------------------------------------------------------- */
#ifdef __cplusplus
#define FIMO_B200_STATIC_ASSERT(c, m) static_assert(c, m)
#else
#define FIMO_B200_STATIC_ASSERT(c, m) _Static_assert(c, m)
#endif
FIMO_B200_STATIC_ASSERT(sizeof(FimoWorldsMeta_SystemDescriptor) == 160, "SystemDescriptor");
FIMO_B200_STATIC_ASSERT(sizeof(FimoWorldsMeta_ResourceDescriptor) == 40, "ResourceDescriptor");
FIMO_B200_STATIC_ASSERT(sizeof(FimoWorldsMeta_SystemDependency) == 16, "SystemDependency");
FIMO_B200_STATIC_ASSERT(sizeof(FimoWorldsMeta_SystemGroupDescriptor) == 40, "SystemGroupDescriptor");
FIMO_B200_STATIC_ASSERT(sizeof(FimoWorldsMeta_WorldDescriptor) == 32, "WorldDescriptor");
FIMO_B200_STATIC_ASSERT(sizeof(FimoWorldsMeta_Fence) == 1, "Fence");
FIMO_B200_STATIC_ASSERT(sizeof(FimoWorldsMeta_TimelineSemaphore) == 8, "TimelineSemaphore");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_CommandBufferEntry) == 24, "CommandBufferEntry");
FIMO_B200_STATIC_ASSERT(offsetof(FimoTasksMeta_CommandBufferEntry, payload) == 8, "CommandBufferEntry.payload");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_Task) == 48, "Task");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_CommandBuffer) == 56, "CommandBuffer");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_Pool) == 16, "Pool");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_PoolVTable) == 80, "PoolVTable");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_CommandBufferHandle) == 16, "Handle");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_CommandBufferHandleVTable) == 32, "HandleVTable");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_PoolConfig) == 64, "PoolConfig");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_PoolConfigStackConfig) == 48, "StackConfig");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_FutexFilter) == 24, "FutexFilter");
FIMO_B200_STATIC_ASSERT(sizeof(FimoTasksMeta_FutexKeyExpect) == 32, "FutexKeyExpect");
FIMO_B200_STATIC_ASSERT(sizeof(FimoStatus) == 4, "FimoStatus");
FIMO_B200_STATIC_ASSERT(sizeof(FimoDuration) == 16 && sizeof(FimoInstant) == 16, "FimoDuration/Instant");
FIMO_B200_STATIC_ASSERT(sizeof(FimoB200_ResourceView) == 64, "ResourceView");

#ifdef __cplusplus
}
#endif

#endif /* FIMO_WORLDS_B200_H */
