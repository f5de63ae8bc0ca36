"""ctypes mirror of include/fimo_worlds_b200.h.

Harness-side only: tests, bench.py and the smoke test drive the C-ABI shared
libraries through this module.  The product is the shared library, not this file.
The same binding serves any implementation of the header (the CUDA product, or —
from tests/ and bench.py only — the CPU oracle), which is what makes the parity
tests read like the reference's own tests (pkgs/fimo_worlds_meta/src/*.zig).
"""
from __future__ import annotations

import ctypes as C
from ctypes import POINTER, c_bool, c_char_p, c_double, c_int32, c_size_t, c_uint8, c_uint32, c_uint64, c_void_p

FimoStatus = c_int32
STATUS_OK = 0

STYPE_WORLD_BATCH = 0xB2000001
STYPE_RESOURCE_DEVICE = 0xB2000002
STYPE_SYSTEM_DEVICE = 0xB2000003
STYPE_GROUP_DEVICE = 0xB2000004

SCOPE_ENTITY, SCOPE_WORLD, SCOPE_ENTITY_BITS, SCOPE_BATCH = 0, 1, 2, 3
BUILTIN_NONE, BUILTIN_INTEGRATE, BUILTIN_PROPAGATE, BUILTIN_CULL = 0, 1, 2, 3
(BUILTIN_FORCES, BUILTIN_INTEGRATE_ANG, BUILTIN_WORLD_AABB, BUILTIN_CULL_WAABB, BUILTIN_BOUNDS_REDUCE, BUILTIN_GROUP_STATS,
 BUILTIN_PREV_COPY, BUILTIN_LOD, BUILTIN_VEL_CLAMP, BUILTIN_SLEEP_FLAGS, BUILTIN_ANGVEL_DAMP, BUILTIN_VIS_BOUNDS) = range(4, 16)
GSTAT_VISIBLE_LO, GSTAT_WORLDS_LO, GSTAT_BOUNDS, GSTAT_WORDS = 0, 2, 4, 16
STAT_VISIBLE, STAT_ENTITIES, STAT_MASK_SUM, STAT_WORLDS, STAT_STEPS, STAT_WORDS = 0, 1, 2, 3, 4, 8
PARENT_ROOT = 0xFFFFFFFF
DEP_WEAK, DEP_IGNORE_DEFERRED = 1, 2
GROUP_FLAG_NO_FUSION, GROUP_FLAG_NO_GRAPH, GROUP_FLAG_STATS_RING = 1, 2, 4
VIEW_FLAG_EXTERNAL = 1

ALLOC_TRANSIENT, ALLOC_SINGLE_GENERATION, ALLOC_MULTI_GENERATION, ALLOC_SYSTEM_PERSISTENT = 0, 1, 2, 3


class Fence(C.Structure):
    """jobs.h:15-17 — one atomic byte."""

    _fields_ = [("state", c_uint8)]


class TimelineSemaphore(C.Structure):
    _fields_ = [("state", c_uint64)]


class FutexFilter(C.Structure):
    """futex.h:36-40; passed BY VALUE to futex_wake."""

    _fields_ = [("op", c_size_t), ("token_mask", c_size_t), ("cmp_arg", c_size_t)]


class Task(C.Structure):
    pass


TaskFn = C.CFUNCTYPE(None, POINTER(Task))
Task._fields_ = [
    ("label", c_char_p),
    ("label_len", c_size_t),
    ("on_start", TaskFn),
    ("on_complete", TaskFn),
    ("on_abort", TaskFn),
    ("on_deinit", TaskFn),
]


class CommandBufferHandle(C.Structure):
    _fields_ = [("data", c_void_p), ("vtable", c_void_p)]


class CommandBufferEntryPayload(C.Union):
    _fields_ = [
        ("abort_on_error", c_bool),
        ("set_min_stack_size", c_size_t),
        ("select_worker", c_size_t),
        ("enqueue_task", POINTER(Task)),
        ("enqueue_command_buffer", c_void_p),
        ("wait_on_command_buffer", CommandBufferHandle),
        ("wait_on_command_indirect", c_size_t),
    ]


class CommandBufferEntry(C.Structure):
    _fields_ = [("type", c_int32), ("payload", CommandBufferEntryPayload)]


# command_buffer.h:44-54
(ENTRY_ABORT_ON_ERROR, ENTRY_SET_MIN_STACK_SIZE, ENTRY_SELECT_WORKER, ENTRY_SELECT_ANY_WORKER, ENTRY_ENQUEUE_TASK,
 ENTRY_ENQUEUE_COMMAND_BUFFER, ENTRY_WAIT_ON_BARRIER, ENTRY_WAIT_ON_COMMAND_BUFFER, ENTRY_WAIT_ON_COMMAND_INDIRECT) = range(9)


class CommandBuffer(C.Structure):
    pass


CommandBufferFn = C.CFUNCTYPE(None, POINTER(CommandBuffer))
CommandBuffer._fields_ = [
    ("label", c_char_p),
    ("label_len", c_size_t),
    ("entries", POINTER(CommandBufferEntry)),
    ("entries_len", c_size_t),
    ("on_complete", CommandBufferFn),
    ("on_abort", CommandBufferFn),
    ("on_deinit", CommandBufferFn),
]


class PoolVTable(C.Structure):
    _fields_ = [
        ("id", C.CFUNCTYPE(c_size_t, c_void_p)),
        ("acquire", C.CFUNCTYPE(None, c_void_p)),
        ("release", C.CFUNCTYPE(None, c_void_p)),
        ("request_close", C.CFUNCTYPE(None, c_void_p)),
        ("accepts_requests", C.CFUNCTYPE(c_bool, c_void_p)),
        ("owns_current_thread", C.CFUNCTYPE(c_bool, c_void_p)),
        ("label", C.CFUNCTYPE(c_void_p, c_void_p, POINTER(c_size_t))),
        ("workers", C.CFUNCTYPE(c_size_t, c_void_p, POINTER(c_size_t), c_size_t)),
        ("stack_sizes", C.CFUNCTYPE(c_size_t, c_void_p, POINTER(c_size_t), c_size_t)),
        ("enqueue_buffer", C.CFUNCTYPE(FimoStatus, c_void_p, POINTER(CommandBuffer), POINTER(CommandBufferHandle))),
    ]


class Pool(C.Structure):
    _fields_ = [("data", c_void_p), ("vtable", POINTER(PoolVTable))]

    def id(self) -> int:
        return self.vtable.contents.id(self.data)

    def unref(self) -> None:
        self.vtable.contents.release(self.data)

    def ref(self) -> "Pool":
        self.vtable.contents.acquire(self.data)
        return self

    def label(self) -> str:
        n = c_size_t(0)
        p = self.vtable.contents.label(self.data, C.byref(n))
        return C.string_at(p, n.value).decode()

    def worker_count(self) -> int:
        return self.vtable.contents.workers(self.data, None, 0)

    def enqueue_buffer(self, buffer: CommandBuffer, handle: CommandBufferHandle | None = None) -> int:
        return self.vtable.contents.enqueue_buffer(self.data, C.byref(buffer), C.byref(handle) if handle else None)


class PoolConfig(C.Structure):
    _fields_ = [
        ("next", c_void_p),
        ("label", c_char_p),
        ("label_len", c_size_t),
        ("stacks", c_void_p),
        ("stacks_len", c_size_t),
        ("default_stack_index", c_size_t),
        ("worker_count", c_size_t),
        ("is_queryable", c_bool),
    ]


class ResourceDescriptor(C.Structure):
    _fields_ = [
        ("next", c_void_p),
        ("label", c_char_p),
        ("label_len", c_size_t),
        ("size", c_size_t),
        ("alignment", c_size_t),
    ]


class SystemDependency(C.Structure):
    _fields_ = [("system", c_void_p), ("flags", c_size_t)]


FactoryDestroyFn = C.CFUNCTYPE(None, c_void_p)
SystemCreateFn = C.CFUNCTYPE(c_bool, c_void_p, c_void_p, c_void_p)
SystemDestroyFn = C.CFUNCTYPE(None, c_void_p)
SystemRunFn = C.CFUNCTYPE(None, c_void_p, POINTER(c_void_p), POINTER(c_void_p), POINTER(Fence))


class SystemDescriptor(C.Structure):
    _fields_ = [
        ("next", c_void_p),
        ("label", c_char_p),
        ("label_len", c_size_t),
        ("exclusive_handles", POINTER(c_void_p)),
        ("exclusive_handles_len", c_size_t),
        ("shared_handles", POINTER(c_void_p)),
        ("shared_handles_len", c_size_t),
        ("before", POINTER(SystemDependency)),
        ("before_len", c_size_t),
        ("after", POINTER(SystemDependency)),
        ("after_len", c_size_t),
        ("factory", c_void_p),
        ("factory_size", c_size_t),
        ("factory_alignment", c_size_t),
        ("factory_destroy", FactoryDestroyFn),
        ("system_size", c_size_t),
        ("system_alignment", c_size_t),
        ("system_create", SystemCreateFn),
        ("system_destroy", SystemDestroyFn),
        ("system_run", SystemRunFn),
    ]


class SystemGroupDescriptor(C.Structure):
    _fields_ = [
        ("next", c_void_p),
        ("label", c_char_p),
        ("label_len", c_size_t),
        ("pool", POINTER(Pool)),
        ("world", c_void_p),
    ]


class WorldDescriptor(C.Structure):
    _fields_ = [("next", c_void_p), ("label", c_char_p), ("label_len", c_size_t), ("pool", POINTER(Pool))]


class ExtHeader(C.Structure):
    _fields_ = [("s_type", c_uint32), ("reserved", c_uint32), ("next", c_void_p)]


class WorldBatchExt(C.Structure):
    _fields_ = [
        ("header", ExtHeader),
        ("n_worlds", c_uint64),
        ("entities_per_world", c_uint64),
        ("device", c_int32),
        ("flags", c_uint32),
    ]


class ResourceDeviceExt(C.Structure):
    _fields_ = [
        ("header", ExtHeader),
        ("scope", c_uint32),
        ("components", c_uint32),
        ("elem_bytes", c_uint32),
        ("flags", c_uint32),
    ]


class ResourceView(C.Structure):
    _fields_ = [
        ("device_ptr", c_void_p),
        ("host_ptr", c_void_p),
        ("n_worlds", c_uint64),
        ("count", c_uint64),
        ("components", c_uint32),
        ("elem_bytes", c_uint32),
        ("comp_stride", c_uint64),
        ("world_stride", c_uint64),
        ("scope", c_uint32),
        ("flags", c_uint32),
    ]


class LaunchInfo(C.Structure):
    _fields_ = [
        ("stream", c_void_p),
        ("n_worlds", c_uint64),
        ("entities_per_world", c_uint64),
        ("device", c_int32),
        ("flags", c_uint32),
        ("reserved", c_uint64),
    ]


DeviceEntryFn = C.CFUNCTYPE(FimoStatus, c_void_p, POINTER(POINTER(ResourceView)), POINTER(POINTER(ResourceView)),
                            POINTER(LaunchInfo))


class SystemDeviceExt(C.Structure):
    _fields_ = [("header", ExtHeader), ("device_entry", DeviceEntryFn), ("builtin_id", c_uint32), ("flags", c_uint32)]


PostStepFn = C.CFUNCTYPE(None, c_void_p, c_void_p, c_uint64)


class GroupDeviceExt(C.Structure):
    _fields_ = [
        ("header", ExtHeader),
        ("flags", c_uint32),
        ("reserved", c_uint32),
        ("post_step", PostStepFn),
        ("post_step_user", c_void_p),
        ("stats_ring", c_void_p),       # with GROUP_FLAG_STATS_RING: device [slots][STAT_WORDS] u64, slot = generation % slots
        ("stats_ring_slots", c_uint32),
        ("reserved2", c_uint32),
    ]


class ExecutorInfo(C.Structure):
    _fields_ = [
        ("systems", c_uint32),
        ("device_systems", c_uint32),
        ("nodes", c_uint32),
        ("fused_world_step", c_uint32),
        ("uses_cuda_graph", c_uint32),
        ("kernel_launches_per_generation", c_uint32),
        ("edges", c_uint64),
        ("algorithmic_bytes_per_step", c_uint64),
        ("max_generations_per_launch", c_uint64),
        ("reserved", c_uint64 * 3),
    ]


class Export(C.Structure):
    _fields_ = [("name", c_char_p), ("name_space", c_char_p), ("function", c_void_p)]


assert C.sizeof(SystemDescriptor) == 160 and C.sizeof(ResourceDescriptor) == 40
assert C.sizeof(SystemGroupDescriptor) == 40 and C.sizeof(WorldDescriptor) == 32
assert C.sizeof(CommandBufferEntry) == 24 and C.sizeof(CommandBuffer) == 56 and C.sizeof(Task) == 48
assert C.sizeof(PoolVTable) == 80 and C.sizeof(Pool) == 16 and C.sizeof(PoolConfig) == 64
assert C.sizeof(ResourceView) == 64

# (name, restype, argtypes) of the 32 `fimo-worlds` symbols, in `all_symbols` order
# (pkgs/fimo_worlds_meta/src/symbols.zig:22-58).
WORLDS_SYMBOLS = [
    ("resource_register", FimoStatus, [POINTER(ResourceDescriptor), POINTER(c_void_p)]),
    ("resource_unregister", None, [c_void_p]),
    ("system_register", FimoStatus, [POINTER(SystemDescriptor), POINTER(c_void_p)]),
    ("system_unregister", None, [c_void_p]),
    ("system_group_create", FimoStatus, [POINTER(SystemGroupDescriptor), POINTER(c_void_p)]),
    ("system_group_destroy", None, [c_void_p]),
    ("system_group_get_world", c_void_p, [c_void_p]),
    ("system_group_get_label", c_void_p, [c_void_p, POINTER(c_size_t)]),
    ("system_group_get_pool", Pool, [c_void_p]),
    ("system_group_add_systems", FimoStatus, [c_void_p, POINTER(c_void_p), c_size_t]),
    ("system_group_remove_system", None, [c_void_p, c_void_p, POINTER(Fence)]),
    ("system_group_schedule", FimoStatus, [c_void_p, POINTER(POINTER(Fence)), c_size_t, POINTER(Fence)]),
    ("system_context_get_group", c_void_p, [c_void_p]),
    ("system_context_get_generation", c_size_t, [c_void_p]),
    ("system_context_allocator_alloc", c_void_p, [c_void_p, c_int32, c_size_t, c_size_t, c_size_t]),
    ("system_context_allocator_resize", c_bool, [c_void_p, c_int32, c_void_p, c_size_t, c_size_t, c_size_t, c_size_t]),
    ("system_context_allocator_remap", c_void_p, [c_void_p, c_int32, c_void_p, c_size_t, c_size_t, c_size_t, c_size_t]),
    ("system_context_allocator_free", None, [c_void_p, c_int32, c_void_p, c_size_t, c_size_t, c_size_t]),
    ("world_create", FimoStatus, [POINTER(WorldDescriptor), POINTER(c_void_p)]),
    ("world_destroy", None, [c_void_p]),
    ("world_get_label", c_void_p, [c_void_p, POINTER(c_size_t)]),
    ("world_get_pool", Pool, [c_void_p]),
    ("world_has_resource", c_bool, [c_void_p, c_void_p]),
    ("world_add_resource", FimoStatus, [c_void_p, c_void_p, c_void_p]),
    ("world_remove_resource", FimoStatus, [c_void_p, c_void_p, c_void_p]),
    ("world_lock_resources", None, [c_void_p, POINTER(c_void_p), c_size_t, POINTER(c_void_p), c_size_t,
                                    POINTER(c_void_p)]),
    ("world_unlock_resource_exclusive", None, [c_void_p, c_void_p]),
    ("world_unlock_resource_shared", None, [c_void_p, c_void_p]),
    ("world_allocator_alloc", c_void_p, [c_void_p, c_size_t, c_size_t, c_size_t]),
    ("world_allocator_resize", c_bool, [c_void_p, c_void_p, c_size_t, c_size_t, c_size_t, c_size_t]),
    ("world_allocator_remap", c_void_p, [c_void_p, c_void_p, c_size_t, c_size_t, c_size_t, c_size_t]),
    ("world_allocator_free", None, [c_void_p, c_void_p, c_size_t, c_size_t, c_size_t]),
]

# plumbing exported by every implementation of the header
COMMON_SYMBOLS = [
    ("fimo_worlds_b200_exports", POINTER(Export), [POINTER(c_size_t)]),
    ("fimo_worlds_b200_last_error", c_char_p, []),
    ("fimo_tasks_futex_wait", c_int32, [c_void_p, c_size_t, c_uint64, c_size_t, c_void_p]),
    ("fimo_tasks_futex_wake", c_size_t, [c_void_p, c_size_t, FutexFilter]),
    ("fimo_worlds_b200_bind_tasks", None, [c_void_p, c_void_p]),
    ("fimo_worlds_b200_fence_wait", None, [POINTER(Fence)]),
    ("fimo_worlds_b200_fence_signal", None, [POINTER(Fence)]),
    ("fimo_tasks_create_worker_pool", FimoStatus, [POINTER(PoolConfig), POINTER(Pool)]),
    ("fimo_tasks_abort", None, []),
    ("fimo_tasks_worker_id", c_bool, [POINTER(c_size_t)]),
]

class TraceEvent(C.Structure):  # FimoB200_TraceEvent
    _fields_ = [("group_label", C.c_void_p), ("group_label_len", c_size_t), ("generation", c_uint64), ("device_ms", c_double),
                ("host_enqueue_ns", C.c_int64), ("host_complete_ns", C.c_int64), ("kernels", c_uint32), ("fused_world_step", c_uint32)]


TraceSinkFn = C.CFUNCTYPE(None, c_void_p, POINTER(TraceEvent))


# fimo tracing subscriber events (fimo_std/tracing.h:35-56,72-204) the bridge emits
class Instant(C.Structure):
    _fields_ = [("secs", c_uint64), ("nanos", c_uint32)]


class TracingEventInfo(C.Structure):
    _fields_ = [("name", c_char_p), ("target", c_char_p), ("scope", c_char_p), ("file_name", c_char_p),
                ("line_number", C.c_int32), ("level", C.c_int32)]


class TracingCallStackEvent(C.Structure):  # create / destroy call stack
    _fields_ = [("event", c_uint32), ("time", Instant), ("stack", c_void_p)]


class TracingEnterSpan(C.Structure):  # also log_message (same layout: info instead of span)
    _fields_ = [("event", c_uint32), ("time", Instant), ("stack", c_void_p), ("span", POINTER(TracingEventInfo)),
                ("message", c_void_p), ("message_length", c_size_t)]


class TracingExitSpan(C.Structure):
    _fields_ = [("event", c_uint32), ("time", Instant), ("stack", c_void_p), ("is_unwinding", C.c_bool)]


class TracingDeclareEventInfo(C.Structure):
    _fields_ = [("event", c_uint32), ("info", POINTER(TracingEventInfo))]


TracingOnEventFn = C.CFUNCTYPE(None, c_void_p, POINTER(c_uint32))


class TracingSubscriber(C.Structure):
    _fields_ = [("ptr", c_void_p), ("on_event", TracingOnEventFn)]


(TRACING_EVENT_CREATE_CALL_STACK, TRACING_EVENT_DESTROY_CALL_STACK, TRACING_EVENT_ENTER_SPAN, TRACING_EVENT_EXIT_SPAN,
 TRACING_EVENT_LOG_MESSAGE, TRACING_EVENT_DECLARE_EVENT_INFO) = 4, 5, 9, 10, 11, 12


def decode_tracing_event(ptr):
    """(kind, dict) of one subscriber event, read through the struct its tag names."""
    kind = ptr[0]
    if kind in (TRACING_EVENT_CREATE_CALL_STACK, TRACING_EVENT_DESTROY_CALL_STACK):
        e = C.cast(ptr, POINTER(TracingCallStackEvent)).contents
        return kind, {"ns": e.time.secs * 10**9 + e.time.nanos, "stack": e.stack}
    if kind in (TRACING_EVENT_ENTER_SPAN, TRACING_EVENT_LOG_MESSAGE):
        e = C.cast(ptr, POINTER(TracingEnterSpan)).contents
        info = e.span.contents
        return kind, {"ns": e.time.secs * 10**9 + e.time.nanos, "stack": e.stack, "info": C.addressof(info),
                      "name": info.name.decode(), "target": info.target.decode(), "scope": info.scope.decode(), "level": info.level,
                      "message": C.string_at(e.message, e.message_length).decode()}
    if kind == TRACING_EVENT_EXIT_SPAN:
        e = C.cast(ptr, POINTER(TracingExitSpan)).contents
        return kind, {"ns": e.time.secs * 10**9 + e.time.nanos, "stack": e.stack, "is_unwinding": bool(e.is_unwinding)}
    if kind == TRACING_EVENT_DECLARE_EVENT_INFO:
        e = C.cast(ptr, POINTER(TracingDeclareEventInfo)).contents
        return kind, {"info": C.addressof(e.info.contents), "name": e.info.contents.name.decode()}
    return kind, {}


# device-side helpers: product library only
PRODUCT_SYMBOLS = [
    ("fimo_b200_builtin_system", FimoStatus, [c_uint32, POINTER(SystemDescriptor), POINTER(SystemDeviceExt)]),
    ("fimo_b200_resource_upload", FimoStatus, [POINTER(ResourceView), c_void_p]),
    ("fimo_b200_resource_download", FimoStatus, [POINTER(ResourceView), c_void_p]),
    ("fimo_b200_group_executor_info", FimoStatus, [c_void_p, POINTER(ExecutorInfo)]),
    ("fimo_b200_group_executor_describe", c_size_t, [c_void_p, c_char_p, c_size_t]),
    ("fimo_b200_group_kernels_launched", c_uint64, [c_void_p]),
    ("fimo_b200_group_stream", c_void_p, [c_void_p]),
    ("fimo_b200_group_wait_idle", None, [c_void_p]),
    ("fimo_b200_group_device_ms", c_double, [c_void_p, POINTER(c_uint64)]),
    ("fimo_b200_group_span_ms", c_double, [c_void_p]),
    ("fimo_b200_group_transfer", FimoStatus, [c_void_p, c_void_p, c_void_p, c_uint32, POINTER(Fence)]),
    ("fimo_b200_group_schedule_io", FimoStatus, [c_void_p, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_size_t, POINTER(Fence)]),
    ("fimo_b200_group_schedule_readback", FimoStatus, [c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_size_t, POINTER(Fence)]),
    ("fimo_worlds_b200_set_trace_sink", None, [TraceSinkFn, c_void_p]),
    ("fimo_worlds_b200_trace_subscribe", FimoStatus, [POINTER(TracingSubscriber), c_size_t]),
    ("fimo_b200_host_alloc", c_void_p, [c_size_t]),
    ("fimo_b200_host_free", None, [c_void_p]),
    ("fimo_b200_cull_pack_static", FimoStatus, [c_void_p, c_void_p, c_void_p, c_uint64, c_uint64, c_uint64, c_uint64, c_void_p, c_uint64, c_void_p]),
    ("fimo_b200_group_stats_peers_open", FimoStatus, [c_void_p, c_uint32, c_void_p]),
    ("fimo_b200_group_stats_peers_bind", FimoStatus, [c_void_p, c_void_p, c_uint32, c_uint32]),
    ("fimo_b200_stats_nccl_unique_id", FimoStatus, [c_void_p]),
    ("fimo_b200_group_stats_nccl_bind", FimoStatus, [c_void_p, c_void_p, c_uint32, c_uint32, c_uint32]),
    ("fimo_b200_group_global_stats", FimoStatus, [c_void_p, c_uint64, c_uint32, POINTER(c_uint64)]),
]
STATS_PEER_HANDLE_BYTES, STATS_NCCL_ID_BYTES, STATS_MAX_RANKS = 64, 128, 16
STATS_MODE_PEERS, STATS_MODE_NCCL = 1, 2
TRANSFER_HOST_TO_DEVICE, TRANSFER_DEVICE_TO_HOST = 0, 1


class StepTransfer(C.Structure):  # FimoB200_StepTransfer
    _fields_ = [("resource", c_void_p), ("host", c_void_p)]


class CullReadback(C.Structure):  # FimoB200_CullReadback
    _fields_ = [("vis_mask", c_void_p), ("vis_list", c_void_p), ("vis_count", c_void_p), ("host", c_void_p), ("host_bytes", c_uint64)]


READBACK_HEADER_WORDS = 8
PACK_STATIC_MAX_BYTES, PACK_STATIC_MAX_WORLDS = 8 << 20, 8192  # fimo_b200_cull_pack_static's fixed staging area

# CPU-oracle-only helpers (tests / bench cpu_baseline only)
ORACLE_SYMBOLS = [
    ("fimo_oracle_builtin_system", FimoStatus, [c_uint32, POINTER(SystemDescriptor)]),
    ("fimo_oracle_step_groups", c_double, [POINTER(c_void_p), c_size_t, c_size_t]),
    ("fimo_oracle_step_groups_queued", c_double, [POINTER(c_void_p), c_size_t, c_size_t]),
    ("fimo_oracle_hardware_threads", c_uint32, []),
]


class FimoError(RuntimeError):
    """A FIMO_STATUS_ERROR; `.name` is the reference's Zig error name (e.g. "Deadlock")."""

    def __init__(self, name: str):
        super().__init__(name)
        self.name = name


class WorldsLib:
    """A loaded implementation of include/fimo_worlds_b200.h."""

    def __init__(self, path: str):
        self.path = path
        self.dll = C.CDLL(path, mode=C.RTLD_LOCAL)
        for name, restype, argtypes in WORLDS_SYMBOLS:
            fn = getattr(self.dll, "fimo_worlds_" + name)
            fn.restype, fn.argtypes = restype, argtypes
            setattr(self, name, fn)
        for table in (COMMON_SYMBOLS, PRODUCT_SYMBOLS, ORACLE_SYMBOLS):
            for name, restype, argtypes in table:
                try:
                    fn = getattr(self.dll, name)
                except AttributeError:
                    if table is COMMON_SYMBOLS:
                        raise
                    continue
                fn.restype, fn.argtypes = restype, argtypes
                setattr(self, name, fn)
        self.is_product = hasattr(self, "fimo_b200_builtin_system")
        self.is_oracle = hasattr(self, "fimo_oracle_builtin_system")

    def last_error(self) -> str:
        return (self.fimo_worlds_b200_last_error() or b"").decode()

    def check(self, status: int) -> None:
        if status != STATUS_OK:
            raise FimoError(self.last_error())

    def exports(self) -> list[tuple[str, str, int]]:
        n = c_size_t(0)
        table = self.fimo_worlds_b200_exports(C.byref(n))
        return [(table[i].name.decode(), table[i].name_space.decode(), table[i].function) for i in range(n.value)]
