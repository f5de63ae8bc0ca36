"""Harness-side wrappers over the C ABI, shaped like the reference's Zig wrappers
(pkgs/fimo_worlds_meta/src/{worlds,systems,resources,Job}.zig) so that tests read like
the reference's own tests:

    World.init / addSystemGroup / getPool            worlds.zig:28-199
    SystemGroup.addSytems / removeSystem / schedule / run   systems.zig:30-203
    TypedResource(T).register / addToWorld / lockInWorldExclusive ...   resources.zig:32-120
    Declaration.initFunctor(...).register            systems.zig:922-1182
    Fence / Job.go                                   Job.zig:20-63, 216-264
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Sequence

from . import abi
from .abi import FimoError, WorldsLib  # noqa: F401  (re-exported)


class Fence:
    """Job.zig:20-63."""

    def __init__(self, lib: WorldsLib):
        self.lib = lib
        self.raw = abi.Fence(0)

    def ptr(self):
        return C.pointer(self.raw)

    def is_signaled(self) -> bool:
        return (self.raw.state & 1) != 0

    def wait(self) -> None:
        self.lib.fimo_worlds_b200_fence_wait(C.byref(self.raw))

    def signal(self) -> None:
        self.lib.fimo_worlds_b200_fence_signal(C.byref(self.raw))

    def reset(self) -> None:
        self.raw.state = 0


def create_pool(lib: WorldsLib, label: str = "pool", workers: int = 0) -> abi.Pool:
    """Pool.init (pkgs/fimo_tasks_meta/src/pool.zig:112-117)."""
    b = label.encode()
    cfg = abi.PoolConfig(None, b, len(b), None, 0, 0, workers, False)
    pool = abi.Pool()
    lib.check(lib.fimo_tasks_create_worker_pool(C.byref(cfg), C.byref(pool)))
    return pool


class Resource:
    """An untyped registered resource (resources.zig:14-30)."""

    def __init__(self, lib: WorldsLib, handle: int, size: int, label: str, keep=()):
        self.lib, self.handle, self.size, self.label = lib, handle, size, label
        self._keep = keep

    @staticmethod
    def register(lib: WorldsLib, label: str, size: int, alignment: int = 8, next_ptr=None, keep=()) -> "Resource":
        b = label.encode()
        desc = abi.ResourceDescriptor(next_ptr, b, len(b), size, alignment)
        h = C.c_void_p()
        lib.check(lib.resource_register(C.byref(desc), C.byref(h)))
        return Resource(lib, h.value, size, label, keep)

    def unregister(self) -> None:
        self.lib.resource_unregister(self.handle)

    def exists_in_world(self, world: "World") -> bool:
        return bool(self.lib.world_has_resource(world.handle, self.handle))

    def add_to_world(self, world: "World", value) -> None:
        self.lib.check(self.lib.world_add_resource(world.handle, self.handle, C.byref(value)))

    def remove_from_world(self, world: "World", out=None):
        if out is None:
            out = (C.c_uint8 * self.size)()
        self.lib.check(self.lib.world_remove_resource(world.handle, self.handle, C.byref(out)))
        return out

    def lock_in_world(self, world: "World", exclusive: bool) -> int:
        arr = (C.c_void_p * 1)(self.handle)
        out = (C.c_void_p * 1)()
        if exclusive:
            self.lib.world_lock_resources(world.handle, arr, 1, None, 0, out)
        else:
            self.lib.world_lock_resources(world.handle, None, 0, arr, 1, out)
        return out[0]

    def unlock_in_world(self, world: "World", exclusive: bool) -> None:
        if exclusive:
            self.lib.world_unlock_resource_exclusive(world.handle, self.handle)
        else:
            self.lib.world_unlock_resource_shared(world.handle, self.handle)


class TypedResource(Resource):
    """TypedResource(T) (resources.zig:32-120) for a ctypes type."""

    @staticmethod
    def register_typed(lib: WorldsLib, ctype, label: str) -> "TypedResource":
        b = label.encode()
        desc = abi.ResourceDescriptor(None, b, len(b), C.sizeof(ctype), C.alignment(ctype))
        h = C.c_void_p()
        lib.check(lib.resource_register(C.byref(desc), C.byref(h)))
        r = TypedResource(lib, h.value, C.sizeof(ctype), label)
        r.ctype = ctype
        return r

    def add_value(self, world: "World", value) -> None:
        self.add_to_world(world, self.ctype(value) if not isinstance(value, self.ctype) else value)

    def remove_value(self, world: "World"):
        out = self.ctype()
        self.remove_from_world(world, out)
        return out

    def lock_exclusive(self, world: "World"):
        return C.cast(self.lock_in_world(world, True), C.POINTER(self.ctype))

    def lock_shared(self, world: "World"):
        return C.cast(self.lock_in_world(world, False), C.POINTER(self.ctype))


class System:
    """A registered system. `run(ctx, exclusive, shared, fence)` receives the SystemContext
    pointer, two lists of resource blob addresses and the deferred fence pointer; like the Zig
    wrappers (systems.zig:1162-1167) the fence is signaled for the user unless `takes_fence`."""

    def __init__(self, lib: WorldsLib, handle: int, keep):
        self.lib, self.handle, self._keep = lib, handle, keep

    @staticmethod
    def register(lib: WorldsLib, label: str, run: Callable | None = None, exclusive: Sequence[Resource] = (),
                 shared: Sequence[Resource] = (), before: Sequence = (), after: Sequence = (),
                 takes_fence: bool = False, init: Callable | None = None, deinit: Callable | None = None,
                 descriptor: abi.SystemDescriptor | None = None, keep=()) -> "System":
        n_ex, n_sh = len(exclusive), len(shared)
        desc = descriptor if descriptor is not None else abi.SystemDescriptor()
        b = label.encode()
        desc.label, desc.label_len = b, len(b)
        ex = (C.c_void_p * max(n_ex, 1))(*[r.handle for r in exclusive])
        sh = (C.c_void_p * max(n_sh, 1))(*[r.handle for r in shared])
        desc.exclusive_handles, desc.exclusive_handles_len = ex, n_ex
        desc.shared_handles, desc.shared_handles_len = sh, n_sh

        def deps(items):
            arr = (abi.SystemDependency * max(len(items), 1))()
            for i, it in enumerate(items):
                s, flags = (it, 0) if isinstance(it, System) else it
                arr[i].system, arr[i].flags = s.handle, flags
            return arr

        bf, af = deps(before), deps(after)
        desc.before, desc.before_len = bf, len(before)
        desc.after, desc.after_len = af, len(after)
        keepalive = [b, ex, sh, bf, af, desc, keep]

        if descriptor is None:
            # Python-callable system: the state is one pointer holding the SystemContext,
            # as Declaration.initFunctor stores it (systems.zig:1100-1133).
            def _create(_factory, ctx, system):
                C.cast(system, C.POINTER(C.c_void_p))[0] = ctx
                if init is not None:
                    return bool(init(ctx))
                return True

            def _destroy(system):
                if deinit is not None:
                    deinit(C.cast(system, C.POINTER(C.c_void_p))[0])

            def _run(system, ex_ptrs, sh_ptrs, fence):
                ctx = C.cast(system, C.POINTER(C.c_void_p))[0]
                e = [ex_ptrs[i] for i in range(n_ex)]
                s = [sh_ptrs[i] for i in range(n_sh)]
                try:
                    run(ctx, e, s, fence)
                finally:
                    if not takes_fence:
                        lib.fimo_worlds_b200_fence_signal(fence)

            c_create, c_destroy, c_run = abi.SystemCreateFn(_create), abi.SystemDestroyFn(_destroy), abi.SystemRunFn(_run)
            desc.system_size, desc.system_alignment = 8, 8
            desc.system_create, desc.system_destroy, desc.system_run = c_create, c_destroy, c_run
            keepalive += [c_create, c_destroy, c_run]
        h = C.c_void_p()
        lib.check(lib.system_register(C.byref(desc), C.byref(h)))
        return System(lib, h.value, keepalive)

    def unregister(self) -> None:
        self.lib.system_unregister(self.handle)


class SystemGroup:
    """systems.zig:30-203."""

    def __init__(self, lib: WorldsLib, handle: int, world: "World", keep=()):
        self.lib, self.handle, self.world, self._keep = lib, handle, world, keep

    def deinit(self) -> None:
        self.lib.system_group_destroy(self.handle)

    def get_world_handle(self) -> int:
        return self.lib.system_group_get_world(self.handle)

    def get_label(self) -> str:
        n = C.c_size_t(0)
        p = self.lib.system_group_get_label(self.handle, C.byref(n))
        return C.string_at(p, n.value).decode()

    def get_pool(self) -> abi.Pool:
        return self.lib.system_group_get_pool(self.handle)

    def add_systems(self, systems: Sequence[System]) -> None:
        arr = (C.c_void_p * max(len(systems), 1))(*[s.handle for s in systems])
        self.lib.check(self.lib.system_group_add_systems(self.handle, arr, len(systems)))

    def remove_system(self, system: System) -> None:
        """removeSystem waits for the removal like the Zig wrapper (systems.zig:170-174)."""
        f = Fence(self.lib)
        self.lib.system_group_remove_system(self.handle, system.handle, f.ptr())
        f.wait()

    def schedule(self, wait_on: Sequence[Fence] = (), signal: Fence | None = None) -> None:
        arr = (C.POINTER(abi.Fence) * max(len(wait_on), 1))(*[f.ptr() for f in wait_on])
        self.lib.check(self.lib.system_group_schedule(self.handle, arr, len(wait_on), signal.ptr() if signal else None))

    def run(self, wait_on: Sequence[Fence] = ()) -> None:
        """schedule + wait (systems.zig:196-202)."""
        f = Fence(self.lib)
        self.schedule(wait_on, f)
        f.wait()


class World:
    """worlds.zig:28-199."""

    def __init__(self, lib: WorldsLib, handle: int, keep=()):
        self.lib, self.handle, self._keep = lib, handle, keep

    @staticmethod
    def init(lib: WorldsLib, label: str = "", pool: abi.Pool | None = None, batch: abi.WorldBatchExt | None = None) -> "World":
        b = label.encode()
        desc = abi.WorldDescriptor(C.cast(C.pointer(batch), C.c_void_p) if batch is not None else None, b, len(b),
                                   C.pointer(pool) if pool is not None else None)
        h = C.c_void_p()
        lib.check(lib.world_create(C.byref(desc), C.byref(h)))
        return World(lib, h.value, (b, batch, pool))

    def deinit(self) -> None:
        self.lib.world_destroy(self.handle)

    def get_label(self) -> str:
        n = C.c_size_t(0)
        p = self.lib.world_get_label(self.handle, C.byref(n))
        return C.string_at(p, n.value).decode()

    def get_pool(self) -> abi.Pool:
        return self.lib.world_get_pool(self.handle)

    def add_system_group(self, label: str = "", pool: abi.Pool | None = None, ext: abi.GroupDeviceExt | None = None) -> SystemGroup:
        b = label.encode()
        desc = abi.SystemGroupDescriptor(C.cast(C.pointer(ext), C.c_void_p) if ext is not None else None, b, len(b),
                                         C.pointer(pool) if pool is not None else None, self.handle)
        h = C.c_void_p()
        self.lib.check(self.lib.system_group_create(C.byref(desc), C.byref(h)))
        return SystemGroup(self.lib, h.value, self, (b, pool, ext))

    def alloc(self, size: int, alignment: int = 8) -> int:
        return self.lib.world_allocator_alloc(self.handle, size, alignment, 0)

    def free(self, ptr: int, size: int, alignment: int = 8) -> None:
        self.lib.world_allocator_free(self.handle, ptr, size, alignment, 0)


class Job:
    """Job.go (Job.zig:216-264): run a Python callable as a one-task detached command buffer
    on `pool`, signalling `signal` when it is done."""

    _live: dict[int, object] = {}

    @staticmethod
    def go(lib: WorldsLib, pool: abi.Pool, fn: Callable[[], None], signal: Fence | None = None) -> None:
        task = abi.Task()
        entries = (abi.CommandBufferEntry * 2)()
        buf = abi.CommandBuffer()
        key = id(buf)

        def _start(_t):
            fn()

        def _deinit(_b):
            # the keep-alive entry is intentionally never dropped: the thunk is still executing here
            if signal is not None:
                signal.signal()

        c_start, c_deinit = abi.TaskFn(_start), abi.CommandBufferFn(_deinit)
        task.label, task.label_len, task.on_start = b"job", 3, c_start
        entries[0].type = abi.ENTRY_ABORT_ON_ERROR
        entries[0].payload.abort_on_error = True
        entries[1].type = abi.ENTRY_ENQUEUE_TASK
        entries[1].payload.enqueue_task = C.pointer(task)
        buf.label, buf.label_len, buf.entries, buf.entries_len, buf.on_deinit = b"job", 3, entries, 2, c_deinit
        Job._live[key] = (task, entries, buf, c_start, c_deinit)
        lib.check(pool.enqueue_buffer(buf))
