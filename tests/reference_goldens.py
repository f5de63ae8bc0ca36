"""The reference's own behavioural tests for the hot path, restated once and run against every
implementation of include/fimo_worlds_b200.h (CPU oracle here; the CUDA product in -m gpu).

Each function cites the Zig test it restates (all under /root/reference/pkgs/fimo_worlds_meta/src/).
The golden values (5, 4000, 10000, generation +1, cycle => error) are the reference's.
"""
from __future__ import annotations

import ctypes as C
import random
import threading

from fimo_b200 import abi
from fimo_b200.worlds import Fence, FimoError, Job, System, TypedResource, World, create_pool


def world_smoke(lib):
    """worlds.zig:201-212 "World: smoke test" — label round trip."""
    world = World.init(lib, "test-world")
    assert world.get_label() == "test-world"
    world.deinit()


def world_custom_pool(lib):
    """worlds.zig:214-234 "World: custom pool" — the world reports the executor it was given."""
    executor = create_pool(lib, "test-pool", 4)
    world = World.init(lib, "test-world", pool=executor)
    ex = world.get_pool()
    assert ex.id() == executor.id()
    ex.unref()
    world.deinit()
    executor.unref()


def world_allocator(lib):
    """worlds.zig:236-261 "World: allocator" — allocations work; leftovers are auto-freed at deinit."""
    world = World.init(lib, "test-world")
    p = world.alloc(100, 8)
    assert p
    C.memset(p, 0xAB, 100)
    world.free(p, 100, 8)
    q = world.alloc(64, 16)  # intentionally leaked: TracingAllocator frees it on world deinit
    assert q and q % 16 == 0
    world.deinit()


def group_label_and_pool(lib):
    """systems.zig:176-211 "SystemGroup: smoke test" / "custom pool"."""
    world = World.init(lib, "test-world")
    executor = create_pool(lib, "test-pool", 2)
    group = world.add_system_group("test-group", pool=executor)
    assert group.get_label() == "test-group"
    assert group.get_world_handle() == world.handle
    ex = group.get_pool()
    assert ex.id() == executor.id()
    ex.unref()
    group.deinit()
    world.deinit()
    executor.unref()


def group_schedule_counts_to_five(lib):
    """systems.zig:214-253 "SystemGroup: schedule": one system `+= 1` on a shared u32;
    4 x schedule + 1 x run  =>  5."""
    world = World.init(lib, "test-world")
    group = world.add_system_group("test-group")
    resource = TypedResource.register_typed(lib, C.c_uint32, "test-resource")
    resource.add_value(world, 0)

    def run(ctx, exclusive, shared, fence):
        C.cast(shared[0], C.POINTER(C.c_uint32))[0] += 1

    sys_ = System.register(lib, "test-system", run, shared=[resource])
    group.add_systems([sys_])
    for _ in range(4):
        group.schedule()
    group.run()
    ptr = resource.lock_shared(world)
    value = ptr[0]
    resource.unlock_in_world(world, False)
    group.remove_system(sys_)
    sys_.unregister()
    resource.remove_value(world)
    resource.unregister()
    group.deinit()
    world.deinit()
    assert value == 5


def context_group_identity(lib):
    """systems.zig:427-459 "SystemContext: group": ctx.getGroup() is the group the system runs in."""
    world = World.init(lib, "test-world")
    group = world.add_system_group("test-group")
    handle = TypedResource.register_typed(lib, C.c_void_p, "group")
    handle.add_value(world, 0)

    def run(ctx, exclusive, shared, fence):
        C.cast(exclusive[0], C.POINTER(C.c_void_p))[0] = lib.system_context_get_group(ctx)

    sys_ = System.register(lib, "test-system", run, exclusive=[handle])
    group.add_systems([sys_])
    group.run()
    ptr = handle.lock_exclusive(world)
    seen = ptr[0]
    handle.unlock_in_world(world, True)
    group.remove_system(sys_)
    sys_.unregister()
    handle.remove_value(world)
    handle.unregister()
    group.deinit()
    world.deinit()
    assert seen == group.handle


def context_generation_increments(lib):
    """systems.zig:461-508 "SystemContext: generation": over 10 schedules + 1 run every run sees
    exactly the previous generation + 1."""
    world = World.init(lib, "test-world")
    group = world.add_system_group("test-group")
    err = TypedResource.register_typed(lib, C.c_int32, "error")
    err.add_value(world, 0)
    seen = []

    def run(ctx, exclusive, shared, fence):
        gen = lib.system_context_get_generation(ctx)
        if seen and seen[-1] + 1 != gen:
            C.cast(exclusive[0], C.POINTER(C.c_int32))[0] = 1
        seen.append(gen)

    sys_ = System.register(lib, "test-system", run, exclusive=[err])
    group.add_systems([sys_])
    for _ in range(10):
        group.schedule()
    group.run()
    ptr = err.lock_exclusive(world)
    failed = ptr[0]
    err.unlock_in_world(world, True)
    group.remove_system(sys_)
    sys_.unregister()
    err.remove_value(world)
    err.unregister()
    group.deinit()
    world.deinit()
    assert failed == 0
    assert seen == list(range(seen[0], seen[0] + 11))


def context_deferred_job(lib):
    """systems.zig:510-577 "SystemContext: deferred": the system spawns a sub-job on the group's pool
    that signals the deferred fence; after 10 schedules + run the sub-job has completed."""
    world = World.init(lib, "test-world")
    group = world.add_system_group("test-group")
    completed = TypedResource.register_typed(lib, C.c_bool, "completed")
    completed.add_value(world, False)
    runs = []

    def run(ctx, exclusive, shared, fence):
        grp = lib.system_context_get_group(ctx)
        executor = lib.system_group_get_pool(grp)
        target = exclusive[0]

        def sub_job():
            C.cast(target, C.POINTER(C.c_bool))[0] = True
            runs.append(1)
            lib.fimo_worlds_b200_fence_signal(fence)

        Job.go(lib, executor, sub_job)
        executor.unref()

    sys_ = System.register(lib, "test-system", run, exclusive=[completed], takes_fence=True)
    group.add_systems([sys_])
    for _ in range(10):
        group.schedule()
    group.run()
    ptr = completed.lock_exclusive(world)
    done = bool(ptr[0])
    completed.unlock_in_world(world, True)
    group.remove_system(sys_)
    sys_.unregister()
    completed.remove_value(world)
    completed.unregister()
    group.deinit()
    world.deinit()
    assert done and len(runs) == 11


def context_allocators(lib):
    """systems.zig:579-749 allocator strategies: each strategy hands out usable, aligned memory
    from inside a running system."""
    world = World.init(lib, "test-world")
    group = world.add_system_group("test-group")
    ok = TypedResource.register_typed(lib, C.c_int32, "ok")
    ok.add_value(world, 0)

    def run(ctx, exclusive, shared, fence):
        good = 0
        for strategy in (abi.ALLOC_TRANSIENT, abi.ALLOC_SINGLE_GENERATION, abi.ALLOC_MULTI_GENERATION,
                         abi.ALLOC_SYSTEM_PERSISTENT):
            p = lib.system_context_allocator_alloc(ctx, strategy, 256, 64, 0)
            if p and p % 64 == 0:
                C.memset(p, 0x5A, 256)
                q = lib.system_context_allocator_remap(ctx, strategy, p, 256, 64, 1024, 0)
                if q and C.string_at(q, 256) == b"\x5a" * 256:
                    good += 1
                    lib.system_context_allocator_free(ctx, strategy, q, 1024, 64, 0)
        C.cast(exclusive[0], C.POINTER(C.c_int32))[0] += good

    sys_ = System.register(lib, "test-system", run, exclusive=[ok])
    group.add_systems([sys_])
    for _ in range(6):  # more than the 4-generation ring
        group.run()
    ptr = ok.lock_exclusive(world)
    total = ptr[0]
    ok.unlock_in_world(world, True)
    group.remove_system(sys_)
    sys_.unregister()
    ok.remove_value(world)
    ok.unregister()
    group.deinit()
    world.deinit()
    assert total == 6 * 4


def system_smoke_and_cycle(lib):
    """systems.zig:1246-1289 "System: smoke test" and "System: cyclic dependency":
    sys1 runs after... (sys1.after = {sys0}); sys2 with before={sys0}, after={sys1} closes a cycle
    and must be rejected."""
    noop = lambda ctx, e, s, f: None  # noqa: E731
    sys0 = System.register(lib, "system-0", noop)
    sys1 = System.register(lib, "system-1", noop, after=[sys0])
    try:
        sys2 = System.register(lib, "system-1", noop, before=[sys0], after=[sys1])
    except FimoError as e:
        assert e.name == "Deadlock"
    else:
        sys2.unregister()
        raise AssertionError("cyclic dependency was accepted")
    finally:
        sys1.unregister()
        sys0.unregister()


def system_registration_errors(lib):
    """Universe.zig:174-219 validation: exclusive dup => Deadlock, shared∩exclusive => Deadlock,
    shared dup => Duplicate, before dup => Duplicate, before∩after => Deadlock."""
    noop = lambda ctx, e, s, f: None  # noqa: E731
    a = TypedResource.register_typed(lib, C.c_int32, "a")
    b = TypedResource.register_typed(lib, C.c_int32, "b")
    s0 = System.register(lib, "s0", noop)

    def expect(name, **kw):
        try:
            s = System.register(lib, "bad", noop, **kw)
        except FimoError as e:
            assert e.name == name, (e.name, name)
        else:
            s.unregister()
            raise AssertionError(f"expected {name}")

    expect("Deadlock", exclusive=[a, a])
    expect("Deadlock", exclusive=[a], shared=[a])
    expect("Duplicate", shared=[b, b])
    expect("Duplicate", before=[s0, s0])
    expect("Deadlock", before=[s0], after=[s0])
    expect("Duplicate", after=[s0, s0])
    s0.unregister()
    a.unregister()
    b.unregister()


def resource_smoke_and_add(lib):
    """resources.zig:122-155 "resource: smoke test" / "add to world"."""
    h1 = TypedResource.register_typed(lib, C.c_int32, "resource-1")
    h2 = TypedResource.register_typed(lib, C.c_int32, "resource-2")
    assert h1.handle != h2.handle
    world = World.init(lib, "test-world")
    assert not h1.exists_in_world(world)
    h1.add_value(world, 5)
    assert h1.exists_in_world(world)
    try:
        h1.add_value(world, 6)
    except FimoError as e:
        assert e.name == "Duplicate"  # World.zig:109
    else:
        raise AssertionError("duplicate add accepted")
    ptr = h1.lock_exclusive(world)
    assert ptr[0] == 5
    h1.unlock_in_world(world, True)
    assert h1.remove_value(world).value == 5
    assert not h1.exists_in_world(world)
    world.deinit()
    h1.unregister()
    h2.unregister()


def resource_unique_lock(lib, use_pool_jobs: bool = True):
    """resources.zig:157-205 "resource: unique lock": 4 jobs x 1000 exclusive increments => 4000."""
    handle = TypedResource.register_typed(lib, C.c_size_t, "resource-1")
    world = World.init(lib, "test-world")
    executor = world.get_pool()
    handle.add_value(world, 0)
    num_jobs, iterations = 4, 1000

    def runner():
        for _ in range(iterations):
            p = handle.lock_exclusive(world)
            p[0] += 1
            handle.unlock_in_world(world, True)

    fences = [Fence(lib) for _ in range(num_jobs)]
    for f in fences:
        Job.go(lib, executor, runner, f)
    for f in fences:
        f.wait()
    p = handle.lock_exclusive(world)
    total = p[0]
    handle.unlock_in_world(world, True)
    handle.remove_value(world)
    executor.unref()
    world.deinit()
    handle.unregister()
    assert total == num_jobs * iterations


def resource_shared_lock(lib, num_writes: int = 2000):
    """resources.zig:207-313 "resource: shared lock": 2 writers / 4 readers; under the lock the
    invariant term_sum == term1 + term2 always holds (the reference uses 10000 writes; the
    count is a parameter here to keep the CPU suite short)."""
    handle = TypedResource.register_typed(lib, C.c_size_t, "resource-1")
    world = World.init(lib, "test-world")
    executor = world.get_pool()
    handle.add_value(world, 0)
    num_reads = num_writes * 2
    state = {"t1": 0, "t2": 0, "sum": 0, "reads": 0, "bad": 0}
    reads_lock = threading.Lock()
    mask = (1 << 64) - 1

    def check():
        s, t2, t1 = state["sum"], state["t2"], state["t1"]
        if s != (t1 + t2) & mask:
            state["bad"] += 1

    def reader():
        while True:
            w = handle.lock_shared(world)
            try:
                if w[0] >= num_writes or state["reads"] >= num_reads:
                    break
                check()
                with reads_lock:
                    state["reads"] += 1
            finally:
                handle.unlock_in_world(world, False)

    def writer(idx):
        rnd = random.Random(idx)
        while True:
            w = handle.lock_exclusive(world)
            try:
                if w[0] >= num_writes:
                    break
                check()
                t1 = rnd.getrandbits(64)
                state["t1"] = t1
                t2 = rnd.getrandbits(64)
                state["t2"] = t2
                state["sum"] = (t1 + t2) & mask
                w[0] += 1
            finally:
                handle.unlock_in_world(world, True)

    fences = [Fence(lib) for _ in range(6)]
    for i in range(2):
        Job.go(lib, executor, (lambda i=i: writer(i)), fences[i])
    for i in range(2, 6):
        Job.go(lib, executor, reader, fences[i])
    for f in fences:
        f.wait()
    w = handle.lock_shared(world)
    writes = w[0]
    handle.unlock_in_world(world, False)
    handle.remove_value(world)
    executor.unref()
    world.deinit()
    handle.unregister()
    assert state["bad"] == 0
    assert writes == num_writes


ALL = [
    world_smoke, world_custom_pool, world_allocator, group_label_and_pool, group_schedule_counts_to_five,
    context_group_identity, context_generation_increments, context_deferred_job, context_allocators,
    system_smoke_and_cycle, system_registration_errors, resource_smoke_and_add, resource_unique_lock,
    resource_shared_lock,
]
