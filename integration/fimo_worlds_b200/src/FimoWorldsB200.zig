//! The `fimo_worlds` module with its system-group step on a B200.
//!
//! Same module name, same imports, same 32 `fimo-worlds` exports as `modules/fimo_worlds/src/FimoWorlds.zig:41-81`;
//! every export points at the C function of libfimo_worlds_b200.so with the identical signature
//! (`include/fimo_worlds_b200.h` cites the `symbols.zig` lines each prototype replaces), so the Zig wrappers of
//! `pkgs/fimo_worlds_meta` and every C caller work unchanged.  There is no Universe on the Zig side: registry,
//! worlds, groups and the device executor live behind the C ABI.
//!
//! What `init` does that the C library cannot do alone:
//!   * binds the pool-aware `futex_wait` / `futex_wake` of the loaded `fimo_tasks` module
//!     (`pkgs/fimo_tasks_meta/src/symbols.zig:132-166`), so a fiber that waits on a fence the GPU path signals parks
//!     through its own pool and is woken by the library's completion thread;
//!   * hands the context's tracing subscribers to the library (`attachSubscribers`), so device generations show up
//!     as spans in the same `.ftrdb` capture as CPU tasks (`pkgs/fimo_std/src/tracing.zig:387-554`).
const std = @import("std");

const fimo_std = @import("fimo_std");
const ctx = fimo_std.ctx;
const modules = fimo_std.modules;
const tracing = fimo_std.tracing;
const fimo_tasks_meta = @import("fimo_tasks_meta");
const fimo_worlds_meta = @import("fimo_worlds_meta");
const symbols = fimo_worlds_meta.symbols;

const c = @cImport(@cInclude("fimo_worlds_b200.h"));

pub const Module = modules.Module(@This());

comptime {
    _ = Module;
}

pub const fimo_module = .{
    .name = .fimo_worlds,
    .author = "fimo",
    .description = "Multi-threaded state processing (B200 system-group step)",
    .license = "MIT OR APACHE 2.0",
};

pub const fimo_imports = .{fimo_tasks_meta.symbols.all_symbols};

/// `@ptrCast`: the C prototypes use the C spellings of the same extern structs (layouts pinned by
/// tests/test_abi_layout.py against the reference headers).
fn export_(comptime symbol: anytype, comptime function: anytype) struct { symbol: @TypeOf(symbol), value: *const symbol.T } {
    return .{ .symbol = symbol, .value = @ptrCast(function) };
}

pub const fimo_exports = .{
    export_(symbols.resource_register, &c.fimo_worlds_resource_register),
    export_(symbols.resource_unregister, &c.fimo_worlds_resource_unregister),
    export_(symbols.system_register, &c.fimo_worlds_system_register),
    export_(symbols.system_unregister, &c.fimo_worlds_system_unregister),
    export_(symbols.system_group_create, &c.fimo_worlds_system_group_create),
    export_(symbols.system_group_destroy, &c.fimo_worlds_system_group_destroy),
    export_(symbols.system_group_get_world, &c.fimo_worlds_system_group_get_world),
    export_(symbols.system_group_get_label, &c.fimo_worlds_system_group_get_label),
    export_(symbols.system_group_get_pool, &c.fimo_worlds_system_group_get_pool),
    export_(symbols.system_group_add_systems, &c.fimo_worlds_system_group_add_systems),
    export_(symbols.system_group_remove_system, &c.fimo_worlds_system_group_remove_system),
    export_(symbols.system_group_schedule, &c.fimo_worlds_system_group_schedule),
    export_(symbols.system_context_get_group, &c.fimo_worlds_system_context_get_group),
    export_(symbols.system_context_get_generation, &c.fimo_worlds_system_context_get_generation),
    export_(symbols.system_context_allocator_alloc, &c.fimo_worlds_system_context_allocator_alloc),
    export_(symbols.system_context_allocator_resize, &c.fimo_worlds_system_context_allocator_resize),
    export_(symbols.system_context_allocator_remap, &c.fimo_worlds_system_context_allocator_remap),
    export_(symbols.system_context_allocator_free, &c.fimo_worlds_system_context_allocator_free),
    export_(symbols.world_create, &c.fimo_worlds_world_create),
    export_(symbols.world_destroy, &c.fimo_worlds_world_destroy),
    export_(symbols.world_get_label, &c.fimo_worlds_world_get_label),
    export_(symbols.world_get_pool, &c.fimo_worlds_world_get_pool),
    export_(symbols.world_has_resource, &c.fimo_worlds_world_has_resource),
    export_(symbols.world_add_resource, &c.fimo_worlds_world_add_resource),
    export_(symbols.world_remove_resource, &c.fimo_worlds_world_remove_resource),
    export_(symbols.world_lock_resources, &c.fimo_worlds_world_lock_resources),
    export_(symbols.world_unlock_resource_exclusive, &c.fimo_worlds_world_unlock_resource_exclusive),
    export_(symbols.world_unlock_resource_shared, &c.fimo_worlds_world_unlock_resource_shared),
    export_(symbols.world_allocator_alloc, &c.fimo_worlds_world_allocator_alloc),
    export_(symbols.world_allocator_resize, &c.fimo_worlds_world_allocator_resize),
    export_(symbols.world_allocator_remap, &c.fimo_worlds_world_allocator_remap),
    export_(symbols.world_allocator_free, &c.fimo_worlds_world_allocator_free),
};

pub const fimo_events = .{
    .init = init,
    .deinit = deinit,
};

fn init(self: *@This()) void {
    _ = self;
    // Fibers of the real fimo_tasks pools must park through their pool, not on an OS futex.
    c.fimo_worlds_b200_bind_tasks(
        @ptrCast(fimo_tasks_meta.symbols.futex_wait.getGlobal().get()),
        @ptrCast(fimo_tasks_meta.symbols.futex_wake.getGlobal().get()),
    );
}

fn deinit(self: *@This()) void {
    _ = c.fimo_worlds_b200_trace_subscribe(null, 0);
    c.fimo_worlds_b200_bind_tasks(null, null); // back to the built-in parking lot
    self.* = undefined;
}

/// Forwards device generations to the given subscribers — the same list the application passed to the context
/// (`tracing.Cfg.subscribers`).  `tracing.Subscriber` is `extern struct { data, on_event }`
/// (`tracing.zig:562-565`), layout-identical to `FimoTracingSubscriber`.
pub fn attachSubscribers(subscribers: []const tracing.Subscriber) error{InvalidArgument}!void {
    if (c.fimo_worlds_b200_trace_subscribe(@ptrCast(subscribers.ptr), subscribers.len) != c.FIMO_STATUS_OK)
        return error.InvalidArgument;
}

/// A failing export returns `.err` and leaves the error's name in a thread-local of the library
/// (`NotFound`, `Deadlock`, `Duplicate`, `InUse`, `InitError`, `OutOfMemory`, `NoDevice`, `DeviceError`); callers
/// that want it in the context result (the reference's `ctx.setResult`, `FimoWorlds.zig:117-120`) call this
/// right after the failing call on the same thread.
pub fn forwardLastError() void {
    const raw = c.fimo_worlds_b200_last_error();
    if (raw == null) return;
    const name = std.mem.span(raw);
    inline for (@typeInfo(Error).error_set.?) |e| {
        if (std.mem.eql(u8, e.name, name)) {
            ctx.setResult(.initErr(.initError(@field(Error, e.name))));
            return;
        }
    }
    ctx.setResult(.initErr(.initError(error.Unknown)));
}

/// Error names the library reports (`fimo_worlds_b200_last_error`); the first five are the reference's
/// (`Universe.zig:174-210`, `SystemGroup.zig:411`, `World.zig:109,121`, `SystemContext.zig:71`).
pub const Error = error{ NotFound, Deadlock, Duplicate, InUse, InitError, OutOfMemory, NoDevice, DeviceError, InvalidArgument };
