//! Root of the `fimo_worlds_b200` module (compare `modules/fimo_worlds/src/root.zig`): one module in the bundle.
const std = @import("std");

pub const fimo_std = @import("fimo_std");
pub const fimo_worlds_meta = @import("fimo_worlds_meta");

pub const FimoWorldsB200 = @import("FimoWorldsB200.zig");

pub const fimo_module_bundle = fimo_std.modules.ModuleBundle(.{
    FimoWorldsB200.Module,
});

comptime {
    _ = fimo_module_bundle;
}
