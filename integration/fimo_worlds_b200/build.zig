//! Build script of the `fimo_worlds_b200` module: the drop-in replacement of `modules/fimo_worlds` whose
//! system-group step runs on a B200 (libfimo_worlds_b200.so).
//!
//! Place this directory at `modules/fimo_worlds_b200` of a fimo checkout and add it to the module list of the
//! top-level build next to `modules/fimo_worlds` (the two export the same `fimo-worlds` symbols: load one).
//! It follows `modules/fimo_worlds/build.zig:5-45` — same packages, same `builder.addModule` — and adds the
//! one step the reference does not have: `make` builds the CUDA library with nvcc for sm_100a and the module
//! links it.  `-Dfimo-worlds-b200=<path>` names the checkout of this repository (default: `third_party/`).
const std = @import("std");

const build_internals = @import("tools/build-internals");

pub fn configure(builder: *build_internals.FimoBuild) void {
    const b = builder.build;
    const fimo_std_pkg = builder.getPackage("fimo_std");
    const fimo_tasks_meta_pkg = builder.getPackage("fimo_tasks_meta");
    const fimo_worlds_meta_pkg = builder.getPackage("fimo_worlds_meta");

    const repo = b.option(
        []const u8,
        "fimo-worlds-b200",
        "Path of the fimo-worlds-b200 checkout (include/, fimo_b200/csrc/)",
    ) orelse "../../third_party/fimo-worlds-b200";

    // nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo + g++ -std=c++17, see fimo_b200/csrc/Makefile.
    // The context version is compiled into the module record the loader checks (pkgs/fimo_std/build.zig:171-194).
    const make = b.addSystemCommand(&.{ "make", "-C" });
    make.addDirectoryArg(b.path(b.pathJoin(&.{ repo, "fimo_b200", "csrc" })));
    make.addArg(b.fmt(
        "CXXFLAGS+=-DFIMO_B200_CONTEXT_VERSION_MAJOR={d} -DFIMO_B200_CONTEXT_VERSION_MINOR={d} -DFIMO_B200_CONTEXT_VERSION_PATCH={d}",
        .{
            build_internals.fimo_version.major,
            build_internals.fimo_version.minor,
            build_internals.fimo_version.patch,
        },
    ));

    const fimo_worlds_b200 = b.addModule("fimo_worlds_b200", .{
        .root_source_file = b.path("src/root.zig"),
        .target = builder.graph.target,
        .optimize = builder.graph.optimize,
        .link_libc = true,
    });
    fimo_worlds_b200.addImport("fimo_std", fimo_std_pkg.root_module);
    fimo_worlds_b200.addImport("fimo_tasks_meta", fimo_tasks_meta_pkg.root_module);
    fimo_worlds_b200.addImport("fimo_worlds_meta", fimo_worlds_meta_pkg.root_module);
    fimo_worlds_b200.addIncludePath(b.path(b.pathJoin(&.{ repo, "include" })));
    fimo_worlds_b200.addLibraryPath(b.path(b.pathJoin(&.{ repo, "fimo_b200" })));
    fimo_worlds_b200.addRPath(b.path(b.pathJoin(&.{ repo, "fimo_b200" })));
    fimo_worlds_b200.linkSystemLibrary("fimo_worlds_b200", .{});

    const module = builder.addModule(.{
        .name = "fimo_worlds_b200",
        .root_module = fimo_worlds_b200,
    });
    // the shared library must exist before the module is linked
    module.step.dependOn(&make.step);

    // The reference's behavioural tests of the path live in the meta package (pkgs/fimo_worlds_meta/src/
    // {systems,worlds,resources}.zig) and run against whichever module exports `fimo-worlds`: with this module in
    // the test context's module list they exercise the CUDA path unchanged.
    _ = module.addTest(.{
        .name = "fimo_worlds_b200_test",
        .step = .{
            .module = blk: {
                const t = b.createModule(.{
                    .root_source_file = b.path("src/root.zig"),
                    .target = builder.graph.target,
                    .optimize = builder.graph.optimize,
                    .link_libc = true,
                });
                t.addImport("fimo_std", fimo_std_pkg.root_module);
                t.addImport("fimo_tasks_meta", fimo_tasks_meta_pkg.root_module);
                t.addImport("fimo_worlds_meta", fimo_worlds_meta_pkg.root_module);
                t.addIncludePath(b.path(b.pathJoin(&.{ repo, "include" })));
                t.addLibraryPath(b.path(b.pathJoin(&.{ repo, "fimo_b200" })));
                t.linkSystemLibrary("fimo_worlds_b200", .{});
                break :blk t;
            },
        },
    });
}

pub fn build(b: *std.Build) void {
    _ = b;
}
