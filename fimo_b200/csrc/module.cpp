// module.cpp — the record that makes libfimo_worlds_b200.so loadable as a fimo module.
//
// The reference's loader opens `module.fimo_module` and calls `fimo_impl_module_export_iterator`
// (pkgs/fimo_std/include/fimo_std/impl/module.h:40,61-62; pkgs/fimo_std/src/context/modules/ModuleHandle.zig:123-223),
// which walks the pointers the linker gathered in the ELF section `fimo_module` and hands every non-NULL
// `FimoModulesExport` to an inspector (pkgs/fimo_std/src/modules/exports.zig:2016-2044).  Here the record is
// plain constant data: name, the "fimo-tasks" namespace + the two futex symbols this path imports
// (pkgs/fimo_tasks_meta/src/symbols.zig:132-160), and the 32 "fimo-worlds" symbols (symbols.zig:22-58) as static
// exports — the same functions `fimo_worlds_b200_exports()` lists.
#include "fimo_worlds_b200.h"

#ifndef FIMO_B200_CONTEXT_VERSION_MAJOR
#define FIMO_B200_CONTEXT_VERSION_MAJOR 0
#define FIMO_B200_CONTEXT_VERSION_MINOR 2
#define FIMO_B200_CONTEXT_VERSION_PATCH 0
#endif

static_assert(sizeof(FimoB200_Version) == 56, "FimoVersion");
static_assert(sizeof(FimoB200_ModulesSymbolImport) == 72, "FimoModulesSymbolImport");
static_assert(sizeof(FimoB200_ModulesSymbolExport) == 88, "FimoModulesSymbolExport");
static_assert(sizeof(FimoB200_ModulesExport) == 208, "FimoModulesExport");

namespace {

#define FB_VERSION {FIMO_B200_CONTEXT_VERSION_MAJOR, FIMO_B200_CONTEXT_VERSION_MINOR, FIMO_B200_CONTEXT_VERSION_PATCH, nullptr, 0, nullptr, 0}
#define FB_SYM(n) {(const void *)&fimo_worlds_##n, 0, FB_VERSION, #n, "fimo-worlds"}

const FimoB200_ModulesSymbolExport kSymbolExports[] = {
        // `all_symbols`, pkgs/fimo_worlds_meta/src/symbols.zig:22-58
        FB_SYM(resource_register), FB_SYM(resource_unregister), FB_SYM(system_register), FB_SYM(system_unregister),
        FB_SYM(system_group_create), FB_SYM(system_group_destroy), FB_SYM(system_group_get_world), FB_SYM(system_group_get_label),
        FB_SYM(system_group_get_pool), FB_SYM(system_group_add_systems), FB_SYM(system_group_remove_system),
        FB_SYM(system_group_schedule), FB_SYM(system_context_get_group), FB_SYM(system_context_get_generation),
        FB_SYM(system_context_allocator_alloc), FB_SYM(system_context_allocator_resize), FB_SYM(system_context_allocator_remap),
        FB_SYM(system_context_allocator_free), FB_SYM(world_create), FB_SYM(world_destroy), FB_SYM(world_get_label),
        FB_SYM(world_get_pool), FB_SYM(world_has_resource), FB_SYM(world_add_resource), FB_SYM(world_remove_resource),
        FB_SYM(world_lock_resources), FB_SYM(world_unlock_resource_exclusive), FB_SYM(world_unlock_resource_shared),
        FB_SYM(world_allocator_alloc), FB_SYM(world_allocator_resize), FB_SYM(world_allocator_remap), FB_SYM(world_allocator_free),
};
const FimoB200_ModulesNamespaceImport kNamespaceImports[] = {{"fimo-tasks"}};
const FimoB200_ModulesSymbolImport kSymbolImports[] = {
        {FB_VERSION, "futex_wait", "fimo-tasks"},
        {FB_VERSION, "futex_wake", "fimo-tasks"},
};

const FimoB200_ModulesExport kModule = {
        nullptr,
        FB_VERSION,
        "fimo_worlds",
        "B200-native fimo_worlds system-group step (CUDA sm_100a device executor behind the fimo_worlds_meta ABI)",
        "fimo_b200",
        "MIT OR Apache-2.0",
        nullptr, 0, // parameters
        nullptr, 0, // resources
        kNamespaceImports, sizeof(kNamespaceImports) / sizeof(kNamespaceImports[0]),
        kSymbolImports, sizeof(kSymbolImports) / sizeof(kSymbolImports[0]),
        kSymbolExports, sizeof(kSymbolExports) / sizeof(kSymbolExports[0]),
        nullptr, 0, // dynamic symbol exports
        nullptr, 0, // modifiers
};

} // namespace

extern "C" {
// One pointer per exported module in section `fimo_module`; `used` + `retain` keep it through --gc-sections.
__attribute__((section("fimo_module"), used, retain)) const FimoB200_ModulesExport *const fimo_b200_module_export_fimo_worlds = &kModule;
// A NULL entry, as the reference always emits (exports.zig:2006 `embedStaticModuleExport(null)`): the iterator skips it.
__attribute__((section("fimo_module"), used, retain)) const FimoB200_ModulesExport *const fimo_b200_module_export_null = nullptr;

extern const FimoB200_ModulesExport *const __start_fimo_module[] __attribute__((visibility("hidden")));
extern const FimoB200_ModulesExport *const __stop_fimo_module[] __attribute__((visibility("hidden")));

void fimo_impl_module_export_iterator(FimoB200_ModuleInspector inspector, void *data) {
    for (const FimoB200_ModulesExport *const *it = __start_fimo_module; it != __stop_fimo_module; ++it) {
        if (!*it) continue;
        if (!inspector(*it, data)) return;
    }
}
}
