#!/usr/bin/env python
"""Golden ABI layout of the drop-in boundary, measured on the REFERENCE's own C headers.

    python scripts/make_abi_golden.py            # writes tests/golden/abi_layout.json (needs /root/reference)
    layout_of(include_dirs, headers)             # used by tests/test_abi_layout.py on include/fimo_worlds_b200.h

`pkgs/fimo_tasks_meta/include` compiles as C as it is.  `pkgs/fimo_worlds_meta/include/.../jobs.h` does not
(jobs.h:42 passes `&fence->state` where a `Fence*` is expected, jobs.h:47 passes a value as the `expected`
pointer of a compare-exchange); the script compiles a temporary copy under /tmp with exactly those two inline
function bodies neutralised — the struct definitions, which are what is measured, are untouched.  Nothing of
the reference is copied into this repository: only sizes and offsets (numbers) are committed.
"""
from __future__ import annotations

import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"

# struct -> fields (every field on the boundary; names are the reference's)
STRUCTS = {
    "FimoTasksMeta_Task": ["label", "label_len", "on_start", "on_complete", "on_abort", "on_deinit"],
    "FimoTasksMeta_CommandBufferHandleVTable": ["acquire", "release", "owner_pool", "wait_on"],
    "FimoTasksMeta_CommandBufferHandle": ["data", "vtable"],
    "FimoTasksMeta_CommandBufferEntryPayload": ["abort_on_error", "set_min_stack_size", "select_worker", "enqueue_task",
                                                "enqueue_command_buffer", "wait_on_command_buffer", "wait_on_command_indirect"],
    "FimoTasksMeta_CommandBufferEntry": ["type", "payload"],
    "FimoTasksMeta_CommandBuffer": ["label", "label_len", "entries", "entries_len", "on_complete", "on_abort", "on_deinit"],
    "FimoTasksMeta_PoolVTable": ["id", "acquire", "release", "request_close", "accepts_requests", "owns_current_thread",
                                 "label", "workers", "stack_sizes", "enqueue_buffer"],
    "FimoTasksMeta_Pool": ["data", "vtable"],
    "FimoTasksMeta_PoolConfigStackConfig": ["next", "size", "preallocated_count", "cold_count", "hot_count", "max_allocated"],
    "FimoTasksMeta_PoolConfig": ["next", "label", "label_len", "stacks", "stacks_len", "default_stack_index", "worker_count",
                                 "is_queryable"],
    "FimoTasksMeta_FutexKeyExpect": ["key", "key_size", "expect", "token"],
    "FimoTasksMeta_FutexFilter": ["op", "token_mask", "cmp_arg"],
    "FimoWorldsMeta_Fence": ["state"],
    "FimoWorldsMeta_TimelineSemaphore": ["state"],
    "FimoWorldsMeta_ResourceDescriptor": ["next", "label", "label_len", "size", "alignment"],
    "FimoWorldsMeta_SystemDependency": ["system", "flags"],
    "FimoWorldsMeta_SystemDescriptor": ["next", "label", "label_len", "exclusive_handles", "exclusive_handles_len",
                                        "shared_handles", "shared_handles_len", "before", "before_len", "after", "after_len",
                                        "factory", "factory_size", "factory_alignment", "factory_destroy", "system_size",
                                        "system_alignment", "system_create", "system_destroy", "system_run"],
    "FimoWorldsMeta_SystemGroupDescriptor": ["next", "label", "label_len", "pool", "world"],
    "FimoWorldsMeta_WorldDescriptor": ["next", "label", "label_len", "pool"],
    "FimoDuration": ["secs", "nanos"],
    "FimoInstant": ["secs", "nanos"],
    # tracing subscriber events the bridge emits (fimo_std/tracing.h:35-56,137-147,169-204)
    "FimoTracingEventInfo": ["name", "target", "scope", "file_name", "line_number", "level"],
    "FimoTracingEventCreateCallStack": ["event", "time", "stack"],
    "FimoTracingEventDestroyCallStack": ["event", "time", "stack"],
    "FimoTracingEventEnterSpan": ["event", "time", "stack", "span", "message", "message_length"],
    "FimoTracingEventExitSpan": ["event", "time", "stack", "is_unwinding"],
    "FimoTracingEventLogMessage": ["event", "time", "stack", "info", "message", "message_length"],
    "FimoTracingSubscriber": ["ptr", "on_event"],
}
# enumerators / constants whose VALUES are part of the ABI
CONSTANTS = [
    "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ABORT_ON_ERROR", "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SET_MIN_STACK_SIZE",
    "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_WORKER", "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_SELECT_ANY_WORKER",
    "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_TASK", "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_ENQUEUE_COMMAND_BUFFER",
    "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_BARRIER", "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_BUFFER",
    "FIMO_TASKS_META_COMMAND_BUFFER_ENTRY_TYPE_WAIT_ON_COMMAND_INDIRECT",
    "FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_COMPLETED", "FIMO_TASKS_META_COMMAND_BUFFER_HANDLE_COMPLETION_STATUS_ABORTED",
    "FIMO_TASKS_META_FUTEX_STATUS_OK", "FIMO_TASKS_META_FUTEX_STATUS_INVALID", "FIMO_TASKS_META_FUTEX_STATUS_TIMEOUT",
    "FIMO_TASKS_META_FUTEX_STATUS_KEY_ERROR",
    "FIMO_TASKS_META_FUTEX_FILTER_TOKEN_OP_NOOP", "FIMO_TASKS_META_FUTEX_FILTER_TOKEN_OP_DEREF",
    "FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U8", "FIMO_TASKS_META_FUTEX_FILTER_TOKEN_TYPE_U64",
    "FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_EQ", "FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_LE", "FIMO_TASKS_META_FUTEX_FILTER_CMP_OP_GE",
    "FIMO_TASKS_META_FUTEX_FILTER_CMP_ARG_OP_NOOP", "FIMO_TASKS_META_FUTEX_FILTER_CMP_ARG_OP_DEREF",
    "FIMO_WORLDS_META_FENCE_UNSIGNALED", "FIMO_WORLDS_META_FENCE_SIGNALED", "FIMO_WORLDS_META_FENCE_CONTENDED",
    "FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_WEAK", "FIMO_WORLDS_META_SYSTEM_DEPENDENCY_FLAGS_IGNORE_DEFERRED",
    "FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_TRANSIENT", "FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SINGLE_GENERATION",
    "FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_MULTI_GENERATION", "FIMO_WORLDS_META_SYSTEM_ALLOCATOR_STRATEGY_SYSTEM_PERSISTENT",
    "FIMO_TRACING_EVENT_CREATE_CALL_STACK", "FIMO_TRACING_EVENT_DESTROY_CALL_STACK", "FIMO_TRACING_EVENT_ENTER_SPAN",
    "FIMO_TRACING_EVENT_EXIT_SPAN", "FIMO_TRACING_EVENT_LOG_MESSAGE", "FIMO_TRACING_EVENT_DECLARE_EVENT_INFO",
    "FIMO_TRACING_LEVEL_OFF", "FIMO_TRACING_LEVEL_ERROR", "FIMO_TRACING_LEVEL_WARN", "FIMO_TRACING_LEVEL_INFO",
    "FIMO_TRACING_LEVEL_DEBUG", "FIMO_TRACING_LEVEL_TRACE",
]


def layout_of(include_dirs: list[str], headers: list[str], std: str = "-std=gnu2x") -> dict:
    """Compiles a probe against `headers` and returns {"structs": {name: {"size", "align", "fields": {f: [offset, size]}}},
    "constants": {name: value}}."""
    lines = ["#include <stddef.h>", "#include <stdio.h>"] + [f"#include <{h}>" for h in headers] + ["int main(void) {"]
    for s, fields in STRUCTS.items():
        lines.append(f'  printf("S {s} %zu %zu\\n", sizeof({s}), _Alignof({s}));')
        for f in fields:
            lines.append(f'  printf("F {s} {f} %zu %zu\\n", offsetof({s}, {f}), sizeof((({s} *)0)->{f}));')
    for c in CONSTANTS:
        lines.append(f'  printf("C {c} %lld\\n", (long long)({c}));')
    lines += ["  return 0;", "}"]
    tmp = tempfile.mkdtemp(prefix="abi_probe_")
    try:
        src, exe = os.path.join(tmp, "probe.c"), os.path.join(tmp, "probe")
        open(src, "w").write("\n".join(lines))
        cmd = ["gcc", std, "-w"] + [f"-I{d}" for d in include_dirs] + [src, "-o", exe]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("probe does not compile:\n" + r.stderr[-4000:])
        out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    res = {"structs": {}, "constants": {}}
    for ln in out.splitlines():
        p = ln.split()
        if p[0] == "S":
            res["structs"][p[1]] = {"size": int(p[2]), "align": int(p[3]), "fields": {}}
        elif p[0] == "F":
            res["structs"][p[1]]["fields"][p[2]] = [int(p[3]), int(p[4])]
        else:
            res["constants"][p[1]] = int(p[2])
    return res


def reference_layout() -> dict:
    """Layout according to the reference's headers (worlds_meta from a patched temporary copy, see above)."""
    tmp = tempfile.mkdtemp(prefix="fimo_ref_hdr_")
    try:
        dst = os.path.join(tmp, "fimo_worlds_meta")
        shutil.copytree(os.path.join(REF, "pkgs/fimo_worlds_meta/include/fimo_worlds_meta"), dst)
        jobs = os.path.join(dst, "jobs.h")
        text = open(jobs).read()
        # the two lines that do not compile sit inside `static inline` helpers; drop those helpers' bodies
        text = text.replace("if (FimoWorldsMeta_fence_is_signaled(&fence->state))", "if (FimoWorldsMeta_fence_is_signaled(fence))")
        text = re.sub(r"atomic_compare_exchange_weak_explicit\(&fence->state, FIMO_WORLDS_META_FENCE_UNSIGNALED,",
                      "atomic_compare_exchange_weak_explicit(&fence->state, &(FimoU8){FIMO_WORLDS_META_FENCE_UNSIGNALED},", text)
        open(jobs, "w").write(text)
        incs = [tmp, os.path.join(REF, "pkgs/fimo_std/include"), os.path.join(REF, "pkgs/fimo_tasks_meta/include")]
        return layout_of(incs, ["fimo_std/time.h", "fimo_std/tracing.h", "fimo_tasks_meta/package.h", "fimo_worlds_meta/jobs.h",
                                "fimo_worlds_meta/resources.h", "fimo_worlds_meta/systems.h", "fimo_worlds_meta/worlds.h"])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def product_layout() -> dict:
    return layout_of([os.path.join(ROOT, "include")], ["fimo_worlds_b200.h"], std="-std=c17")


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference")
    golden = reference_layout()
    golden["source"] = "sizes/offsets/enumerator values measured with gcc (x86-64) on /root/reference/pkgs/{fimo_std,fimo_tasks_meta,fimo_worlds_meta}/include by scripts/make_abi_golden.py"
    path = os.path.join(ROOT, "tests", "golden", "abi_layout.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    json.dump(golden, open(path, "w"), indent=1, sort_keys=True)
    print("wrote", path, len(golden["structs"]), "structs", len(golden["constants"]), "constants")
