// tracing.cpp — device generations as fimo tracing events.
//
// The reference routes spans and messages to *subscribers*: `{data, on_event(data, *const Event)}` records that
// receive extern structs beginning with an `Event` tag (pkgs/fimo_std/src/tracing.zig:387-554, 562-565;
// pkgs/fimo_std/include/fimo_std/tracing.h:72-204).  Spans are entered and exited on a *call stack*; the CPU path
// creates one per task (modules/fimo_tasks/src/Worker.zig:225-248).  Device work has no fiber, so the bridge gives
// every system group one call stack of its own (token = the group's address) and brackets each finished generation
// with enter_span / exit_span on it, timed by the CUDA events around the generation.
#include <chrono>
#include <cstring>
#include <mutex>
#include <unordered_set>

#include "rt.h"

namespace fb {
namespace {

constexpr size_t kMaxSubscribers = 8;
std::mutex g_trace_m; // subscribers, known stacks, and the delivery itself (events of one group stay in order)
FimoTracingSubscriber g_subs[kMaxSubscribers];
size_t g_n_subs = 0;
std::unordered_set<const void *> g_stacks; // groups whose call stack has been created
bool g_declared = false;
std::atomic<bool> g_any{false};

const FimoTracingEventInfo kGenerationSpan = {"system_group generation", "fimo_worlds_b200", "device", __FILE__, __LINE__, FIMO_TRACING_LEVEL_TRACE};
const FimoTracingEventInfo kKernelInfo = {"device launch", "fimo_worlds_b200", "device", __FILE__, __LINE__, FIMO_TRACING_LEVEL_TRACE};

FimoInstant instant_of(int64_t monotonic_ns) {
    FimoInstant t;
    t.secs = (uint64_t)(monotonic_ns / 1000000000ll), t.nanos = (uint32_t)(monotonic_ns % 1000000000ll);
    return t;
}

template <class E> void emit(const E &ev) { // g_trace_m held
    for (size_t i = 0; i < g_n_subs; i++) g_subs[i].on_event(g_subs[i].ptr, &ev.event);
}

void destroy_stack_locked(const void *stack, int64_t now_ns) {
    FimoTracingEventDestroyCallStack ev{};
    ev.event = FIMO_TRACING_EVENT_DESTROY_CALL_STACK, ev.time = instant_of(now_ns), ev.stack = const_cast<void *>(stack);
    emit(ev);
}

int64_t now_ns() {
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

} // namespace

bool trace_bridge_active() { return g_any.load(std::memory_order_acquire); }

void trace_bridge_generation(const void *group, const FimoB200_TraceEvent &t) {
    std::lock_guard<std::mutex> lock(g_trace_m);
    if (!g_n_subs) return;
    void *stack = const_cast<void *>(group);
    const int64_t end_ns = t.host_complete_ns;
    int64_t begin_ns = end_ns - (int64_t)(t.device_ms * 1.0e6);
    if (begin_ns > end_ns) begin_ns = end_ns;
    if (!g_declared) {
        FimoTracingEventDeclareEventInfo d{};
        d.event = FIMO_TRACING_EVENT_DECLARE_EVENT_INFO, d.info = &kGenerationSpan;
        emit(d);
        d.info = &kKernelInfo;
        emit(d);
        g_declared = true;
    }
    if (g_stacks.insert(group).second) {
        FimoTracingEventCreateCallStack c{};
        c.event = FIMO_TRACING_EVENT_CREATE_CALL_STACK, c.time = instant_of(begin_ns), c.stack = stack;
        emit(c);
    }
    char msg[256];
    int n = std::snprintf(msg, sizeof msg, "%.*s generation %llu: %u kernels, %.6f ms %s", (int)t.group_label_len,
                          t.group_label ? t.group_label : "", (unsigned long long)t.generation, t.kernels, t.device_ms,
                          t.kernels ? "on device" : "on the host pool");
    if (n < 0) n = 0;
    if ((size_t)n >= sizeof msg) n = (int)sizeof msg - 1;
    FimoTracingEventEnterSpan enter{};
    enter.event = FIMO_TRACING_EVENT_ENTER_SPAN, enter.time = instant_of(begin_ns), enter.stack = stack;
    enter.span = &kGenerationSpan, enter.message = msg, enter.message_length = (FimoUSize)n;
    emit(enter);
    char note[160];
    int m = std::snprintf(note, sizeof note, "%s, enqueued %.3f us before completion",
                          t.fused_world_step ? "fused world-step executor" : (t.kernels ? "one kernel per system" : "host systems buffer"), (double)(end_ns - t.host_enqueue_ns) / 1.0e3);
    if (m < 0) m = 0;
    if ((size_t)m >= sizeof note) m = (int)sizeof note - 1;
    FimoTracingEventLogMessage log{};
    log.event = FIMO_TRACING_EVENT_LOG_MESSAGE, log.time = instant_of(begin_ns), log.stack = stack;
    log.info = &kKernelInfo, log.message = note, log.message_length = (FimoUSize)m;
    emit(log);
    FimoTracingEventExitSpan exit_{};
    exit_.event = FIMO_TRACING_EVENT_EXIT_SPAN, exit_.time = instant_of(end_ns), exit_.stack = stack, exit_.is_unwinding = false;
    emit(exit_);
}

void trace_bridge_group_destroyed(const void *group) {
    if (!trace_bridge_active()) return;
    std::lock_guard<std::mutex> lock(g_trace_m);
    if (g_stacks.erase(group)) destroy_stack_locked(group, now_ns());
}

} // namespace fb

extern "C" FIMO_B200_EXPORT FimoStatus fimo_worlds_b200_trace_subscribe(const FimoTracingSubscriber *subscribers, FimoUSize count) {
    using namespace fb;
    if (count > kMaxSubscribers || (count && !subscribers)) {
        set_last_error("InvalidArgument");
        return FIMO_STATUS_ERROR;
    }
    for (FimoUSize i = 0; i < count; i++)
        if (!subscribers[i].on_event) {
            set_last_error("InvalidArgument");
            return FIMO_STATUS_ERROR;
        }
    std::lock_guard<std::mutex> lock(g_trace_m);
    if (g_n_subs) { // the old subscribers see their stacks end
        const int64_t t = now_ns();
        for (const void *stack : g_stacks) destroy_stack_locked(stack, t);
    }
    g_stacks.clear();
    g_declared = false;
    g_n_subs = count;
    for (FimoUSize i = 0; i < count; i++) g_subs[i] = subscribers[i];
    g_any.store(count != 0, std::memory_order_release);
    return FIMO_STATUS_OK;
}
