"""generations of a system group as fimo tracing events.

A fimo tracing subscriber is `{data, on_event(data, *const Event)}` and receives extern structs that begin with
the event tag (pkgs/fimo_std/src/tracing.zig:387-554, 562-565; include/fimo_std/tracing.h:72-204).  The layout of
the structs the bridge emits is pinned against the reference header in tests/test_abi_layout.py; here the
*sequence* is checked, decoded through ctypes mirrors written from the header independently of the C++ side:
declare_event_info, create_call_stack (one per group), then per generation enter_span / log_message / exit_span on
that stack, and destroy_call_stack when the group dies or the subscribers are detached."""
import ctypes as C

import pytest

from fimo_b200 import abi, library
from fimo_b200.worlds import System, TypedResource, World


class Recorder:
    def __init__(self, lib):
        self.lib, self.events = lib, []
        self.cb = abi.TracingOnEventFn(lambda _d, ev: self.events.append(abi.decode_tracing_event(ev)))
        self.sub = (abi.TracingSubscriber * 1)(abi.TracingSubscriber(None, self.cb))

    def __enter__(self):
        assert self.lib.fimo_worlds_b200_trace_subscribe(self.sub, 1) == 0
        return self

    def __exit__(self, *exc):
        assert self.lib.fimo_worlds_b200_trace_subscribe(None, 0) == 0


def check_sequence(events, label, generations, kernels_positive):
    kinds = [k for k, _ in events]
    declares = [e for k, e in events if k == abi.TRACING_EVENT_DECLARE_EVENT_INFO]
    assert {d["name"] for d in declares} == {"system_group generation", "device launch"}
    creates = [e for k, e in events if k == abi.TRACING_EVENT_CREATE_CALL_STACK]
    assert len(creates) == 1
    stack = creates[0]["stack"]
    assert stack
    body = [(k, e) for k, e in events if k in (abi.TRACING_EVENT_ENTER_SPAN, abi.TRACING_EVENT_LOG_MESSAGE, abi.TRACING_EVENT_EXIT_SPAN)]
    assert [k for k, _ in body] == [abi.TRACING_EVENT_ENTER_SPAN, abi.TRACING_EVENT_LOG_MESSAGE, abi.TRACING_EVENT_EXIT_SPAN] * len(generations)
    assert kinds.index(abi.TRACING_EVENT_CREATE_CALL_STACK) < kinds.index(abi.TRACING_EVENT_ENTER_SPAN)
    span_ids = {d["info"] for d in declares}
    last_exit = 0
    for i, gen in enumerate(generations):
        (_, enter), (_, log), (_, exit_) = body[3 * i:3 * i + 3]
        assert enter["stack"] == log["stack"] == exit_["stack"] == stack
        assert enter["info"] in span_ids and log["info"] in span_ids  # declared before use, static lifetime
        assert enter["name"] == "system_group generation" and enter["target"] == "fimo_worlds_b200" and enter["level"] == 5
        assert enter["message"].startswith(f"{label} generation {gen}: ")
        assert ("0 kernels" not in enter["message"]) == kernels_positive
        assert enter["ns"] <= exit_["ns"] and not exit_["is_unwinding"]
        assert exit_["ns"] >= last_exit  # one group's generations are delivered in order
        last_exit = exit_["ns"]
    return stack


def test_host_generations_become_spans():
    lib = library.load()
    world = World.init(lib, "trace-world")
    group = world.add_system_group("trace-group")
    resource = TypedResource.register_typed(lib, C.c_uint32, "trace-resource")
    resource.add_value(world, 0)
    sys_ = System.register(lib, "trace-system", lambda ctx, ex, sh, fence: None, shared=[resource])
    group.add_systems([sys_])
    group.run()  # before subscribing: not reported
    rec = Recorder(lib)
    with rec:
        for _ in range(3):
            group.run()
        n_before_destroy = len(rec.events)
        group.remove_system(sys_)
        group.deinit()
        destroys = [e for k, e in rec.events[n_before_destroy:] if k == abi.TRACING_EVENT_DESTROY_CALL_STACK]
        assert len(destroys) == 1
    stack = check_sequence(rec.events, "trace-group", [1, 2, 3], kernels_positive=False)
    assert destroys[0]["stack"] == stack
    sys_.unregister()
    resource.remove_value(world)
    resource.unregister()
    world.deinit()
    assert lib.fimo_worlds_b200_trace_subscribe(None, 9) != 0  # more than 8 subscribers: InvalidArgument


@pytest.mark.gpu
def test_device_generations_become_spans():
    from fimo_b200 import synth

    lib = library.load()
    rig = synth.WorldStepRig(lib, 64, 1024, 4)
    rec = Recorder(lib)
    try:
        rig.step(2)
        with rec:
            rig.step(4)
            events = list(rec.events)  # detach destroys the live stack
    finally:
        rig.close()
    check_sequence(events, "world-step", [2, 3, 4, 5], kernels_positive=True)
    assert [k for k, _ in rec.events][-1] == abi.TRACING_EVENT_DESTROY_CALL_STACK
    spans = [e for k, e in events if k == abi.TRACING_EVENT_ENTER_SPAN]
    exits = [e for k, e in events if k == abi.TRACING_EVENT_EXIT_SPAN]
    assert all(0 < x["ns"] - s["ns"] < 50_000_000 for s, x in zip(spans, exits))  # device time between the CUDA events
