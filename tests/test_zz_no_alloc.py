"""fimo_b200_cull_pack_static — the packed cull read-back for hosts that must not allocate device memory.

Some users of this library cannot let it call cudaMalloc: a frame loop that cannot afford the device-wide stall of
cudaMalloc / cudaFree, an engine that gave all of HBM to its own pool at start-up, a plug-in host that forbids driver
allocations after initialisation.  fimo_b200_group_schedule_readback allocates its HBM staging slots itself, so those
users get a second entry point that packs through a fixed 8 MiB array inside the library's device image instead.

The GPU test drives both entry points over the same cull outputs and demands the same words from both (integer data:
bit for bit, the bar of every parity test here) on several calls in a row — the array is shared by all calls and keeps
the previous call's contents, so the later calls are the ones that would expose stale data.

This file sorts last on purpose (`pytest --collect-only` lists it after every other test): with `pytest -x` a failure
here cannot hide another test."""
import ctypes as C

import numpy as np
import pytest

from fimo_b200 import abi, synth


def _reject(lib, *args) -> str:
    assert lib.fimo_b200_cull_pack_static(*args) != abi.STATUS_OK
    return lib.last_error()


def _check_limits(lib):
    """Sizes beyond the fixed array are refused before anything touches the device (so this part needs no GPU)."""
    dummy = (C.c_uint32 * 64)()
    p = C.addressof(dummy)
    words = (abi.PACK_STATIC_MAX_WORLDS + 1) * 3 + 8
    # more worlds than one call packs
    assert _reject(lib, p, p, p, abi.PACK_STATIC_MAX_WORLDS + 1, 1, 1, 0, p, 4 * words, None) == "TooLarge"
    # a list that does not fit: 4096 x 1024 entities all visible is 16 MiB of indices
    assert _reject(lib, p, p, p, 4096, 1024, 32, 0, p, 64 << 20, None) == "TooLarge"
    # the masks alone do not fit
    assert _reject(lib, p, p, p, 8192, 1 << 20, 1 << 15, 0, p, 1 << 40, None) == "TooLarge"
    # `out` too small for counts + offsets + masks, or missing
    assert _reject(lib, p, p, p, 16, 64, 2, 0, p, 4 * (8 + 2 * 16 + 16 * 2) - 4, None) == "InvalidArgument"
    assert _reject(lib, p, p, p, 16, 64, 2, 0, None, 1 << 20, None) == "InvalidArgument"


def test_static_pack_refuses_what_does_not_fit():
    from fimo_b200 import library

    _check_limits(library.load())


@pytest.mark.gpu
@pytest.mark.parametrize("W,E,D,margin", [(24, 1024, 4, 1.0), (700, 96, 4, 1.0), (24, 1024, 4, 0.5)])
def test_static_pack_equals_schedule_readback_call_after_call(product, W, E, D, margin):
    """Four steps with different frusta (so the visible sets, and their sizes, change from call to call).  After each
    step the existing entry point has delivered its packed buffer; the new one then packs the very same resources
    (device pointers taken from locked resource views — the caller's inputs; a pinned host buffer — the caller's
    output) on a caller-owned stream.  Both buffers must agree word for word over everything that is defined: header,
    counts, offsets, masks and the list up to min(total, capacity).  Past the total the new entry point promises zeros
    (never the previous call's indices), which the second and later calls put to the test.
    margin < 1: the list does not fit and both must cut it at the same place and say so."""
    import torch

    from fimo_b200.worlds import Fence

    _check_limits(product)
    data = synth.generate(W, E, D)
    frusta = [synth.generate(W, E, D, world0=100 * (i + 1))["frustum"] for i in (0, 1, 2, 0)]
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()})
    locked = []

    def pinned(n_words):
        ptr = product.fimo_b200_host_alloc(4 * n_words)  # the CALLER's output buffer, made before the calls under test
        assert ptr
        locked.append(ptr)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint32)), shape=(n_words,))

    try:
        mask_words = (E + 31) // 32
        capacity = int(margin * W * E) if margin == 1.0 else None
        stream = torch.cuda.Stream()
        totals = []
        for i, frustum in enumerate(frusta):
            rig.upload("frustum", frustum)
            if capacity is None:  # a first plain step tells how much is visible; cut the list at half of that
                rig.step(1)
                capacity = int(margin * int(rig.download(("vis_count",))["vis_count"].sum()))
                rig.upload("frustum", frustum)
            words = rig.readback_words(capacity)
            a, b = pinned(words), pinned(words)
            a[...] = 0xDEADBEEF
            b[...] = 0xDEADBEEF
            fence = Fence(product)
            up, desc = rig.io_list({}), rig.readback_desc(a)
            rig.schedule_readback(up, desc, fence)
            fence.wait()
            want = rig.parse_readback(a)

            names = ("vis_count", "vis_mask", "vis_list")
            views = {n: rig._locked_view(n, exclusive=False) for n in names}  # shared locks: nobody rewrites them meanwhile
            try:
                product.check(product.fimo_b200_cull_pack_static(
                    views["vis_count"].device_ptr, views["vis_mask"].device_ptr, views["vis_list"].device_ptr, W, E, mask_words,
                    want["generation"], b.ctypes.data, b.nbytes, stream.cuda_stream))
                stream.synchronize()
            finally:
                for n in names:
                    rig.resources[n].unlock_in_world(rig.world, False)
            got = rig.parse_readback(b)

            defined = words - capacity + min(want["total"], capacity)
            assert want["capacity"] == capacity and want["truncated"] == (1 if want["total"] > capacity else 0)
            assert np.array_equal(b[:defined], a[:defined]), f"call {i}: packed words differ"
            assert got["total"] == want["total"] and np.array_equal(got["list"], want["list"])
            assert not b[defined:].any(), f"call {i}: words past the packed total are not zero"
            totals.append(want["total"])
        assert len(set(totals)) > 1  # the calls really packed different amounts
    finally:
        rig.close()
        for ptr in locked:
            product.fimo_b200_host_free(ptr)
