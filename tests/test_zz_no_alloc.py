"""rsb_jet_render_static: the jet renderer for callers that may not allocate.

Some hosts of this library cannot call cudaMalloc once they are running -- a plugin
inside a process whose framework has already claimed the device for its own pool, or
a service whose device-memory budget is fixed at start-up.  For them the scratch that
``rsb_jet_render`` takes from the caller (needed when the image is too large for
shared memory, e.g. 384x384) is replaced by a 32 MiB array built into the library.

The test drives both entry points through the C ABI on the same inputs, twice in a
row with different batches (the shared array keeps the first call's contents, which
must not leak into the second), and asks for the same images, margins and peaks.  The
static entry is also held against the float64 oracle directly (images within 1e-5 of the
peak, margins exact), so the check does not rest on the other CUDA entry point alone.
It is collected last on purpose: a run that stops at the first failure has already
run every other test.
"""

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIZE = 384   # too large for shared memory -> both entry points use global scratch
SLOTS = 19


def _render(L, fn_name, params, zoom_pad, use_workspace):
    lib = L.lib()
    n = params.shape[0]
    d_p, d_z = L.dev_from(params), L.dev_from(zoom_pad)
    d_img = L.DeviceBuffer(n * SIZE * SIZE * 4)
    d_m, d_k = L.DeviceBuffer(n * 4), L.DeviceBuffer(n * 8)
    L.check(lib.rsb_memset(d_img.ptr, 0xFF, d_img.nbytes, None))
    if use_workspace:
        need = C.c_size_t(0)
        L.check(lib.rsb_jet_workspace_bytes(n, SLOTS, SIZE, C.byref(need)))
        assert need.value > 0, "expected a scratch requirement at this size"
        d_w = L.DeviceBuffer(need.value)
        rc = lib.rsb_jet_render(d_p.ptr, d_z.ptr, n, SLOTS, SIZE, L.JET_NORMALIZE, d_img.ptr,
                                d_m.ptr, d_k.ptr, d_w.ptr, d_w.nbytes, None)
    else:
        rc = lib.rsb_jet_render_static(d_p.ptr, d_z.ptr, n, SLOTS, SIZE, L.JET_NORMALIZE,
                                       d_img.ptr, d_m.ptr, d_k.ptr, None)
    L.check(rc, fn_name)
    L.device_sync()
    return (d_img.download((n, SIZE, SIZE), np.float32), d_m.download((n,), np.int32),
            d_k.download((n,), np.float64))


def test_static_scratch_entry_matches_workspace_entry():
    from radiosim_b200 import _lib as L
    from radiosim_b200.jets.packer import pack_jets

    L.require_device()
    assert hasattr(L.lib(), "rsb_jet_render_static")
    # more images than the 56 slots the built-in array offers at 384x384
    for call, (seed, n) in enumerate([(101, 150), (202, 90)]):
        np.random.seed(seed)
        b = pack_jets(n, SIZE, (3, 10), "normalize")
        assert (b.zoom_pad >= 0).any() and (b.zoom_pad < 0).any()
        ref_img, ref_m, ref_k = _render(L, "rsb_jet_render", b.params, b.zoom_pad, True)
        img, m, k = _render(L, "rsb_jet_render_static", b.params, b.zoom_pad, False)
        assert np.isfinite(img).all(), f"call {call}"
        assert np.array_equal(m, ref_m), f"call {call}"                 # integers: exact
        assert np.allclose(k, ref_k, rtol=1e-6, atol=0), f"call {call}"
        # the tolerance of the other jet tests: 1e-5 of the (normalised) image peak
        assert np.abs(img - ref_img).max() <= 1e-5, f"call {call}"


def test_static_scratch_entry_matches_oracle():
    """384x384 jets through the no-allocation entry against ``oracle.jet`` (float64)."""
    from oracle import jet as ojet
    from radiosim_b200 import _lib as L
    from radiosim_b200.jets.packer import pack_jets

    L.require_device()
    n = 12
    np.random.seed(303)
    trace = []
    ox, _ = ojet.create_jet((n, 3, SIZE, SIZE), (3, 10), "normalize", "clean", trace=trace)
    np.random.seed(303)
    b = pack_jets(n, SIZE, (3, 10), "normalize")
    for call in range(2):
        img, m, k = _render(L, "rsb_jet_render_static", b.params, b.zoom_pad, False)
        assert np.array_equal(m, np.array([t["margin"] for t in trace])), f"call {call}"
        assert np.allclose(k, np.array([t["jet_max"] for t in trace]), rtol=1e-6, atol=0), f"call {call}"
        assert np.abs(img.astype(np.float64) - ox[:, 0]).max() <= 1e-5, f"call {call}"


def test_static_scratch_entry_rejects_oversize():
    from radiosim_b200 import _lib as L

    L.require_device()
    rc = L.lib().rsb_jet_render_static(None, None, 1, SLOTS, 600, 0, None, None, None, None)
    assert rc > 0
    assert b"512" in L.lib().rsb_last_error()
