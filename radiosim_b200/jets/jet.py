"""Drop-in ``create_jet`` / ``component_from_list`` / ``apply_train_type``.

Same signatures, return shapes and dtypes as the reference's
``src/radiosim/jets/jet.py:14-261``; the pixel work runs in ``librsb200.so``
(``rsb_jet_render`` / ``rsb_jet_planes``), the RNG chain in :mod:`.packer`.
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _lib as L
from .packer import JetBatch, pack_jets, source_lists

__all__ = ["apply_train_type", "component_from_list", "create_jet", "render_batch"]


def render_batch(batch: JetBatch, normalize: bool, planes: bool = False, f64: bool = True, planes_out=None,
                 out=None):
    """Render a packed batch on the current device through the host-buffer C ABI.

    Returns ``(images[n, N, N], margin[n] int32, peak[n] float64, planes or None)``.
    ``planes_out``: a host array ``[n, P >= slots, N, N]`` of the output dtype (best page-locked)
    whose first ``slots`` planes per image receive the component planes directly.
    ``out``: ``(images, margin, peak)`` host arrays (contiguous, of the right dtypes) to fill
    instead of fresh ones -- slices of a larger page-locked block when a call is pipelined.
    """
    L.require_device()
    n, size, slots = batch.n, batch.size, batch.slots
    flags = (L.JET_NORMALIZE if normalize else 0) | (L.JET_OUT_F64 if f64 else 0)
    dt = np.float64 if f64 else np.float32
    if out is None:
        img = L.pinned_empty((n, size, size), dt)
        margin = L.pinned_empty((n,), np.int32)
        peak = L.pinned_empty((n,), np.float64)
    else:
        img, margin, peak = out
        assert img.shape == (n, size, size) and img.dtype == dt and img.flags.c_contiguous
        assert margin.dtype == np.int32 and peak.dtype == np.float64 and len(margin) == len(peak) == n
    p = np.ascontiguousarray(batch.params, dtype=np.float64)
    zp = np.ascontiguousarray(batch.zoom_pad, dtype=np.int32)
    if n:
        L.check(
            L.lib().rsb_jet_render_host(
                p.ctypes.data, zp.ctypes.data, n, slots, size, flags,
                img.ctypes.data, margin.ctypes.data, peak.ctypes.data,
            ),
            "rsb_jet_render_host",
        )
    pl = None
    if planes:
        pl = planes_out if planes_out is not None else L.pinned_empty((n, slots, size, size), dt)
        assert pl.dtype == dt and pl.shape[0] == n and pl.shape[1] >= slots and pl.flags.c_contiguous
        img_bytes = slots * size * size * pl.itemsize        # one image's planes on the device
        host_pitch = pl.shape[1] * size * size * pl.itemsize  # ... and where the next image starts on the host
        step = max(1, (256 << 20) // img_bytes)
        for i0 in range(0, n, step):
            i1 = min(n, i0 + step)
            d_p = L.dev_from(p[i0:i1])
            d_z = L.dev_from(zp[i0:i1])
            d_m = L.dev_from(np.asarray(margin[i0:i1]))
            d_k = L.dev_from(np.asarray(peak[i0:i1]))
            d_o = L.DeviceBuffer((i1 - i0) * img_bytes)
            L.check(
                L.lib().rsb_jet_planes(
                    d_p.ptr, d_z.ptr, d_m.ptr, d_k.ptr, i1 - i0, slots, size, flags, d_o.ptr, None
                ),
                "rsb_jet_planes",
            )
            L.check(L.lib().rsb_memcpy2d_d2h(pl[i0:i1].ctypes.data, host_pitch, d_o.ptr, img_bytes, img_bytes,
                                             i1 - i0, None), "rsb_memcpy2d_d2h")
            L.device_sync()
    return img, np.asarray(margin), np.asarray(peak), pl


def apply_train_type(train_type, jet_img, jet_comp, source_list):
    """``src/radiosim/jets/jet.py:200-231``."""
    if train_type == "list":
        return source_list
    if train_type == "gauss":
        size = jet_comp.shape[-1]
        tail = np.full((1, size, size), np.nan)
        tail[0, : source_list.shape[0], : source_list.shape[1]] = source_list
        return np.concatenate((jet_comp, tail))
    if train_type == "clean":
        return jet_img[None]
    raise UnboundLocalError(f"unknown training type {train_type!r}")


_PIPE_CHUNK = 512


def _pack_and_render_pipelined(n, size, comps, scaling, normalize):
    """Large batches: the RNG chain of chunk i+1 is drawn on a helper thread while chunk i is
    rendered and copied back (both release the GIL).  The chunks are packed strictly one after the
    other, so the global ``np.random`` stream is consumed exactly as by one ``pack_jets(n, ...)``."""
    from concurrent.futures import ThreadPoolExecutor

    img = L.pinned_empty((n, size, size), np.float64)
    margin = L.pinned_empty((n,), np.int32)
    peak = L.pinned_empty((n,), np.float64)
    parts = []
    with ThreadPoolExecutor(1) as ex:
        fut = ex.submit(pack_jets, min(_PIPE_CHUNK, n), size, comps, scaling)
        i0 = 0
        while i0 < n:
            b = fut.result()
            i1 = i0 + b.n
            if i1 < n:
                fut = ex.submit(pack_jets, min(_PIPE_CHUNK, n - i1), size, comps, scaling)
            render_batch(b, normalize, out=(img[i0:i1], margin[i0:i1], peak[i0:i1]))
            parts.append(b)
            i0 = i1
    batch = JetBatch(size, parts[0].slots, np.concatenate([b.params for b in parts]),
                     np.concatenate([b.extra for b in parts]), np.concatenate([b.zoom_pad for b in parts]),
                     np.concatenate([b.zoom_out for b in parts]))
    return batch, img, np.asarray(margin), np.asarray(peak)


def create_jet(grid, conf):
    """Drop-in for ``radiosim.jets.create_jet`` (``jets/jet.py:14-197``).

    ``grid`` is only inspected for its shape (as in the reference, ``:46-53``);
    ``conf`` needs ``conf.jet.num_jet_components / scaling / training_type``.
    """
    shape = tuple(grid.shape)
    n = 1 if len(shape) == 3 else shape[0]
    size = shape[-1]
    scaling = conf.jet.scaling
    ttype = conf.jet.training_type
    normalize = scaling == "normalize"
    gauss_targets = None
    if ttype != "gauss" and n >= 2 * _PIPE_CHUNK:
        batch, img, margin, peak = _pack_and_render_pipelined(n, size, conf.jet.num_jet_components, scaling, normalize)
    else:
        batch = pack_jets(n, size, conf.jet.num_jet_components, scaling)
        if ttype == "gauss":  # planes, background and list plane side by side: the planes land there directly
            gauss_targets = L.pinned_empty((n, batch.slots + 2, size, size), np.float64)
        img, margin, peak, _ = render_batch(batch, normalize, planes=(ttype == "gauss"), planes_out=gauss_targets)
    slist = source_lists(batch, margin, peak, normalize)
    jets = img[:, None, :, :]  # (page-locked block owned by the array; recycled when it is released)
    if ttype == "list":
        targets = slist
    elif ttype == "clean":
        targets = L.pinned_empty(jets.shape, jets.dtype)
        L.parallel_copy(targets, jets)
    elif ttype == "gauss":
        targets = gauss_targets
        np.subtract(1, jets[:, 0], out=targets[:, batch.slots])
        targets[:, batch.slots + 1] = np.nan
        targets[:, batch.slots + 1, : batch.slots, :8] = slist
    else:
        raise UnboundLocalError(f"unknown training type {ttype!r}")
    return jets, targets


def component_from_list(size, amp, x, y, sx, sy, rotation):
    """``src/radiosim/jets/jet.py:234-261``: one float64 plane per slot (zeros for
    ``amp == 0``), rendered on the device."""
    from ..utils.gauss import gaussian_planes

    p = np.stack([np.asarray(v, dtype=np.float64) for v in (amp, x, y, sx, sy, rotation)], axis=-1)
    return list(gaussian_planes(p, size))
