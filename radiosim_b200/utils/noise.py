"""Drop-in ``add_noise`` (reference ``src/radiosim/utils/utils.py:259-315``).

Host: the reference's draws on the legacy global ``np.random`` stream, in its order -- the white
field, then for each of the two filtered fields ``uniform(0, 1, n)`` and a normal field, then the
per-image noise levels -- so the stream ends where the reference's does.  The ``3 n N^2`` normals
are drawn natively (``rsb_legacy_normal_fill``) straight into page-locked float32 blocks.
Device: ``rsb_add_noise`` (two separable Gaussian filters, bundle-wide and per-image maxima, scaling).
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _lib as L

__all__ = ["add_noise"]


def _normal_field(count: int) -> np.ndarray:
    """``np.random.normal(size=...)`` (flattened) as page-locked float32, stream handed over and back."""
    out = L.pinned_empty((count,), np.float32)
    name, key, pos, has_gauss, cached = np.random.get_state()
    key = np.ascontiguousarray(key, dtype=np.uint32).copy()
    c_pos, c_has, c_cached = C.c_int32(int(pos)), C.c_int32(int(has_gauss)), C.c_double(float(cached))
    L.check(L.lib().rsb_legacy_normal_fill(key.ctypes.data, C.byref(c_pos), C.byref(c_has), C.byref(c_cached),
                                           count, out.ctypes.data), "rsb_legacy_normal_fill")
    np.random.set_state((name, key, c_pos.value, c_has.value, c_cached.value))
    return out


def add_noise(image, noise_level):
    """``image`` ``[n, 1, size, size]`` (any float dtype) -> noised bundle, float64, same shape.

    ``noise_level``: ``(low, high)`` of the maximum noise intensity in percent of the image peak."""
    L.require_device()
    lib = L.lib()
    image = np.asarray(image)
    shape = image.shape
    n, size = shape[0], shape[-1]
    if image.ndim != 4 or shape[1] != 1 or shape[2] != size:
        raise ValueError(f"add_noise expects a bundle [n, 1, size, size], got {shape}")
    if n == 0:
        return np.array(image, dtype=np.float64)
    count = n * size * size
    white = _normal_field(count)                                   # utils.py:291
    amp_a = np.random.uniform(0, 1, n).astype(np.float32)          # :276 (kernel size / 32)
    field_a = _normal_field(count)                                 # :277
    amp_b = np.random.uniform(0, 1, n).astype(np.float32)          # :276 (kernel size / 8)
    field_b = _normal_field(count)
    level = np.random.uniform(*noise_level, size=n) / 100          # :306-308
    need = C.c_size_t(0)
    L.check(lib.rsb_noise_workspace_bytes(n, size, C.byref(need)))
    img64 = np.ascontiguousarray(image, dtype=np.float64)
    d_small, (p_aa, p_ab, p_lv), _stage = L.upload_blob([amp_a, amp_b, level])
    d_img, d_w = L.DeviceBuffer(count * 8), L.DeviceBuffer(need.value)
    d_f = [L.DeviceBuffer(count * 4) for _ in range(3)]
    L.check(lib.rsb_memcpy_h2d(d_img.ptr, img64.ctypes.data, count * 8, None), "h2d")
    for d, f in zip(d_f, (white, field_a, field_b)):
        L.check(lib.rsb_memcpy_h2d(d.ptr, f.ctypes.data, count * 4, None), "h2d")
    L.check(lib.rsb_add_noise(d_img.ptr, d_f[0].ptr, d_f[1].ptr, d_f[2].ptr, p_aa, p_ab, p_lv, n, size,
                              d_img.ptr, d_w.ptr, d_w.nbytes, None), "rsb_add_noise")
    return d_img.download(shape, np.float64, pinned=True)
