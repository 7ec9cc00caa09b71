"""Drop-ins for ``src/radiosim/utils/gauss.py``: ``twodgaussian`` and ``skewed_gauss`` rendered
on the device (float32 arithmetic, float64 arrays returned like the reference)."""

from __future__ import annotations

import numpy as np

from .. import _lib as L

__all__ = ["gaussian_planes", "skewed_gauss", "twodgaussian"]


def gaussian_planes(params, size: int) -> np.ndarray:
    """One ``twodgaussian`` plane per row ``[amp, x, y, sx, sy, rot]`` -> ``[k, size, size]``
    float64 (zeros for ``amp == 0``).  Device: ``rsb_jet_planes`` without zoom / scaling."""
    L.require_device()
    p = np.ascontiguousarray(np.asarray(params, dtype=np.float64).reshape(-1, 6))
    k = p.shape[0]
    out = np.empty((k, size, size))
    for i0 in range(0, k, 64):
        part = p[i0 : i0 + 64]
        d_p = L.dev_from(part[None])
        d_z = L.dev_from(np.array([-1], np.int32))
        d_m = L.dev_from(np.zeros(1, np.int32))
        d_k = L.dev_from(np.ones(1))
        d_o = L.DeviceBuffer(len(part) * size * size * 8)
        L.check(L.lib().rsb_jet_planes(d_p.ptr, d_z.ptr, d_m.ptr, d_k.ptr, 1, len(part), size,
                                       L.JET_OUT_F64, d_o.ptr, None), "rsb_jet_planes")
        out[i0 : i0 + len(part)] = d_o.download((len(part), size, size), np.float64)
    return out


def twodgaussian(params, size: int) -> np.ndarray:
    """``src/radiosim/utils/gauss.py:50-97``."""
    return gaussian_planes([list(params)], size)[0]


def skewed_gauss(size, x, y, amp, width, length, a) -> np.ndarray:
    """``src/radiosim/utils/gauss.py:100-150``."""
    from ..mojave.mojave import _Blob, _Jet, _Source, render_sources

    j = _Jet(size)
    j.blobs.append(_Blob(size, x, y, amp, width, length, a))
    j.strength, j.radius = 0.0, 100.0   # identity swirl
    s = _Source(size)
    s.jets = [j]
    return render_sources([s], size)[0]
