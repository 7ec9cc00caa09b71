"""Host-side component-parameter packer for jets.

Consumes the legacy global ``np.random`` stream in exactly the order of the
reference's ``create_jet`` (``src/radiosim/jets/jet.py:54-125,168``) so that every
image is comparable draw for draw, and emits the struct-of-arrays the CUDA
renderer takes: ``params[n, slots, 6]`` float64 (amp, x, y, sx, sy, rot) plus
``zoom_pad[n]`` int32.

The only image-dependent draw -- ``uniform(1/2, 1)`` taken iff the rendered
image's border maximum is below 0.01 (``jet.py:149-168``) -- needs the 4(N-1)
border pixels only; they are evaluated here in float64 with the very operation
order of ``twodgaussian`` (``src/radiosim/utils/gauss.py:77-88``) so the branch is
bit-identical to the reference's.  Nothing else of the image is computed on the
host.
"""

from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np

EDGE_THRESHOLD = 0.01
COLS = ("amp", "x", "y", "sx", "sy", "rot", "z_rot", "beta")


@dataclass
class JetBatch:
    size: int
    slots: int
    params: np.ndarray      # [n, slots, 6] float64
    extra: np.ndarray       # [n, slots, 2] float64: z_rotation, beta
    zoom_pad: np.ndarray    # [n] int32, -1 = no zoom
    zoom_out: np.ndarray    # [n] float64 zoom_out_factor (1 when not zoomed)

    @property
    def n(self) -> int:
        return self.params.shape[0]


def _border_coords(size: int):
    idx = np.arange(size)
    last = np.full(size, size - 1)
    zero = np.zeros(size, dtype=idx.dtype)
    rows = np.concatenate((zero, last, idx, idx))
    cols = np.concatenate((idx, idx, zero, last))
    return rows, cols


def border_max(p6: np.ndarray, size: int, coords=None) -> float:
    """Maximum of the summed components over the image border, float64."""
    rows, cols = coords if coords is not None else _border_coords(size)
    planes = np.zeros((p6.shape[0], rows.size))
    for k in range(p6.shape[0]):
        amp, cx, cy, wx, wy, rot = p6[k]
        if amp == 0:
            continue
        cs, sn = np.cos(rot), np.sin(rot)
        ra = cy * cs - cx * sn
        rb = cy * sn + cx * cs
        ua = rows * cs - cols * sn
        ub = rows * sn + cols * cs
        planes[k] = amp * np.exp(-(((ra - ua) / wy) ** 2 + ((rb - ub) / wx) ** 2) / 2.0)
    return np.sum(planes, axis=0).max()


def draw_from_mojave_dist():
    """``jets/flux_scaling.py:12-20``: one draw (legacy global stream, through scipy) of the beta fit to
    the MOJAVE archive's peak fluxes."""
    from scipy import stats

    return stats.beta(0.8639672251677816, 47.64189171625089, 0.09163404776954732, 1892.3881692069087).rvs(1)


def get_start_amp(scale_type):
    """``jets/flux_scaling.py:6-9`` (any other ``scale_type`` raises ``UnboundLocalError`` there too)."""
    if scale_type == "mojave":
        amp_start = draw_from_mojave_dist()
    return amp_start / 1e3


def _mojave_flux():
    return get_start_amp("mojave")


def pack_jets(n: int, size: int, num_jet_components=(3, 10), scaling="normalize") -> JetBatch:
    """Parameters of ``n`` jets from the global legacy ``np.random`` stream.

    The draws are made by the native packer (``rsb_jet_pack``: the stream's state is handed over
    and written back) unless ``scaling="mojave"`` needs scipy's beta draw, the component range
    exceeds the native limit, or ``RSB_PY_PACKER=1`` asks for :func:`pack_jets_py`.
    """
    lo, hi = int(num_jet_components[0]), int(num_jet_components[1])
    if scaling != "mojave" and 2 <= lo <= hi <= 32 and os.environ.get("RSB_PY_PACKER") != "1":
        return _pack_jets_native(n, size, lo, hi)
    return pack_jets_py(n, size, num_jet_components, scaling)


def _pack_jets_native(n: int, size: int, lo: int, hi: int) -> JetBatch:
    import ctypes as C

    from .. import _lib as L

    slots = 2 * hi - 1
    params = np.empty((n, slots, 6))
    extra = np.empty((n, slots, 2))
    zoom_pad = np.empty(n, dtype=np.int32)
    zoom_out = np.empty(n)
    name, key, pos, has_gauss, cached = np.random.get_state()
    assert name == "MT19937"
    key = np.ascontiguousarray(key, dtype=np.uint32).copy()
    c_pos, c_has, c_cached = C.c_int32(int(pos)), C.c_int32(int(has_gauss)), C.c_double(float(cached))
    counts = np.zeros(3, dtype=np.int64)
    L.check(
        L.lib().rsb_jet_pack(
            key.ctypes.data, C.byref(c_pos), C.byref(c_has), C.byref(c_cached), n, size, lo, hi,
            params.ctypes.data, extra.ctypes.data, zoom_pad.ctypes.data, zoom_out.ctypes.data,
            counts.ctypes.data,
        ),
        "rsb_jet_pack",
    )
    np.random.set_state((name, key, c_pos.value, c_has.value, c_cached.value))
    batch = JetBatch(size, slots, params, extra, zoom_pad, zoom_out)
    batch.border_counts = counts
    return batch


def pack_jets_py(n: int, size: int, num_jet_components=(3, 10), scaling="normalize") -> JetBatch:
    """Pure-numpy restatement of the same draws (the native packer's check, and the mojave path)."""
    lo, hi = int(num_jet_components[0]), int(num_jet_components[1])
    slots = 2 * hi - 1
    params = np.zeros((n, slots, 6))
    extra = np.zeros((n, slots, 2))
    zoom_pad = np.full(n, -1, dtype=np.int32)
    zoom_out = np.ones(n)
    coords = _border_coords(size)
    rnd = np.random
    mid = size // 2
    for b in range(n):
        k = rnd.randint(lo, hi + 1)
        col = np.zeros((7, hi))  # amp x y sx sy rot beta
        amp, x, y, sx, sy, rot, beta = col
        rot[0] = rnd.uniform(0, np.pi)
        beta[1] = rnd.uniform(0, 1)
        bend = rnd.uniform(0, np.pi)
        tilt = rnd.uniform(0, np.pi / 2)
        sin_tilt = np.sin(tilt)
        for i in range(k):
            amp[i] = np.exp(-np.sqrt(i) * rnd.normal(1.3, 0.2))
            if i >= 1 and rnd.rand() < 0.1:
                amp[i] = 0
            if i >= 2:
                beta[i] = beta[1] * np.exp(-np.sqrt(i - 1) * rnd.normal(0.5, 0.1))
            bend += rnd.normal(0, np.pi / 24)
            cb, sb = np.cos(bend), np.sin(bend)
            acb, asb = np.abs(cb), np.abs(sb)
            fill = 1 / asb if acb < asb else (1 / acb if acb > asb else np.sqrt(2))
            dist = i / (k - 1) * size / 2 * fill * sin_tilt * 0.7
            x[i] = dist * cb + mid
            y[i] = dist * sb + mid
            pair = size / k * fill * np.sqrt(i + 1) / rnd.uniform(3, 8, size=2)
            if pair[0] >= pair[1]:
                sx[i], sy[i] = pair[0], pair[1]
            else:
                sx[i], sy[i] = pair[1], pair[0]
            if i >= 1:
                rot[i] = rot[i - 1] + rnd.normal(0, np.pi / 18)
        lorentz = 1 / np.sqrt(1 - beta**2)
        mu = np.cos(tilt)
        toward = 1 / (lorentz * (1 - beta * mu))
        away = 1 / (lorentz * (1 + beta * mu))
        off_x = rnd.uniform(-size / 20, size / 20)
        off_y = rnd.uniform(-size / 20, size / 20)
        if scaling == "mojave":
            amp *= _mojave_flux()
        p = params[b]
        p[:hi, 0] = amp * toward
        p[:hi, 1] = x + off_x
        p[:hi, 2] = y + off_y
        p[:hi, 3] = sx
        p[:hi, 4] = sy
        p[:hi, 5] = rot
        extra[b, :, 0] = tilt
        extra[b, :hi, 1] = beta
        if rnd.rand() < 0.3:  # counter jet: mirror slots 1.. through the image centre
            p[hi:, 0] = amp[1:] * away[1:]
            p[hi:, 1] = size - x[1:] + off_x
            p[hi:, 2] = size - y[1:] + off_y
            p[hi:, 3] = sx[1:]
            p[hi:, 4] = sy[1:]
            p[hi:, 5] = rot[1:]
            extra[b, hi:, 1] = beta[1:]
        if border_max(p, size, coords) < EDGE_THRESHOLD:
            zo = rnd.uniform(1 / 2, 1)
            zoom_out[b] = zo
            zoom_pad[b] = np.int64((1 / zo - 1) * size / 2)
    return JetBatch(size, slots, params, extra, zoom_pad, zoom_out)


def source_lists(batch: JetBatch, margin: np.ndarray, peak: np.ndarray, normalize: bool):
    """The ``[n, slots, 8]`` list targets (``jet.py:162-174,181,186``) from the device's
    ``margin`` (idx of zoom_on_source) and ``peak`` (jet_max)."""
    n, size = batch.n, batch.size
    p = batch.params
    out = np.empty((n, batch.slots, 8))
    amp, x, y, sx, sy = (p[:, :, i].copy() for i in range(5))
    zoomed = batch.zoom_pad >= 0
    zf = np.where(zoomed, size / (size - 2 * margin.astype(np.float64)), 1.0)[:, None]
    zo = batch.zoom_out[:, None]
    z = zoomed[:, None]
    half = size / 2
    x1 = half + (x - half) * zf
    y1 = half + (y - half) * zf
    x2 = half + (x1 - half) * zo
    y2 = half + (y1 - half) * zo
    out[:, :, 1] = np.where(z, x2, x)
    out[:, :, 2] = np.where(z, y2, y)
    out[:, :, 3] = np.where(z, sx * zf * zo, sx)
    out[:, :, 4] = np.where(z, sy * zf * zo, sy)
    out[:, :, 0] = amp / peak[:, None] if normalize else amp
    out[:, :, 5] = p[:, :, 5]
    out[:, :, 6:8] = batch.extra
    return out
