"""Drop-in ``generate_proto_set`` / ``create_proto`` (reference
``src/radiosim/ppdisks/ppdisks.py:21-185``).

The host replays the legacy global ``np.random`` chain (``:67,74-76,135,137,305,360-361,
474,483``) and packs, per image, the ellipses, smoothing widths and amplitudes; the 900 x 900
tensor work (masks, circular Gaussian smoothing, crop, antialiased resize, paste) runs in
``librsb200.so`` (``rsb_ppdisk_render``).  Unlike the reference this path does not need torch;
``create_proto`` returns a CPU ``torch.Tensor`` when torch is importable (as the reference
does) and a numpy array otherwise.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .. import _lib as L

__all__ = ["create_proto", "generate_proto_set", "pack_proto", "render_protos"]

FIX = 900
_FITS = {  # ppdisks.py:467-472
    "average_1": (-0.08908080536281453, -4.808388735775005),
    "average_2": (-0.04053517757866287, -5.718461569072661),
    "average_3": (-0.023989028595164514, -6.22105762345096),
    "average_4": (-0.011198837787046627, -6.701733964123176),
}
COMP = np.dtype([("ca", "<f4"), ("sa", "<f4"), ("a2_in", "<f4"), ("b2_in", "<f4"),
                 ("a2_out", "<f4"), ("b2_out", "<f4"), ("scale", "<f4"), ("sigma", "<i4"),
                 ("image", "<i4"), ("pad", "<i4", (3,))])
IMAGE = np.dtype([("comp0", "<i4"), ("ncomp", "<i4"), ("lo", "<i4"), ("hi", "<i4"),
                  ("new_size", "<i4"), ("pad", "<i4", (3,))])
assert COMP.itemsize == 48 and IMAGE.itemsize == 32
_HTAB = None


def _smooth_factor(a):
    """``_get_smooth_factor`` ``:380-394``."""
    return 2 if a < 15 else 4 if a < 20 else 6 if a < 30 else 9 if a < 40 else 12


def _scale(d, m, n):
    """``_get_scale`` ``:477-483``: one ``normal`` draw."""
    e = np.exp(m * d + n)
    return e + np.random.normal(0, e * 0.05)


def pack_proto(img_size, alpha, ratio, size_ratio):
    """RNG chain of one ``create_proto`` call -> (component rows, image row)."""
    key = np.random.choice(list(_FITS.keys()))           # :474
    m, n = _FITS[key]
    ca = np.float32(np.cos(np.deg2rad(alpha)))
    sa = np.float32(np.sin(np.deg2rad(alpha)))
    comps = []
    a = np.float64(np.random.randint(3, 15))              # :305
    b = a / ratio
    d_core = (a + b) / 2
    sc = np.random.randint(5, 10)                         # :135
    comps.append((ca, sa, np.float32(-1), np.float32(1), np.float32(a**2), np.float32(b**2),
                  _scale(d_core, m, n), sc))              # :136
    num_circ = int(np.random.normal(3, 0.5) // 1)         # :137
    a_max = d_core
    for i in range(num_circ):                             # :143-149 -> _sim_ring :310-377
        ra = np.random.randint(10, 35) + (i + 1) * 30
        ex = np.random.randint(5, 10)
        d_ring = ((ra + ex + (ra + ex) / ratio) / 2 + (ra + ra / ratio) / 2) / 2
        s = _smooth_factor(ra)
        comps.append((ca, sa, np.float32(ra**2), np.float32((ra / ratio) ** 2),
                      np.float32((ra + ex) ** 2), np.float32(((ra + ex) / ratio) ** 2),
                      _scale(d_ring, m, n), s))
        extent = np.max([ra, ra + ex, ra / ratio, (ra + ex) / ratio]) + 6 * s
        a_max = np.max([extent, a_max])
    half = FIX // 2
    a_max = np.min([half, a_max])                         # :156
    lo, hi = int(half - a_max), int(half + a_max)         # :158-161
    new = int(size_ratio * img_size)                      # :166-167
    new = new if new % 2 == 0 else new - 1
    return comps, (lo, hi, new)


def _htab():
    global _HTAB
    if _HTAB is None:
        tab = np.zeros((17, FIX), dtype=np.float32)
        t = np.arange(FIX) - (FIX // 2 - 0.5)
        for s in range(1, 17):
            g = np.exp(-(t**2) / (2.0 * s * s))
            tab[s] = (g / g.sum()).astype(np.float32)
        _HTAB = tab
    return _HTAB


def _records(part):
    ncomp = sum(len(c) for c, _ in part)
    comps = np.zeros(ncomp, dtype=COMP)
    images = np.zeros(len(part), dtype=IMAGE)
    k = 0
    for j, (cl, (lo, hi, new)) in enumerate(part):
        images[j] = (k, len(cl), lo, hi, new, (0, 0, 0))
        for (ca, sa, a2i, b2i, a2o, b2o, scale, sigma) in cl:
            comps[k] = (ca, sa, a2i, b2i, a2o, b2o, np.float32(scale), int(sigma), j, (0, 0, 0))
            k += 1
    return comps, images


class DeviceProtos:
    """Packed protos resident on the device; :meth:`enqueue` issues ``rsb_ppdisk_render`` on the
    default stream without synchronising (device-side timing of the kernels alone)."""

    def __init__(self, img_size, packed):
        L.require_device()
        self.lib, self.img_size, self.n = L.lib(), img_size, len(packed)
        comps, images = _records(packed)
        self.ncomp = len(comps)
        need = C.c_size_t(0)
        L.check(self.lib.rsb_ppdisk_workspace_bytes(self.n, self.ncomp, C.byref(need)))
        self.d_h = L.dev_from(_htab())
        self.d_c, self.d_i = L.dev_from(comps), L.dev_from(images)
        self.d_w = L.DeviceBuffer(need.value)
        self.d_o = L.DeviceBuffer(self.n * img_size * img_size * 4)

    def enqueue(self):
        L.check(self.lib.rsb_ppdisk_render(self.d_c.ptr, self.ncomp, self.d_i.ptr, self.n, self.d_h.ptr,
                                           self.img_size, self.d_o.ptr, self.d_w.ptr, self.d_w.nbytes, None),
                "rsb_ppdisk_render")


def pack_protos_native(n, img_size, alpha_range, ratio_range, size_ratio_range):
    """The RNG chain of ``generate_proto_set`` for ``n`` images natively (``rsb_ppdisk_pack``): component
    and image records, with the global ``np.random`` stream handed over and back.  ``None`` when a range
    is not made of integers (numpy then truncates in its own way: the Python packer handles that)."""
    try:
        ints = [float(v).is_integer() for v in (*alpha_range, *ratio_range)]
    except (TypeError, ValueError):
        return None
    if not all(ints) or len(alpha_range) != 2 or len(ratio_range) != 2 or len(size_ratio_range) != 2:
        return None
    a_lo, a_hi, r_lo, r_hi = (int(v) for v in (*alpha_range, *ratio_range))
    if not (0 <= a_lo < a_hi <= 100000 and 0 < r_lo < r_hi) or os.environ.get("RSB_PY_PACKER") == "1":
        return None
    angles = np.deg2rad(np.arange(a_hi, dtype=np.float64))
    cos_tab = np.array([np.float32(np.cos(t)) for t in angles], dtype=np.float32)   # scalar calls, as pack_proto
    sin_tab = np.array([np.float32(np.sin(t)) for t in angles], dtype=np.float32)
    cap = 12 * max(n, 1) + 64
    comps = np.zeros(cap, dtype=COMP)
    images = np.zeros(max(n, 1), dtype=IMAGE)
    nc = C.c_int32(0)
    name, key, pos, has_gauss, cached = np.random.get_state()
    key = np.ascontiguousarray(key, dtype=np.uint32).copy()
    c_pos, c_has, c_cached = C.c_int32(int(pos)), C.c_int32(int(has_gauss)), C.c_double(float(cached))
    rc = L.lib().rsb_ppdisk_pack(key.ctypes.data, C.byref(c_pos), C.byref(c_has), C.byref(c_cached), n, img_size,
                                 a_lo, a_hi, r_lo, r_hi, float(size_ratio_range[0]), float(size_ratio_range[1]),
                                 cos_tab.ctypes.data, sin_tab.ctypes.data, comps.ctypes.data, cap,
                                 images.ctypes.data, C.byref(nc))
    if rc != 0:   # (more components than the buffer holds: left to the Python packer, stream untouched)
        return None
    np.random.set_state((name, key, c_pos.value, c_has.value, c_cached.value))
    return comps[: nc.value], images[:n]


def render_records(img_size, comps, images, chunk=64):
    """Packed records of many images -> float32 ``[n, img_size, img_size]``; chunks of ``chunk`` images,
    a chunk's copy to the host (second stream) under the next chunk's kernels."""
    L.require_device()
    lib = L.lib()
    n = len(images)
    out = L.pinned_empty((n, img_size, img_size), np.float32)
    d_h = L.dev_from(_htab())
    copier = L.CopyStream.get()
    keep = []
    for i0 in range(0, n, chunk):
        im = images[i0 : i0 + chunk].copy()
        c0, c1 = int(im["comp0"][0]), int(im["comp0"][-1] + im["ncomp"][-1])
        cm = comps[c0:c1].copy()
        cm["image"] -= i0
        im["comp0"] -= c0
        need = C.c_size_t(0)
        L.check(lib.rsb_ppdisk_workspace_bytes(len(im), len(cm), C.byref(need)))
        d_in, (p_c, p_i), stage = L.upload_blob([cm, im])
        d_w = L.DeviceBuffer(need.value)
        d_o = L.DeviceBuffer(len(im) * img_size * img_size * 4)
        L.check(lib.rsb_ppdisk_render(p_c, len(cm), p_i, len(im), d_h.ptr, img_size, d_o.ptr, d_w.ptr,
                                      d_w.nbytes, None), "rsb_ppdisk_render")
        copier.after_default()
        copier.d2h(out[i0 : i0 + len(im)], d_o.ptr)
        keep.append((d_in, stage, d_w, d_o))
        if len(keep) > 3:
            copier.sync()
            del keep[:-1]
    copier.sync()
    L.device_sync()
    return out


def render_protos(img_size, packed, chunk=32):
    """``packed``: list of ``pack_proto`` results -> float32 ``[n, img_size, img_size]``."""
    L.require_device()
    lib = L.lib()
    n = len(packed)
    out = L.pinned_empty((n, img_size, img_size), np.float32)
    d_h = L.dev_from(_htab())
    for i0 in range(0, n, chunk):
        part = packed[i0 : i0 + chunk]
        comps, images = _records(part)
        ncomp = len(comps)
        need = C.c_size_t(0)
        L.check(lib.rsb_ppdisk_workspace_bytes(len(part), ncomp, C.byref(need)))
        d_c, d_i = L.dev_from(comps), L.dev_from(images)
        d_w = L.DeviceBuffer(need.value)
        d_o = L.DeviceBuffer(len(part) * img_size * img_size * 4)
        L.check(lib.rsb_ppdisk_render(d_c.ptr, ncomp, d_i.ptr, len(part), d_h.ptr, img_size,
                                      d_o.ptr, d_w.ptr, d_w.nbytes, None), "rsb_ppdisk_render")
        dst = out[i0 : i0 + len(part)]
        L.check(lib.rsb_memcpy_d2h(dst.ctypes.data, d_o.ptr, dst.nbytes, None), "d2h")
        L.device_sync()
    return out


def ellipse_masks(comps: np.ndarray) -> np.ndarray:
    """Bit-exact ``_create_ellipse`` / ring XOR masks for packed component rows (uint8)."""
    L.require_device()
    d_c = L.dev_from(comps)
    d_m = L.DeviceBuffer(len(comps) * FIX * FIX)
    L.check(L.lib().rsb_ppdisk_masks(d_c.ptr, len(comps), d_m.ptr, None), "rsb_ppdisk_masks")
    return d_m.download((len(comps), FIX, FIX), np.uint8)


def create_proto(img_size, alpha, ratio, size_ratio, device="cuda", seed=None):
    """Drop-in for ``create_proto`` (``:95-185``); ``device`` is accepted and ignored (the
    work always runs on the current CUDA device)."""
    if seed is not None:
        np.random.seed(seed)
    img = render_protos(img_size, [pack_proto(img_size, alpha, ratio, size_ratio)])[0]
    try:
        import torch

        return torch.from_numpy(img)
    except ImportError:  # pragma: no cover
        return img


def generate_proto_set(img_size, size, alpha_range=(0, 180), ratio_range=(3, 15),
                       size_ratio_range=(0.1, 1), device="cuda", seed=1337, verbose=False):
    """Drop-in for ``generate_proto_set`` (``:21-92``): ``(list of float32 arrays, params)``."""
    np.random.seed(seed)                                             # :67
    params = dict(seed=seed, N=size, img_size=img_size, alpha_range=alpha_range,
                  ratio_range=ratio_range, size_ratio_range=size_ratio_range)
    native = pack_protos_native(size, img_size, alpha_range, ratio_range, size_ratio_range) if size else None
    if native is not None:
        imgs = render_records(img_size, *native)
        return [imgs[i] for i in range(size)], params
    packed = []
    for _ in range(size):
        alpha = np.random.randint(*alpha_range)                      # :74
        ratio = np.random.randint(*ratio_range) / np.random.randint(*ratio_range)  # :75
        size_ratio = np.random.uniform(*size_ratio_range)            # :76
        packed.append(pack_proto(img_size, alpha, ratio, size_ratio))
    imgs = render_protos(img_size, packed)
    return [imgs[i] for i in range(size)], params
