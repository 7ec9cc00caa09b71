"""Drop-in ``create_survey`` (reference ``src/radiosim/survey/survey.py:9-132``).

Host: the RNG chain of the reference, draw for draw (:func:`pack_survey`), including the
nested ``create_jet`` chain of class-0 sources.  Device: sub-jets through ``rsb_jet_render``
(one launch per distinct sub-jet size), everything else through ``rsb_survey_render``.
"""

from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

from .. import _lib as L
from ..jets.packer import JetBatch, pack_jets, pack_jets_py

__all__ = ["create_survey", "pack_survey", "render_survey"]

JET_RECT = np.dtype([("pix", "<i8"), ("size", "<i4"), ("row0", "<i4"), ("col0", "<i4"), ("pad", "<i4")])


@dataclass
class SurveyBatch:
    size: int
    n: int
    src_params: np.ndarray          # [total, 6] float64, per image: class 1 rows then class 2 rows
    src_offs: np.ndarray            # [n, 3] int32
    jets: list = field(default_factory=list)        # JetBatch (one jet each), in draw order (numpy packer)
    jet_image: np.ndarray = None    # [njets] image index
    jet_row0: np.ndarray = None
    jet_col0: np.ndarray = None
    # the same sub-jets as flat arrays (native packer fills these directly)
    jet_params: np.ndarray = None   # [njets, slots, 6] float64
    jet_zoom_pad: np.ndarray = None  # [njets] int32
    jet_sizes: np.ndarray = None    # [njets] int64

    @property
    def njets(self) -> int:
        return len(self.jet_image)

    def flat_jets(self):
        if self.jet_params is None:
            slots = self.jets[0].slots if self.jets else 1
            self.jet_params = (np.concatenate([j.params for j in self.jets]) if self.jets
                               else np.zeros((0, slots, 6)))
            self.jet_zoom_pad = (np.concatenate([j.zoom_pad for j in self.jets]) if self.jets
                                 else np.zeros(0, np.int32))
            self.jet_sizes = np.array([j.size for j in self.jets], dtype=np.int64)
        return self.jet_params, self.jet_zoom_pad, self.jet_sizes


def pack_survey(n: int, size: int, num_sources: int, class_distribution, scale_sources: bool,
                num_jet_components=(3, 10), jet_scaling="normalize") -> SurveyBatch:
    """Consume the global ``np.random`` stream exactly like ``survey.py:49-107``: natively
    (``rsb_survey_pack``) unless the sub-jets need scipy's draw (``jet_scaling="mojave"``), the
    component range exceeds the native limit, or ``RSB_PY_PACKER=1`` asks for :func:`pack_survey_py`."""
    lo, hi = int(num_jet_components[0]), int(num_jet_components[1])
    if (jet_scaling != "mojave" and 2 <= lo <= hi <= 32 and len(class_distribution) == 3
            and os.environ.get("RSB_PY_PACKER") != "1"):
        return _pack_survey_native(n, size, int(num_sources), class_distribution, bool(scale_sources), lo, hi)
    return pack_survey_py(n, size, num_sources, class_distribution, scale_sources, num_jet_components, jet_scaling)


def _pack_survey_native(n, size, num_sources, class_distribution, scale_sources, lo, hi) -> SurveyBatch:
    slots = 2 * hi - 1
    cap = max(n * num_sources, 1)
    src = np.empty((cap, 6))
    offs = np.zeros((n, 3), dtype=np.int32)
    jp = np.empty((cap, slots, 6))
    jz, jsz, jimg, jr0, jc0 = (np.empty(cap, dtype=np.int32) for _ in range(5))
    nj = C.c_int32(0)
    dist = np.asarray(class_distribution, dtype=np.float64)
    name, key, pos, has_gauss, cached = np.random.get_state()
    key = np.ascontiguousarray(key, dtype=np.uint32).copy()
    c_pos, c_has, c_cached = C.c_int32(int(pos)), C.c_int32(int(has_gauss)), C.c_double(float(cached))
    L.check(
        L.lib().rsb_survey_pack(
            key.ctypes.data, C.byref(c_pos), C.byref(c_has), C.byref(c_cached), n, size, num_sources,
            dist.ctypes.data, int(scale_sources), lo, hi, src.ctypes.data, offs.ctypes.data, jp.ctypes.data,
            jz.ctypes.data, jsz.ctypes.data, jimg.ctypes.data, jr0.ctypes.data, jc0.ctypes.data, C.byref(nj)),
        "rsb_survey_pack",
    )
    np.random.set_state((name, key, c_pos.value, c_has.value, c_cached.value))
    k = nj.value
    total = int(offs[-1, 2]) if n else 0
    return SurveyBatch(size, n, src[:total], offs, [], jimg[:k].astype(np.int64), jr0[:k].astype(np.int64),
                       jc0[:k].astype(np.int64), jp[:k], jz[:k], jsz[:k].astype(np.int64))


def pack_survey_py(n: int, size: int, num_sources: int, class_distribution, scale_sources: bool,
                   num_jet_components=(3, 10), jet_scaling="normalize") -> SurveyBatch:
    """Pure-numpy restatement of the same draws (the native packer's check, and the mojave path)."""
    rnd = np.random
    total = sum(class_distribution)
    rows, offs = [], np.zeros((n, 3), dtype=np.int32)
    jets, jimg, jr0, jc0 = [], [], [], []
    count = 0
    for b in range(n):
        wide, narrow = [], []
        for _ in range(num_sources):
            u = rnd.uniform(0, total)
            if u < class_distribution[0]:
                if scale_sources:
                    js = rnd.randint(size / 10.24, size / 5.12)
                else:
                    js = rnd.randint(100, 200)
                pa, pb = rnd.rand(2) * size
                jets.append(pack_jets_py(1, js, num_jet_components, jet_scaling))
                jimg.append(b)
                jr0.append(int(np.floor(pa - js / 2)))
                jc0.append(int(np.floor(pb - js / 2)))
            else:
                is_wide = u < class_distribution[0] + class_distribution[1]
                if is_wide:
                    if scale_sources:
                        sx = size * rnd.uniform(1 / 40.96, 1 / 20.48)
                        sy = sx + rnd.normal(scale=size / 204.8)
                    else:
                        sx = rnd.uniform(25, 50)
                        sy = sx + rnd.normal(scale=5)
                else:
                    if scale_sources:
                        sx = size * rnd.uniform(1 / 204.8, 1 / 102.4)
                        sy = sx + rnd.normal(scale=size / 1024)
                    else:
                        sx = rnd.uniform(5, 10)
                        sy = sx + rnd.normal(scale=1)
                px, py = rnd.rand(2) * size
                amp = rnd.rand()
                rot = rnd.uniform(0, np.pi)
                (wide if is_wide else narrow).append((amp, px, py, sx, sy, rot))
        offs[b] = (count, count + len(wide), count + len(wide) + len(narrow))
        rows.extend(wide)
        rows.extend(narrow)
        count += len(wide) + len(narrow)
    params = np.array(rows, dtype=np.float64).reshape(-1, 6)
    return SurveyBatch(size, n, params, offs, jets, np.array(jimg, dtype=np.int64),
                       np.array(jr0, dtype=np.int64), np.array(jc0, dtype=np.int64))


def _render_subjets(batch: SurveyBatch, normalize: bool):
    """All sub-jets, grouped by size, into one packed float32 device buffer.

    One upload of the parameters (sorted by size), one launch per distinct size on slices of it."""
    lib = L.lib()
    all_params, all_zpad, sizes = batch.flat_jets()
    nj = batch.njets
    order = np.argsort(sizes, kind="stable")
    pix_off = np.zeros(nj, dtype=np.int64)
    ssz = sizes[order]
    starts = np.flatnonzero(np.r_[True, ssz[1:] != ssz[:-1]]) if nj else np.zeros(0, np.int64)
    ends = np.r_[starts[1:], nj] if nj else np.zeros(0, np.int64)
    total_pix = int((sizes.astype(np.int64) ** 2).sum()) + 4 * (len(starts) + 1)
    d_pix = L.DeviceBuffer(max(total_pix * 4, 16))
    if nj == 0:
        return d_pix, pix_off, sizes
    slots = all_params.shape[1]
    d_p = L.dev_from(all_params[order])
    d_z = L.dev_from(all_zpad[order])
    d_m, d_k = L.DeviceBuffer(nj * 4), L.DeviceBuffer(nj * 8)
    need_max = 0
    need = C.c_size_t(0)
    for a, b in zip(starts, ends):
        L.check(lib.rsb_jet_workspace_bytes(int(b - a), slots, int(ssz[a]), C.byref(need)))
        need_max = max(need_max, need.value)
    d_w = L.DeviceBuffer(max(need_max, 16))
    flags = L.JET_NORMALIZE if normalize else 0
    cursor = 0
    for a, b in zip(starts, ends):
        a, b = int(a), int(b)
        js, cnt = int(ssz[a]), b - a
        pix_off[order[a:b]] = cursor + np.arange(cnt, dtype=np.int64) * (js * js)
        L.check(lib.rsb_jet_render(d_p.ptr + a * slots * 48, d_z.ptr + a * 4, cnt, slots, js, flags,
                                   d_pix.ptr + cursor * 4, d_m.ptr + a * 4, d_k.ptr + a * 8, d_w.ptr,
                                   d_w.nbytes, None),
                "rsb_jet_render (survey sub-jets)")
        cursor += cnt * js * js
        cursor = (cursor + 3) & ~3  # keep every group 16-byte aligned (128-bit stores)
    L.device_sync()
    return d_pix, pix_off, sizes


def render_survey(batch: SurveyBatch, jet_normalize: bool = True, f64: bool = True):
    L.require_device()
    lib = L.lib()
    n, size = batch.n, batch.size
    d_pix, pix_off, sizes = _render_subjets(batch, jet_normalize)
    nj = batch.njets
    rects = np.zeros(max(nj, 1), dtype=JET_RECT)
    jet_offs = np.zeros(n + 1, dtype=np.int32)
    if nj:
        by_img = np.argsort(batch.jet_image, kind="stable")
        rects["pix"][:nj] = pix_off[by_img]
        rects["size"][:nj] = sizes[by_img]
        rects["row0"][:nj] = batch.jet_row0[by_img]
        rects["col0"][:nj] = batch.jet_col0[by_img]
        jet_offs[1:] = np.cumsum(np.bincount(batch.jet_image, minlength=n))
    total = batch.src_params.shape[0]
    need = C.c_size_t(0)
    L.check(lib.rsb_survey_workspace_bytes(n, size, total, C.byref(need)))
    dt = np.float64 if f64 else np.float32
    d_src = L.dev_from(batch.src_params if total else np.zeros((1, 6)))
    d_offs = L.dev_from(batch.src_offs)
    d_rects = L.dev_from(rects)
    d_joffs = L.dev_from(jet_offs)
    d_w = L.DeviceBuffer(need.value)
    d_x = L.DeviceBuffer(n * size * size * np.dtype(dt).itemsize)
    d_y = L.DeviceBuffer(n * 4 * size * size * np.dtype(dt).itemsize)
    L.check(lib.rsb_survey_render(d_src.ptr, d_offs.ptr, total, d_rects.ptr, d_joffs.ptr, d_pix.ptr,
                                  n, size, L.JET_OUT_F64 if f64 else 0, d_x.ptr, d_y.ptr, d_w.ptr,
                                  d_w.nbytes, None), "rsb_survey_render")
    x = d_x.download((n, 1, size, size), dt, pinned=True)
    y = d_y.download((n, 4, size, size), dt, pinned=True)
    return x, y


def create_survey(grid, conf):
    """Drop-in for ``radiosim.survey.create_survey``: ``(survey[n,1,N,N], comps[n,4,N,N])`` float64."""
    shape = tuple(grid.shape)
    n = 1 if len(shape) == 3 else shape[0]
    size = shape[-1]
    batch = pack_survey(n, size, conf.survey.num_sources, conf.survey.class_distribution,
                        conf.survey.scale_sources, conf.jet.num_jet_components, conf.jet.scaling)
    return render_survey(batch, conf.jet.scaling == "normalize", f64=True)
