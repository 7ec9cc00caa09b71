"""Drop-in ``create_survey`` (reference ``src/radiosim/survey/survey.py:9-132``).

Host: the RNG chain of the reference, draw for draw (:func:`pack_survey`), including the
nested ``create_jet`` chain of class-0 sources.  Device: sub-jets through ``rsb_jet_render``
(one launch per distinct sub-jet size), everything else through ``rsb_survey_render``.
"""

from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from .. import _lib as L
from ..jets.packer import JetBatch, pack_jets

__all__ = ["create_survey", "pack_survey", "render_survey"]

JET_RECT = np.dtype([("pix", "<i8"), ("size", "<i4"), ("row0", "<i4"), ("col0", "<i4"), ("pad", "<i4")])


@dataclass
class SurveyBatch:
    size: int
    n: int
    src_params: np.ndarray          # [total, 6] float64, per image: class 1 rows then class 2 rows
    src_offs: np.ndarray            # [n, 3] int32
    jets: list = field(default_factory=list)        # JetBatch (one jet each), in draw order
    jet_image: np.ndarray = None    # [njets] image index
    jet_row0: np.ndarray = None
    jet_col0: np.ndarray = None


def pack_survey(n: int, size: int, num_sources: int, class_distribution, scale_sources: bool,
                num_jet_components=(3, 10), jet_scaling="normalize") -> SurveyBatch:
    """Consume the global ``np.random`` stream exactly like ``survey.py:49-107``."""
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
                jets.append(pack_jets(1, js, num_jet_components, jet_scaling))
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
    """All sub-jets, grouped by size, into one packed float32 device buffer."""
    lib = L.lib()
    nj = len(batch.jets)
    sizes = np.array([j.size for j in batch.jets], dtype=np.int64)
    order = np.argsort(sizes, kind="stable")
    pix_off = np.zeros(nj, dtype=np.int64)
    total_pix = int((sizes.astype(np.int64) ** 2).sum()) + 4 * (len(np.unique(sizes)) + 1)
    d_pix = L.DeviceBuffer(max(total_pix * 4, 16))
    cursor = 0
    flags = L.JET_NORMALIZE if normalize else 0
    keep = []
    i = 0
    while i < nj:
        js = int(sizes[order[i]])
        j = i
        while j < nj and sizes[order[j]] == js:
            j += 1
        group = order[i:j]
        params = np.concatenate([batch.jets[g].params for g in group])
        zpad = np.concatenate([batch.jets[g].zoom_pad for g in group])
        cnt = len(group)
        for q, g in enumerate(group):
            pix_off[g] = cursor + q * js * js
        need = C.c_size_t(0)
        L.check(lib.rsb_jet_workspace_bytes(cnt, params.shape[1], js, C.byref(need)))
        d_p, d_z = L.dev_from(params), L.dev_from(zpad)
        d_m, d_k = L.DeviceBuffer(cnt * 4), L.DeviceBuffer(cnt * 8)
        d_w = L.DeviceBuffer(max(need.value, 16))
        L.check(lib.rsb_jet_render(d_p.ptr, d_z.ptr, cnt, params.shape[1], js, flags,
                                   d_pix.ptr + cursor * 4, d_m.ptr, d_k.ptr, d_w.ptr, d_w.nbytes, None),
                "rsb_jet_render (survey sub-jets)")
        keep.append((d_p, d_z, d_m, d_k, d_w))
        cursor += cnt * js * js
        cursor = (cursor + 3) & ~3  # keep every group 16-byte aligned (128-bit stores)
        i = j
    L.device_sync()
    return d_pix, pix_off, sizes


def render_survey(batch: SurveyBatch, jet_normalize: bool = True, f64: bool = True):
    L.require_device()
    lib = L.lib()
    n, size = batch.n, batch.size
    d_pix, pix_off, sizes = _render_subjets(batch, jet_normalize)
    nj = len(batch.jets)
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
    x = d_x.download((n, 1, size, size), dt)
    y = d_y.download((n, 4, size, size), dt)
    return x, y


def create_survey(grid, conf):
    """Drop-in for ``radiosim.survey.create_survey``: ``(survey[n,1,N,N], comps[n,4,N,N])`` float64."""
    shape = tuple(grid.shape)
    n = 1 if len(shape) == 3 else shape[0]
    size = shape[-1]
    batch = pack_survey(n, size, conf.survey.num_sources, conf.survey.class_distribution,
                        conf.survey.scale_sources, conf.jet.num_jet_components, conf.jet.scaling)
    return render_survey(batch, conf.jet.scaling == "normalize", f64=True)
