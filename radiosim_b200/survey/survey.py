"""Drop-in ``create_survey`` (reference ``src/radiosim/survey/survey.py:9-132``).

Host: the RNG chain of the reference, draw for draw (:func:`pack_survey`), including the
nested ``create_jet`` chain of class-0 sources.  Device: all sub-jets of the batch through
``rsb_jet_render_var`` (every sub-jet has its own size; at most two launches), everything else
through ``rsb_survey_render`` (one fused kernel: splat, paste, per-image maximum, normalise).
"""

from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

from .. import _lib as L
from ..jets.packer import JetBatch, pack_jets, pack_jets_py

__all__ = ["DeviceSurvey", "create_survey", "pack_survey", "render_survey"]

JET_RECT = np.dtype([("pix", "<i8"), ("size", "<i4"), ("row0", "<i4"), ("col0", "<i4"), ("pitch", "<i4")])


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


class DeviceSurvey:
    """A packed batch resident on the device: everything is uploaded in one asynchronous copy,
    :meth:`enqueue` issues the kernels (sub-jets in at most two variable-size launches, then
    ``rsb_survey_render``) on the default stream without synchronising, :meth:`download` copies the
    two results to host arrays."""

    def __init__(self, batch: SurveyBatch, jet_normalize: bool = True, f64: bool = True):
        L.require_device()
        self.lib = L.lib()
        self.n, self.size, self.f64 = batch.n, batch.size, f64
        n, size = batch.n, batch.size
        all_params, all_zpad, sizes = batch.flat_jets()
        nj = self.nj = batch.njets
        self.flags_jet = L.JET_NORMALIZE if jet_normalize else 0
        # sub-jets sorted by size (CTAs of one wave equally busy); sizes <= 128 take one CTA per
        # image, larger ones a 2-CTA cluster -> two launches at most
        # ... and anything above 256 pixels (surveys larger than 1310^2) one launch per distinct size
        # through rsb_jet_render, rows packed without padding
        order = np.argsort(sizes, kind="stable")
        ssz = sizes[order].astype(np.int64)
        slots = all_params.shape[1] if nj else 1
        var_ok = slots <= 32
        n_var = int(np.searchsorted(ssz, 256, side="right")) if var_ok else 0
        pitch = np.where(np.arange(nj) < n_var, (ssz + 3) & ~3, ssz)
        pix_sorted = np.zeros(nj, dtype=np.int64)
        self.groups, self.big_groups = [], []
        cursor = 0
        if n_var:
            pix_sorted[1:n_var] = np.cumsum((ssz * pitch)[:n_var])[:-1]   # (multiples of 4: every pitch is)
            cursor = int((ssz * pitch)[:n_var].sum())
            split = int(np.searchsorted(ssz[:n_var], 128, side="right"))
            for a, b in ((0, split), (split, n_var)):
                if b > a:
                    self.groups.append((a, b, int(ssz[b - 1])))
        need_ws = 0
        if nj > n_var:
            big = ssz[n_var:]
            starts = n_var + np.flatnonzero(np.r_[True, big[1:] != big[:-1]])
            need = C.c_size_t(0)
            for a, b in zip(starts, np.r_[starts[1:], nj]):
                a, b, js = int(a), int(b), int(ssz[a])
                pix_sorted[a:b] = cursor + np.arange(b - a, dtype=np.int64) * (js * js)
                self.big_groups.append((a, b, js, cursor))
                cursor = (cursor + (b - a) * js * js + 3) & ~3      # every group 16-byte aligned
                L.check(self.lib.rsb_jet_workspace_bytes(b - a, slots, js, C.byref(need)))
                need_ws = max(need_ws, need.value)
        total_pix = cursor
        self.d_jw = L.DeviceBuffer(max(need_ws, 16))
        self.slots = slots
        rects = np.zeros(max(nj, 1), dtype=JET_RECT)
        jet_offs = np.zeros(n + 1, dtype=np.int32)
        if nj:
            inv = np.empty(nj, dtype=np.int64)
            inv[order] = np.arange(nj)                        # draw order -> sorted position
            by_img = np.argsort(batch.jet_image, kind="stable")
            pos = inv[by_img]
            rects["pix"][:nj] = pix_sorted[pos]
            rects["size"][:nj] = ssz[pos]
            rects["pitch"][:nj] = pitch[pos]
            rects["row0"][:nj] = batch.jet_row0[by_img]
            rects["col0"][:nj] = batch.jet_col0[by_img]
            jet_offs[1:] = np.cumsum(np.bincount(batch.jet_image, minlength=n))
        self.total = batch.src_params.shape[0]
        need = C.c_size_t(0)
        L.check(self.lib.rsb_survey_workspace_bytes(n, size, self.total, C.byref(need)))
        self.dt = np.float64 if f64 else np.float32
        none = np.zeros(1)
        self.d_in, ptrs, self._stage = L.upload_blob([
            all_params[order] if nj else none, all_zpad[order] if nj else none,
            ssz.astype(np.int32) if nj else none, pix_sorted if nj else none, rects, jet_offs,
            batch.src_params if self.total else np.zeros((1, 6)),
            batch.src_offs if n else np.zeros((1, 3), np.int32)])
        (self.p_params, self.p_zpad, self.p_sizes, self.p_offs, self.p_rects, self.p_joffs,
         self.p_src, self.p_soffs) = ptrs
        self.d_pix = L.DeviceBuffer(max(total_pix * 4, 16))
        self.d_mk = L.DeviceBuffer(max(nj, 1) * 12 + 256)     # margins and peaks of the sub-jets (unused here)
        self.d_w = L.DeviceBuffer(max(need.value, 16))
        esz = np.dtype(self.dt).itemsize
        self.d_x = L.DeviceBuffer(max(n * size * size * esz, 16))
        self.d_y = L.DeviceBuffer(max(n * 4 * size * size * esz, 16))

    def enqueue(self):
        lib = self.lib
        p_margin = self.d_mk.ptr
        p_peak = (self.d_mk.ptr + max(self.nj, 1) * 4 + 255) & ~255
        for a, b, max_size in self.groups:
            L.check(lib.rsb_jet_render_var(self.p_params + a * self.slots * 48, self.p_zpad + a * 4,
                                           self.p_sizes + a * 4, self.p_offs + a * 8, b - a,
                                           self.slots, max_size, self.flags_jet, self.d_pix.ptr,
                                           p_margin + a * 4, p_peak + a * 8, None),
                    "rsb_jet_render_var (survey sub-jets)")
        for a, b, js, cursor in self.big_groups:
            L.check(lib.rsb_jet_render(self.p_params + a * self.slots * 48, self.p_zpad + a * 4, b - a,
                                       self.slots, js, self.flags_jet, self.d_pix.ptr + cursor * 4,
                                       p_margin + a * 4, p_peak + a * 8, self.d_jw.ptr, self.d_jw.nbytes, None),
                    "rsb_jet_render (survey sub-jets above 256 pixels)")
        if self.n:
            L.check(lib.rsb_survey_render(self.p_src, self.p_soffs, self.total, self.p_rects,
                                          self.p_joffs, self.d_pix.ptr, self.n, self.size,
                                          L.JET_OUT_F64 if self.f64 else 0, self.d_x.ptr, self.d_y.ptr,
                                          self.d_w.ptr, self.d_w.nbytes, None), "rsb_survey_render")

    def download_into(self, x: np.ndarray, y: np.ndarray, copier=None):
        """Asynchronous copies into (page-locked, contiguous) host arrays; ``L.device_sync()`` or a
        later synchronising call completes them.  Keep this object alive until then.  With a
        ``copier`` (:class:`_lib.CopyStream`) the copies run on its stream, behind everything enqueued
        so far, and overlap whatever the default stream is given next."""
        n, size = self.n, self.size
        assert x.shape == (n, 1, size, size) and y.shape == (n, 4, size, size)
        assert x.dtype == self.dt and y.dtype == self.dt and x.flags.c_contiguous and y.flags.c_contiguous
        if not n:
            return
        if copier is not None:
            copier.after_default()
            copier.d2h(x, self.d_x.ptr)
            copier.d2h(y, self.d_y.ptr)
            return
        L.check(self.lib.rsb_memcpy_d2h(x.ctypes.data, self.d_x.ptr, x.nbytes, None), "d2h")
        L.check(self.lib.rsb_memcpy_d2h(y.ctypes.data, self.d_y.ptr, y.nbytes, None), "d2h")

    def download(self):
        n, size = self.n, self.size
        x = L.pinned_empty((n, 1, size, size), self.dt)
        y = L.pinned_empty((n, 4, size, size), self.dt)
        self.download_into(x, y)
        L.device_sync()
        return x, y


def render_survey(batch: SurveyBatch, jet_normalize: bool = True, f64: bool = True):
    dev = DeviceSurvey(batch, jet_normalize, f64)
    dev.enqueue()
    return dev.download()


def _chunk_images(n: int, size: int) -> int:
    """Images per pipeline step: about 64 MB of float64 results."""
    return max(1, min(n, (64 << 20) // (5 * size * size * 8)))


def create_survey(grid, conf):
    """Drop-in for ``radiosim.survey.create_survey``: ``(survey[n,1,N,N], comps[n,4,N,N])`` float64.

    The legacy RNG chain (sequential by nature) is drawn chunk after chunk on a helper thread while
    the device renders the previous chunk and copies it to the host; the chunks are packed strictly
    in order, so the global ``np.random`` stream is consumed exactly as by one pass."""
    from concurrent.futures import ThreadPoolExecutor

    shape = tuple(grid.shape)
    n = 1 if len(shape) == 3 else shape[0]
    size = shape[-1]
    sconf, jconf = conf.survey, conf.jet
    args = (size, sconf.num_sources, sconf.class_distribution, sconf.scale_sources,
            jconf.num_jet_components, jconf.scaling)
    normalize = jconf.scaling == "normalize"
    step = _chunk_images(n, size)
    if n <= step:
        return render_survey(pack_survey(n, *args), normalize, f64=True)
    L.require_device()
    x = L.pinned_empty((n, 1, size, size), np.float64)
    y = L.pinned_empty((n, 4, size, size), np.float64)
    in_flight = []
    copier = L.CopyStream.get()
    with ThreadPoolExecutor(1) as ex:
        fut = ex.submit(pack_survey, min(step, n), *args)
        i0 = 0
        while i0 < n:
            batch = fut.result()
            i1 = i0 + batch.n
            if i1 < n:
                fut = ex.submit(pack_survey, min(step, n - i1), *args)
            dev = DeviceSurvey(batch, normalize, f64=True)
            dev.enqueue()
            dev.download_into(x[i0:i1], y[i0:i1], copier)
            in_flight.append(dev)
            if len(in_flight) > 3:      # bound the device memory held by unfinished chunks
                copier.sync()
                del in_flight[:-1]
            i0 = i1
    copier.sync()
    L.device_sync()
    return x, y
