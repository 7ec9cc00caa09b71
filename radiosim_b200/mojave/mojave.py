"""Drop-in ``create_mojave`` / ``gen_compact`` / ``gen_one_jet`` / ``gen_two_jet``
(reference ``src/radiosim/mojave/mojave.py``).

Host: every child generator's draws are replayed in the reference's order (scipy's
``expon`` / ``skewnorm`` variates are drawn by calling scipy exactly as the reference does)
and turned into a *recipe* -- skewed-Gaussian blobs as pairs of 1-D profiles, swirl
parameters, central Gaussians, rotation angle.  Device: ``rsb_mojave_render`` evaluates the
recipe (``radiosim_b200/csrc/mojave.cu``).  The only pixel the host ever evaluates is the one
a shock's amplitude is read from (``mojave.py:190``), from the 1-D profiles.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np
from scipy.special import ndtr
from scipy.stats import expon, skewnorm

from .. import _lib as L

__all__ = ["add_swirl", "create_mojave", "gen_compact", "gen_jet", "gen_one_jet", "gen_shocks",
           "gen_two_jet"]

GAUSS = np.dtype([("amp", "<f8"), ("x", "<f8"), ("y", "<f8"), ("sx", "<f8"), ("sy", "<f8"), ("rot", "<f8")])
RECIPE = np.dtype([("g", GAUSS, (3,)), ("strength", "<f8", (2,)), ("radius", "<f8", (2,)),
                   ("angle", "<f8"), ("amp", "<f8"), ("blob0", "<i4", (2,)), ("nblob", "<i4", (2,)),
                   ("flip1", "<i4"), ("compact", "<i4"), ("pad", "<i4", (2,))])
BLOB = np.dtype([("amp", "<f4"), ("pad", "<i4"), ("prof", "<i8")])
BLOBP = np.dtype([("x", "<f8"), ("y", "<f8"), ("amp", "<f8"), ("width", "<f8"), ("length", "<f8"), ("a", "<f8"),
                  ("ref_r", "<i4"), ("ref_c", "<i4"), ("pad", "<i4", (2,))])
assert RECIPE.itemsize == 224 and BLOB.itemsize == 16 and BLOBP.itemsize == 64


def _as_rng(rng):
    return np.random.default_rng(seed=rng) if isinstance(rng, (int, np.integer)) else rng


def _expon_rvs(rng, loc, scale):
    """``scipy.stats.expon.rvs(loc, scale, random_state=rng)`` without its argument checking
    (same draw, same arithmetic: ``standard_exponential * scale + loc``)."""
    return rng.standard_exponential(()) * scale + loc


def _skewnorm_rvs(rng, a, loc, scale):
    """``scipy.stats.skewnorm.rvs(a, loc, scale, random_state=rng)``: scipy's two-normal
    construction (``skewnorm_gen._rvs``), then ``* scale + loc``."""
    a = np.float64(a)
    u0 = rng.normal(size=())
    v = rng.normal(size=())
    d = a / np.sqrt(1 + a**2)
    u1 = d * u0 + v * np.sqrt(1 - d**2)
    return (u1 if u0 >= 0 else -u1) * scale + loc


class _Blob:
    """amp * normal[row] * skew[col] (``src/radiosim/utils/gauss.py:131-148``)."""

    __slots__ = ("amp", "normal", "skew", "args", "factor", "ref")

    def __init__(self, size, x, y, amp, width, length, a):
        self.args = (x, y, amp, width, length, a)   # as drawn (compared with the native packer)
        self.factor, self.ref = None, None          # shocks: amplitude = jet pixel `ref` x `factor`
        vals = np.arange(size)
        t = (vals - y) / width
        normal = np.exp(-(t**2) / 2.0) / np.sqrt(2 * np.pi) / width
        self.normal = normal / normal.max()
        u = (vals - x) / length
        skew = 2.0 * (np.exp(-(u**2) / 2.0) / np.sqrt(2 * np.pi)) * ndtr(a * u) / length
        self.skew = skew / skew.max()
        self.amp = amp

    def at(self, r, c):
        v = (self.normal[r] * self.skew[c]) * self.amp
        return 0.0 if abs(v) <= 1e-8 else v  # np.isclose(v, 0) with its default atol, rtol * 0 = 0


class _Jet:
    def __init__(self, size):
        self.size = size
        self.blobs = []
        self.strength = 0.0
        self.radius = 100.0

    def pixel(self, r, c):
        """Value of the un-swirled jet image at one pixel, summed in build order."""
        s = 0.0
        for b in self.blobs:
            s = s + b.at(r, c)
        return s


def _base_jet(size, amp, width, length, a):
    j = _Jet(size)
    j.blobs.append(_Blob(size, size / 2, size / 2, amp, width, length, a))
    return j


def _add_shocks(jet, n_shock, length, width, sx, rng):
    """``gen_shocks`` ``:124-219``."""
    size = jet.size
    step = (2 * length - sx) / n_shock
    dist = step
    for _ in range(n_shock):
        if rng.uniform(0, 1) > 2 / 3:
            continue
        px, py = size / 2 + dist, size / 2
        r, c = int(py), int(px)
        if not (-size <= r < size and -size <= c < size):   # the reference's IndexError
            continue
        factor = rng.uniform(0.03, 0.6)
        s_amp = jet.pixel(r % size, c % size) * factor
        s_sx = rng.uniform(*np.sort([2, width]))
        s_sy = rng.uniform(*np.sort([s_sx, length / n_shock]))
        s_a = rng.uniform(5, 9)
        jet.blobs.append(_Blob(size, int(px), int(py), s_amp, s_sx, s_sy, s_a))
        jet.blobs[-1].factor, jet.blobs[-1].ref = factor, (r % size, c % size)
        dist += step


def _swirl(jet, rng, first=None):
    """``add_swirl`` ``:222-261`` (draws only)."""
    if first is None:
        jet.strength = rng.normal(loc=0, scale=0.2)
        jet.radius = rng.uniform(50, 200)
    else:
        dev = 1 + rng.normal(loc=0, scale=0.1, size=2)
        jet.strength = first.strength * dev[0]
        jet.radius = first.radius * dev[1]


def _core_widths(rng, dims):
    sx = rng.uniform(dims.min(), dims.max()) * 1.5
    sy = rng.uniform(dims.min(), dims.max()) * 1.5
    while sx / sy < 1 / 2 or sx / sy > 2:
        sy = rng.uniform(dims.min(), dims.max()) * 1.5
    return sx, sy, rng.uniform(0, 2 * np.pi)


class _Source:
    def __init__(self, size):
        self.size = size
        self.jets = []       # up to two _Jet
        self.flip1 = False
        self.gauss = []      # [amp, x, y, sx, sy, rot]
        self.angle = 0.0
        self.compact = False
        self.amp = 0.0


def _recipe_two_jet(rng, size):
    """``gen_two_jet`` ``:264-398``."""
    dev = 1 + rng.normal(loc=0, scale=0.1, size=5)
    amp = _expon_rvs(rng, 1.715e-4, 1.985e-2)
    r_w = _skewnorm_rvs(rng, 4.37, 22.71, 16.85) / 10
    r_l = _skewnorm_rvs(rng, 8.62, 69.60, 54.73) / 4
    r_amp = amp * rng.power(0.7)
    l_w, l_l, l_amp = r_w * dev[2], r_l * dev[1], r_amp * dev[0]
    dims = np.sort([r_w / 2, r_l / 5, l_w / 2, l_l / 5])
    sx, sy, rot = _core_widths(rng, dims)
    r_jet = _base_jet(size, r_amp, r_w, r_l, 4)
    l_jet = _base_jet(size, l_amp, l_w, l_l, 4)
    n_r = int(np.round(rng.power(0.5), decimals=1) * 10)
    n_l = int(np.round(rng.power(0.5), decimals=1) * 10)
    if n_r > 0:
        _add_shocks(r_jet, n_r, r_l, r_w, sx, rng)
    if n_l > 0:
        _add_shocks(l_jet, n_l, l_l, l_w, sx, rng)
    _swirl(r_jet, rng)
    _swirl(l_jet, rng, r_jet)
    s = _Source(size)
    s.jets = [r_jet, l_jet]
    s.flip1 = True
    s.gauss.append([amp, size / 2, size / 2, sx, sy, rot])
    s.angle = rng.uniform(0, 360)
    return s


def _recipe_one_jet(rng, size):
    """``gen_one_jet`` ``:401-492``."""
    amp = _expon_rvs(rng, 1.715e-4, 1.985e-2)
    width = _skewnorm_rvs(rng, 4.37, 22.71, 16.85) / 12
    length = _skewnorm_rvs(rng, 8.62, 69.60, 54.73) / 6.5
    jet_amp = amp * rng.power(0.7)
    dims = np.sort([width / 2, length / 5])
    sx, sy, rot = _core_widths(rng, dims)
    jet = _base_jet(size, jet_amp, width, length, 4)
    n_s = int(np.round(rng.power(0.5), decimals=1) * 10)
    if n_s > 0:
        _add_shocks(jet, n_s, length, width, sx, rng)
    _swirl(jet, rng)
    s = _Source(size)
    s.jets = [jet]
    s.gauss.append([amp, size / 2, size / 2, sx, sy, rot])
    s.angle = rng.uniform(0, 360)
    return s


def _bounded_width(rng, dims):
    v = rng.uniform(*dims) * 1.7
    tries = 0
    while v > 15:
        if tries > 10:
            v /= 5
            break
        v = rng.uniform(*dims) * 1.7
        tries += 1
    return v


def _recipe_compact(rng, size):
    """``gen_compact`` ``:495-662``."""
    ncomps = rng.integers(1, 4)
    outflow = rng.uniform() > 0.5
    amp = _expon_rvs(rng, 5.54e-05, 0.02)
    jet_amp = amp * rng.normal(2 / 3, 0.05)
    width = _skewnorm_rvs(rng, 8.41, 3.04, 3.46) * 0.8
    length = _skewnorm_rvs(rng, 6.04, 3.02, 2.59) * 1
    s = _Source(size)
    if outflow:
        jet = _base_jet(size, jet_amp, width, length, 4)
        _swirl(jet, rng)
        s.jets = [jet]
    dims = np.sort([width / 2, length / 4])
    dims[0] *= 1.2
    dims = np.sort(dims)
    c_sx, c_sy, c_rot = [], [], []
    vertical = True
    for i in range(int(ncomps)):
        dev = 1 + rng.normal(loc=0, scale=0.2, size=2)
        if i != 0:
            ox, oy = rng.uniform(-5, 5), rng.uniform(-5, 5)
        else:
            ox = oy = 0
        if i == 0:
            sx = _bounded_width(rng, dims)
            sy = _bounded_width(rng, dims)
            tries = 0
            while (sx / sy < 0.33 or sx / sy > 3) or (0.75 < sx / sy < 1.33):
                if tries > 10:
                    sy = sx * rng.choice([rng.uniform(0.4, 0.8), rng.uniform(1.2, 2.5)])
                    if sy > 15:
                        sy /= 5
                    break
                sy = _bounded_width(rng, dims)
                tries += 1
            vertical = sx > sy
        else:
            sx, sy = c_sx[-1] * dev[0], c_sy[-1] * dev[1]
        if c_rot:
            rot = rng.uniform(c_rot[-1] - 0.1, c_rot[-1] + 0.1)
        elif vertical:
            rot = rng.uniform(-np.pi / 4, np.pi / 4)
        else:
            rot = rng.uniform(-np.pi / 4 + np.pi / 2, np.pi / 4 + np.pi / 2)
        c_amp = amp * rng.uniform(0.5, 1.2)
        s.gauss.append([c_amp, size / 2 - ox, size / 2 - oy, sx, sy, rot])
        c_sx.append(sx)
        c_sy.append(sy)
        c_rot.append(rot)
    s.angle = rng.uniform(0, 360)
    s.compact = True
    s.amp = amp
    return s


def pack_recipes(part, size, pad=True):
    """Recipes -> the three arrays ``rsb_mojave_render`` takes (records, blobs, 1-D profiles).
    ``pad=False`` leaves the blob / profile arrays empty when there is no blob (for merging)."""
    rec = np.zeros(len(part), dtype=RECIPE)
    blobs, profs = [], []
    for k, s in enumerate(part):
        r = rec[k]
        for q, g in enumerate(s.gauss):
            r["g"][q] = tuple(g)
        for q, j in enumerate(s.jets):
            r["blob0"][q] = len(blobs)
            r["nblob"][q] = len(j.blobs)
            r["strength"][q], r["radius"][q] = j.strength, j.radius
            for b in j.blobs:
                blobs.append((np.float32(b.amp), 0, 2 * size * len(profs)))
                profs.append(np.concatenate((b.normal, b.skew)).astype(np.float32))
        r["angle"], r["amp"] = s.angle, s.amp
        r["flip1"], r["compact"] = int(s.flip1), int(s.compact)
    blob_arr = np.array(blobs, dtype=BLOB) if blobs else np.zeros(1 if pad else 0, dtype=BLOB)
    prof_arr = np.concatenate(profs) if profs else np.zeros(2 * size if pad else 0, dtype=np.float32)
    return rec, blob_arr, prof_arr


def merge_packed(chunks, size):
    """Concatenate packed chunks, re-basing blob and profile offsets."""
    recs, blobs, profs = [], [], []
    nb = npf = 0
    for rec, b, p in chunks:
        rec = rec.copy()
        rec["blob0"] = np.where(rec["nblob"] > 0, rec["blob0"] + nb, 0)  # (unused entries stay 0)
        b = b.copy()
        b["prof"] += npf
        recs.append(rec); blobs.append(b); profs.append(p)
        nb += len(b)
        npf += len(p)
    rec = np.concatenate(recs) if recs else np.zeros(0, dtype=RECIPE)
    blob_arr = np.concatenate(blobs) if nb else np.zeros(1, dtype=BLOB)
    prof_arr = np.concatenate(profs) if npf else np.zeros(2 * size, dtype=np.float32)
    return rec, blob_arr, prof_arr


def render_packed(rec, blob_arr, prof_arr, size, f64=True, chunk=2048):
    """Render packed recipes: slices of the records against the whole blob / profile arrays."""
    L.require_device()
    lib = L.lib()
    n = len(rec)
    dt = np.float64 if f64 else np.float32
    out = L.pinned_empty((n, size, size), dt)
    max_imgs = max(1, min(chunk, (1 << 30) // (size * size * 8)))
    d_b, d_p = L.dev_from(blob_arr), L.dev_from(prof_arr)
    for i0 in range(0, n, max_imgs):
        cnt = min(max_imgs, n - i0)
        need = C.c_size_t(0)
        L.check(lib.rsb_mojave_workspace_bytes(cnt, size, C.byref(need)))
        d_r = L.dev_from(rec[i0 : i0 + cnt])
        d_w = L.DeviceBuffer(need.value)
        d_o = L.DeviceBuffer(cnt * size * size * np.dtype(dt).itemsize)
        L.check(lib.rsb_mojave_render(d_r.ptr, cnt, d_b.ptr, d_p.ptr, size,
                                      L.JET_OUT_F64 if f64 else 0, d_o.ptr, d_w.ptr, d_w.nbytes, None),
                "rsb_mojave_render")
        dst = out[i0 : i0 + cnt]
        L.check(lib.rsb_memcpy_d2h(dst.ctypes.data, d_o.ptr, dst.nbytes, None), "d2h")
        L.device_sync()
    return out


def _words(v: int) -> np.ndarray:
    out = []
    while True:
        out.append(v & 0xFFFFFFFF)
        v >>= 32
        if v == 0:
            return np.array(out, dtype=np.uint32)


def native_recipes(rng, amount, size):
    """The host half of ``create_mojave`` (``mojave.py:56-64`` + the ``gen_*`` draws) natively
    (``rsb_mojave_pack``): the children the three ``rng.spawn`` calls would hand out are derived from
    the parent's seed sequence in C++, their draws replayed, and the parent's spawn counter moved on
    as ``spawn`` would.  Returns ``(recipes, blob parameters)`` or ``None`` when the parent is not a
    PCG64 generator on a plain integer-seeded ``SeedSequence`` (then the Python packer runs)."""
    bg = rng.bit_generator
    ss = getattr(bg, "seed_seq", None)
    if (type(bg).__name__ != "PCG64" or type(ss) is not np.random.SeedSequence or ss.pool_size != 4
            or not isinstance(ss.entropy, (int, np.integer)) or ss.entropy < 0
            or not all(isinstance(k, (int, np.integer)) and 0 <= k for k in ss.spawn_key)):
        return None
    n = int(np.sum(amount))
    first = ss.n_children_spawned
    if first + n >= 2**32 or len(ss.spawn_key) > 8:
        return None
    entropy = _words(int(ss.entropy))
    key = (np.concatenate([_words(int(k)) for k in ss.spawn_key]) if ss.spawn_key else np.zeros(0, np.uint32))
    kinds = np.repeat(np.arange(3, dtype=np.int32), np.asarray(amount, dtype=np.int64))
    rec = np.zeros(n, dtype=RECIPE)
    cap = 24 * max(n, 1)          # <= 2 jets x (1 + 10 shocks) per image
    bp = np.zeros(cap, dtype=BLOBP)
    nb = C.c_int32(0)
    L.check(L.lib().rsb_mojave_pack(entropy.ctypes.data, len(entropy), key.ctypes.data if len(key) else None,
                                    len(key), first, n, kinds.ctypes.data, size, rec.ctypes.data, bp.ctypes.data,
                                    cap, C.byref(nb)), "rsb_mojave_pack")
    # the three spawn() calls of the reference leave the parent's child counter n further on
    moved = np.random.SeedSequence(ss.entropy, spawn_key=ss.spawn_key, pool_size=ss.pool_size,
                                   n_children_spawned=first + n)
    bg.__setstate__((bg.state, moved))
    return rec, bp[: nb.value]


def render_native(rec, bp, size, f64=True, chunk=2048):
    """Recipes + blob parameters (:func:`native_recipes`) -> images; profiles and shock amplitudes are
    finished on the device (``rsb_mojave_blobs``)."""
    L.require_device()
    lib = L.lib()
    n = len(rec)
    dt = np.float64 if f64 else np.float32
    out = L.pinned_empty((n, size, size), dt)
    nb = max(len(bp), 1)
    d_bp = L.dev_from(bp if len(bp) else np.zeros(1, dtype=BLOBP))
    d_blobs, d_prof = L.DeviceBuffer(nb * 16), L.DeviceBuffer(nb * 2 * size * 4)
    # chunks of ~64 MB of results: a chunk's copy to the host (second stream) runs under the next
    # chunk's kernels (default stream)
    max_imgs = max(1, min(chunk, (64 << 20) // (size * size * np.dtype(dt).itemsize)))
    copier = L.CopyStream.get()
    keep = []
    for i0 in range(0, n, max_imgs):
        cnt = min(max_imgs, n - i0)
        need = C.c_size_t(0)
        L.check(lib.rsb_mojave_workspace_bytes(cnt, size, C.byref(need)))
        d_r, (p_r,), stage = L.upload_blob([rec[i0 : i0 + cnt]])
        d_w = L.DeviceBuffer(need.value)
        d_o = L.DeviceBuffer(cnt * size * size * np.dtype(dt).itemsize)
        L.check(lib.rsb_mojave_blobs(p_r, cnt, d_bp.ptr, len(bp), size, d_blobs.ptr, d_prof.ptr, None),
                "rsb_mojave_blobs")
        L.check(lib.rsb_mojave_render(p_r, cnt, d_blobs.ptr, d_prof.ptr, size, L.JET_OUT_F64 if f64 else 0,
                                      d_o.ptr, d_w.ptr, d_w.nbytes, None), "rsb_mojave_render")
        copier.after_default()
        copier.d2h(out[i0 : i0 + cnt], d_o.ptr)
        keep.append((d_r, stage, d_w, d_o))
        if len(keep) > 3:               # bound the blocks held by chunks still being copied
            copier.sync()
            del keep[:-1]
    copier.sync()
    L.device_sync()
    return out


def render_sources(sources, size, f64=True, chunk=2048):
    """Evaluate recipes on the device -> ``[n, size, size]`` float64 (or float32)."""
    return render_packed(*pack_recipes(sources, size), size, f64=f64, chunk=chunk)


def gen_two_jet(rng, size=1024, printout=False):
    return render_sources([_recipe_two_jet(_as_rng(rng), size)], size)[0]


def gen_one_jet(rng, size=1024, printout=False):
    return render_sources([_recipe_one_jet(_as_rng(rng), size)], size)[0]


def gen_compact(rng, size=1024, printout=False):
    return render_sources([_recipe_compact(_as_rng(rng), size)], size)[0]


def gen_jet(size, amplitude, width, length, a):
    """``gen_jet`` ``:84-121`` as an image (device render of a single un-swirled blob)."""
    s = _Source(size)
    j = _base_jet(size, amplitude, width, length, a)
    j.strength, j.radius = 0.0, 100.0
    s.jets = [j]
    return render_sources([s], size)[0]


def gen_shocks(jet, shock_comp, length, width, sx, rng):
    """``gen_shocks`` ``:124-219`` on a host image, as the reference's signature has it: equally spaced
    shocks (each a device-rendered ``skewed_gauss``) added to ``jet`` IN PLACE; returns ``(jet,
    printout_information, skipped)``.  Same draws in the same order from ``rng``; an out-of-range
    position is passed over before its amplitude factor is drawn, as the reference's ``IndexError``
    handler does.  (The bundle path does not come through here: there the shocks are blobs of the
    recipe, ``_add_shocks``.)"""
    from ..utils.gauss import skewed_gauss

    info = ([], [], [], [], [], [], [], [], [])
    size = len(jet)
    step = (2 * length - sx) / shock_comp
    dist = step
    skipped = 0
    for _ in range(shock_comp):
        if rng.uniform(0, 1) > 2 / 3:
            skipped += 1
            continue
        s_x, s_y = size / 2 + dist, size / 2
        try:
            at = jet[int(s_y), int(s_x)]
        except IndexError:
            continue
        s_amp = at * rng.uniform(0.03, 0.6)
        s_sx = rng.uniform(*np.sort([2, width]))
        s_sy = rng.uniform(*np.sort([s_sx, length / shock_comp]))
        s_a = rng.uniform(5, 9)
        jet += skewed_gauss(size, int(s_x), int(s_y), s_amp, s_sx, s_sy, s_a)   # (float32 render: values below 1e-8 come back as 0)
        for lst, v in zip(info, ("r", s_amp, int(s_x), int(s_y), s_sx, s_sy, 0, dist, s_a)):
            lst.append(v)
        dist += step
    return jet, info, skipped


def swirl_image(image, center, strength, radius):
    """``skimage.transform.swirl(image, center=center, strength=strength, radius=radius)`` (all other
    arguments at their defaults) on the device: ``rsb_swirl``."""
    from .. import _lib as L

    L.require_device()
    img = np.ascontiguousarray(image, dtype=np.float64)
    if img.ndim != 2:
        raise ValueError("swirl_image: a 2-D image is expected")
    rows, cols = img.shape
    d_in = L.dev_from(img)
    d_out = L.DeviceBuffer(max(img.nbytes, 16))
    L.check(L.lib().rsb_swirl(d_in.ptr, rows, cols, float(center[0]), float(center[1]), float(strength),
                              float(radius), float(img.min()), float(img.max()), d_out.ptr, None), "rsb_swirl")
    return np.array(d_out.download((rows, cols), np.float64))


def add_swirl(jet, rng, first_jet_params=None):
    """``add_swirl`` ``:222-261``: draws the swirl (or varies the first jet's by ~10 %) and applies it;
    returns ``(swirled_jet, (strength, radius, center))``."""
    size = len(jet)
    if not first_jet_params:
        strength = rng.normal(loc=0, scale=0.2)
        radius = rng.uniform(50, 200)
        center = np.array([size / 2, size / 2])
    else:
        deviation = 1 + rng.normal(loc=0, scale=0.1, size=2)
        strength = first_jet_params[0] * deviation[0]
        radius = first_jet_params[1] * deviation[1]
        center = first_jet_params[2]
    return swirl_image(jet, center, strength, radius), (strength, radius, center)


def class_counts(class_ratio, bundle_size):
    """``create_mojave`` ``:47-53``."""
    ratio = np.array(class_ratio)
    amount = np.array((ratio / ratio.sum()) * bundle_size).astype(int)
    while amount.sum() < bundle_size:
        amount[amount.argmax()] += 1
    return amount


def _recipe_chunk(kind, children, size):
    fn = (_recipe_compact, _recipe_one_jet, _recipe_two_jet)[kind]
    return [fn(c, size) for c in children]


def _packed_chunk(kind, children, size):
    return pack_recipes(_recipe_chunk(kind, children, size), size, pad=False)


def make_recipes(rng, amount, size, threads=1):
    """The host half of ``create_mojave`` ``:56-64``: one recipe per spawned child generator, class
    after class, returned packed (:func:`pack_recipes`).  The children are independent streams,
    so ``threads`` > 1 deals them to that many worker processes (the reference hands them to
    joblib workers); the result does not depend on it."""
    kids = [rng.spawn(int(a)) for a in amount]  # same spawn order as the reference
    threads = int(threads)
    if threads <= 1 or sum(len(k) for k in kids) < 8 * threads:
        return merge_packed([_packed_chunk(kind, ch, size) for kind, ch in enumerate(kids)], size)
    import multiprocessing as mp

    jobs = []
    for kind, ch in enumerate(kids):
        step = max(1, -(-len(ch) // (4 * threads)))
        jobs += [(kind, ch[i : i + step], size) for i in range(0, len(ch), step)]
    # "forkserver": the parent may hold a CUDA context and writer threads by now, and forking such a
    # process is unsafe; the workers only need numpy / scipy and never touch the GPU
    with mp.get_context("forkserver").Pool(threads) as pool:
        parts = pool.starmap(_packed_chunk, jobs)
    return merge_packed(parts, size)


def create_mojave(conf, rng, today=None):
    """Drop-in for ``create_mojave`` (``:23-81``): ``(data, data_name)``.

    ``general.threads`` is validated like the reference (``:38-42``); the recipes of the child
    generators are drawn on that many host processes and rendered in one device batch.
    """
    threads = conf.general.threads
    if not threads:
        threads = 1
    elif not isinstance(threads, int) or threads <= 0:
        raise ValueError("threads has to be int > 0 or None")
    size = conf.dataset.img_size
    bundle_size = conf.dataset.bundle_size
    amount = class_counts(conf.mojave.class_ratio, bundle_size)
    native = None if os.environ.get("RSB_PY_PACKER") == "1" else native_recipes(rng, amount, size)
    if native is None:
        rec, blob_arr, prof_arr = make_recipes(rng, amount, size, threads)
    classes = np.repeat(np.arange(3), amount)
    order = np.arange(bundle_size)
    rng.shuffle(order)
    # rendered in shuffled order, so the images need no host-side permutation afterwards
    if native is not None:
        jets_shuffled = render_native(native[0][order], native[1], size)
    else:
        jets_shuffled = render_packed(rec[order], blob_arr, prof_arr, size)
    ra = rng.uniform(0, 360, size=bundle_size)      # utils.py:386-390
    dec = rng.uniform(-30, 87, size=bundle_size)
    start = np.datetime64("1995-01-01")             # utils.py:393-406
    if today is None:
        from datetime import datetime

        today = datetime.now().date()
    delta = np.datetime64(today) - start
    dates = start + rng.integers(delta.astype(int), size=bundle_size)
    data = [jets_shuffled, classes[order], ra[order], dec[order], dates.astype("S10")]
    return data, ["galaxies", "galaxy_classes", "RA", "DEC", "date"]
