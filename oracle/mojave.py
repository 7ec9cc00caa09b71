"""CPU restatement of the MOJAVE-like source generator
(reference ``src/radiosim/mojave/mojave.py``).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

PARITY UNPINNED for the warp steps (``oracle/warps.py`` restates scikit-image,
which is not available here).  Everything else -- draw order on the numpy
``Generator`` children, the scipy ``expon``/``skewnorm`` variates (scipy is called,
not restated), skewed Gaussians, shocks, cores, isclose zeroing, recentring -- is
pinned against the unmodified reference run with the same restated warps plugged
in (``tests/golden/mojave_*.npz``).
"""

from __future__ import annotations

import numpy as np
from scipy.stats import expon, skewnorm

from . import warps
from .gauss import skewed_gauss, twodgaussian


def _as_rng(rng):
    return np.random.default_rng(seed=rng) if isinstance(rng, int) else rng


def base_jet(size, amp, width, length, a):
    """``gen_jet`` ``:84-121``: skewed Gaussian at the image centre."""
    j = skewed_gauss(size, size / 2, size / 2, amp, width, length, a)
    j[np.isclose(j, 0)] = 0
    return j


def add_shocks(jet, n_shock, length, width, sx, rng):
    """``gen_shocks`` ``:124-219`` (amplitude reads the image built so far, ``:190``)."""
    size = len(jet)
    step = (2 * length - sx) / n_shock
    dist = step
    for _ in range(n_shock):
        if rng.uniform(0, 1) > 2 / 3:  # :182
            continue
        px = size / 2 + dist
        py = size / 2
        try:
            s_amp = jet[int(py), int(px)] * rng.uniform(0.03, 0.6)  # :190
        except IndexError:
            continue
        s_sx = rng.uniform(*np.sort([2, width]))  # :194
        s_sy = rng.uniform(*np.sort([s_sx, length / n_shock]))  # :195
        s_a = rng.uniform(5, 9)  # :199
        jet += skewed_gauss(size, int(px), int(py), s_amp, s_sx, s_sy, s_a)  # :201
        dist += step
    return jet


def swirl_jet(jet, rng, first=None):
    """``add_swirl`` ``:222-261``."""
    size = len(jet)
    if not first:
        strength = rng.normal(loc=0, scale=0.2)
        radius = rng.uniform(50, 200)
        center = np.array([size / 2, size / 2])
    else:
        dev = 1 + rng.normal(loc=0, scale=0.1, size=2)
        strength = first[0] * dev[0]
        radius = first[1] * dev[1]
        center = first[2]
    return warps.swirl(jet, center=center, strength=strength, radius=radius), (
        strength,
        radius,
        center,
    )


def _core_widths(rng, dims):
    """sx, sy, rot of the central Gaussian, ``:307-315`` / ``:436-444``."""
    sx = rng.uniform(dims.min(), dims.max()) * 1.5
    sy = rng.uniform(dims.min(), dims.max()) * 1.5
    while sx / sy < 1 / 2 or sx / sy > 2:
        sy = rng.uniform(dims.min(), dims.max()) * 1.5
    rot = rng.uniform(0, 2 * np.pi)
    return sx, sy, rot


def gen_two_jet(rng, size=1024):
    """``gen_two_jet`` ``:264-398``."""
    rng = _as_rng(rng)
    dev = 1 + rng.normal(loc=0, scale=0.1, size=5)  # :289-291
    amp = expon.rvs(1.715e-4, 1.985e-2, random_state=rng)  # :293
    r_w = skewnorm.rvs(4.37, 22.71, 16.85, random_state=rng) / 10  # :295
    r_l = skewnorm.rvs(8.62, 69.60, 54.73, random_state=rng) / 4  # :296
    r_amp = amp * rng.power(0.7)  # :297
    l_w, l_l, l_amp = r_w * dev[2], r_l * dev[1], r_amp * dev[0]  # :299-301
    dims = np.sort([r_w / 2, r_l / 5, l_w / 2, l_l / 5])  # :303
    sx, sy, rot = _core_widths(rng, dims)
    r_jet = base_jet(size, r_amp, r_w, r_l, 4)  # :319-320
    l_jet = base_jet(size, l_amp, l_w, l_l, 4)
    n_r = int(np.round(rng.power(0.5), decimals=1) * 10)  # :322-323
    n_l = int(np.round(rng.power(0.5), decimals=1) * 10)
    if n_r > 0:
        r_jet = add_shocks(r_jet, n_r, r_l, r_w, sx, rng)
    if n_l > 0:
        l_jet = add_shocks(l_jet, n_l, l_l, l_w, sx, rng)
    r_jet, first = swirl_jet(r_jet, rng)  # :335-336
    l_jet, _ = swirl_jet(l_jet, rng, first)
    glx = r_jet + np.flip(l_jet)  # :337
    glx += twodgaussian([amp, size / 2, size / 2, sx, sy, rot], size)  # :341-342
    glx = warps.rotate(glx, rng.uniform(0, 360))  # :344-346
    glx[np.isclose(glx, 0)] = 0  # :348
    return glx


def gen_one_jet(rng, size=1024):
    """``gen_one_jet`` ``:401-492``."""
    rng = _as_rng(rng)
    amp = expon.rvs(1.715e-4, 1.985e-2, random_state=rng)  # :426
    width = skewnorm.rvs(4.37, 22.71, 16.85, random_state=rng) / 12  # :428
    length = skewnorm.rvs(8.62, 69.60, 54.73, random_state=rng) / 6.5  # :429
    jet_amp = amp * rng.power(0.7)  # :430
    dims = np.sort([width / 2, length / 5])  # :432
    sx, sy, rot = _core_widths(rng, dims)
    jet = base_jet(size, jet_amp, width, length, 4)  # :448
    n_s = int(np.round(rng.power(0.5), decimals=1) * 10)  # :450
    if n_s > 0:
        jet = add_shocks(jet, n_s, length, width, sx, rng)
    jet, _ = swirl_jet(jet, rng)  # :456
    glx = jet + twodgaussian([amp, size / 2, size / 2, sx, sy, rot], size)  # :458-459
    glx = warps.rotate(glx, rng.uniform(0, 360))  # :461-463
    glx[np.isclose(glx, 0)] = 0  # :465
    return glx


def _bounded_width(rng, dims):
    """One of the ``while s > 15`` redraw loops, ``:560-577``."""
    s = rng.uniform(*dims) * 1.7
    tries = 0
    while s > 15:
        if tries > 10:
            s /= 5
            break
        s = rng.uniform(*dims) * 1.7
        tries += 1
    return s


def gen_compact(rng, size=1024):
    """``gen_compact`` ``:495-662``."""
    rng = _as_rng(rng)
    ncomps = rng.integers(1, 4)  # :519
    outflow = rng.uniform() > 0.5  # :520
    amp = expon.rvs(5.54e-05, 0.02, random_state=rng)  # :522
    jet_amp = amp * rng.normal(2 / 3, 0.05)  # :524
    width = skewnorm.rvs(8.41, 3.04, 3.46, random_state=rng) * 0.8  # :528
    length = skewnorm.rvs(6.04, 3.02, 2.59, random_state=rng) * 1  # :529
    if outflow:  # :531-536
        glx = base_jet(size, jet_amp, width, length, 4)
        glx, _ = swirl_jet(glx, rng)
    else:
        glx = np.zeros((size, size))
    dims = np.sort([width / 2, length / 4])  # :538-540
    dims[0] *= 1.2
    dims = np.sort(dims)
    c_sx, c_sy, c_rot = [], [], []
    vertical = True
    for i in range(int(ncomps)):  # :550-623
        dev = 1 + rng.normal(loc=0, scale=0.2, size=2)
        if i != 0:
            ox = rng.uniform(-5, 5)
            oy = rng.uniform(-5, 5)
        else:
            ox = oy = 0
        if i == 0:
            sx = _bounded_width(rng, dims)
            sy = _bounded_width(rng, dims)
            tries = 0
            while (sx / sy < 0.33 or sx / sy > 3) or (0.75 < sx / sy < 1.33):  # :580-596
                if tries > 10:
                    sy = sx * rng.choice([rng.uniform(0.4, 0.8), rng.uniform(1.2, 2.5)])
                    if sy > 15:
                        sy /= 5
                    break
                sy = _bounded_width(rng, dims)
                tries += 1
            vertical = sx > sy
        else:
            sx = c_sx[-1] * dev[0]
            sy = c_sy[-1] * dev[1]
        if c_rot:  # :604-610
            rot = rng.uniform(c_rot[-1] - 0.1, c_rot[-1] + 0.1)
        elif vertical:
            rot = rng.uniform(-np.pi / 4, np.pi / 4)
        else:
            rot = rng.uniform(-np.pi / 4 + np.pi / 2, np.pi / 4 + np.pi / 2)
        c_amp = amp * rng.uniform(0.5, 1.2)  # :612
        glx += twodgaussian([c_amp, size / 2 - ox, size / 2 - oy, sx, sy, rot], size)
        c_sx.append(sx)
        c_sy.append(sy)
        c_rot.append(rot)
    glx = warps.rotate(glx, rng.uniform(0, 360))  # :625-626
    shift = size // 2 - np.argwhere(glx == glx.max())[0]  # :628
    glx = np.roll(glx, shift=shift, axis=[0, 1])  # :630
    glx[np.isclose(glx, 0)] = 0  # :632
    glx /= glx.max()  # :634-635
    glx *= amp
    return glx


def class_counts(class_ratio, bundle_size):
    """``create_mojave`` ``:47-53``."""
    ratio = np.array(class_ratio)
    amount = np.array((ratio / ratio.sum()) * bundle_size).astype(int)
    while amount.sum() < bundle_size:
        amount[amount.argmax()] += 1
    return amount


def create_mojave(img_size, bundle_size, class_ratio, rng, today=None):
    """``create_mojave`` ``:23-81`` (single process; ``joblib`` only changes where
    each independent child generator is consumed).

    ``today`` freezes ``datetime.now().date()`` of ``_gen_date``
    (``src/radiosim/utils/utils.py:393-406``).
    """
    amount = class_counts(class_ratio, bundle_size)
    compact = [gen_compact(c, img_size) for c in rng.spawn(amount[0])]
    one = [gen_one_jet(c, img_size) for c in rng.spawn(amount[1])]
    two = [gen_two_jet(c, img_size) for c in rng.spawn(amount[2])]
    jets = np.array([*compact, *one, *two])
    classes = np.repeat(np.arange(3), amount)
    order = np.arange(bundle_size)
    rng.shuffle(order)
    ra = rng.uniform(0, 360, size=bundle_size)
    dec = rng.uniform(-30, 87, size=bundle_size)
    start = np.datetime64("1995-01-01")
    if today is None:
        from datetime import datetime

        today = datetime.now().date()
    delta = np.datetime64(today) - start
    dates = start + rng.integers(delta.astype(int), size=bundle_size)
    data = [jets[order], classes[order], ra[order], dec[order], dates.astype("S10")]
    return data, ["galaxies", "galaxy_classes", "RA", "DEC", "date"]
