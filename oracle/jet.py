"""CPU restatement of the jet generator (reference ``src/radiosim/jets/jet.py``).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The restatement is split where the B200 path is split -- parameter draws
(:func:`draw_jet`), rendering (:func:`render_slots`), the data-dependent zoom
(:func:`finish_jet`) -- but consumes the legacy global ``np.random`` stream in
exactly the reference's order, so ``create_jet`` here is draw-for-draw and
bit-for-bit comparable with the reference (pinned in ``tests/test_oracle.py``
against ``tests/golden/jet_*.npz``, which ``oracle/make_golden.py`` produced from
the unmodified reference).
"""

from __future__ import annotations

import numpy as np

from .gauss import twodgaussian
from .resize import resize_linear

EDGE_THRESHOLD = 0.01  # src/radiosim/jets/jet.py:156


def boosting(theta, beta):
    """``src/radiosim/utils/utils.py:64-87``."""
    gamma = 1 / np.sqrt(1 - beta**2)
    mu = np.cos(theta)
    return 1 / (gamma * (1 - beta * mu)), 1 / (gamma * (1 + beta * mu))


def draw_jet(img_size: int, n_lo: int, n_hi: int, scaling: str):
    """Parameter chain of one jet, ``src/radiosim/jets/jet.py:54-142``.

    Returns the eight per-slot columns (``2*n_hi-1`` slots each), already boosted,
    shifted and mirrored / zero-filled.
    """
    rnd = np.random
    half = img_size // 2
    k = rnd.randint(n_lo, n_hi + 1)  # :54
    amp, x, y, sx, sy, rot, beta = (np.zeros(n_hi) for _ in range(7))
    rot[0] = rnd.uniform(0, np.pi)  # :62
    beta[1] = rnd.uniform(0, 1)  # :66
    phi = rnd.uniform(0, np.pi)  # :67  (y_rotation)
    incl = rnd.uniform(0, np.pi / 2)  # :68  (z_rotation)
    for i in range(k):
        amp[i] = np.exp(-np.sqrt(i) * rnd.normal(1.3, 0.2))  # :72
        if i >= 1 and rnd.rand() < 0.1:  # :73
            amp[i] = 0
        if i >= 2:  # :77-78
            beta[i] = beta[1] * np.exp(-np.sqrt(i - 1) * rnd.normal(0.5, 0.1))
        phi += rnd.normal(0, np.pi / 24)  # :81
        ac, as_ = np.abs(np.cos(phi)), np.abs(np.sin(phi))  # :84-91
        if ac < as_:
            stretch = 1 / as_
        elif ac > as_:
            stretch = 1 / ac
        else:
            stretch = np.sqrt(2)
        radius = i / (k - 1) * img_size / 2 * stretch * np.sin(incl) * 0.7  # :94
        x[i] = radius * np.cos(phi) + half  # :97 via pol2cart utils.py:433-454
        y[i] = radius * np.sin(phi) + half
        w = img_size / k * stretch * np.sqrt(i + 1) / rnd.uniform(3, 8, size=2)  # :101-107
        sx[i], sy[i] = w.max(), w.min()
        if i >= 1:  # :110-111
            rot[i] = rot[i - 1] + rnd.normal(0, np.pi / 18)

    b_app, b_rec = boosting(incl, beta)  # :114
    dx = rnd.uniform(-img_size / 20, img_size / 20)  # :116
    dy = rnd.uniform(-img_size / 20, img_size / 20)  # :117
    if scaling == "mojave":  # :119-120, flux_scaling.py:6-20
        from scipy import stats

        amp *= (
            stats.beta(
                0.8639672251677816,
                47.64189171625089,
                0.09163404776954732,
                1892.3881692069087,
            ).rvs(1)
            / 1e3
        )
    pad = np.zeros(n_hi - 1)
    if rnd.rand() < 0.3:  # :125-133 counter jet
        cols = dict(
            amp=np.concatenate((amp * b_app, amp[1:] * b_rec[1:])),
            x=np.concatenate((x + dx, img_size - x[1:] + dx)),
            y=np.concatenate((y + dy, img_size - y[1:] + dy)),
            sx=np.concatenate((sx, sx[1:])),
            sy=np.concatenate((sy, sy[1:])),
            rot=np.concatenate((rot, rot[1:])),
            beta=np.concatenate((beta, beta[1:])),
        )
    else:  # :134-142
        cols = dict(
            amp=np.concatenate((amp * b_app, pad)),
            x=np.concatenate((x + dx, pad)),
            y=np.concatenate((y + dy, pad)),
            sx=np.concatenate((sx, pad)),
            sy=np.concatenate((sy, pad)),
            rot=np.concatenate((rot, pad)),
            beta=np.concatenate((beta, pad)),
        )
    cols["z_rot"] = np.repeat(incl, 2 * n_hi - 1)
    return cols


def render_slots(size, amp, x, y, sx, sy, rot):
    """``component_from_list``, ``src/radiosim/jets/jet.py:234-261``."""
    planes = np.zeros((len(amp), size, size))
    for i in range(len(amp)):
        if amp[i] != 0:
            planes[i] = twodgaussian([amp[i], x[i], y[i], sx[i], sy[i], rot[i]], size)
    return planes


def perimeter_max(img):
    """``src/radiosim/jets/jet.py:149-155``: max over the 4(N-1) border pixels."""
    return max(img[0, :].max(), img[-1, :].max(), img[:, 0].max(), img[:, -1].max())


def source_margin(img, thr=EDGE_THRESHOLD):
    """``idx`` of ``zoom_on_source``, ``src/radiosim/utils/utils.py:111-119``.

    Smallest distance from any pixel ``> thr`` to the image border (0 if none).
    """
    mask = img > thr
    cols = mask.any(axis=0)
    rows = mask.any(axis=1)
    return int(
        min(np.argmax(cols), np.argmax(cols[::-1]), np.argmax(rows), np.argmax(rows[::-1]))
    )


def zoom_in(img, planes, thr=EDGE_THRESHOLD):
    """``zoom_on_source``, ``src/radiosim/utils/utils.py:90-138``."""
    n = img.shape[0]
    m = source_margin(img, thr)
    factor = n / (n - 2 * m)
    out = resize_linear(img[m : n - m, m : n - m], n)
    out_planes = None
    if planes is not None:
        out_planes = np.stack([resize_linear(p[m : n - m, m : n - m], n) for p in planes])
    return out, out_planes, factor, m


def zoom_away(img, pad_value):
    """``zoom_out``, ``src/radiosim/utils/utils.py:141-177``.

    ``pad_value`` is truncated to int64 (``:161-162``).  The reference's component
    loop assigns to a loop local (``:169-175``), so the planes come back untouched;
    only the image is resampled here.
    """
    p = int(np.int64(pad_value))
    n = img.shape[0]
    return resize_linear(np.pad(img, p), n), p


def finish_jet(cols, img_size, scaling, training_type, trace=None, planes_needed=True):
    """Render + edge test + zoom + normalise + targets, ``jets/jet.py:144-192``."""
    amp, x, y, sx, sy, rot = (cols[k] for k in ("amp", "x", "y", "sx", "sy", "rot"))
    planes = render_slots(img_size, amp, x, y, sx, sy, rot)  # :145
    img = np.sum(planes, axis=0)  # :146
    info = dict(zoomed=False, margin=0, pad=0, zoom_out_factor=1.0, jet_max=1.0)
    if perimeter_max(img) < EDGE_THRESHOLD:  # :157
        img, planes, zf, m = zoom_in(img, planes)  # :159-161
        x = img_size / 2 + (x - img_size / 2) * zf  # :162-165
        y = img_size / 2 + (y - img_size / 2) * zf
        sx = sx * zf
        sy = sy * zf
        zo = np.random.uniform(1 / 2, 1)  # :168
        pad_value = (1 / zo - 1) * img_size / 2  # :169
        img, p = zoom_away(img, pad_value)  # :170
        x = img_size / 2 + (x - img_size / 2) * zo  # :171-174
        y = img_size / 2 + (y - img_size / 2) * zo
        sx = sx * zo
        sy = sy * zo
        info.update(zoomed=True, margin=m, pad=p, zoom_out_factor=zo)
    if scaling == "normalize":  # :177-181
        peak = img.max()
        img = img / peak
        planes = planes / peak
        amp = amp / peak
        info["jet_max"] = peak
    slist = np.array([amp, x, y, sx, sy, rot, cols["z_rot"], cols["beta"]]).T  # :186
    if training_type == "list":  # :200-231
        target = slist
    elif training_type == "gauss":
        full = np.concatenate((planes, (1 - img)[None]))  # :185
        extra = np.full((1, img_size, img_size), np.nan)
        extra[0, : slist.shape[0], : slist.shape[1]] = slist
        target = np.concatenate((full, extra))
    elif training_type == "clean":
        target = img[None]
    else:
        raise ValueError(training_type)
    if trace is not None:
        trace.append(info)
    return img, target


def create_jet(grid_shape, num_jet_components=(3, 10), scaling="normalize",
               training_type="list", trace=None):
    """``create_jet`` (``src/radiosim/jets/jet.py:14-197``) for ``len(grid)`` images.

    ``grid_shape`` is the shape of the reference's ``grid`` argument (only its
    length and last axis are ever read, ``:46-53``).
    """
    shape = tuple(grid_shape)
    n_img = 1 if len(shape) == 3 else shape[0]
    size = shape[-1]
    lo, hi = num_jet_components
    jets, targets = [], []
    for _ in range(n_img):
        cols = draw_jet(size, lo, hi, scaling)
        img, tgt = finish_jet(cols, size, scaling, training_type, trace)
        jets.append(img)
        targets.append(tgt)
    return np.array(jets)[:, None], np.array(targets)
