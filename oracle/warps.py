"""CPU restatement of the two scikit-image warps the mojave path calls.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

PARITY UNPINNED: the arithmetic lives in scikit-image (pinned 0.25.2 in the
reference's ``uv.lock``), which is a third-party dependency that is neither
vendored under the reference tree nor installed on this image.  The functions
below restate its published algorithm; the anchor is the reference's call sites
``src/radiosim/mojave/mojave.py:259`` (``swirl(jet, center=, strength=, radius=)``)
and ``src/radiosim/mojave/mojave.py:346,463,626`` (``rotate(glx, angle=)``), i.e.
all other arguments are skimage defaults (``order=1`` for float images,
``clip=True``, ``swirl: mode='reflect', rotation=0``, ``rotate: mode='constant',
cval=0, resize=False, center=None``).

Published algorithm (skimage/transform/_warps.py):

* ``swirl``: inverse map of an output pixel (x=col, y=row):
  ``rho = hypot(x-x0, y-y0)``; ``r' = radius/5*ln 2``;
  ``theta = rotation + strength*exp(-rho/r') + atan2(y-y0, x-x0)``;
  sample the input at ``(x0 + rho cos theta, y0 + rho sin theta)``; bilinear
  interpolation; ``mode='reflect'`` is handed to ndimage as ``'mirror'``.
* ``rotate``: centre ``c = (cols/2-0.5, rows/2-0.5)``; inverse map
  ``p_in = c + R(angle)(p_out - c)`` with ``R = [[cos,-sin],[sin,cos]]`` on
  (x, y), angle in degrees; bilinear; neighbours outside the image contribute
  ``cval=0`` (ndimage ``'grid-constant'``).
* both clip the result to the input's [min, max] (a no-op up to one rounding for
  bilinear interpolation of a non-negative image whose minimum is 0).
"""

from __future__ import annotations

import numpy as np
from scipy import ndimage as ndi


def _clip_like_input(out, src):
    lo, hi = src.min(), src.max()
    return np.clip(out, lo, hi, out=out)


def swirl(image, center=None, strength=1, radius=100, rotation=0):
    img = np.asarray(image, dtype=np.float64)
    rows, cols = img.shape
    if center is None:
        center = np.array([cols, rows]) / 2
    x0, y0 = float(center[0]), float(center[1])
    yy, xx = np.mgrid[0:rows, 0:cols].astype(np.float64)
    dx = xx - x0
    dy = yy - y0
    rho = np.sqrt(dx**2 + dy**2)
    r_eff = radius / 5 * np.log(2)
    theta = rotation + strength * np.exp(-rho / r_eff) + np.arctan2(dy, dx)
    src_x = x0 + rho * np.cos(theta)
    src_y = y0 + rho * np.sin(theta)
    out = ndi.map_coordinates(
        img, [src_y, src_x], order=1, mode="mirror", prefilter=False
    )
    return _clip_like_input(out, img)


def rotate(image, angle):
    img = np.asarray(image, dtype=np.float64)
    rows, cols = img.shape
    cx = cols / 2.0 - 0.5
    cy = rows / 2.0 - 0.5
    a = np.deg2rad(angle)
    ca, sa = np.cos(a), np.sin(a)
    yy, xx = np.mgrid[0:rows, 0:cols].astype(np.float64)
    # composed homogeneous matrix, as skimage builds it: T(+c) R T(-c)
    m02 = cx - ca * cx + sa * cy
    m12 = cy - sa * cx - ca * cy
    src_x = ca * xx - sa * yy + m02
    src_y = sa * xx + ca * yy + m12
    out = ndi.map_coordinates(
        img, [src_y, src_x], order=1, mode="grid-constant", cval=0.0, prefilter=False
    )
    lo, hi = min(img.min(), 0.0), max(img.max(), 0.0)
    return np.clip(out, lo, hi, out=out)
