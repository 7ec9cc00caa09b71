"""CPU restatement of the Gaussian primitives (reference ``src/radiosim/utils/gauss.py``).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).
"""

from __future__ import annotations

import numpy as np
from scipy.special import ndtr


def twodgaussian(params, size: int) -> np.ndarray:
    """Rotated elliptical Gaussian on the full ``size x size`` grid.

    Follows ``src/radiosim/utils/gauss.py:50-97`` operation for operation (so the
    float64 result is bit-identical): ``params = [amp, p1, p2, p3, p4, rot]`` where
    the reference's argument swap makes ``p1`` the COLUMN centre, ``p2`` the ROW
    centre, ``p4`` the width dividing the first rotated coordinate and ``p3`` the
    width dividing the second (``:77-79``).
    """
    amp, p1, p2, p3, p4, rot = (params[i] for i in range(6))
    c_row, c_col = p2, p1
    w_a, w_b = p4, p3
    cr, sr = np.cos(rot), np.sin(rot)
    ra = c_row * cr - c_col * sr  # :80
    rb = c_row * sr + c_col * cr  # :81
    r, c = np.indices((size, size))
    ua = r * cr - c * sr  # :84
    ub = r * sr + c * cr  # :85
    return amp * np.exp(-(((ra - ua) / w_a) ** 2 + ((rb - ub) / w_b) ** 2) / 2.0)  # :86-88


def twodgaussian_at(params, rows, cols) -> np.ndarray:
    """Same formula as :func:`twodgaussian`, evaluated only at the given pixels."""
    amp, p1, p2, p3, p4, rot = (params[i] for i in range(6))
    cr, sr = np.cos(rot), np.sin(rot)
    ra = p2 * cr - p1 * sr
    rb = p2 * sr + p1 * cr
    ua = rows * cr - cols * sr
    ub = rows * sr + cols * cr
    return amp * np.exp(-(((ra - ua) / p4) ** 2 + ((rb - ub) / p3) ** 2) / 2.0)


def _norm_pdf(t):
    return np.exp(-(t**2) / 2.0) / np.sqrt(2 * np.pi)


def skew_profiles(size: int, x, y, width, length, a):
    """The two max-normalised 1-D factors of ``skewed_gauss``.

    ``src/radiosim/utils/gauss.py:131-143``: ``norm.pdf(vals, loc=y, scale=width)``
    (rows) and ``skewnorm.pdf(vals, a, loc=x, scale=length)`` (columns), each divided
    by its maximum over the INTEGER samples.  scipy's ``skewnorm.pdf`` is
    ``2 phi(t) Phi(a t) / scale``.
    """
    vals = np.arange(size)
    normal = _norm_pdf((vals - y) / width) / width
    normal = normal / normal.max()
    t = (vals - x) / length
    skew = 2.0 * _norm_pdf(t) * ndtr(a * t) / length
    skew = skew / skew.max()
    return normal, skew


def skewed_gauss(size: int, x, y, amp, width, length, a) -> np.ndarray:
    """``src/radiosim/utils/gauss.py:100-150``: outer product * amp, isclose-zeroed."""
    normal, skew = skew_profiles(size, x, y, width, length, a)
    out = normal.reshape(-1, 1) * skew
    out *= amp
    out[np.isclose(out, 0)] = 0
    return out
