"""CPU restatement of the two resamplers the hot path calls in third-party code.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

* :func:`resize_linear` -- ``cv2.resize(src, (D, D), interpolation=INTER_LINEAR)`` on
  float64, as called from ``src/radiosim/utils/utils.py:124-127,133-135,164-166``
  (opencv-python pinned 4.11.0.86 in ``uv.lock``).  Pinned: ``tests/test_oracle.py``
  checks it against the installed cv2 to <= 1e-14 when cv2 imports.
* :func:`resize_bilinear_aa` -- ``torchvision.transforms.functional.resize`` on a
  float tensor = ATen ``_upsample_bilinear2d_aa`` (``align_corners=False``), as
  called from ``src/radiosim/ppdisks/ppdisks.py:171``.  Pinned against torch itself
  in ``tests/test_oracle.py`` and through the ppdisk fixture.
"""

from __future__ import annotations

import numpy as np


def linear_taps(S: int, D: int):
    """Per destination sample: left source index, right source index, right weight.

    cv2 INTER_LINEAR: ``f = (d+0.5)*S/D - 0.5``; ``s = floor(f)``; ``w = f - s``;
    ``s < 0 -> s = 0, w = 0``; ``s >= S-1 -> s = S-1, w = 0``.
    """
    d = np.arange(D, dtype=np.float64)
    f = (d + 0.5) * (S / D) - 0.5
    s = np.floor(f)
    w = f - s
    lo = s < 0
    hi = s >= S - 1
    s = np.where(lo, 0, np.where(hi, S - 1, s)).astype(np.int64)
    w = np.where(lo | hi, 0.0, w)
    s1 = np.minimum(s + 1, S - 1)
    return s, s1, w


def resize_linear(src: np.ndarray, D: int) -> np.ndarray:
    """Square-output bilinear resize without antialiasing (cv2 semantics)."""
    src = np.asarray(src, dtype=np.float64)
    R, C = src.shape
    s, s1, w = linear_taps(C, D)
    tmp = src[:, s] * (1.0 - w)[None, :] + src[:, s1] * w[None, :]
    s, s1, w = linear_taps(R, D)
    return tmp[s, :] * (1.0 - w)[:, None] + tmp[s1, :] * w[:, None]


def aa_taps(S: int, D: int):
    """Antialiased triangle-filter taps of ATen's ``_upsample_bilinear2d_aa``.

    Returns ``(xmin[D], count[D], weights[D, K])`` with rows normalised to 1.
    """
    scale = S / D
    support = scale if scale >= 1.0 else 1.0
    invscale = 1.0 / scale if scale >= 1.0 else 1.0
    K = int(np.ceil(support)) * 2 + 1
    xmin = np.zeros(D, dtype=np.int64)
    cnt = np.zeros(D, dtype=np.int64)
    wts = np.zeros((D, K), dtype=np.float64)
    for d in range(D):
        center = scale * (d + 0.5)
        lo = max(0, int(center - support + 0.5))
        hi = min(S, int(center + support + 0.5))
        j = np.arange(hi - lo, dtype=np.float64)
        t = (j + lo - center + 0.5) * invscale
        w = np.maximum(0.0, 1.0 - np.abs(t))
        tot = w.sum()
        if tot != 0.0:
            w = w / tot
        xmin[d], cnt[d] = lo, hi - lo
        wts[d, : hi - lo] = w
    return xmin, cnt, wts


def resize_bilinear_aa(src: np.ndarray, D: int) -> np.ndarray:
    """Square-output antialiased bilinear resize (torch semantics), float64.

    ATen runs the horizontal pass first, then the vertical one.
    """
    src = np.asarray(src, dtype=np.float64)
    R, C = src.shape
    xmin, cnt, wts = aa_taps(C, D)
    tmp = np.zeros((R, D))
    for d in range(D):
        tmp[:, d] = src[:, xmin[d] : xmin[d] + cnt[d]] @ wts[d, : cnt[d]]
    xmin, cnt, wts = aa_taps(R, D)
    out = np.zeros((D, D))
    for d in range(D):
        out[d, :] = wts[d, : cnt[d]] @ tmp[xmin[d] : xmin[d] + cnt[d], :]
    return out
