"""CPU restatement of ``add_noise`` (reference ``src/radiosim/utils/utils.py:259-315``).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Same legacy global ``np.random`` draw
order as the reference; pinned against ``tests/golden/noise_*.npz`` (made from the unmodified
reference; its ``astropy.convolution.Gaussian2DKernel`` is the restatement in
``oracle/refstub.py``: default size = 8 sigma rounded up to an odd integer, samples at integer
offsets, normalised to sum 1 -- astropy itself is absent from the image).

The reference convolves with ``scipy.signal.fftconvolve(mode="same")``; the kernel is an outer
product, so the same numbers come from two 1-D zero-padded convolutions, which is what this
restatement (and the CUDA path) does.
"""

from __future__ import annotations

import math

import numpy as np

KERNEL_FRACTIONS = (1 / 32, 1 / 8)      # kernels = [1, N/32, N/8]           (:301)
STRENGTHS = (0.2, 0.3, 0.5)             # white, N/32, N/8                    (:302)


def kernel_1d(kernel: float) -> np.ndarray:
    """1-D factor of ``Gaussian2DKernel(kernel / 2).array`` (:281)."""
    sigma = kernel / 2
    size = int(math.ceil(8 * sigma))
    if size % 2 == 0:
        size += 1
    h = (size - 1) // 2
    g = np.exp(-np.arange(-h, h + 1, dtype=np.float64) ** 2 / (2 * sigma**2))
    return g / g.sum()


def blur_same(field: np.ndarray, g: np.ndarray) -> np.ndarray:
    """``fftconvolve(field, outer(g, g)[None, None], mode="same")`` as two 1-D passes."""
    h = (len(g) - 1) // 2
    n = field.shape[-1]
    pad = np.pad(field, [(0, 0)] * (field.ndim - 2) + [(h, h), (h, h)])
    tmp = np.zeros(field.shape[:-1] + (n + 2 * h,))
    for k in range(len(g)):  # rows
        tmp += g[k] * pad[..., k : k + n, :]
    out = np.zeros(field.shape)
    for k in range(len(g)):  # columns
        out += g[k] * tmp[..., :, k : k + n]
    return out


def add_noise(image: np.ndarray, noise_level) -> np.ndarray:
    shape = image.shape
    n = shape[0]
    size = shape[-1]
    total = np.zeros(shape)
    white = np.random.normal(size=shape)                              # :291
    total += white / (np.abs(white).max() / STRENGTHS[0])            # :294 (maximum over the whole bundle)
    for frac, strength in zip(KERNEL_FRACTIONS, STRENGTHS[1:]):
        max_noise = np.random.uniform(0, 1, n)                        # :276
        field = np.random.normal(0, 1, size=shape) * max_noise[:, None, None, None]   # :277-279
        field = blur_same(field, kernel_1d(size * frac))              # :281-282
        total += field / (np.abs(field).max() / strength)
    level = np.random.uniform(*noise_level, size=n)[:, None, None, None] / 100        # :306-308
    total = total / total.max(axis=(1, 2, 3))[:, None, None, None]                      # :310
    total = total * (image.max(axis=(1, 2, 3))[:, None, None, None] * level)           # :311
    return total + image                                                                # :312
