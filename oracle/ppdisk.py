"""CPU restatement of the protoplanetary-disk generator
(reference ``src/radiosim/ppdisks/ppdisks.py``).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Pinned: reproduces the reference's own fixture ``tests/ppdisks_128.h5``
(``tests/test_ppdisks_simulation.py:1-34``) -- see ``tests/test_oracle.py``.

Two dtype facts of the reference that this restatement keeps, because they decide
mask bits and rounding (both probed with torch 2.11 in the build container):

* ``_create_ellipse`` (``:225-234``) multiplies an int64 tensor by numpy float64
  scalars.  ``np.float64`` is a Python ``float`` subclass, so torch promotes to its
  DEFAULT dtype: the whole membership test runs in **float32** (scalars rounded to
  float32 first, true division, no FMA).
* ``els_full`` (``:139``) is a float32 tensor, so every ``els_full += ring`` rounds
  the running ring sum to float32; the final canvas (``:178``) is float32 too.

``_smooth`` (``:426-451``) is a circular convolution of the mask with a
sum-normalised Gaussian sampled at half-integer offsets; after the two rolls of
``_convolve2d`` (``:420-421``) the net effect is
``out[i,j] = sum_kl mask[k,l] h[(i-k+449)%900] h[(j-l+449)%900]`` with
``h[t] = g(t-449.5)/sum g``.  The reference evaluates it with a complex64 FFT of
the mask (noise ~1e-10 absolute); here it is the exact float64 sum.
"""

from __future__ import annotations

import numpy as np

from .resize import resize_bilinear_aa

FIX = 900  # ppdisks.py:131
_FITS = (  # ppdisks.py:467-472
    (-0.08908080536281453, -4.808388735775005),
    (-0.04053517757866287, -5.718461569072661),
    (-0.023989028595164514, -6.22105762345096),
    (-0.011198837787046627, -6.701733964123176),
)
_FIT_KEYS = ["average_1", "average_2", "average_3", "average_4"]


def ellipse_mask(a, b, alpha_deg, n=FIX, c0=FIX // 2):
    """``_create_ellipse`` ``:188-236`` with its float32 arithmetic (see header)."""
    f32 = np.float32
    ca = f32(np.cos(np.deg2rad(alpha_deg)))
    sa = f32(np.sin(np.deg2rad(alpha_deg)))
    a2 = f32(a**2)
    b2 = f32(b**2)
    d = (np.arange(n) - c0).astype(f32)
    dr = d[:, None]
    dc = d[None, :]
    t1 = dr * ca + dc * sa
    t1 = t1 * t1
    t2 = dr * sa - dc * ca
    t2 = t2 * t2
    with np.errstate(divide="ignore", invalid="ignore"):
        q = t1 / a2 + t2 / b2
    assert q.dtype == np.float32
    return q <= f32(1)


def smooth_matrix(sigma, n=FIX):
    """Circulant matrix ``H[i,k] = h[(i-k+449)%900]`` of the 1-D factor of ``_smooth``.

    Kernel samples: astropy ``Gaussian2DKernel(x_size=900)`` evaluates at
    ``-449.5 ... 449.5`` (even size) and divides by the 2-D sum (``:446-450``);
    the 2-D kernel is the outer product of ``h`` with itself.
    """
    t = np.arange(n) - (n // 2 - 0.5)
    g = np.exp(-(t**2) / (2.0 * sigma**2))
    h = g / g.sum()
    i = np.arange(n)
    return h[(i[:, None] - i[None, :] + (n // 2 - 1)) % n]


def smooth(mask, sigma):
    H = smooth_matrix(sigma, mask.shape[0])
    return H @ mask.astype(np.float64) @ H.T


def smooth_factor(a):
    """``_get_smooth_factor`` ``:380-394``."""
    if a < 15:
        return 2
    if a < 20:
        return 4
    if a < 30:
        return 6
    if a < 40:
        return 9
    return 12


def _scale(d, m, n):
    """``_get_scale`` ``:477-483`` (one ``normal`` draw)."""
    e = np.exp(m * d + n)
    return e + np.random.normal(0, e * 0.05)


def create_proto(img_size, alpha, ratio, size_ratio, seed=None, trace=None):
    """``create_proto`` ``:95-185``."""
    if seed is not None:
        np.random.seed(seed)
    key = np.random.choice(_FIT_KEYS)  # :474
    m, n = _FITS[_FIT_KEYS.index(key)]
    a = np.float64(np.random.randint(3, 15))  # :305
    core_mask = ellipse_mask(a, a / ratio, alpha)
    d_core = (a + a / ratio) / 2
    sc = np.random.randint(5, 10)  # :135
    core = smooth(core_mask, sc) * _scale(d_core, m, n)  # :136
    num_circ = int(np.random.normal(3, 0.5) // 1)  # :137
    rings = np.zeros((FIX, FIX), dtype=np.float32)  # :139
    a_max = d_core
    ring_log = []
    for i in range(num_circ):  # :143-149 -> _sim_ring :310-377
        ra = np.random.randint(10, 35) + (i + 1) * 30
        ex = np.random.randint(5, 10)
        inner = ellipse_mask(ra, ra / ratio, alpha)
        outer = ellipse_mask(ra + ex, (ra + ex) / ratio, alpha)
        ring_mask = outer ^ inner  # :269
        d_ring = ((ra + ex + (ra + ex) / ratio) / 2 + (ra + ra / ratio) / 2) / 2
        s = smooth_factor(ra)
        ring = smooth(ring_mask, s) * _scale(d_ring, m, n)
        extent = np.max([ra, ra + ex, ra / ratio, (ra + ex) / ratio]) + 6 * s
        a_max = np.max([extent, a_max])
        rings = (rings.astype(np.float64) + ring).astype(np.float32)  # float32 accumulator
        ring_log.append((ra, ex, s))
    proto = np.abs(core + rings.astype(np.float64))  # :151-152
    half = FIX // 2
    a_max = np.min([half, a_max])
    lo, hi = int(half - a_max), int(half + a_max)
    snip = proto[lo:hi, lo:hi]
    if snip.max() < 1e-10:  # :163-164
        return create_proto(img_size, alpha, ratio, size_ratio)
    new = int(size_ratio * img_size)
    new = new if new % 2 == 0 else new - 1
    snip = resize_bilinear_aa(snip, new)  # :171
    h2 = new // 2
    c = img_size // 2
    canvas = np.ones((img_size, img_size), dtype=np.float32) * np.float32(1e-10)  # :178
    win = canvas[c - h2 : c + h2, c - h2 : c + h2]
    canvas[c - h2 : c + h2, c - h2 : c + h2] = (win.astype(np.float64) + snip).astype(
        np.float32
    )
    if trace is not None:
        trace.append(dict(a=a, sc=sc, num_circ=num_circ, rings=ring_log, lo=lo, hi=hi,
                          new=new, fit=key))
    return canvas


def generate_proto_set(img_size, size, alpha_range=(0, 180), ratio_range=(3, 15),
                       size_ratio_range=(0.1, 1), seed=1337, trace=None):
    """``generate_proto_set`` ``:21-92``."""
    np.random.seed(seed)  # :67
    protos = []
    for _ in range(size):
        alpha = np.random.randint(*alpha_range)  # :74
        ratio = np.random.randint(*ratio_range) / np.random.randint(*ratio_range)  # :75
        size_ratio = np.random.uniform(*size_ratio_range)  # :76
        protos.append(create_proto(img_size, alpha, ratio, size_ratio, trace=trace))
    params = dict(seed=seed, N=size, img_size=img_size, alpha_range=alpha_range,
                  ratio_range=ratio_range, size_ratio_range=size_ratio_range)
    return protos, params
