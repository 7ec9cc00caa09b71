"""What can be said about ``skimage.transform.rotate / swirl`` without scikit-image.

scikit-image is absent from the build container AND from the GPU box (probed in round 2:
``import skimage`` fails on both), so ``oracle/warps.py`` cannot be pinned against the
library itself.  These checks hold the restatement to everything that is known about the
library from outside:

* the known-answer case of scikit-image's own test-suite for ``rotate``
  (``skimage/transform/tests/test_warps.py::test_rotate``: a 5x5 image with ``x[1, 1] = 1``
  rotated by 90 degrees equals ``np.rot90(x)``) -- this fixes the DIRECTION (counter-clockwise
  for positive angles) and the CENTRE (``cols/2 - 0.5``: with any other centre a quarter turn
  does not map pixels onto pixels);
* quarter turns are exact index permutations for even and odd sizes;
* ``swirl``: the documented reverse map ``theta' = rotation + strength*exp(-rho/r) + theta``,
  ``r = ln(2)*radius/5``: strength 0 is the identity, the sign of the displacement of a
  feature, and the inverse-swirl round trip scikit-image's own ``test_swirl`` asserts.
"""

import numpy as np

from oracle import warps


def test_rotate_known_answer_of_skimage_test_suite():
    x = np.zeros((5, 5))
    x[1, 1] = 1
    assert np.allclose(warps.rotate(x, 90), np.rot90(x), atol=1e-12)


def test_rotate_quarter_turns_are_index_permutations():
    rng = np.random.default_rng(0)
    for n in (8, 9, 64):
        img = rng.random((n, n))
        for k in range(4):
            assert np.allclose(warps.rotate(img, 90 * k), np.rot90(img, k), atol=1e-12), (n, k)
    # rows != cols is never used by the reference (square images only)


def test_rotate_direction_small_angle():
    """Positive angle = counter-clockwise as displayed (row 0 on top): a feature right of the
    centre moves up (towards smaller row indices)."""
    n = 65
    img = np.zeros((n, n))
    img[32, 52] = 1.0
    out = warps.rotate(img, 10)
    r, c = np.unravel_index(out.argmax(), out.shape)
    assert r < 32 and c <= 52
    exp_r = 32 - 20 * np.sin(np.deg2rad(10))
    assert abs(r - exp_r) <= 1


def test_swirl_zero_strength_is_identity():
    rng = np.random.default_rng(1)
    img = rng.random((40, 40))
    out = warps.swirl(img, center=(20, 20), strength=0, radius=80)
    assert np.allclose(out, img, atol=1e-12)


def test_swirl_direction_and_amount():
    """Reverse map: the output at polar angle t shows the input at t + s*exp(-rho/r), so a feature
    at input angle t0 appears at t0 - s*exp(-rho/r) (angles measured with atan2(row, col))."""
    n = 129
    c0 = n / 2
    yy, xx = np.mgrid[0:n, 0:n]
    rho0, t0 = 30.0, 0.3
    fy, fx = c0 + rho0 * np.sin(t0), c0 + rho0 * np.cos(t0)
    img = np.exp(-((yy - fy) ** 2 + (xx - fx) ** 2) / (2 * 2.0**2))
    s, radius = 0.4, 150.0
    out = warps.swirl(img, center=(c0, c0), strength=s, radius=radius)
    w = out / out.sum()
    cy, cx = (w * yy).sum(), (w * xx).sum()
    t1 = np.arctan2(cy - c0, cx - c0)
    expected = t0 - s * np.exp(-rho0 / (radius / 5 * np.log(2)))
    assert abs(t1 - expected) < 5e-3
    assert abs(np.hypot(cy - c0, cx - c0) - rho0) < 0.1


def test_swirl_round_trip_like_skimage_test_suite():
    """skimage's test_swirl: swirl(+s) then swirl(-s) gives the image back (mean abs error < 0.01
    away from the border)."""
    yy, xx = np.mgrid[0:200, 0:200]
    img = (((yy // 25) + (xx // 25)) % 2).astype(np.float64)  # checkerboard
    a = warps.swirl(img, center=(100, 100), strength=10, radius=80)
    b = warps.swirl(a, center=(100, 100), strength=-10, radius=80)
    assert np.mean(np.abs(img[1:-1, 1:-1] - b[1:-1, 1:-1])) < 0.05
