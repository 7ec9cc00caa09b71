"""Import the UNMODIFIED reference (``/root/reference/src``) in the build container.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Used by
``oracle/make_golden.py`` and by the container-only tests that pin the numpy
restatement against the real reference.  ``/root/reference`` does not exist on
the GPU box, so nothing that runs there may call :func:`load_reference`.

Four packages the reference imports at module level are absent on this image
(astropy, scikit-image, h5py, rich_click; no network).  They are replaced by
minimal stand-ins in ``sys.modules``:

* ``rich_click``      -> ``click`` plus a ``rich_click`` attribute bag
  (``src/radiosim/__init__.py:1-17`` only sets attributes on it);
* ``h5py``            -> raise-on-use (the generators never touch it);
* ``astropy.nddata.utils.Cutout2D`` -> raise-on-use (only dead code calls it);
* ``astropy.convolution.Gaussian2DKernel`` -> :class:`Gaussian2DKernel` below, a
  restatement of astropy 7.1 (third-party, pinned in ``uv.lock``; call site
  ``src/radiosim/ppdisks/ppdisks.py:446-450``).  It is pinned indirectly: with it
  the reference reproduces its own fixture ``tests/ppdisks_128.h5``;
* ``skimage.transform.swirl/rotate`` -> ``oracle.warps`` (PARITY UNPINNED).
"""

from __future__ import annotations

import os
import sys
import types

import numpy as np

REFERENCE_SRC = "/root/reference/src"


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "radiosim"))


class Gaussian2DKernel:
    """astropy.convolution.Gaussian2DKernel, restated.

    Model ``1/(2 pi sx sy) exp(-(a x^2 + b x y + c y^2))`` sampled at pixel centres
    (``mode='center'``): integer offsets for an odd size, half-integer offsets
    ``-(n/2)+0.5 ... n/2-0.5`` for an even size; then divided by the array sum.
    Default size is ``8*max(sigma)`` rounded up to an odd integer.
    """

    def __init__(self, x_stddev, y_stddev=None, theta=0.0, x_size=None, y_size=None):
        if y_stddev is None:
            y_stddev = x_stddev
        default = int(np.ceil(8 * max(x_stddev, y_stddev)))
        if default % 2 == 0:
            default += 1
        if x_size is None:
            x_size = default
        if y_size is None:
            y_size = x_size

        def axis(n):
            n = int(n)
            if n % 2 == 0:
                return np.arange(-(n // 2) + 0.5, n // 2 + 0.5)
            return np.arange(-(n - 1) // 2, (n - 1) // 2 + 1).astype(np.float64)

        xs, ys = np.meshgrid(axis(x_size), axis(y_size))
        ct, st = np.cos(theta), np.sin(theta)
        a = 0.5 * (ct**2 / x_stddev**2 + st**2 / y_stddev**2)
        b = 0.5 * np.sin(2 * theta) * (1 / x_stddev**2 - 1 / y_stddev**2)
        c = 0.5 * (st**2 / x_stddev**2 + ct**2 / y_stddev**2)
        arr = np.exp(-(a * xs**2 + b * xs * ys + c * ys**2)) / (
            2 * np.pi * x_stddev * y_stddev
        )
        self.array = arr / arr.sum()

    def __array__(self, dtype=None, copy=None):
        return self.array if dtype is None else self.array.astype(dtype)

    @property
    def shape(self):
        return self.array.shape


def _raiser(name):
    def _f(*a, **k):
        raise RuntimeError(f"{name} is a stub: not available on this image")

    return _f


def install_stubs() -> None:
    import click

    from oracle import warps

    if "rich_click" not in sys.modules:
        rc = types.ModuleType("rich_click")
        for k in dir(click):
            if not k.startswith("__"):
                setattr(rc, k, getattr(click, k))
        rc.rich_click = types.SimpleNamespace()
        sys.modules["rich_click"] = rc

    if "h5py" not in sys.modules:
        h5 = types.ModuleType("h5py")
        h5.File = _raiser("h5py.File")
        h5.string_dtype = _raiser("h5py.string_dtype")
        sys.modules["h5py"] = h5

    if "astropy" not in sys.modules:
        ap = types.ModuleType("astropy")
        conv = types.ModuleType("astropy.convolution")
        conv.Gaussian2DKernel = Gaussian2DKernel
        ndd = types.ModuleType("astropy.nddata")
        nddu = types.ModuleType("astropy.nddata.utils")
        nddu.Cutout2D = _raiser("astropy Cutout2D")
        ap.convolution, ap.nddata, ndd.utils = conv, ndd, nddu
        sys.modules.update(
            {
                "astropy": ap,
                "astropy.convolution": conv,
                "astropy.nddata": ndd,
                "astropy.nddata.utils": nddu,
            }
        )

    if "skimage" not in sys.modules:
        sk = types.ModuleType("skimage")
        tr = types.ModuleType("skimage.transform")
        tr.swirl = warps.swirl
        tr.rotate = warps.rotate
        sk.transform = tr
        sys.modules.update({"skimage": sk, "skimage.transform": tr})


def load_reference():
    """Return the imported reference package ``radiosim`` (container only)."""
    if not reference_available():
        raise RuntimeError(
            "/root/reference is not present (GPU box?) - use tests/golden fixtures"
        )
    install_stubs()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import radiosim  # noqa: F401
    import radiosim.jets
    import radiosim.mojave
    import radiosim.ppdisks
    import radiosim.survey

    return radiosim
