"""Drop-ins for the hot-path helpers of ``src/radiosim/utils/utils.py``."""

from __future__ import annotations

import numpy as np

from .. import _lib as L
from ..h5min import write_h5

__all__ = ["cart2pol", "create_grid", "load_data", "pol2cart", "relativistic_boosting",
           "save_sky_distribution_bundle", "zoom_on_source", "zoom_out", "_gen_date",
           "_gen_vlba_obs_position", "_save_mojave_bundle"]


class _ShapeOnly:
    """What ``create_grid`` returns: the generators only ever read ``len`` and ``.shape`` of the
    grid (``jets/jet.py:46-53``), so the 3 x N x N float64 planes per image of the reference
    (``utils.py:52-61``, 25 MB per 1024^2 image) are not materialised."""

    def __init__(self, shape):
        self.shape = tuple(shape)

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, item):
        return np.broadcast_to(np.zeros(1), self.shape)[item]


def create_grid(pixel, bundle_size):
    """``utils.py:36-61`` (shape-only stand-in, see :class:`_ShapeOnly`)."""
    return _ShapeOnly((bundle_size, 3, pixel, pixel))


def relativistic_boosting(theta, beta):
    """``utils.py:64-87``."""
    gamma = 1 / np.sqrt(1 - beta**2)
    mu = np.cos(theta)
    return 1 / (gamma * (1 - beta * mu)), 1 / (gamma * (1 + beta * mu))


def pol2cart(r, phi):
    """``utils.py:433-454``."""
    return (r * np.cos(phi), r * np.sin(phi))


def cart2pol(x, y):
    """``utils.py:409-430``."""
    return (np.sqrt(x**2 + y**2), np.arctan2(y, x))


def _gen_vlba_obs_position(rng, size):
    """``utils.py:386-390``: RA in [0, 360), DEC in [-30, 87), in this order from ``rng``."""
    ra = rng.uniform(0, 360, size=size)
    dec = rng.uniform(-30, 87, size=size)
    return ra, dec


def _gen_date(rng, start_date, size, today=None):
    """``utils.py:393-406``: uniformly drawn days between ``start_date`` and today as fixed-length
    10-byte strings (what h5py's ``string_dtype("utf-8", 10)`` stores).  ``today`` can be frozen for
    reproducible bundles."""
    start = np.datetime64(start_date)
    if today is None:
        from datetime import datetime

        today = datetime.now().date()
    delta = np.datetime64(today) - start
    dates = start + rng.integers(delta.astype(int), size=size)
    return dates.astype("S10")


def _resize(planes: np.ndarray, crop: int, dst: int, pad: int) -> np.ndarray:
    """planes [k, S, S] float64 -> resize(pad(planes[:, crop:S-crop, crop:S-crop])) on the device."""
    k, S, _ = planes.shape
    src = np.ascontiguousarray(planes, dtype=np.float64)
    d_s = L.dev_from(src)
    d_o = L.DeviceBuffer(k * dst * dst * 8)
    off = (crop * S + crop) * 8
    L.check(L.lib().rsb_resize_linear(d_s.ptr + off, S - 2 * crop, S, S * S, pad, k, d_o.ptr, dst, None),
            "rsb_resize_linear")
    return d_o.download((k, dst, dst), np.float64)


def zoom_on_source(img, comp=None, max_amp=0.01):
    """``utils.py:90-138``."""
    L.require_device()
    img = np.ascontiguousarray(img, dtype=np.float64)
    size = img.shape[0]
    d_i = L.dev_from(img)
    d_m = L.DeviceBuffer(16)
    L.check(L.lib().rsb_source_margin(d_i.ptr, size, float(max_amp), d_m.ptr, None), "rsb_source_margin")
    idx = int(d_m.download((1,), np.int32)[0])
    idx = 0 if idx >= size else idx
    zoom_factor = size / (size - 2 * idx)
    zoomed = _resize(img[None], idx, size, 0)[0]
    if comp is not None:
        return zoomed, _resize(np.asarray(comp, dtype=np.float64), idx, size, 0), zoom_factor
    return zoomed, zoom_factor


def zoom_out(img, comp=None, pad_value=0):
    """``utils.py:141-177`` -- including its quirk: the components come back unchanged."""
    L.require_device()
    if not isinstance(pad_value, int):
        pad_value = int(np.int64(pad_value))
    img = np.ascontiguousarray(img, dtype=np.float64)
    out = _resize(img[None], 0, img.shape[0], pad_value)[0]
    if comp is not None:
        return out, comp
    return out


def save_sky_distribution_bundle(path, x, y, name_x="x", name_y="y"):
    """``utils.py:344-364`` through the built-in HDF5 writer (no h5py on this image)."""
    write_h5(path, {name_x: np.asarray(x), name_y: np.asarray(y)})


def _save_mojave_bundle(path, data, data_name):
    """``utils.py:367-383``: one dataset per (array, name) pair."""
    assert len(data) == len(data_name)
    write_h5(path, {name: np.asarray(dat) for dat, name in zip(data, data_name)})


def load_data(conf_path, data_type="train", key="x"):
    """``utils.py:457-470``: every image of dataset ``key`` from the bundles of one split under the
    configuration's ``paths.outpath`` (the reference's pattern ``"data*" + data_type`` as a regular
    expression on the file name -- which, as in the reference, matches ``datatrain...`` but not the
    writer's own ``data_jet_train_0.h5``), through the built-in HDF5 reader."""
    import re
    import tomllib
    from pathlib import Path

    from ..h5min import read_h5

    with open(conf_path, "rb") as f:
        config = tomllib.load(f)
    path = Path(config["paths"]["outpath"])
    data = []
    for p in path.iterdir():
        if re.findall("data*" + data_type, p.name):
            data.extend(read_h5(p)[key])
    return data
