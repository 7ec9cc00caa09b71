"""GPU parity tests of the stand-alone primitives (utils drop-ins)."""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_twodgaussian_and_component_from_list():
    from oracle.gauss import twodgaussian as o_g
    from radiosim_b200.jets import component_from_list
    from radiosim_b200.utils import twodgaussian

    p = [0.7, 40.3, 55.6, 9.0, 4.0, 1.1]
    ref = o_g(p, 96)
    out = twodgaussian(p, 96)
    assert out.dtype == np.float64 and np.abs(out - ref).max() <= 1e-5 * ref.max()
    amp = np.array([1.0, 0.0, 0.4]); x = np.array([30.0, 0.0, 50.5]); y = np.array([31.0, 0.0, 20.25])
    sx = np.array([8.0, 0.0, 5.0]); sy = np.array([4.0, 0.0, 3.0]); rot = np.array([0.3, 0.0, 2.0])
    planes = component_from_list(64, amp, x, y, sx, sy, rot)
    assert len(planes) == 3 and not planes[1].any()
    for k in (0, 2):
        r = o_g([amp[k], x[k], y[k], sx[k], sy[k], rot[k]], 64)
        assert np.abs(planes[k] - r).max() <= 1e-5 * r.max()


def test_skewed_gauss():
    from oracle.gauss import skewed_gauss as o_s
    from radiosim_b200.utils import skewed_gauss

    ref = o_s(128, 64, 64, 0.02, 3.0, 20.0, 4)
    out = skewed_gauss(128, 64, 64, 0.02, 3.0, 20.0, 4)
    assert np.abs(out - ref).max() <= 1e-5 * ref.max()
    assert np.array_equal(out == 0, ref == 0) or np.abs(out - ref).max() <= 2e-8


@pytest.mark.parametrize("size", [96, 128, 151])
def test_zoom_on_source_and_zoom_out(size):
    from oracle import jet as ojet
    from radiosim_b200.utils import zoom_on_source, zoom_out

    rng = np.random.default_rng(size)
    yy, xx = np.mgrid[0:size, 0:size]
    img = np.exp(-((xx - size / 2 - 3) ** 2 / 60.0 + (yy - size / 2 + 2) ** 2 / 30.0))
    comp = np.stack([img * 0.25, img * 0.75])
    z, zc, zf, m = ojet.zoom_in(img, comp)
    out, outc, f = zoom_on_source(img, comp)
    assert f == zf                                   # integer idx -> identical factor
    assert np.abs(out - z).max() <= 1e-12 and np.abs(outc - zc).max() <= 1e-12
    out2, f2 = zoom_on_source(img)
    assert f2 == zf and np.abs(out2 - z).max() <= 1e-12
    pad_value = (1 / 0.7 - 1) * size / 2
    zo, p = ojet.zoom_away(z, pad_value)
    got, same = zoom_out(z, zc, pad_value=pad_value)
    assert same is zc                                # the reference returns comp untouched
    assert np.abs(got - zo).max() <= 1e-12
    # empty mask -> idx 0 -> identity resize
    e, fe = zoom_on_source(img * 1e-3)
    assert fe == 1.0 and np.abs(e - img * 1e-3).max() <= 1e-15
