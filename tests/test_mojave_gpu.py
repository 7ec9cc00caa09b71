"""GPU parity tests of the mojave path.

The oracle's swirl / rotate are restatements of scikit-image (PARITY UNPINNED, see
oracle/warps.py); everything else is pinned through the reference (tests/test_oracle.py).
"""

import datetime
import os
from types import SimpleNamespace

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5  # of the image peak


@pytest.mark.parametrize("kind", ["compact", "one_jet", "two_jet"])
def test_generators_match_reference_golden(golden_dir, kind):
    from radiosim_b200 import mojave as m

    g = np.load(os.path.join(golden_dir, "mojave_128_classes.npz"))
    fn = {"compact": m.gen_compact, "one_jet": m.gen_one_jet, "two_jet": m.gen_two_jet}[kind]
    for k, s in enumerate(g["seeds"]):
        out = fn(int(s), size=int(g["size"]))
        ref = g[kind][k]
        assert out.dtype == np.float64 and out.shape == ref.shape
        assert np.abs(out - ref).max() <= TOL * ref.max(), (kind, k)


@pytest.mark.parametrize("size", [96, 256])
def test_generators_match_oracle_many_seeds(size):
    from oracle import mojave as om
    from radiosim_b200 import mojave as m

    for fn, ofn in [(m.gen_compact, om.gen_compact), (m.gen_one_jet, om.gen_one_jet),
                    (m.gen_two_jet, om.gen_two_jet)]:
        for seed in range(10, 22):
            ref = ofn(np.random.default_rng(seed), size=size)
            out = fn(np.random.default_rng(seed), size=size)
            assert np.abs(out - ref).max() <= TOL * ref.max(), (fn.__name__, seed)


def test_create_mojave_bundle_matches_reference_golden(golden_dir):
    from radiosim_b200.mojave import create_mojave

    g = np.load(os.path.join(golden_dir, "mojave_64_bundle.npz"))
    conf = SimpleNamespace(
        general=SimpleNamespace(threads=None),
        dataset=SimpleNamespace(img_size=int(g["size"]), bundle_size=int(g["bundle_size"])),
        mojave=SimpleNamespace(class_ratio=list(g["class_ratio"])))
    today = datetime.date.fromisoformat(str(g["today"]))
    data, names = create_mojave(conf, np.random.default_rng(int(g["seed"])), today=today)
    d = dict(zip(names, data))
    assert np.array_equal(d["galaxy_classes"], g["galaxy_classes"])           # labels: exact
    assert np.array_equal(d["RA"], g["RA"]) and np.array_equal(d["DEC"], g["DEC"])
    assert [bytes(x) for x in d["date"]] == [bytes(x) for x in g["date"]]
    for i in range(len(g["galaxies"])):
        assert np.abs(d["galaxies"][i] - g["galaxies"][i]).max() <= TOL * g["galaxies"][i].max(), i
    with pytest.raises(ValueError):
        conf.general.threads = -2
        create_mojave(conf, np.random.default_rng(0))


# ---- the warps of the kernel against what is known of scikit-image from outside -------------
# (scikit-image exists neither in the build container nor on the GPU box; see
# tests/test_warps_cpu.py for the sources of these facts)
def _one_gauss_source(size, g, angle, jet=None, strength=0.0, radius=100.0):
    from radiosim_b200.mojave import mojave as mm

    s = mm._Source(size)
    if g is not None:
        s.gauss.append(list(g))
    if jet is not None:
        j = mm._base_jet(size, *jet)
        j.strength, j.radius = strength, radius
        s.jets = [j]
    s.angle = angle
    return s


@pytest.mark.parametrize("size", [64, 97])
def test_kernel_rotate_quarter_turns_equal_rot90(size):
    """skimage's own known answer (rotate(x, 90) == np.rot90(x)): direction and the
    (N/2 - 0.5) centre.  A quarter turn maps pixels onto pixels, so the match is to float32
    rounding of the same values."""
    from radiosim_b200.mojave.mojave import render_sources

    g = [1.0, size / 2 + 11.3, size / 2 - 6.7, 5.0, 2.5, 0.4]
    imgs = render_sources([_one_gauss_source(size, g, 90.0 * k) for k in range(4)], size)
    base = imgs[0]
    assert base.max() > 0.9
    for k in range(1, 4):
        assert np.abs(imgs[k] - np.rot90(base, k)).max() <= 1e-6, k


def test_kernel_swirl_identity_and_direction():
    from radiosim_b200.mojave.mojave import gen_jet, render_sources

    size = 128
    jet = (1.0, 4.0, 25.0, 4)   # amplitude, width, length, skew: extends to the right of the centre
    plain = gen_jet(size, *jet)
    # strength 0 -> the swirl is the identity (up to the bilinear taps at integer positions)
    same = render_sources([_one_gauss_source(size, None, 0.0, jet, 0.0, 120.0)], size)[0]
    assert np.abs(same - plain).max() <= 1e-6
    # reverse map theta' = theta + s*exp(-rho/r): the content moves by -s*exp(-rho/r) in atan2(row, col)
    s, radius = 0.5, 150.0
    sw = render_sources([_one_gauss_source(size, None, 0.0, jet, s, radius)], size)[0]
    yy, xx = np.mgrid[0:size, 0:size]
    sel = (xx - size / 2) > 20    # the part of the jet well away from the centre
    def ang(img):
        w = img * sel
        return np.arctan2((w * (yy - size / 2)).sum(), (w * (xx - size / 2)).sum())
    rho = 30.0
    expect = -s * np.exp(-rho / (radius / 5 * np.log(2)))
    assert ang(plain) == pytest.approx(0.0, abs=2e-2)
    assert ang(sw) < -0.05 and ang(sw) == pytest.approx(expect, abs=0.08)   # (a flux-weighted mean over rho)


def test_native_recipes_render_like_the_python_packer():
    """create_mojave through the native recipe generator (rsb_mojave_pack + rsb_mojave_blobs: profiles and
    shock amplitudes finished on the device) against the Python packer path on the same generator."""
    from radiosim_b200.mojave import create_mojave

    conf = SimpleNamespace(general=SimpleNamespace(threads=None),
                           dataset=SimpleNamespace(img_size=128, bundle_size=90),
                           mojave=SimpleNamespace(class_ratio=[1, 1, 1]))
    today = datetime.date(2026, 1, 1)
    a, names = create_mojave(conf, np.random.default_rng(77), today=today)
    os.environ["RSB_PY_PACKER"] = "1"
    try:
        b, _ = create_mojave(conf, np.random.default_rng(77), today=today)
    finally:
        del os.environ["RSB_PY_PACKER"]
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])
    for i in range(90):
        assert np.abs(a[0][i] - b[0][i]).max() <= TOL * b[0][i].max(), i


def test_config2_subset_matches_oracle():
    """BASELINE config 2 (mojave 256^2, class_ratio [1,1,1]): a 300-image bundle, 100 per class, through
    create_mojave (native recipes) against the oracle's create_mojave on the same generator: labels,
    coordinates and dates exact, images within 1e-5 of their peak."""
    from oracle import mojave as om
    from radiosim_b200.mojave import create_mojave

    conf = SimpleNamespace(general=SimpleNamespace(threads=None),
                           dataset=SimpleNamespace(img_size=256, bundle_size=300),
                           mojave=SimpleNamespace(class_ratio=[1, 1, 1]))
    today = datetime.date(2026, 1, 1)
    ref, names = om.create_mojave(256, 300, [1, 1, 1], np.random.default_rng(1337), today=today)
    got, gnames = create_mojave(conf, np.random.default_rng(1337), today=today)
    assert names == gnames
    assert np.bincount(ref[1]).tolist() == [100, 100, 100]
    for k in (1, 2, 3):
        assert np.array_equal(ref[k], got[k]), names[k]
    assert [bytes(x) for x in ref[4]] == [bytes(x) for x in got[4]]
    worst = 0.0
    for i in range(300):
        e = np.abs(got[0][i] - ref[0][i]).max() / ref[0][i].max()
        worst = max(worst, e)
        assert e <= TOL, (i, int(ref[1][i]), e)
    print(f"mojave 256^2, 300 images: worst error {worst:.2e} of the image peak")
