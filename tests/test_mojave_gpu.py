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


@pytest.mark.parametrize("size", [96, 256, 1024])   # 1024: the reference's default size (mojave.py:265,402,496)
def test_generators_match_oracle_many_seeds(size):
    from oracle import mojave as om
    from radiosim_b200 import mojave as m

    for fn, ofn in [(m.gen_compact, om.gen_compact), (m.gen_one_jet, om.gen_one_jet),
                    (m.gen_two_jet, om.gen_two_jet)]:
        for seed in range(10, 22 if size < 1024 else 13):
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


def test_gen_jet_and_gen_shocks_standalone_match_oracle():
    """``gen_jet`` / ``gen_shocks`` with the reference's signatures on host images
    (mojave.py:84-219): same draws, images within 1e-5 of the jet's peak."""
    from oracle import mojave as om
    from radiosim_b200 import mojave as m

    for seed, size, comp in ((3, 256, 5), (4, 200, 8), (5, 128, 3), (6, 96, 6)):
        amp, width, length, a, sx = 3.0, 6.0 + seed, size / 5, 4.0, 5.0
        ref = om.base_jet(size, amp, width, length, a)
        jet = m.gen_jet(size, amp, width, length, a)
        assert jet.shape == ref.shape and jet.dtype == np.float64
        assert np.abs(jet - ref).max() <= TOL * ref.max()
        r_o, r_g = np.random.default_rng(seed), np.random.default_rng(seed)
        ref2 = om.add_shocks(ref.copy(), comp, length, width, sx, r_o)
        out, info, skipped = m.gen_shocks(ref.copy(), comp, length, width, sx, r_g)
        assert r_o.random() == r_g.random()                       # same number of draws
        assert np.abs(out - ref2).max() <= TOL * ref2.max()
        assert len(info) == 9 and all(len(col) == len(info[0]) for col in info)
        assert 0 <= skipped <= comp and len(info[0]) + skipped <= comp
        assert all(o == "r" for o in info[0]) and all(rot == 0 for rot in info[6])


def test_gen_shocks_out_of_range_position_draws_nothing_more():
    """A shock position outside the image is passed over before its amplitude factor is drawn
    (the reference's ``IndexError`` handler, mojave.py:189-192)."""
    from oracle import mojave as om
    from radiosim_b200 import mojave as m

    size = 64
    jet = om.base_jet(size, 2.0, 4.0, 20.0, 3.0)
    for seed in range(4):
        r_o, r_g = np.random.default_rng(seed), np.random.default_rng(seed)
        ref = om.add_shocks(jet.copy(), 4, 200.0, 5.0, 3.0, r_o)      # step ~ 100 px: beyond the image
        out, info, skipped = m.gen_shocks(jet.copy(), 4, 200.0, 5.0, 3.0, r_g)
        assert r_o.random() == r_g.random()
        assert np.abs(out - ref).max() <= TOL * ref.max()
        assert len(info[0]) == 0


@pytest.mark.parametrize("shape", [(256, 256), (128, 200), (97, 61)])
def test_swirl_image_matches_oracle_and_known_answers(shape):
    """``rsb_swirl`` against the restated scikit-image swirl (PARITY UNPINNED like the fused path)
    plus what needs no pin: strength 0 is the identity, and the result stays inside the input's range."""
    from oracle import warps
    from radiosim_b200.mojave import swirl_image

    rng = np.random.default_rng(8)
    img = rng.random(shape) ** 3
    rows, cols = shape
    for center, strength, radius in (((cols / 2, rows / 2), 0.3, 120.0), ((cols / 3, rows / 2 + 5.5), -0.8, 60.0),
                                     ((cols / 2, rows / 2), 2.5, 200.0)):
        ref = warps.swirl(img, center=np.array(center), strength=strength, radius=radius)
        out = swirl_image(img, center, strength, radius)
        assert out.shape == img.shape and out.dtype == np.float64
        assert np.abs(out - ref).max() <= 1e-12
        assert out.min() >= img.min() and out.max() <= img.max()
    same = swirl_image(img, (cols / 2, rows / 2), 0.0, 100.0)
    assert np.abs(same - img).max() <= 1e-15


def test_add_swirl_standalone_matches_oracle():
    """``add_swirl`` (mojave.py:222-261): first jet draws strength and radius, the second jet varies the
    first one's by ~10 %; parameters equal, images within 1e-12."""
    from oracle import mojave as om
    from radiosim_b200 import mojave as m

    size = 192
    jet = om.base_jet(size, 2.0, 7.0, 40.0, 4.0)
    r_o, r_g = np.random.default_rng(12), np.random.default_rng(12)
    ref1, p_o = om.swirl_jet(jet, r_o)
    out1, p_g = m.add_swirl(jet, r_g)
    assert p_o[0] == p_g[0] and p_o[1] == p_g[1] and np.array_equal(p_o[2], p_g[2])
    assert np.abs(out1 - ref1).max() <= 1e-12
    ref2, q_o = om.swirl_jet(jet[:, ::-1].copy(), r_o, first=p_o)
    out2, q_g = m.add_swirl(jet[:, ::-1].copy(), r_g, first_jet_params=p_g)
    assert q_o[0] == q_g[0] and q_o[1] == q_g[1]
    assert np.abs(out2 - ref2).max() <= 1e-12
    assert r_o.random() == r_g.random()
