"""CPU tests: the oracle (numpy restatement) against the reference's golden vectors.

The vectors under tests/golden/ were produced by oracle/make_golden.py from the
unmodified reference; ppdisks_128.h5 is the reference's own test fixture.
"""

import datetime
import os

import numpy as np
import pytest

from oracle import jet as ojet
from oracle import mojave as omoj
from oracle import ppdisk as opp
from oracle import survey as osur
from oracle.resize import resize_bilinear_aa, resize_linear


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


@pytest.mark.parametrize(
    "name", ["jet_128_list", "jet_256_list", "jet_64_gauss", "jet_100_clean_mojave"]
)
def test_jet_oracle_matches_reference(golden_dir, name):
    g = _load(golden_dir, name + ".npz")
    np.random.seed(int(g["seed"]))
    n, size = int(g["n"]), int(g["size"])
    x, y = ojet.create_jet(
        (n, 3, size, size), tuple(g["num_jet_components"]), str(g["scaling"]),
        str(g["training_type"]),
    )
    assert np.random.rand() == float(g["next_rand"])  # RNG stream stayed aligned
    assert x.shape == g["x"].shape and y.shape == g["y"].shape
    assert np.abs(x - g["x"]).max() <= 1e-12 * g["x"].max()
    assert np.array_equal(np.isnan(y), np.isnan(g["y"]))
    assert np.nanmax(np.abs(y - g["y"])) <= 1e-12 * np.nanmax(np.abs(g["y"]))


@pytest.mark.parametrize("name", ["survey_128", "survey_160"])
def test_survey_oracle_matches_reference(golden_dir, name):
    g = _load(golden_dir, name + ".npz")
    np.random.seed(int(g["seed"]))
    n, size = int(g["n"]), int(g["size"])
    x, y = osur.create_survey(
        (n, 3, size, size), int(g["num_sources"]), tuple(g["class_distribution"]),
        bool(g["scale_sources"]),
    )
    assert np.random.rand() == float(g["next_rand"])
    assert np.abs(x - g["x"]).max() <= 1e-12
    assert np.abs(y - g["y"]).max() <= 1e-12


def test_ppdisk_oracle_matches_reference_fixture(golden_dir):
    """The recipe of the reference's tests/test_ppdisks_simulation.py:1-34 at 128."""
    from radiosim_b200.h5min import read_h5

    ref = read_h5(os.path.join(golden_dir, "ppdisks_128.h5"))["y"]
    assert ref.shape == (9, 128, 128) and ref.dtype == np.float32
    imgs = []
    for size in (1, 2):
        for srr in [(0.1, 1), (0.1, 0.5), (0.5, 1)]:
            p, _ = opp.generate_proto_set(128, size, size_ratio_range=srr, seed=42)
            imgs.extend(p)
    imgs = np.array(imgs)
    assert imgs.dtype == np.float32
    assert np.allclose(ref, imgs, rtol=1e-5, atol=1e-8)  # the reference test's own bar
    assert np.abs(ref - imgs).max() <= 1e-6 * ref.max()


def test_ppdisk_oracle_512(golden_dir):
    g = _load(golden_dir, "ppdisks_512_first3.npz")
    for i, srr in enumerate(g["size_ratio_ranges"]):
        p, _ = opp.generate_proto_set(512, 1, size_ratio_range=tuple(srr), seed=int(g["seed"]))
        assert np.allclose(g["y"][i], p[0], rtol=1e-5, atol=1e-8)


def test_mojave_oracle_matches_reference_with_restated_warps(golden_dir):
    g = _load(golden_dir, "mojave_128_classes.npz")
    size = int(g["size"])
    for key, fn in [("compact", omoj.gen_compact), ("one_jet", omoj.gen_one_jet),
                    ("two_jet", omoj.gen_two_jet)]:
        for k, s in enumerate(g["seeds"]):
            out = fn(np.random.default_rng(int(s)), size=size)
            assert np.abs(out - g[key][k]).max() <= 1e-12 * g[key][k].max()


def test_mojave_bundle_labels(golden_dir):
    g = _load(golden_dir, "mojave_64_bundle.npz")
    today = datetime.date.fromisoformat(str(g["today"]))
    data, names = omoj.create_mojave(
        int(g["size"]), int(g["bundle_size"]), list(g["class_ratio"]),
        np.random.default_rng(int(g["seed"])), today=today,
    )
    d = dict(zip(names, data))
    assert np.array_equal(d["galaxy_classes"], g["galaxy_classes"])
    assert np.array_equal(d["RA"], g["RA"]) and np.array_equal(d["DEC"], g["DEC"])
    assert [bytes(x) for x in d["date"]] == [bytes(x) for x in g["date"]]
    assert np.abs(d["galaxies"] - g["galaxies"]).max() <= 1e-12 * g["galaxies"].max()


def test_resize_linear_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    for S, D in [(100, 128), (128, 128), (37, 256), (300, 128), (511, 256), (129, 128)]:
        a = rng.random((S, S))
        ref = cv2.resize(a, dsize=(D, D), interpolation=cv2.INTER_LINEAR)
        assert np.abs(ref - resize_linear(a, D)).max() <= 1e-14


def test_resize_aa_matches_torch():
    torch = pytest.importorskip("torch")
    import torch.nn.functional as F

    rng = np.random.default_rng(1)
    for S, D in [(651, 102), (322, 234), (900, 504), (242, 400), (65, 64), (64, 64)]:
        a = rng.random((S, S))
        ref = F.interpolate(
            torch.from_numpy(a)[None, None], size=(D, D), mode="bilinear",
            antialias=True, align_corners=False,
        )[0, 0].numpy()
        assert np.abs(ref - resize_bilinear_aa(a, D)).max() <= 1e-13


def test_ellipse_mask_matches_torch_semantics():
    """The float32 promotion of ppdisks.py:225-234, replayed with torch itself."""
    torch = pytest.importorskip("torch")
    xx, yy = torch.from_numpy(np.mgrid[:900, :900])
    for a, ratio, alpha in [(7.0, 3 / 14, 17), (44, 14 / 3, 133), (101, 1.0, 90), (12.0, 2.5, 0)]:
        a = np.float64(a) if isinstance(a, float) else a
        b = a / ratio
        t1 = ((xx - 450) * np.cos(np.deg2rad(alpha)) + (yy - 450) * np.sin(np.deg2rad(alpha))) ** 2
        t2 = ((xx - 450) * np.sin(np.deg2rad(alpha)) - (yy - 450) * np.cos(np.deg2rad(alpha))) ** 2
        ref = ((t1 / a**2) + (t2 / b**2)) <= 1
        assert np.array_equal(ref.numpy(), opp.ellipse_mask(a, b, alpha))


@pytest.mark.parametrize("name", ["noise_64", "noise_100"])
def test_add_noise_oracle_matches_reference_golden(golden_dir, name):
    """oracle.noise.add_noise against the unmodified reference's add_noise (utils/utils.py:259-315)."""
    from oracle import noise as on

    g = np.load(os.path.join(golden_dir, name + ".npz"))
    np.random.seed(int(g["seed"]))
    out = on.add_noise(g["image"], list(g["level"]))
    assert np.abs(out - g["out"]).max() <= 1e-12
    assert np.random.rand() == float(g["next_rand"])
