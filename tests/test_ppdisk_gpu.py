"""GPU parity tests of the ppdisk path against the reference's own fixture and the oracle."""

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _recipe(gen, img_size):
    """The recipe of the reference's tests/test_ppdisks_simulation.py:1-34."""
    imgs = []
    for size in (1, 2):
        for srr in [(0.1, 1), (0.1, 0.5), (0.5, 1)]:
            p, _ = gen(img_size=img_size, size=size, size_ratio_range=srr, seed=42)
            imgs.extend(p)
    return np.array(imgs)


def test_generate_proto_set_matches_reference_fixture(golden_dir):
    from radiosim_b200.h5min import read_h5
    from radiosim_b200.ppdisks import generate_proto_set

    ref = read_h5(os.path.join(golden_dir, "ppdisks_128.h5"))["y"]
    imgs = _recipe(generate_proto_set, 128)
    assert imgs.dtype == np.float32 and imgs.shape == ref.shape
    assert np.allclose(ref, imgs, rtol=1e-5, atol=1e-8)   # the reference test's own criterion
    assert np.abs(ref - imgs).max() <= 1e-5 * ref.max()   # and ours: 1e-5 of the peak


def test_generate_proto_set_512(golden_dir):
    from radiosim_b200.ppdisks import generate_proto_set

    g = np.load(os.path.join(golden_dir, "ppdisks_512_first3.npz"))
    for i, srr in enumerate(g["size_ratio_ranges"]):
        p, _ = generate_proto_set(512, 1, size_ratio_range=tuple(srr), seed=int(g["seed"]))
        assert np.allclose(g["y"][i], p[0], rtol=1e-5, atol=1e-8), i


def test_masks_bit_exact():
    """Masks are integer results: bit-exact against the oracle's float32 replay of torch."""
    from oracle import ppdisk as opp
    from radiosim_b200.ppdisks.ppdisks import COMP, ellipse_masks

    rows, ref = [], []
    for a, ratio, alpha, ex in [(7.0, 3 / 14, 17, 0), (44, 14 / 3, 133, 6), (101, 1.0, 90, 9),
                                (12.0, 2.5, 0, 0), (130, 0.25, 45, 8)]:
        ca = np.float32(np.cos(np.deg2rad(alpha)))
        sa = np.float32(np.sin(np.deg2rad(alpha)))
        r = np.zeros(1, dtype=COMP)[0]
        if ex:
            r["a2_in"], r["b2_in"] = np.float32(a**2), np.float32((a / ratio) ** 2)
            r["a2_out"], r["b2_out"] = np.float32((a + ex) ** 2), np.float32(((a + ex) / ratio) ** 2)
            ref.append(opp.ellipse_mask(a + ex, (a + ex) / ratio, alpha) ^ opp.ellipse_mask(a, a / ratio, alpha))
        else:
            r["a2_in"], r["b2_in"] = -1, 1
            r["a2_out"], r["b2_out"] = np.float32(a**2), np.float32((a / ratio) ** 2)
            ref.append(opp.ellipse_mask(a, a / ratio, alpha))
        r["ca"], r["sa"], r["sigma"] = ca, sa, 2
        rows.append(r)
    masks = ellipse_masks(np.array(rows, dtype=COMP))
    for k in range(len(rows)):
        assert np.array_equal(masks[k].astype(bool), ref[k]), k


def test_create_proto_corner_grid_matches_oracle():
    """The corner grid of tests/test_ppdisks_simulation.py:37-55, checked (the reference only
    smoke-runs it)."""
    from oracle import ppdisk as opp
    from radiosim_b200.ppdisks import create_proto

    k = 0
    for img_size in (128, 256, 512):
        for alpha in (0.0, 180.0):
            for ratio in (3, 15):
                for size_ratio in (0.1, 1.0):
                    k += 1
                    o = opp.create_proto(img_size, alpha, ratio, size_ratio, seed=100 + k)
                    g = np.asarray(create_proto(img_size, alpha, ratio, size_ratio, seed=100 + k))
                    assert np.allclose(o, g, rtol=1e-5, atol=1e-8), (img_size, alpha, ratio, size_ratio)


def test_create_proto_no_seed_runs_like_the_reference_test():
    """tests/test_ppdisks_simulation.py:37-55 as the reference runs it (``seed=None``: the global stream
    is used as it stands): every corner renders, finite, non-negative, on the 1e-10 background."""
    from radiosim_b200.ppdisks import create_proto

    np.random.seed(99)
    for img_size in (128, 512):
        for alpha in (0.0, 180.0):
            for ratio in (3, 15):
                for size_ratio in (0.1, 1.0):
                    before = np.random.get_state()[2]
                    g = np.asarray(create_proto(img_size=img_size, alpha=alpha, ratio=ratio, size_ratio=size_ratio,
                                                seed=None))
                    assert g.shape == (img_size, img_size) and g.dtype == np.float32
                    assert np.isfinite(g).all() and g.min() >= 1e-10 * (1 - 1e-6) and g.max() > 1e-10
                    assert np.random.get_state()[2] != before       # (draws came from the global stream)
