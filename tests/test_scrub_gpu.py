"""No kernel may depend on what its predecessor left in shared memory.

Every drop-in is run three times on the same inputs, each time behind ``rsb_debug_scrub_shared``
with a different bit pattern (zeros, quiet NaN / -1, +infinity): the results must be finite and
bit-identical.  (Found with it: zero-weight resampling taps of the jet kernel that touched tiles
the renderer skips -- 0 x NaN.)
"""

from types import SimpleNamespace

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
PATTERNS = (0x0, 0xFFFFFFFF, 0x7F800000)


def _scrub(pattern):
    from radiosim_b200 import _lib as L

    L.require_device()
    L.check(L.lib().rsb_debug_scrub_shared(pattern, None), "rsb_debug_scrub_shared")
    L.device_sync()


def _same(runs, nan_by_design=False):
    first = runs[0]
    if not nan_by_design:
        for arrs in runs:
            for a in arrs:
                assert np.isfinite(a).all()
    for arrs in runs[1:]:
        for a, b in zip(first, arrs):
            assert np.array_equal(a, b, equal_nan=nan_by_design)


@pytest.mark.parametrize("size,ns,dist", [(150, 25, (1, 1, 1)), (1024, 400, (2, 1, 2)), (1600, 60, (2, 1, 2))])
def test_survey(size, ns, dist):
    from radiosim_b200.survey import create_survey

    conf = SimpleNamespace(jet=SimpleNamespace(training_type="list", scaling="normalize", num_jet_components=[3, 10]),
                           survey=SimpleNamespace(num_sources=ns, class_distribution=list(dist), scale_sources=True))
    runs = []
    for p in PATTERNS:
        _scrub(p)
        np.random.seed(21)
        x, y = create_survey(np.broadcast_to(np.zeros(1), (2, 3, size, size)), conf)
        runs.append((np.array(x), np.array(y)))
    _same(runs)


@pytest.mark.parametrize("training_type", ["list", "gauss", "clean"])
def test_create_jet(training_type):
    from radiosim_b200.jets import create_jet

    conf = SimpleNamespace(jet=SimpleNamespace(training_type=training_type, num_jet_components=[3, 10],
                                               scaling="normalize"))
    runs = []
    for p in PATTERNS:
        _scrub(p)
        np.random.seed(22)
        x, y = create_jet(np.broadcast_to(np.zeros(1), (40, 3, 96, 96)), conf)
        runs.append((np.array(x), np.array(y)))
        assert np.isfinite(x).all()
    _same(runs, nan_by_design=(training_type == "gauss"))   # (the list plane is NaN-padded, jet.py:217-221)


def test_mojave():
    from radiosim_b200 import mojave as m

    runs = []
    for p in PATTERNS:
        _scrub(p)
        runs.append(tuple(fn(np.random.default_rng(5), size=200) for fn in (m.gen_compact, m.gen_one_jet, m.gen_two_jet)))
    _same(runs)


def test_ppdisks():
    from radiosim_b200.ppdisks import generate_proto_set

    runs = []
    for p in PATTERNS:
        _scrub(p)
        protos, _ = generate_proto_set(256, 6, seed=11)
        runs.append((np.array(protos),))
    _same(runs)


def test_add_noise():
    from radiosim_b200.utils import add_noise

    img = np.random.default_rng(3).random((3, 1, 200, 200)) ** 8
    runs = []
    for p in PATTERNS:
        _scrub(p)
        np.random.seed(23)
        runs.append((np.array(add_noise(img, (0, 15))),))
    _same(runs)
