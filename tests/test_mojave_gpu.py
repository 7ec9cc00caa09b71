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
