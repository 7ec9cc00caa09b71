"""GPU tests of the bundle drivers: BASELINE config 1 (default TOML, jets 128^2) end to end."""

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_config1_jet_bundles_match_oracle(tmp_path):
    """rc/default_simulation.toml (seed 1337, 128^2, bundle 20, list targets), first bundles of
    every split: images <= 1e-5 of peak, list targets rtol 1e-6, written as .h5 x / y."""
    from oracle import jet as ojet
    from radiosim_b200.h5min import read_h5
    from radiosim_b200.io import Config
    from radiosim_b200.simulations import simulate_sky_distributions

    conf = Config.from_toml(os.path.join(ROOT, "rc", "default_simulation.toml"))
    conf.paths.outpath = tmp_path
    conf.dataset.bundles_train, conf.dataset.bundles_valid, conf.dataset.bundles_test = 3, 2, 1
    written = simulate_sky_distributions(conf, "jet", rank=0, world=1)
    assert [p.name for p in written] == ["data_jet_train_0.h5", "data_jet_train_1.h5", "data_jet_train_2.h5",
                                         "data_jet_valid_0.h5", "data_jet_valid_1.h5", "data_jet_test_0.h5"]
    np.random.seed(1337)
    for p in written:
        x, y = ojet.create_jet((20, 3, 128, 128), (3, 10), "normalize", "list")
        d = read_h5(p)
        assert d["x"].dtype == np.float64 and d["x"].shape == (20, 1, 128, 128) and d["y"].shape == (20, 19, 8)
        assert np.abs(d["x"] - x).max() <= 1e-5
        assert np.allclose(d["y"], y, rtol=1e-6, atol=1e-12)


def test_survey_mojave_and_ppdisk_drivers(tmp_path):
    from radiosim_b200.h5min import read_h5
    from radiosim_b200.io import Config
    from radiosim_b200.simulations import simulate_ppdisks, simulate_sky_distributions

    conf = Config.from_dict({"general": {"seed": 5}, "paths": {"outpath": str(tmp_path / "s")},
                             "dataset": dict(bundles_train=1, bundles_valid=0, bundles_test=1,
                                             bundle_size=3, img_size=128, noise=False),
                             "survey": {"num_sources": 10}})
    for mode, names in [("survey", {"x", "y"}), ("mojave", {"galaxies", "galaxy_classes", "RA", "DEC", "date"})]:
        written = simulate_sky_distributions(conf, mode, rank=0, world=1)
        assert len(written) == 2
        d = read_h5(written[0])
        assert set(d) == names
    assert read_h5(written[0])["galaxies"].shape == (3, 128, 128)
    pp = {"metadata": dict(img_size=128, alpha_range=[0, 180], ratio_range=[3, 15], size_ratio_range=[0.1, 1.0], seed=7),
          "dataset": dict(file_prefix="ppdisks", outpath=str(tmp_path / "p"), batch_size=2, batches=10,
                          batches_train_ratio=0.2, batches_valid_ratio=0.1, batches_test_ratio=0.1)}
    written = simulate_ppdisks(pp, rank=0, world=1)
    assert [p.name for p in written] == ["ppdisks_train_0.h5", "ppdisks_train_1.h5", "ppdisks_valid_0.h5", "ppdisks_test_0.h5"]
    d = read_h5(written[0])
    assert d["y"].shape == (2, 128, 128) and d["y"].dtype == np.float32
    assert d["params"].startswith("{'seed':") and "src_ra" in d["metadata"]
