"""GPU tests of the bundle drivers: BASELINE config 1 (default TOML, jets 128^2) end to end."""

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_config1_jet_bundles_match_oracle(tmp_path):
    """rc/default_simulation.toml as shipped (seed 1337, 128^2, bundle 20, list targets): ALL 40
    bundles (20 train / 10 valid / 10 test = 800 images, BASELINE config 1) against the oracle:
    images <= 1e-5 of peak, list targets rtol 1e-6, written as .h5 x / y."""
    from oracle import jet as ojet
    from radiosim_b200.h5min import read_h5
    from radiosim_b200.io import Config
    from radiosim_b200.simulations import simulate_sky_distributions

    conf = Config.from_toml(os.path.join(ROOT, "rc", "default_simulation.toml"))
    conf.paths.outpath = tmp_path
    assert (conf.dataset.bundles_train, conf.dataset.bundles_valid, conf.dataset.bundles_test) == (20, 10, 10)
    written = simulate_sky_distributions(conf, "jet", rank=0, world=1)
    assert [p.name for p in written] == ([f"data_jet_train_{i}.h5" for i in range(20)]
                                         + [f"data_jet_valid_{i}.h5" for i in range(10)]
                                         + [f"data_jet_test_{i}.h5" for i in range(10)])
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


def test_config3_default_ppdisk_toml_matches_oracle(tmp_path):
    """BASELINE config 3: rc/default_ppdisks_simulation.toml as shipped (512^2, seed 1337, batch 5,
    10 batches split 7/2/1): the first two training shards (10 images) against the float64 oracle,
    with the per-shard seeds the driver drew (they are recorded in the shard's ``params``)."""
    import ast
    import tomllib

    from oracle import ppdisk as opp
    from radiosim_b200.h5min import read_h5
    from radiosim_b200.simulations import simulate_ppdisks

    with open(os.path.join(ROOT, "rc", "default_ppdisks_simulation.toml"), "rb") as f:
        cfg = tomllib.load(f)
    cfg["dataset"]["outpath"] = str(tmp_path)
    written = simulate_ppdisks(cfg, rank=0, world=1)
    assert [p.name for p in written] == ([f"ppdisks_train_{i}.h5" for i in range(7)]
                                         + ["ppdisks_valid_0.h5", "ppdisks_valid_1.h5", "ppdisks_test_0.h5"])
    meta = cfg["metadata"]
    for p in written[:2]:
        d = read_h5(p)
        params = ast.literal_eval(d["params"])
        assert d["y"].shape == (5, 512, 512) and d["y"].dtype == np.float32
        ref, _ = opp.generate_proto_set(512, 5, meta["alpha_range"], meta["ratio_range"],
                                        meta["size_ratio_range"], seed=params["seed"])
        ref = np.array(ref)
        assert np.allclose(ref, d["y"], rtol=1e-5, atol=1e-8)     # the reference test's criterion
        assert np.abs(ref - d["y"]).max() <= 1e-5 * ref.max()     # 1e-5 of the peak
        assert d["metadata"].count("'index'") == 5
