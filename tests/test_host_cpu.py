"""CPU tests: C-ABI exports, config, sharder, drivers (pixel work replaced by the oracle),
including the 2-rank gloo path."""

import ctypes
import os
import re
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "radiosim_b200.h")).read()
    declared = set(re.findall(r"\b(rsb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = ctypes.CDLL(os.path.join(ROOT, "radiosim_b200", "librsb200.so"))
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, missing
    from radiosim_b200 import _lib as L

    assert set(L.EXPORTS) <= declared
    assert lib.rsb_abi_version() == 1


def test_no_cpu_fallback_without_device():
    from radiosim_b200 import _lib as L

    if L.device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(L.RsbError, match="no CPU fallback"):
        L.require_device()


def test_config_loader_and_required_fields(tmp_path):
    from radiosim_b200.io import Config, ConfigError

    c = Config.from_toml(os.path.join(ROOT, "rc", "default_simulation.toml"))
    assert c.jet.num_jet_components == [3, 10] and c.general.threads is None
    assert c.dataset.img_size == 128 and c.survey.class_distribution == [2, 1, 2]
    with pytest.raises(ConfigError, match="field required"):
        Config.from_dict({"dataset": {"img_size": 128}})
    with pytest.raises(ConfigError):
        Config.from_dict({"dataset": dict(bundles_train=1, bundles_valid=1, bundles_test=1,
                                          bundle_size=2, img_size=128), "jet": {"training_type": "x"}})


def test_shard_ranges_partition():
    from radiosim_b200.sharder import rank_seed, shard_bundles, shard_name

    for total in (0, 1, 7, 20, 1000003):
        for world in (1, 2, 4, 8):
            parts = [shard_bundles(total, r, world) for r in range(world)]
            assert sum(len(p) for p in parts) == total
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
            flat = [i for p in parts for i in (p if total < 100 else [p.start, p.stop - 1] if len(p) else [])]
            assert flat == sorted(flat)
    assert rank_seed(1337, 3) == 1340 and rank_seed(None, 3) is None
    assert shard_name("jet", "train", 4, 0, 1) == "data_jet_train_4.h5"
    assert shard_name("jet", "train", 4, 1, 2) == "data_jet_train_r1_4.h5"


def _oracle_render(batch, normalize, planes=False, f64=True, planes_out=None):
    """Stand-in for the CUDA renderer in CPU tests: the oracle on the packed parameters."""
    from oracle import jet as ojet

    imgs, margins, peaks = [], [], []
    for i in range(batch.n):
        p = batch.params[i]
        planes_ = ojet.render_slots(batch.size, *[p[:, k] for k in range(6)])
        img = planes_.sum(axis=0)
        m, peak = 0, 1.0
        if batch.zoom_pad[i] >= 0:
            img, _, _, m = ojet.zoom_in(img, None)
            img, _ = ojet.zoom_away(img, int(batch.zoom_pad[i]))
        if normalize:
            peak = img.max()
            img = img / peak
        imgs.append(img); margins.append(m); peaks.append(peak)
    return np.array(imgs), np.array(margins, np.int32), np.array(peaks), None


def _worker(rank, world, port, outdir):
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    import radiosim_b200.jets.jet as jj
    from radiosim_b200.io import Config
    from radiosim_b200.simulations import simulate_sky_distributions

    jj.render_batch = _oracle_render
    conf = Config.from_dict({"general": {"seed": 11}, "paths": {"outpath": outdir},
                             "dataset": dict(bundles_train=3, bundles_valid=1, bundles_test=0,
                                             bundle_size=2, img_size=64, noise=False)})
    written = simulate_sky_distributions(conf, "jet")
    dist.barrier()  # plumbing only: nothing of the data path is communicated
    assert len(written) == (3 if rank == 0 else 1)
    dist.destroy_process_group()


def test_two_rank_gloo_shards(tmp_path):
    import torch.multiprocessing as mp

    from oracle import jet as ojet
    from radiosim_b200.h5min import read_h5

    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    names = sorted(os.listdir(tmp_path))
    assert names == ["data_jet_train_r0_0.h5", "data_jet_train_r0_1.h5", "data_jet_train_r1_2.h5",
                     "data_jet_valid_r0_0.h5"]
    # rank r's shards equal a single-process reference run seeded with seed + r
    for rank, files in [(0, ["data_jet_train_r0_0.h5", "data_jet_train_r0_1.h5", "data_jet_valid_r0_0.h5"]),
                        (1, ["data_jet_train_r1_2.h5"])]:
        np.random.seed(11 + rank)
        for f in files:
            x, y = ojet.create_jet((2, 3, 64, 64), (3, 10), "normalize", "list")
            d = read_h5(os.path.join(tmp_path, f))
            assert d["x"].shape == (2, 1, 64, 64) and d["y"].shape == (2, 19, 8)
            assert np.abs(d["x"] - x).max() <= 1e-12 and np.abs(d["y"] - y).max() <= 1e-12


def test_mojave_recipes_do_not_depend_on_worker_count():
    """create_mojave's host half (mojave.py:56-64): the spawned child generators are independent,
    so dealing them to worker processes must give the very same packed recipes."""
    from radiosim_b200.mojave.mojave import class_counts, make_recipes

    amount = class_counts([1, 1, 1], 96)
    a = make_recipes(np.random.default_rng(5), amount, 128, 1)
    b = make_recipes(np.random.default_rng(5), amount, 128, 3)
    assert len(a[0]) == 96
    for u, v in zip(a, b):
        assert u.dtype == v.dtype and u.tobytes() == v.tobytes()


def test_mojave_direct_draws_equal_scipy():
    """The recipe code draws expon / skewnorm variates without scipy's rvs wrapper (it costs more
    than the draw); the values and the stream position must be those of scipy's."""
    from scipy.stats import expon, skewnorm
    from radiosim_b200.mojave.mojave import _expon_rvs, _skewnorm_rvs

    r1, r2 = np.random.default_rng(7), np.random.default_rng(7)
    for _ in range(300):
        for a, loc, scale in ((4.37, 22.71, 16.85), (8.62, 69.60, 54.73), (8.41, 3.04, 3.46), (6.04, 3.02, 2.59)):
            assert skewnorm.rvs(a, loc, scale, random_state=r1) == _skewnorm_rvs(r2, a, loc, scale)
        assert expon.rvs(1.715e-4, 1.985e-2, random_state=r1) == _expon_rvs(r2, 1.715e-4, 1.985e-2)
    assert r1.random() == r2.random()


def test_small_helpers_of_the_reference_utils(tmp_path):
    """cart2pol / pol2cart, the VLBA position and date draws (utils.py:386-454), get_start_amp
    (jets/flux_scaling.py) and load_data (utils.py:457-470) through the built-in HDF5 reader."""
    import datetime

    from radiosim_b200 import utils as u
    from radiosim_b200.h5min import write_h5
    from radiosim_b200.jets import draw_from_mojave_dist, get_start_amp
    from radiosim_b200.utils.utils import _gen_date, _gen_vlba_obs_position, _save_mojave_bundle

    r, phi = u.cart2pol(3.0, -4.0)
    assert r == 5.0 and phi == np.arctan2(-4.0, 3.0)
    x, y = u.pol2cart(r, phi)
    assert abs(x - 3.0) < 1e-15 and abs(y + 4.0) < 1e-15

    a, b = np.random.default_rng(5), np.random.default_rng(5)
    ra, dec = _gen_vlba_obs_position(a, 7)
    assert np.array_equal(ra, b.uniform(0, 360, size=7)) and np.array_equal(dec, b.uniform(-30, 87, size=7))
    today = datetime.date(2024, 5, 17)
    d = _gen_date(a, "1995-01-01", 4, today=today)
    days = (np.datetime64(today) - np.datetime64("1995-01-01")).astype(int)
    want = (np.datetime64("1995-01-01") + b.integers(days, size=4)).astype("S10")
    assert d.dtype == np.dtype("S10") and np.array_equal(d, want)

    from scipy import stats
    np.random.seed(11)
    amp = get_start_amp("mojave")
    np.random.seed(11)
    ref = stats.beta(0.8639672251677816, 47.64189171625089, 0.09163404776954732, 1892.3881692069087).rvs(1) / 1e3
    assert np.array_equal(amp, ref) and draw_from_mojave_dist().shape == (1,)
    with pytest.raises(UnboundLocalError):
        get_start_amp("normalize")

    out = tmp_path / "build"
    out.mkdir()
    rng = np.random.default_rng(2)
    bundles = {}
    for split, count in (("train", 2), ("valid", 1)):
        for i in range(count):
            xs = rng.random((3, 1, 8, 8))
            bundles[(split, i)] = xs
            write_h5(out / f"data_jet_{split}_{i}.h5", {"x": xs, "y": rng.random((3, 19, 8))})
    _save_mojave_bundle(out / "other.h5", (np.arange(3.0),), ("galaxies",))
    conf = tmp_path / "c.toml"
    conf.write_text(f'[paths]\noutpath = "{out}"\n')
    # the reference's pattern "data*" + data_type is a REGULAR EXPRESSION ("dat", any number of "a", then the
    # split name), so it does not match the writer's own "data_jet_train_0.h5"; replicated as it is
    assert u.load_data(conf, "train", "x") == []
    for (split, i), xs in bundles.items():
        write_h5(out / f"data{split}_{i}.h5", {"x": xs})
    got = u.load_data(conf, "train", "x")
    assert len(got) == 6
    assert any(np.array_equal(got[0], w[0]) for w in (bundles[("train", 0)], bundles[("train", 1)]))
    assert len(u.load_data(conf, "valid", "x")) == 3
