"""Generate ``tests/golden/*.npz`` from the UNMODIFIED reference.

TEST INFRASTRUCTURE ONLY.  Runs only in the build container (needs
``/root/reference``); the GPU box uses the committed outputs.

    python -m oracle.make_golden

Each fixture stores the exact call (seed, sizes, config) next to the reference's
float64 outputs, so both the oracle (``tests/test_oracle.py``) and the CUDA path
(``tests/test_*_gpu.py``) can replay it.
"""

from __future__ import annotations

import datetime as _dt
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden")
FROZEN_TODAY = _dt.date(2025, 1, 1)


def _conf(ref, training_type="list", scaling="normalize", num_sources=20):
    from radiosim.io import Config

    conf = Config.from_toml("/root/reference/rc/default_simulation.toml")
    conf.jet.training_type = training_type
    conf.jet.scaling = scaling
    conf.survey.num_sources = num_sources
    return conf


def jets(ref):
    from radiosim.jets import create_jet

    cases = [
        ("jet_128_list", 1337, 8, 128, "list", "normalize"),
        ("jet_256_list", 11, 2, 256, "list", "normalize"),
        ("jet_64_gauss", 5, 2, 64, "gauss", "normalize"),
        ("jet_100_clean_mojave", 9, 3, 100, "clean", "mojave"),
    ]
    for name, seed, n, size, tt, sc in cases:
        np.random.seed(seed)
        x, y = create_jet(np.zeros((n, 3, size, size)), _conf(ref, tt, sc))
        nxt = np.random.rand()
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"), x=x, y=y, seed=seed, n=n, size=size,
            training_type=tt, scaling=sc, num_jet_components=np.array([3, 10]),
            next_rand=nxt,
        )
        print(name, x.shape, y.shape)


def surveys(ref):
    from radiosim.survey import create_survey

    for name, seed, n, size, ns in [("survey_128", 3, 1, 128, 20), ("survey_160", 4, 1, 160, 12)]:
        np.random.seed(seed)
        x, y = create_survey(np.zeros((n, 3, size, size)), _conf(ref, num_sources=ns))
        nxt = np.random.rand()
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"), x=x, y=y, seed=seed, n=n, size=size,
            num_sources=ns, class_distribution=np.array([2, 1, 2]), scale_sources=True,
            next_rand=nxt,
        )
        print(name, x.shape, y.shape)


def ppdisks(ref):
    from radiosim.ppdisks import generate_proto_set

    shutil.copyfile(
        "/root/reference/tests/ppdisks_128.h5", os.path.join(OUT, "ppdisks_128.h5")
    )
    # the reference's 512 fixture is a missing large blob: regenerate the first
    # three images of its recipe (size=1, the three size_ratio ranges, seed 42)
    imgs = []
    for srr in [(0.1, 1), (0.1, 0.5), (0.5, 1)]:
        p, _ = generate_proto_set(img_size=512, size=1, size_ratio_range=srr, seed=42)
        imgs.extend(p)
    np.savez_compressed(
        os.path.join(OUT, "ppdisks_512_first3.npz"), y=np.array(imgs), seed=42,
        size_ratio_ranges=np.array([(0.1, 1), (0.1, 0.5), (0.5, 1)]),
    )
    print("ppdisks_512_first3", np.array(imgs).shape)


def mojave(ref):
    import radiosim.utils.utils as ru
    from radiosim.mojave import create_mojave, gen_compact, gen_one_jet, gen_two_jet

    size = 128
    out = {}
    for cname, fn in [("compact", gen_compact), ("one_jet", gen_one_jet), ("two_jet", gen_two_jet)]:
        out[cname] = np.array([fn(np.random.default_rng(s), size=size) for s in (0, 1, 2)])
    np.savez_compressed(os.path.join(OUT, "mojave_128_classes.npz"), size=size,
                        seeds=np.array([0, 1, 2]), **out)

    # full bundle with a frozen "today" (utils.py:402 reads datetime.now())
    class _FrozenDT(_dt.datetime):
        @classmethod
        def now(cls, tz=None):
            return cls(FROZEN_TODAY.year, FROZEN_TODAY.month, FROZEN_TODAY.day)

    import h5py

    h5py.string_dtype = lambda enc, n: f"S{n}"
    real = _dt.datetime
    _dt.datetime = _FrozenDT
    try:
        conf = _conf(ref)
        conf.dataset.img_size = 64
        conf.dataset.bundle_size = 7
        conf.mojave.class_ratio = [1, 2, 1]
        conf.general.threads = None
        data, names = create_mojave(conf, np.random.default_rng(1337))
    finally:
        _dt.datetime = real
    np.savez_compressed(
        os.path.join(OUT, "mojave_64_bundle.npz"), seed=1337, size=64, bundle_size=7,
        class_ratio=np.array([1, 2, 1]), today=str(FROZEN_TODAY),
        **{n: d for n, d in zip(names, data)},
    )
    print("mojave ok", [d.shape for d in data])


def noise(ref):
    """add_noise (utils/utils.py:259-315) on small bundles: input images, result, RNG position."""
    from radiosim.utils.utils import add_noise

    for name, seed, n, size, level in [("noise_64", 11, 3, 64, [0, 15]), ("noise_100", 12, 2, 100, [5, 10])]:
        rng = np.random.default_rng(seed)            # (input images only: any positive pattern will do)
        yy, xx = np.mgrid[0:size, 0:size]
        img = np.stack([np.exp(-((yy - rng.uniform(20, size - 20)) ** 2 + (xx - rng.uniform(20, size - 20)) ** 2)
                               / (2 * rng.uniform(3, 9) ** 2)) * rng.uniform(0.01, 1.0) for _ in range(n)])[:, None]
        np.random.seed(seed)
        out = add_noise(img, level)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), seed=seed, n=n, size=size, level=np.array(level),
                            image=img, out=out, next_rand=np.random.rand())
        print(name, "ok", out.shape, float(np.abs(out - img).max()))


def main():
    sys.path.insert(0, ROOT)
    from oracle.refstub import load_reference

    ref = load_reference()
    os.makedirs(OUT, exist_ok=True)
    jets(ref)
    surveys(ref)
    ppdisks(ref)
    mojave(ref)
    noise(ref)


if __name__ == "__main__":
    main()
