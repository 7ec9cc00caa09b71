"""Bundle drivers (reference ``src/radiosim/simulations/simulations.py``,
``src/radiosim/simulations/ppdisks.py``): loop over train/valid/test bundles, call the
generator, write ``.h5`` shards with the reference's dataset names."""

from __future__ import annotations

from pathlib import Path

import numpy as np

from ..h5min import write_h5
from ..sharder import rank_info, rank_seed, shard_bundles, shard_name

__all__ = ["create_sky_distribution", "simulate_ppdisks", "simulate_sky_distributions"]

# resolution and pointings of the DSHARP observations the reference attaches as metadata
# (src/radiosim/simulations/ppdisks.py:33-56)
_DSHARP_CELL = 0.00299999999999988
_DSHARP_POINTINGS = [
    (252.1901583333, -14.27663888889), (242.8806334167, -18.64062026556),
    (246.6002916667, -24.27041666667), (269.088663625, -21.9562676325),
    (236.6862890833, -34.51002105694), (246.912515125, -23.97199338639),
    (246.6875916667, -24.38563138889), (239.1667003167, -22.02789646667),
    (246.484, -24.34672222222), (246.5986583334, -24.72051666667),
    (246.578625, -24.47213611111), (246.9142916667, -24.65431666667),
    (239.653748625, -22.95433409167), (252.313727625, -14.36918004583),
    (239.0382866667, -37.93514649278), (236.3035304583, -34.29194894083),
    (239.1762312125, -37.82110333889), (239.8185583333, -41.95297045306),
    (242.2576419583, -39.08689564417), (240.1854510899, -41.92538824062),
]


class _ShardWriter:
    """Writes shards on a few background threads while the next bundle is being generated (the
    file write releases the GIL).  At most ``depth`` bundles are in flight, so their page-locked
    blocks return to the pool at the rate they are taken."""

    def __init__(self, workers=3, depth=3):
        from concurrent.futures import ThreadPoolExecutor

        self.pool = ThreadPoolExecutor(workers)
        self.depth = depth
        self.pending = []

    def submit(self, path, datasets):
        while len(self.pending) >= self.depth:
            self.pending.pop(0).result()
        self.pending.append(self.pool.submit(write_h5, path, datasets))

    def close(self):
        try:
            for f in self.pending:
                f.result()  # re-raises a writer's exception
        finally:
            self.pending = []
            self.pool.shutdown()


def _noised_sky(sky, conf):
    """jet / survey bundles with ``dataset.noise`` (``simulations.py:119-125``).  The reference's
    condition ``noise_level > 0`` compares a list with an int and raises ``TypeError`` on its HEAD;
    its intent -- add noise when a level above zero is configured, then bring every image back to
    [0, 1] -- is what runs here."""
    level = conf.dataset.noise_level
    if not conf.dataset.noise or not max(level) > 0:
        return sky
    from ..utils import add_noise

    out = add_noise(sky, level)
    for img in out:
        img -= img.min()
        img /= img.max()
    return out


def _bundles_per_call(conf, mode, remaining):
    """How many bundles one generator call renders.  The jet / survey generators consume the legacy
    RNG stream image after image, so one call for k * bundle_size images gives exactly the images of k
    calls for bundle_size -- small bundles (the default TOML: 20 images of 128^2) are rendered several
    at a time and cut into their shards afterwards.  Not with noise: the reference draws the noise of a
    bundle before the next bundle's parameters."""
    if mode not in ("jet", "survey") or (conf.dataset.noise and max(conf.dataset.noise_level) > 0):
        return 1
    n, b = conf.dataset.img_size, conf.dataset.bundle_size
    planes = 1 + (5 if mode == "survey" else 22 if conf.jet.training_type == "gauss" else
                  1 if conf.jet.training_type == "clean" else 0)
    per_bundle = b * n * n * 8 * planes
    return int(max(1, min(remaining, (256 << 20) // max(per_bundle, 1), 4096 // max(b, 1))))


def create_sky_distribution(conf, bundles, mode, opt, rng=None, rank=0, world=1):
    """One split: the bundles in ``bundles`` (an iterable of indices) -> one shard each."""
    from ..utils import create_grid

    out = Path(conf.paths.outpath)
    out.mkdir(parents=True, exist_ok=True)
    written = []
    sky = target = data = None
    bundles = list(bundles)
    b = conf.dataset.bundle_size
    # small shards (the default TOML: 2.6 MB each) are bound by per-file latency, not by bandwidth: more
    # writer threads; large ones keep three, so that at most three bundles' page-locked blocks are held
    shard_bytes = b * conf.dataset.img_size**2 * 8 * 2
    lanes = int(max(3, min(8, (64 << 20) // max(shard_bytes, 1))))
    writer = _ShardWriter(workers=lanes, depth=lanes)
    try:
        pos = 0
        while pos < len(bundles):
            k = _bundles_per_call(conf, mode, len(bundles) - pos)
            paths = [out / shard_name(mode, opt, i, rank, world) for i in bundles[pos : pos + k]]
            grid = create_grid(conf.dataset.img_size, b * k)
            if mode in ("jet", "survey"):
                if mode == "jet":
                    from ..jets import create_jet as generate
                else:
                    from ..survey import create_survey as generate
                sky, target = generate(grid, conf)
                for q, path in enumerate(paths):  # (views of the same page-locked blocks)
                    writer.submit(path, {"x": _noised_sky(sky[q * b : (q + 1) * b], conf),
                                         "y": target[q * b : (q + 1) * b]})
            elif mode == "mojave":
                from ..mojave import create_mojave

                data, names = create_mojave(conf, rng)
                if conf.dataset.noise:  # simulations.py:101-108
                    from ..utils import add_noise

                    data[0] = add_noise(data[0][:, None], conf.dataset.noise_level)[:, 0]
                writer.submit(paths[0], dict(zip(names, data)))
            else:
                raise ValueError(f"Mode {mode} not found. Choose 'survey', 'jet' or 'mojave' in config file")
            sky = target = data = None  # (the writer holds the only references now)
            written += paths
            pos += k
    finally:
        writer.close()
    return written


def simulate_sky_distributions(conf, mode, rank=None, world=None):
    """``simulate_sky_distributions`` (``simulations.py:46-79``), sharded over ranks."""
    if rank is None or world is None:
        rank, world, _ = rank_info()
    seed = rank_seed(conf.general.seed, rank)
    if seed is not None:
        np.random.seed(seed)  # the legacy global stream, as tools/simulation_cli.py:38-40 does
    rng = np.random.default_rng(seed)
    written = []
    for opt, total in (("train", conf.dataset.bundles_train), ("valid", conf.dataset.bundles_valid),
                       ("test", conf.dataset.bundles_test)):
        written += create_sky_distribution(conf, shard_bundles(total, rank, world), mode, opt, rng, rank, world)
    return written


def simulate_ppdisks(config, rank=None, world=None):
    """``simulate_ppdisks`` re-created from its intent (it raises ``AttributeError`` on the
    reference's HEAD): accepts the flat ``[metadata]/[simulation]/[dataset]`` layout as a dict."""
    from ..ppdisks import generate_proto_set

    if rank is None or world is None:
        rank, world, _ = rank_info()
    meta, ds = config["metadata"], config["dataset"]
    rng = np.random.default_rng(seed=rank_seed(meta.get("seed", 1337), rank))
    out = Path(ds["outpath"])
    out.mkdir(parents=True, exist_ok=True)
    n, bs = ds["batches"], ds["batch_size"]
    counts = {"train": int(n * ds["batches_train_ratio"]), "valid": int(n * ds["batches_valid_ratio"]),
              "test": int(n * ds["batches_test_ratio"])}
    written = []
    for kind, num in counts.items():
        for i in shard_bundles(num, rank, world):
            protos, params = generate_proto_set(
                img_size=meta["img_size"], size=bs, alpha_range=meta["alpha_range"],
                ratio_range=meta["ratio_range"], size_ratio_range=meta["size_ratio_range"],
                seed=int(rng.integers(low=0, high=2**32 - 1)))
            metadata = []
            for j in range(bs):
                p = rng.choice(_DSHARP_POINTINGS)
                metadata.append(dict(index=j, cell_size=_DSHARP_CELL, src_ra=p[0], src_dec=p[1]))
            suffix = f"_r{rank}" if world > 1 else ""
            path = out / f"{ds['file_prefix']}_{kind}{suffix}_{i}.h5"
            write_h5(path, {"y": np.array(protos), "metadata": str(metadata), "params": str(params)})
            written.append(path)
    return written
