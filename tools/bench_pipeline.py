"""Wall-clock of the whole `simulate` path (RNG chain, render, .h5 shards) for the reference's default jet config."""
import sys, time, tempfile, glob, os
sys.path.insert(0, ".")
from radiosim_b200.io import Config
from radiosim_b200.simulations import simulate_sky_distributions
import radiosim_b200, pathlib

toml = next(pathlib.Path(radiosim_b200.__file__).resolve().parent.parent.rglob("default_simulation.toml"))
for size, bundle in ((128, 20), (256, 500)):
    with tempfile.TemporaryDirectory() as d:
        conf = Config.from_toml(str(toml))
        conf.paths.outpath = d
        conf.general.seed = 1337
        conf.dataset.img_size = size
        conf.dataset.bundle_size = bundle
        t = time.time()
        files = simulate_sky_distributions(conf, "jet", 0, 1)
        dt = time.time() - t
        n = bundle * len(files)
        nbytes = sum(os.path.getsize(f) for f in files)
        print(f"simulate jet {size}x{size}: {len(files)} shards, {n} images, {nbytes / 1e6:.0f} MB in {dt:.2f} s = {n / dt:.0f} images/s")
