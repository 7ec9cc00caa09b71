"""Where the 17 ms of the default-TOML run (C1) go: cProfile of simulate_sky_distributions."""
import cProfile, pstats, os, sys, tempfile, shutil, time
sys.path.insert(0, ".")
import numpy as np
from radiosim_b200.io import Config
from radiosim_b200.simulations import simulate_sky_distributions
conf = Config.from_toml("rc/default_simulation.toml")
def run():
    tmp = tempfile.mkdtemp(prefix="rsb_c1_")
    conf.paths.outpath = tmp
    t = time.perf_counter(); files = simulate_sky_distributions(conf, "jet", rank=0, world=1); dt = time.perf_counter() - t
    shutil.rmtree(tmp, ignore_errors=True)
    return dt
print("warm", run(), run(), run())
pr = cProfile.Profile(); pr.enable(); dt = run(); pr.disable()
print("profiled run", dt)
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
