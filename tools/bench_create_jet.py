"""Throughput of the drop-in create_jet (host RNG chain + render + float64 results on the host)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from types import SimpleNamespace as NS
from radiosim_b200.jets import create_jet

for ttype in ("list", "clean"):
    for n, size in ((2048, 256), (800, 128)):
        conf = NS(jet=NS(num_jet_components=[3, 10], scaling="normalize", training_type=ttype))
        grid = np.empty((n, 3, size, size), dtype=np.float32)
        np.random.seed(1337)
        create_jet(grid, conf)  # warm-up (library load, first launch)
        t = time.time()
        reps = 5
        x = y = None
        for _ in range(reps):
            del x, y  # a caller that has written the previous bundle; its page-locked blocks are reused
            x, y = create_jet(grid, conf)
        dt = (time.time() - t) / reps
        print(f"create_jet {ttype:5s} n={n} {size}x{size}: {n / dt:9.0f} images/s ({dt * 1e3:.1f} ms per call), x {x.dtype} {x.shape}")
