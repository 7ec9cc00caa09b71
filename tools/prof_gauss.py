import sys, time
import numpy as np
sys.path.insert(0, ".")
from types import SimpleNamespace as NS
from radiosim_b200.jets import create_jet
for n, size in ((20, 128), (200, 256)):
    conf = NS(jet=NS(num_jet_components=[3, 10], scaling="normalize", training_type="gauss"))
    grid = np.empty((n, 3, size, size), dtype=np.float32)
    np.random.seed(1)
    create_jet(grid, conf)
    x = y = None
    t = time.time()
    for _ in range(3):
        x = y = None
        x, y = create_jet(grid, conf)
    dt = (time.time() - t) / 3
    print(f"create_jet gauss n={n} {size}^2: {n/dt:.0f} images/s ({dt*1e3:.1f} ms/call, {y.nbytes/1e6:.0f} MB targets)")
