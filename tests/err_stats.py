"""Error statistics of the CUDA jets against the float64 oracle (images are normalised: peak = 1).

A test-side script (it imports oracle/, which only tests may do): python tests/err_stats.py on a GPU box."""
import sys, numpy as np
sys.path.insert(0, '.')
from oracle import jet as ojet
from radiosim_b200.jets.jet import render_batch
from radiosim_b200.jets.packer import pack_jets
worst = 0.0
import os
SCALE = int(os.environ.get("RSB_ERR_SCALE", "1"))  # more jets per case
COMPS = tuple(int(v) for v in os.environ.get("RSB_ERR_COMPS", "3,10").split(","))  # num_jet_components
for size, n, seed in [(128, 48, 1), (200, 32, 2), (256, 64, 3), (256, 64, 11), (64, 48, 5)]:
    n *= SCALE
    np.random.seed(seed); tr = []
    ox, _ = ojet.create_jet((n, 3, size, size), COMPS, "normalize", "clean", trace=tr)
    np.random.seed(seed); b = pack_jets(n, size, COMPS, "normalize")
    img, margin, peak, _ = render_batch(b, True, f64=False)
    e = np.abs(img - ox[:, 0]).reshape(n, -1).max(axis=1)
    pk = np.array([t["jet_max"] for t in tr]); mg = np.array([t["margin"] for t in tr])
    rel = np.abs(peak - pk) / pk
    worst = max(worst, e.max())
    print(f"size {size} n {n}: image err max {e.max():.2e} median {np.median(e):.2e} | peak rel err max {rel.max():.2e} | margins equal {np.array_equal(margin, mg)}")
print("worst image error / peak: %.2e (tolerance 1e-5)" % worst)
