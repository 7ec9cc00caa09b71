"""add_noise throughput (drop-in: host image bundle in, noised float64 bundle out)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200.utils import add_noise
size, n = (int(sys.argv[1]) if len(sys.argv) > 1 else 256), (int(sys.argv[2]) if len(sys.argv) > 2 else 200)
img = np.random.default_rng(1).random((n, 1, size, size)) ** 8
np.random.seed(3)
add_noise(img[:8], (0, 15))
best = 1e9
for rep in range(3):
    t0 = time.perf_counter(); out = add_noise(img, (0, 15)); best = min(best, time.perf_counter() - t0)
print(f"add_noise {size}^2 x {n}: {n / best:.0f} images/s ({best / n / (3 * size * size) * 1e9:.2f} ns per normal drawn)")
