"""End-to-end timing of the ppdisk and mojave drop-ins (host recipe + device + results on the host)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from types import SimpleNamespace as NS
from radiosim_b200.ppdisks import generate_proto_set
from radiosim_b200.mojave import create_mojave

for rep in range(3):
    t = time.time()
    imgs, params = generate_proto_set(512, 50, (0, 180), (3, 15), (0.1, 1.0), "cuda", 1337 + rep, False)
    dt = time.time() - t
    print(f"generate_proto_set 512^2 x50: {1e3 * dt / 50:.2f} ms/image")
import os
threads = int(os.environ.get("THREADS", "0")) or None
conf = NS(general=NS(threads=threads, seed=1337), dataset=NS(bundle_size=1000, img_size=256), mojave=NS(class_ratio=[1, 1, 1]))
for rep in range(3):
    data = names = None  # (release the previous bundle: its page-locked block is reused)
    rng = np.random.default_rng(1337 + rep)
    t = time.time()
    data, names = create_mojave(conf, rng, today="2026-01-01")
    dt = time.time() - t
    print(f"create_mojave (threads={threads}) 256^2 x1000: {1e3 * dt / 1000:.3f} ms/image ({1000 / dt:.0f} images/s)")
