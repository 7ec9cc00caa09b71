"""ncu targets for the mojave and ppdisk kernels: python tools/run_dropins.py mojave|ppdisk"""
import sys
import numpy as np
sys.path.insert(0, ".")
from types import SimpleNamespace as NS
which = sys.argv[1]
if which == "mojave":
    from radiosim_b200.mojave import create_mojave
    conf = NS(general=NS(threads=None, seed=1337), dataset=NS(bundle_size=1000, img_size=256), mojave=NS(class_ratio=[1, 1, 1]))
    for rep in range(2):
        data, names = create_mojave(conf, np.random.default_rng(1337 + rep), today="2026-01-01")
    print("ok", data[0].shape)
else:
    from radiosim_b200.ppdisks import generate_proto_set
    for rep in range(2):
        imgs, params = generate_proto_set(512, 32, (0, 180), (3, 15), (0.1, 1.0), "cuda", 1337 + rep, False)
    print("ok", len(imgs))
