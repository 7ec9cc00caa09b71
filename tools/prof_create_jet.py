import sys, time
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200.jets.packer import pack_jets, source_lists
from radiosim_b200.jets.jet import render_batch
n, size = 2048, 256
np.random.seed(1)
for rep in range(5):
    t0 = time.time(); b = pack_jets(n, size, (3, 10), "normalize")
    t1 = time.time(); img, margin, peak, _ = render_batch(b, True)
    t2 = time.time(); sl = source_lists(b, margin, peak, True)
    t3 = time.time()
    print(f"rep {rep}: pack {1e3*(t1-t0):.1f} ms, render_batch {1e3*(t2-t1):.1f} ms, lists {1e3*(t3-t2):.1f} ms")
    del img
