"""Render one batch of 256x256 jets twice (profiling target for ncu)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L
from radiosim_b200.jets.packer import pack_jets
lib = L.lib()
n, size = 148 * 16, 256
np.random.seed(1)
b = pack_jets(n, size, (3, 10), "normalize")
d_p, d_z = L.dev_from(b.params), L.dev_from(b.zoom_pad)
d_img = L.DeviceBuffer(n * size * size * 4); d_m = L.DeviceBuffer(n * 4); d_k = L.DeviceBuffer(n * 8)
for rep in range(2):
    L.check(lib.rsb_jet_render(d_p.ptr, d_z.ptr, n, 19, size, 1, d_img.ptr, d_m.ptr, d_k.ptr, None, 0, None))
    L.device_sync()
print("ok", d_m.download((n,), np.int32)[:8])
