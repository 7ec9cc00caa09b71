"""Splat vs resampling share of the scratch-image kernel (sizes above 256): all-plain vs all-zoom batches."""
import sys, ctypes as C
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L
from radiosim_b200.jets.packer import pack_jets
lib = L.lib()
for size, n in ((384, 4096), (512, 2048)):
    np.random.seed(1)
    b = pack_jets(n, size, (3, 10), "normalize")
    for label, zp in (("as drawn", b.zoom_pad), ("all plain", np.full(n, -1, np.int32)),
                      ("all zoom", np.where(b.zoom_pad >= 0, b.zoom_pad, size // 8).astype(np.int32))):
        d_p, d_z = L.dev_from(b.params), L.dev_from(np.ascontiguousarray(zp))
        d_img = L.DeviceBuffer(n * size * size * 4); d_m = L.DeviceBuffer(n * 4); d_k = L.DeviceBuffer(n * 8)
        need = C.c_size_t(0)
        L.check(lib.rsb_jet_workspace_bytes(n, 19, size, C.byref(need)))
        d_w = L.DeviceBuffer(max(need.value, 16))
        e0, e1 = C.c_void_p(), C.c_void_p()
        L.check(lib.rsb_event_create(C.byref(e0))); L.check(lib.rsb_event_create(C.byref(e1)))
        for rep in range(3):
            if rep == 1:
                L.check(lib.rsb_event_record(e0, None))
            L.check(lib.rsb_jet_render(d_p.ptr, d_z.ptr, n, 19, size, 1, d_img.ptr, d_m.ptr, d_k.ptr, d_w.ptr, d_w.nbytes, None))
        L.check(lib.rsb_event_record(e1, None)); L.device_sync()
        ms = C.c_float(0); L.check(lib.rsb_event_elapsed_ms(e0, e1, C.byref(ms)))
        print(f"{size}^2 {label:10s}: {2 * n / (ms.value * 1e-3) / 1e6:6.3f} M images/s  ({ms.value / (2 * n) * 1e3:.2f} us/image, zoom fraction {(zp >= 0).mean():.2f})")
