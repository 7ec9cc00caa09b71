"""Device-resident jet throughput at other sizes (which kernel serves them: jet2 on-chip up to 256, jet.cu beyond)."""
import sys, ctypes as C
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L
from radiosim_b200.jets.packer import pack_jets
lib = L.lib()
CASES = ((64, 32768), (100, 16384), (128, 16384), (150, 8192), (200, 8192), (256, 8192), (384, 4096), (512, 2048), (1024, 512))
want = [int(a) for a in sys.argv[1:]]
known = dict(CASES)
cases = [(w, known.get(w, max(256, (1 << 29) // (w * w) // 256 * 256))) for w in want] if want else list(CASES)
for size, n in cases:
    np.random.seed(1)
    b = pack_jets(n, size, (3, 10), "normalize")
    d_p, d_z = L.dev_from(b.params), L.dev_from(b.zoom_pad)
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
    live = (b.params[:, :, 0] != 0).sum() / n
    rate = 2 * n / (ms.value * 1e-3)
    print(f"{size:5d}^2: {rate/1e6:7.3f} M images/s, {rate*size*size*live/1e12:6.2f}e12 algorithmic exp/s (workspace {need.value>>20} MiB)")
