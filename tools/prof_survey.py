"""Device time of the survey pipeline's parts (CUDA events): variable-size sub-jet launches and the
fused tile kernel, for BASELINE config 4 shapes."""
import ctypes as C, json, sys, time
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L
from radiosim_b200.survey import survey as sv

lib = L.lib()
L.require_device()
ev = [C.c_void_p(), C.c_void_p()]
for e in ev:
    L.check(lib.rsb_event_create(C.byref(e)))


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    L.device_sync()
    L.check(lib.rsb_event_record(ev[0], None))
    for _ in range(reps):
        fn()
    L.check(lib.rsb_event_record(ev[1], None))
    ms = C.c_float()
    L.check(lib.rsb_event_elapsed_ms(ev[0], ev[1], C.byref(ms)))
    return ms.value / reps


n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
for ns, dist in ((2000, (2, 1, 2)), (2000, (0, 1, 2)), (5000, (2, 1, 2)), (5000, (0, 1, 2))):
    np.random.seed(1337)
    b = sv.pack_survey(n, 1024, ns, dist, True)
    d = sv.DeviceSurvey(b, True, False)
    groups = d.groups
    full = timed(d.enqueue)
    d.groups = []
    tile = timed(d.enqueue)
    d.groups = groups
    jets = []
    for g in groups:
        d2 = sv.DeviceSurvey.__new__(sv.DeviceSurvey)
        d2.__dict__.update(d.__dict__)
        d2.groups, d2.n = [g], 0
        jets.append((g[1] - g[0], g[2], timed(d2.enqueue)))
    print(json.dumps(dict(n=n, ns=ns, dist=dist, ms_per_image=full / n, tile_ms_per_image=tile / n,
                          jets=[dict(count=c, max_size=m, ms=t, us_per_jet=1e3 * t / c) for c, m, t in jets],
                          gaussians=b.src_params.shape[0] / n, subjets=b.njets / n)))
