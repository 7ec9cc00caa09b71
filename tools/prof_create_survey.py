"""Where the time of the drop-in create_survey goes (host packer / uploads / kernels / download)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L
from radiosim_b200.survey import survey as sv
n, size, nsrc = 4, 1024, int(sys.argv[1]) if len(sys.argv) > 1 else 2000
np.random.seed(1337)
for rep in range(4):
    t0 = time.perf_counter(); b = sv.pack_survey(n, size, nsrc, [2, 1, 2], True)
    t1 = time.perf_counter(); d = sv.DeviceSurvey(b, True, True)
    t2 = time.perf_counter(); d.enqueue(); L.device_sync()
    t3 = time.perf_counter(); out = d.download()
    t4 = time.perf_counter()
    del out, d
    f = 1e3 / n
    print(f"rep {rep}: pack {f*(t1-t0):.2f}  setup+upload {f*(t2-t1):.2f}  kernels {f*(t3-t2):.2f}  download(f64) {f*(t4-t3):.2f}  ms/image "
          f"({b.njets/n:.0f} sub-jets/image)")
