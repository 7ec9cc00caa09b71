"""One survey batch through the device pipeline (for ncu): python tools/run_survey.py N_IMAGES NS D0 D1 D2 [reps]"""
import sys
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L
from radiosim_b200.survey import survey as sv
n, ns = int(sys.argv[1]), int(sys.argv[2])
dist = tuple(int(v) for v in sys.argv[3:6])
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 2
np.random.seed(1337)
b = sv.pack_survey(n, 1024, ns, dist, True)
d = sv.DeviceSurvey(b, True, False)
for _ in range(reps):
    d.enqueue()
L.device_sync()
x, y = d.download()
print("ok", float(x.max()), b.njets, b.src_params.shape)
