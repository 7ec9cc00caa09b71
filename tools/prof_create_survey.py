"""Where the time of the drop-in create_survey goes (host packer / sub-jets / splat)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200.survey.survey import pack_survey, render_survey
n, size, nsrc = 4, 1024, 2000
np.random.seed(1337)
for rep in range(3):
    t0 = time.time(); b = pack_survey(n, size, nsrc, [2, 1, 2], True)
    t1 = time.time(); out = render_survey(b, True)
    t2 = time.time()
    del out
    print(f"rep {rep}: pack {1e3*(t1-t0)/n:.1f} ms/image ({b.njets/n:.0f} sub-jets/image), render {1e3*(t2-t1)/n:.1f} ms/image")
