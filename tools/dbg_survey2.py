import sys
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200.survey.survey import render_survey, pack_survey
from oracle import survey as osur
N = int(sys.argv[1]) if len(sys.argv) > 1 else 128
for seed, ns, dist in ((3, 20, (0, 1, 2)), (3, 20, (0, 1, 0)), (3, 20, (0, 0, 1)), (3, 40, (0, 1, 0)), (3, 6, (0, 1, 0)), (3, 20, (2, 1, 2))):
    np.random.seed(seed)
    ox, oy = osur.create_survey((1, 3, N, N), ns, dist, True)
    np.random.seed(seed)
    b = pack_survey(1, N, ns, dist, True)
    x, y = render_survey(b, f64=False)
    errs = [float(np.abs(y[0, k] - oy[0, k]).max()) for k in range(4)]
    print(f"N {N} seed {seed} ns {ns} dist {dist}: plane errs {['%.2e' % e for e in errs]} offs {b.src_offs.tolist()} njets {b.njets}")
