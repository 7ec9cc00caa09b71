import sys
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200.survey.survey import render_survey, pack_survey
from oracle import survey as osur
from oracle.gauss import twodgaussian
N = 128
for seed, ns, dist in ((3, 6, (0, 1, 0)),):
    np.random.seed(seed)
    ox, oy = osur.create_survey((1, 3, N, N), ns, dist, True)
    np.random.seed(seed)
    b = pack_survey(1, N, ns, dist, True)
    print(b.src_params.round(3).tolist())
    x, y = render_survey(b, f64=False)
    peak = sum(twodgaussian(p, N) for p in b.src_params).max()
    for k in (1, 2):
        got = y[0, k] * 1.0
        print("plane", k, "max", got.max(), "argmax", np.unravel_index(got.argmax(), got.shape))
    # which sources are in plane 1 / plane 2? least squares on the individual renders
    basis = np.stack([twodgaussian(p, N).ravel() for p in b.src_params], 1)
    for k in (1, 2):
        coef, *_ = np.linalg.lstsq(basis, y[0, k].ravel().astype(np.float64), rcond=None)
        print("plane", k, "coefficients of the sources (x peak):", (coef * peak).round(3).tolist())
