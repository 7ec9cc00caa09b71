"""Debug aid: single-source survey renders against the oracle, error maps summarised."""
import sys
import numpy as np
sys.path.insert(0, ".")
from oracle.gauss import twodgaussian
from radiosim_b200.survey.survey import SurveyBatch, render_survey, pack_survey
from oracle import survey as osur


def one(N, p, cls):
    p = np.array([p], dtype=np.float64)
    offs = np.array([[0, 1, 1]] if cls == 1 else [[0, 0, 1]], np.int32)
    b = SurveyBatch(N, 1, p, offs, [], np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0, np.int64))
    x, y = render_survey(b, f64=False)
    ref = twodgaussian(p[0], N)
    ref = ref / ref.max()
    err = np.abs(x[0, 0] - ref)
    r, c = np.unravel_index(err.argmax(), err.shape)
    print(f"N={N} cls={cls} p={p[0].tolist()} max err {err.max():.3e} at {(r, c)} got {x[0,0,r,c]:.4e} ref {ref[r,c]:.4e}"
          f" rows with err>1e-5: {np.unique(np.nonzero(err > 1e-5)[0])[:20]} cols: {np.unique(np.nonzero(err > 1e-5)[1])[:20]}")


for N in (128, 200, 1024):
    one(N, [0.8, N * 0.3, N * 0.6, N / 30, N / 40, 0.6], 1)
    one(N, [0.5, N * 0.7, N * 0.2, N / 150, N / 160, 2.1], 2)
    one(N, [0.9, 3.2, N - 2.5, N / 30, N / 25, 1.0], 1)
for seed, ns, dist in ((1, 3, (0, 1, 1)), (2, 3, (1, 0, 0)), (3, 20, (2, 1, 2))):
    N = 128
    np.random.seed(seed)
    ox, oy = osur.create_survey((1, 3, N, N), ns, dist, True)
    np.random.seed(seed)
    b = pack_survey(1, N, ns, dist, True)
    x, y = render_survey(b, f64=False)
    for k in range(4):
        e = np.abs(y[0, k] - oy[0, k])
        print(f"seed {seed} ns {ns} dist {dist} plane {k}: max err {e.max():.3e} at {np.unravel_index(e.argmax(), e.shape)}"
              f" rows>1e-5 {np.unique(np.nonzero(e > 1e-5)[0])[:12]} cols {np.unique(np.nonzero(e > 1e-5)[1])[:12]}")
    print("  sources:", b.src_params.round(2).tolist()[:6], "jets", b.njets, b.jet_sizes, b.jet_row0, b.jet_col0)
