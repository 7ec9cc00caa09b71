"""CPU restatement of the survey generator (reference ``src/radiosim/survey/survey.py``).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Same legacy global
``np.random`` draw order as the reference; pinned against
``tests/golden/survey_*.npz`` (made from the unmodified reference).
"""

from __future__ import annotations

import numpy as np

from .gauss import twodgaussian
from .jet import create_jet


def _sum_gaussians(rows, size, workers):
    """Sum of ``twodgaussian`` over ``rows`` in draw order.  ``workers > 1`` splits the rows
    over threads (numpy releases the GIL in the big ufuncs) and adds the partial sums --
    float64 rounding order changes at the 1e-16 level, nothing else does.  Only there so
    that the C4-sized parity cases (thousands of full-grid Gaussians at 1024^2) finish in
    about a minute on the test box."""
    out = np.zeros((size, size))
    if workers <= 1 or len(rows) < 2 * workers:
        for r in rows:
            out += twodgaussian(r, size)
        return out
    from concurrent.futures import ThreadPoolExecutor

    def part(chunk):
        acc = np.zeros((size, size))
        for r in chunk:
            acc += twodgaussian(r, size)
        return acc

    step = -(-len(rows) // workers)
    with ThreadPoolExecutor(workers) as ex:
        for acc in ex.map(part, [rows[i : i + step] for i in range(0, len(rows), step)]):
            out += acc
    return out


def create_survey(grid_shape, num_sources=20, class_distribution=(2, 1, 2),
                  scale_sources=True, jet_kwargs=None, trace=None, workers=1):
    """``create_survey``, ``src/radiosim/survey/survey.py:9-132``."""
    shape = tuple(grid_shape)
    n_img = 1 if len(shape) == 3 else shape[0]
    size = shape[-1]
    jet_kwargs = dict(jet_kwargs or {})
    rnd = np.random
    total = sum(class_distribution)
    survey = np.zeros((n_img, 1, size, size))
    planes = np.zeros((n_img, len(class_distribution) + 1, size, size))
    for b in range(n_img):
        pending = ([], [])  # Gaussians of planes 1 / 2, evaluated after the draws (same order)
        for _ in range(num_sources):
            u = rnd.uniform(0, total)  # :51
            if u < class_distribution[0]:  # jets :54-76
                if scale_sources:
                    js = rnd.randint(size / 10.24, size / 5.12)
                else:
                    js = rnd.randint(100, 200)
                pa, pb = rnd.rand(2) * size
                jet, _ = create_jet((1, js, js), **jet_kwargs)
                jet = np.squeeze(jet)
                # literal transcription of the slicing semantics of :61-76
                r0 = int(np.floor(pa - js / 2))
                r1 = int(np.floor(pa + js / 2))
                c0 = int(np.floor(pb - js / 2))
                c1 = int(np.floor(pb + js / 2))
                if r0 < 0:
                    jet = jet[-r0:js, :]
                    r0 = 0
                if c0 < 0:
                    jet = jet[:, -c0:js]
                    c0 = 0
                if r1 > size:
                    jet = jet[0 : size - r1 + js, :]
                if c1 > size:
                    jet = jet[:, 0 : size - c1 + js]
                planes[b, 0, r0:r1, c0:c1] += jet
                if trace is not None:
                    trace.append(("jet", js, pa, pb))
            else:
                wide = u < class_distribution[0] + class_distribution[1]
                if wide:  # :79-92
                    if scale_sources:
                        sx = size * rnd.uniform(1 / 40.96, 1 / 20.48)
                        sy = sx + rnd.normal(scale=size / 204.8)
                    else:
                        sx = rnd.uniform(25, 50)
                        sy = sx + rnd.normal(scale=5)
                else:  # :95-107
                    if scale_sources:
                        sx = size * rnd.uniform(1 / 204.8, 1 / 102.4)
                        sy = sx + rnd.normal(scale=size / 1024)
                    else:
                        sx = rnd.uniform(5, 10)
                        sy = sx + rnd.normal(scale=1)
                px, py = rnd.rand(2) * size
                amp = rnd.rand()
                rot = rnd.uniform(0, np.pi)
                pending[0 if wide else 1].append([amp, px, py, sx, sy, rot])  # :90-91, :105-106
                if trace is not None:
                    trace.append(("gauss" if wide else "point", amp, px, py, sx, sy, rot))
        planes[b, 1] += _sum_gaussians(pending[0], size, workers)
        planes[b, 2] += _sum_gaussians(pending[1], size, workers)
        survey[b] = np.sum(planes[b], axis=0)  # :109
        peak = survey[b].max()  # :112-114
        survey[b] /= peak
        planes[b] /= peak
        planes[b, -1] = 1 - survey[b]  # :122
    return survey, planes
