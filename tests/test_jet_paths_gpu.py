"""Hand-made component sets that steer the jet kernels through their less travelled paths
(one exponential per pixel for narrow / non-positive components, ragged tiles, sizes whose rows
are not 16-byte multiples, the slot limit of the on-chip kernel, centres outside the image),
each against the oracle's finish_jet (reference jets/jet.py:144-192) on the same parameters."""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_IMG = 1e-5  # of the image peak (BASELINE north_star)


def _cols(rng, size, slots, kind):
    amp = rng.uniform(0.05, 1.0, slots)
    x = rng.uniform(0.3 * size, 0.7 * size, slots)
    y = rng.uniform(0.3 * size, 0.7 * size, slots)
    sx = rng.uniform(size / 40, size / 12, slots)
    sy = rng.uniform(size / 60, size / 16, slots)
    rot = rng.uniform(0, np.pi, slots)
    amp[0] = 1.0
    if kind == "negative":
        amp[1::3] *= -0.3
    elif kind == "narrow":
        sx[1::2] = rng.uniform(0.35, 0.9, sx[1::2].size)
        sy[1::2] = rng.uniform(0.3, 0.6, sy[1::2].size)
    elif kind == "wide":
        sx[:] = rng.uniform(size / 6, size / 3, slots)
        sy[:] = rng.uniform(size / 8, size / 4, slots)
    elif kind == "outside":
        x[1::2] = rng.uniform(-0.3 * size, 1.3 * size, x[1::2].size)
        y[1::2] = rng.uniform(-0.3 * size, 1.3 * size, y[1::2].size)
    elif kind == "sparse":
        amp[rng.uniform(size=slots) < 0.6] = 0.0
        amp[0] = 1.0
    elif kind == "faint":
        amp[1:] *= 1e-9  # below the footprint cut: dropped by the kernel, < 2e-8 of the peak each
    return dict(amp=amp, x=x, y=y, sx=sx, sy=sy, rot=rot, z_rot=np.zeros(slots), beta=np.zeros(slots))


CASES = [
    (256, 19, "plain", True), (256, 19, "negative", True), (256, 19, "narrow", True), (256, 12, "wide", True),
    (256, 19, "outside", False), (256, 32, "plain", True), (256, 33, "plain", True), (256, 19, "sparse", True),
    (256, 19, "faint", True), (250, 19, "plain", True), (250, 9, "narrow", False), (130, 19, "negative", False),
    (128, 19, "narrow", True), (100, 7, "plain", True), (36, 5, "plain", True), (65, 5, "plain", True),
    (254, 19, "outside", True), (200, 32, "sparse", True),
    # the band renderer for 401 .. 512 pixels (jet3.cu) and, at 33 slots / 384 pixels, the scratch-image kernel
    (512, 19, "negative", True), (512, 19, "narrow", False), (480, 32, "outside", True), (430, 12, "wide", True),
    (512, 19, "faint", True), (418, 19, "sparse", False), (512, 33, "plain", True), (384, 19, "narrow", True),
]


@pytest.mark.parametrize("size,slots,kind,normalize", CASES)
@pytest.mark.parametrize("f64", [False, True])
def test_paths_match_oracle(size, slots, kind, normalize, f64):
    from oracle import jet as ojet
    from radiosim_b200.jets.jet import render_batch
    from radiosim_b200.jets.packer import JetBatch

    n = 6
    rng = np.random.default_rng(size * 1000 + slots * 10 + len(kind))
    params = np.zeros((n, slots, 6))
    zoom_pad = np.full(n, -1, np.int32)
    zoom_out = np.ones(n)
    refs, traces = [], []
    np.random.seed(size + slots)
    for i in range(n):
        cols = _cols(rng, size, slots, kind)
        if i % 3 == 2:  # push the source away from the border so that the zoom branch runs
            for k in ("x", "y"):
                cols[k] = size / 2 + (cols[k] - size / 2) * 0.4
            cols["sx"] *= 0.4
            cols["sy"] *= 0.4
            if kind == "narrow":
                cols["sx"][1::2] = np.maximum(cols["sx"][1::2], 0.35)
                cols["sy"][1::2] = np.maximum(cols["sy"][1::2], 0.3)
        for j, k in enumerate(("amp", "x", "y", "sx", "sy", "rot")):
            params[i, :, j] = cols[k]
        trace = []
        with np.errstate(all="ignore"):
            img, _ = ojet.finish_jet({k: v.copy() for k, v in cols.items()}, size,
                                     "normalize" if normalize else "none", "clean", trace=trace,
                                     planes_needed=False)
        refs.append(img)
        traces.append(trace[0])
        if trace[0]["zoomed"]:
            zoom_pad[i] = trace[0]["pad"]
            zoom_out[i] = trace[0]["zoom_out_factor"]
    b = JetBatch(size, slots, params, np.zeros((n, slots, 2)), zoom_pad, zoom_out)
    img, margin, peak, _ = render_batch(b, normalize, f64=f64)
    assert any(t["zoomed"] for t in traces) or kind in ("wide", "outside")
    for i in range(n):
        ref = refs[i]
        scale = np.abs(ref).max()
        assert np.isfinite(img[i]).all(), (i, kind)
        assert np.abs(img[i] - ref).max() <= TOL_IMG * scale, (i, kind, traces[i])
        if traces[i]["zoomed"]:
            assert margin[i] == traces[i]["margin"], (i, kind)
        if normalize:
            assert np.isclose(peak[i], traces[i]["jet_max"], rtol=1e-6, atol=0), (i, kind)
