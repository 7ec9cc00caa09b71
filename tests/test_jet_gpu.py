"""GPU parity tests of the jet path: CUDA (through the C ABI) vs oracle / golden."""

import os
from types import SimpleNamespace

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_IMG = 1e-5      # max abs error relative to the image peak (BASELINE north_star)
TOL_LIST = 1e-6     # relative tolerance of the float list targets


def _conf(training_type="list", scaling="normalize", comps=(3, 10)):
    return SimpleNamespace(jet=SimpleNamespace(
        training_type=training_type, scaling=scaling, num_jet_components=list(comps)))


def _grid(n, size):
    return np.broadcast_to(np.zeros(1), (n, 3, size, size))


@pytest.mark.parametrize(
    "name", ["jet_128_list", "jet_256_list", "jet_64_gauss", "jet_100_clean_mojave"])
def test_create_jet_matches_reference_golden(golden_dir, name):
    from radiosim_b200.jets import create_jet

    g = np.load(os.path.join(golden_dir, name + ".npz"))
    n, size = int(g["n"]), int(g["size"])
    np.random.seed(int(g["seed"]))
    x, y = create_jet(_grid(n, size), _conf(str(g["training_type"]), str(g["scaling"])))
    assert np.random.rand() == float(g["next_rand"])
    assert x.shape == g["x"].shape and x.dtype == np.float64
    assert y.shape == g["y"].shape and y.dtype == np.float64
    for i in range(n):
        peak = g["x"][i].max()
        assert np.abs(x[i] - g["x"][i]).max() <= TOL_IMG * peak, i
    assert np.array_equal(np.isnan(y), np.isnan(g["y"]))
    if str(g["training_type"]) == "list":
        assert np.allclose(y, g["y"], rtol=TOL_LIST, atol=1e-12)
    else:
        for i in range(n):
            peak = np.nanmax(g["y"][i, :-1]) if y.ndim == 4 and y.shape[1] > 1 else g["y"][i].max()
            assert np.nanmax(np.abs(y[i] - g["y"][i])) <= max(TOL_IMG * peak, 1e-12), i


@pytest.mark.parametrize("size,n,seed,f64", [
    (128, 40, 21, True), (64, 30, 22, False), (101, 12, 23, True), (150, 12, 24, False),
    (256, 24, 25, False), (300, 6, 26, True), (66, 10, 27, True),
    # odd sizes above 128: two CTAs per image with uneven bands (ceil / floor of N/2 rows)
    (151, 12, 28, False), (199, 10, 29, True), (255, 8, 30, False), (130, 8, 31, False),
    # 401 .. 512: eight CTAs per image, 8-row groups dealt round-robin, V pass through distributed shared
    # memory (jet3.cu); 384: the scratch-image kernel; odd / not a multiple of 4 / of 8 among them
    (512, 12, 32, False), (512, 6, 33, True), (404, 8, 34, False), (449, 8, 35, True), (470, 6, 36, False),
    (401, 6, 37, False), (384, 6, 38, False)])
def test_render_matches_oracle(size, n, seed, f64):
    """Same seed -> same parameters -> images within 1e-5 of peak, margins exact."""
    from oracle import jet as ojet
    from radiosim_b200.jets.jet import render_batch
    from radiosim_b200.jets.packer import pack_jets

    np.random.seed(seed)
    trace = []
    ox, _ = ojet.create_jet((n, 3, size, size), (3, 10), "normalize", "clean", trace=trace)
    np.random.seed(seed)
    b = pack_jets(n, size, (3, 10), "normalize")
    img, margin, peak, _ = render_batch(b, True, f64=f64)
    assert img.dtype == (np.float64 if f64 else np.float32)
    o_margin = np.array([t["margin"] for t in trace])
    o_peak = np.array([t["jet_max"] for t in trace])
    assert np.array_equal(margin, o_margin)          # integer: bit-exact
    assert np.allclose(peak, o_peak, rtol=1e-6, atol=0)
    err = np.abs(img.astype(np.float64) - ox[:, 0]).reshape(n, -1).max(axis=1)
    assert err.max() <= TOL_IMG, err
    assert np.abs(img.reshape(n, -1).max(axis=1) - 1.0).max() <= 2e-6


def test_unnormalised_mojave_scaling_matches_oracle():
    from oracle import jet as ojet
    from radiosim_b200.jets.jet import render_batch
    from radiosim_b200.jets.packer import pack_jets

    n, size = 16, 128
    np.random.seed(31)
    ox, _ = ojet.create_jet((n, 3, size, size), (3, 10), "mojave", "clean")
    np.random.seed(31)
    b = pack_jets(n, size, (3, 10), "mojave")
    img, margin, peak, _ = render_batch(b, False)
    assert np.all(peak == 1.0)
    for i in range(n):
        assert np.abs(img[i] - ox[i, 0]).max() <= TOL_IMG * ox[i, 0].max()


def test_gauss_planes_match_oracle():
    from oracle import jet as ojet
    from radiosim_b200.jets import create_jet

    n, size = 6, 96
    np.random.seed(41)
    ox, oy = ojet.create_jet((n, 3, size, size), (3, 10), "normalize", "gauss")
    np.random.seed(41)
    x, y = create_jet(_grid(n, size), _conf("gauss"))
    assert y.shape == oy.shape == (n, 21, size, size)
    assert np.array_equal(np.isnan(y), np.isnan(oy))
    assert np.nanmax(np.abs(y[:, :20] - oy[:, :20])) <= 2 * TOL_IMG
    lst, olst = y[:, 20, :19, :8], oy[:, 20, :19, :8]
    assert np.allclose(lst, olst, rtol=TOL_LIST, atol=1e-12)


def test_edge_cases():
    from radiosim_b200 import _lib as L
    from radiosim_b200.jets.jet import render_batch
    from radiosim_b200.jets.packer import JetBatch

    # empty batch
    b = JetBatch(64, 19, np.zeros((0, 19, 6)), np.zeros((0, 19, 2)),
                 np.zeros(0, np.int32), np.zeros(0))
    img, margin, peak, _ = render_batch(b, True)
    assert img.shape == (0, 64, 64)

    # one live slot, no zoom: analytic check
    p = np.zeros((1, 1, 6))
    p[0, 0] = [2.0, 20.25, 30.5, 6.0, 3.0, 0.7]
    b = JetBatch(64, 1, p, np.zeros((1, 1, 2)), np.full(1, -1, np.int32), np.ones(1))
    img, margin, peak, _ = render_batch(b, False)
    from oracle.gauss import twodgaussian

    ref = twodgaussian(p[0, 0], 64)
    assert np.abs(img[0] - ref).max() <= TOL_IMG * ref.max()

    # argument errors surface as exceptions with a message
    with pytest.raises(L.RsbError, match="slots"):
        b = JetBatch(64, 65, np.zeros((1, 65, 6)), np.zeros((1, 65, 2)),
                     np.full(1, -1, np.int32), np.ones(1))
        render_batch(b, True)


def test_full_size_properties():
    """BASELINE config 5 sizes (256^2): properties that need no oracle."""
    from radiosim_b200.jets.jet import render_batch
    from radiosim_b200.jets.packer import pack_jets

    np.random.seed(5)
    b = pack_jets(600, 256, (3, 10), "normalize")
    img, margin, peak, _ = render_batch(b, True, f64=False)
    flat = img.reshape(len(img), -1)
    assert np.isfinite(flat).all() and (flat >= 0).all()
    assert np.abs(flat.max(axis=1) - 1.0).max() <= 2e-6
    zoomed = b.zoom_pad >= 0
    assert np.all(margin[~zoomed] == 0) and np.all(margin >= 0) and np.all(margin < 128)
    # scale invariance: multiplying every amplitude by 3 leaves the normalised image unchanged
    # for images that do not zoom (the 0.01 threshold is absolute)
    b3 = pack_jets.__globals__["JetBatch"](b.size, b.slots, b.params.copy(), b.extra, b.zoom_pad, b.zoom_out)
    b3.params[:, :, 0] *= 3.0
    img3, _, peak3, _ = render_batch(b3, True, f64=False)
    nz = ~zoomed
    assert np.abs(img3[nz] - img[nz]).max() <= 2e-6
    assert np.allclose(peak3[nz], 3.0 * peak[nz], rtol=1e-12)
    # determinism
    img2, margin2, peak2, _ = render_batch(b, True, f64=False)
    assert np.array_equal(img, img2) and np.array_equal(margin, margin2) and np.array_equal(peak, peak2)


def test_pipelined_create_jet_equals_single_shot():
    """Large batches are packed in chunks on a helper thread while the previous chunk renders:
    same stream consumption, same images, same targets as one pack + one render."""
    from radiosim_b200.jets import create_jet
    from radiosim_b200.jets.jet import _PIPE_CHUNK, render_batch
    from radiosim_b200.jets.packer import pack_jets, source_lists

    n, size = 2 * _PIPE_CHUNK + 77, 64
    np.random.seed(99)
    x, y = create_jet(_grid(n, size), _conf("list"))
    end_a = np.random.rand()
    np.random.seed(99)
    b = pack_jets(n, size, (3, 10), "normalize")
    end_b = np.random.rand()
    img, margin, peak, _ = render_batch(b, True)
    assert end_a == end_b
    assert x.shape == (n, 1, size, size) and np.array_equal(x[:, 0], img)
    assert np.array_equal(y, source_lists(b, margin, peak, True))


@pytest.mark.parametrize("scrub", [None, 0xFFFFFFFF])
def test_variable_size_launch_matches_oracle(scrub):
    """rsb_jet_render_var: jets of many sizes (the sub-jets of a 1024^2 survey: 100..199 px, plus
    small and 256 px ones) in one launch per cluster shape, each against the oracle.  Second pass:
    behind a kernel that leaves NaN bit patterns in every SM's shared memory (tiles the renderer
    skips are never written; a zero-weight tap on one of them must not turn into NaN)."""
    import ctypes as C

    from oracle import jet as ojet
    from radiosim_b200 import _lib as L
    from radiosim_b200.jets.packer import pack_jets

    L.require_device()
    lib = L.lib()
    rng = np.random.default_rng(5)
    for sizes in (np.sort(rng.integers(100, 200, size=70)), np.array([8, 31, 64, 100, 101, 127, 128]),
                  np.array([129, 130, 131, 197, 199, 200, 255, 256, 256])):
        n = len(sizes)
        refs, margins, peaks, params, zpads = [], [], [], [], []
        for k, sz in enumerate(sizes):
            np.random.seed(1000 + k)
            trace = []
            ox, _ = ojet.create_jet((1, 3, int(sz), int(sz)), (3, 10), "normalize", "clean", trace=trace)
            np.random.seed(1000 + k)
            b = pack_jets(1, int(sz), (3, 10), "normalize")
            refs.append(ox[0, 0]); margins.append(trace[0]["margin"]); peaks.append(trace[0]["jet_max"])
            params.append(b.params[0]); zpads.append(b.zoom_pad[0])
        pitch = (sizes + 3) & ~3
        offs = np.zeros(n, dtype=np.int64)
        offs[1:] = np.cumsum(sizes * pitch)[:-1]
        total = int((sizes * pitch).sum())
        d_p, d_z = L.dev_from(np.array(params)), L.dev_from(np.array(zpads, dtype=np.int32))
        d_s, d_o = L.dev_from(sizes.astype(np.int32)), L.dev_from(offs)
        d_img, d_m, d_k = L.DeviceBuffer(total * 4), L.DeviceBuffer(n * 4), L.DeviceBuffer(n * 8)
        if scrub is not None:
            L.check(lib.rsb_debug_scrub_shared(scrub, None))
        L.check(lib.rsb_jet_render_var(d_p.ptr, d_z.ptr, d_s.ptr, d_o.ptr, n, 19, int(sizes.max()),
                                       L.JET_NORMALIZE, d_img.ptr, d_m.ptr, d_k.ptr, None), "rsb_jet_render_var")
        L.device_sync()
        flat = d_img.download((total,), np.float32)
        assert np.array_equal(d_m.download((n,), np.int32), np.array(margins))
        assert np.allclose(d_k.download((n,), np.float64), np.array(peaks), rtol=1e-6, atol=0)
        for k, sz in enumerate(sizes):
            img = flat[offs[k] : offs[k] + sz * pitch[k]].reshape(sz, pitch[k])[:, :sz]
            assert np.abs(img.astype(np.float64) - refs[k]).max() <= TOL_IMG, (int(sz), k)


@pytest.mark.parametrize("size", [64, 100, 128, 157, 200, 224, 255, 256, 300, 384, 512])
def test_shared_memory_leftovers_do_not_matter(size):
    """Fixed-size launches behind a kernel that fills shared memory with zeros, NaN and infinity
    bit patterns: bit-identical images and labels."""
    from radiosim_b200 import _lib as L
    from radiosim_b200.jets.packer import pack_jets
    from radiosim_b200.jets.jet import render_batch

    L.require_device()
    np.random.seed(77)
    b = pack_jets(300 if size <= 256 else 60, size, (3, 10), "normalize")
    got = []
    for pattern in (0x0, 0xFFFFFFFF, 0x7F800000):
        L.check(L.lib().rsb_debug_scrub_shared(pattern, None))
        img, margin, peak, _ = render_batch(b, True, f64=False)
        got.append((np.array(img), np.array(margin), np.array(peak)))
    for img, margin, peak in got:
        assert np.isfinite(img).all()
    for img, margin, peak in got[1:]:
        assert np.array_equal(img, got[0][0]) and np.array_equal(margin, got[0][1]) and np.array_equal(peak, got[0][2])
