"""CPU tests of the host-side jet packer: draw-for-draw alignment with the oracle."""

import numpy as np
import pytest

from oracle import jet as ojet
from radiosim_b200.jets.packer import pack_jets, pack_jets_py, source_lists


@pytest.mark.parametrize("impl", ["native", "python"])
@pytest.mark.parametrize("size,scaling,seed", [(128, "normalize", 1337), (64, "normalize", 3),
                                               (200, "mojave", 5), (256, "normalize", 11)])
def test_packer_matches_oracle_stream(size, scaling, seed, impl):
    """pack_jets (native rsb_jet_pack; numpy restatement for "mojave") and pack_jets_py against the
    oracle: same draws in the same order, same zoom decisions; list targets bit-identical for the
    numpy restatement and within 2 ulp for the native packer (libm vs numpy's SIMD exp/sin/cos)."""
    n = 24
    packer = pack_jets if impl == "native" else pack_jets_py
    np.random.seed(seed)
    trace = []
    _, y = ojet.create_jet((n, 3, size, size), (3, 10), scaling, "list", trace=trace)
    end_oracle = np.random.rand()

    np.random.seed(seed)
    b = packer(n, size, (3, 10), scaling)
    end_packer = np.random.rand()
    assert end_packer == end_oracle  # same number of draws, same order

    zoomed = np.array([t["zoomed"] for t in trace])
    assert np.array_equal(zoomed, b.zoom_pad >= 0)
    assert np.array_equal(np.array([t["pad"] for t in trace])[zoomed], b.zoom_pad[zoomed])
    if scaling == "normalize":
        assert zoomed.any() and (~zoomed).any()

    # with the oracle's margin and peak the list targets are bit-identical
    margin = np.array([t["margin"] for t in trace])
    peak = np.array([t["jet_max"] for t in trace])
    out = source_lists(b, margin, peak, scaling == "normalize")
    if impl == "python" or scaling == "mojave":
        assert np.array_equal(out, y)
    else:
        assert hasattr(b, "border_counts")  # the native path really ran
        assert np.allclose(out, y, rtol=1e-15, atol=0) and np.array_equal(out == 0, y == 0)


@pytest.mark.parametrize("size,n,seed,comps", [(256, 300, 1, (3, 10)), (128, 300, 7, (3, 10)), (64, 200, 3, (2, 6)),
                                               (101, 100, 11, (3, 10)), (32, 100, 4, (4, 4))])
def test_native_packer_continues_the_numpy_stream(size, n, seed, comps):
    """rsb_jet_pack against the numpy restatement: generator state (key, position, cached gaussian)
    identical afterwards, every branch identical, floats within 2 ulp."""
    np.random.seed(seed)
    np.random.normal()  # leave a cached second gaussian in the state that is handed over
    a = pack_jets(n, size, comps)
    sa = np.random.get_state()
    np.random.seed(seed)
    np.random.normal()
    b = pack_jets_py(n, size, comps)
    sb = np.random.get_state()
    assert sa[0] == sb[0] and np.array_equal(sa[1], sb[1]) and sa[2:] == sb[2:]
    assert np.array_equal(a.zoom_pad, b.zoom_pad) and np.array_equal(a.zoom_out, b.zoom_out)
    for u, v in ((a.params, b.params), (a.extra, b.extra)):
        assert np.array_equal(u == 0, v == 0)
        assert np.allclose(u, v, rtol=1e-15, atol=0)
    assert a.border_counts.sum() == n


def test_native_packer_rejects_bad_arguments():
    from radiosim_b200 import _lib as L
    import ctypes as C
    key = np.zeros(624, dtype=np.uint32)
    pos, has, cached = C.c_int32(700), C.c_int32(0), C.c_double(0.0)
    out = np.zeros(64)
    rc = L.lib().rsb_jet_pack(key.ctypes.data, C.byref(pos), C.byref(has), C.byref(cached), 1, 64, 3, 10,
                              out.ctypes.data, out.ctypes.data, out.ctypes.data, out.ctypes.data, None)
    assert rc > 0 and b"position" in L.lib().rsb_last_error()


def test_packer_slot_layout():
    np.random.seed(0)
    b = pack_jets(5, 128, (3, 6))
    assert b.params.shape == (5, 11, 6) and b.zoom_pad.dtype == np.int32
    assert np.all(b.params[:, 0, 0] == 1.0)  # core: amp 1, beta 0 -> boost 1


@pytest.mark.parametrize("n,size,nsrc,dist,scale,seed", [
    (2, 1024, 300, [2, 1, 2], True, 7), (3, 160, 40, [1, 1, 1], True, 8), (2, 512, 100, [2, 1, 2], False, 9),
    (2, 128, 25, [0, 1, 1], True, 10), (1, 256, 30, [3, 0, 0], True, 11)])
def test_native_survey_packer_continues_the_numpy_stream(n, size, nsrc, dist, scale, seed):
    """rsb_survey_pack (survey.py:49-107 incl. the nested create_jet chain) against the numpy
    restatement: generator state identical afterwards, same sources and sub-jets in the same
    order, integers exact, floats within 2 ulp."""
    from radiosim_b200.survey.survey import pack_survey, pack_survey_py

    np.random.seed(seed)
    np.random.normal()
    a = pack_survey(n, size, nsrc, dist, scale)
    sa = np.random.get_state()
    np.random.seed(seed)
    np.random.normal()
    b = pack_survey_py(n, size, nsrc, dist, scale)
    sb = np.random.get_state()
    assert a.jets == [] and a.jet_params is not None  # the native path really ran
    assert np.array_equal(sa[1], sb[1]) and sa[2:] == sb[2:]
    assert np.array_equal(a.src_offs, b.src_offs) and a.src_params.shape == b.src_params.shape
    assert np.allclose(a.src_params, b.src_params, rtol=1e-15, atol=0)
    bp, bz, bs = b.flat_jets()
    assert a.njets == b.njets and np.array_equal(a.jet_sizes, bs) and np.array_equal(a.jet_zoom_pad, bz)
    assert np.allclose(a.jet_params, bp, rtol=1e-15, atol=0)
    for u, v in ((a.jet_image, b.jet_image), (a.jet_row0, b.jet_row0), (a.jet_col0, b.jet_col0)):
        assert np.array_equal(u, v)


def test_native_ppdisk_packer_matches_python():
    """rsb_ppdisk_pack against the Python replay of generate_proto_set's chain (ppdisks.py:74-76 +
    create_proto's draws): stream position equal, integer fields and mask parameters bit-equal, scales
    to float32 rounding of a last-bit difference in exp()."""
    from radiosim_b200.ppdisks import ppdisks as pp

    for seed, n, ar, rr, sr in [(42, 300, (0, 180), (3, 15), (0.1, 1.0)), (7, 200, (0.0, 180.0), (3.0, 15.0), (0.5, 1.0)),
                                (9, 100, (10, 90), (2, 5), (0.1, 0.5))]:
        np.random.seed(seed)
        packed = []
        for _ in range(n):
            alpha = np.random.randint(*ar)
            ratio = np.random.randint(*rr) / np.random.randint(*rr)
            size_ratio = np.random.uniform(*sr)
            packed.append(pp.pack_proto(512, alpha, ratio, size_ratio))
        ref_next = np.random.rand()
        rc, ri = pp._records(packed)
        np.random.seed(seed)
        comps, images = pp.pack_protos_native(n, 512, ar, rr, sr)
        assert np.random.rand() == ref_next
        assert len(comps) == len(rc) and len(images) == n
        for f in ("comp0", "ncomp", "lo", "hi", "new_size"):
            assert np.array_equal(images[f], ri[f]), f
        for f in ("ca", "sa", "a2_in", "b2_in", "a2_out", "b2_out", "sigma", "image"):
            assert np.array_equal(comps[f], rc[f]), f
        assert np.allclose(comps["scale"], rc["scale"], rtol=2e-7, atol=0)
    assert pp.pack_protos_native(3, 128, (0.5, 180), (3, 15), (0.1, 1)) is None


@pytest.mark.parametrize("threads,factor", [("1", None), ("3", None), ("8", None), ("4", "1.05"), ("8", "1.3")])
def test_legacy_normal_fill_equals_numpy(threads, factor, monkeypatch):
    """``rsb_legacy_normal_fill`` (add_noise's ``np.random.normal(size=...)``, utils.py:277,291): the
    generator and the polar method's rejection run on the calling thread, the log / sqrt on workers --
    values (as float32), cached second variate and stream position equal numpy's for even and odd
    counts, with and without a cached variate at entry, across block boundaries."""
    import ctypes as C

    from radiosim_b200 import _lib as L

    monkeypatch.setenv("RSB_HOST_THREADS", threads)
    if factor is not None:      # blocks sized above the expected requirement: the last one overshoots and
        monkeypatch.setenv("RSB_DEBUG_FILL_FACTOR", factor)   # the generator is rewound behind its last used candidate
    for count, pre in ((0, 0), (1, 1), (5, 0), (65537, 1), (2 * (1 << 18) + 70001, 0), (3 * (1 << 18) * 2 + 7, 3)):
        np.random.seed(7)
        for _ in range(pre):
            np.random.normal()
        name, key, pos, hg, cg = np.random.get_state()
        key = np.ascontiguousarray(key, dtype=np.uint32).copy()
        c_pos, c_has, c_c = C.c_int32(int(pos)), C.c_int32(int(hg)), C.c_double(float(cg))
        out = np.empty(max(count, 1), np.float32)
        L.check(L.lib().rsb_legacy_normal_fill(key.ctypes.data, C.byref(c_pos), C.byref(c_has), C.byref(c_c),
                                               count, out.ctypes.data), "rsb_legacy_normal_fill")
        ref = np.random.normal(size=count).astype(np.float32)
        after = np.random.get_state()
        assert np.array_equal(out[:count], ref), (count, pre)
        assert c_pos.value == after[2] and c_has.value == after[3] and c_c.value == after[4]
        assert np.array_equal(key, after[1])
