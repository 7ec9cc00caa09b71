"""Throughput of the other BASELINE.json configs (not the headline bench): device time of the
render entry points on pre-packed inputs, CUDA events on the launching stream, plus the
oracle on a small sample of the same workload.  One JSON line per config."""
import ctypes as C, json, sys, time
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L

lib = L.lib()


def timed(fn, reps=5, warm=2):
    ev0, ev1 = C.c_void_p(), C.c_void_p()
    L.check(lib.rsb_event_create(C.byref(ev0))); L.check(lib.rsb_event_create(C.byref(ev1)))
    for _ in range(warm):
        fn()
    L.device_sync()
    L.check(lib.rsb_event_record(ev0, None))
    for _ in range(reps):
        fn()
    L.check(lib.rsb_event_record(ev1, None))
    ms = C.c_float()
    L.check(lib.rsb_event_elapsed_ms(ev0, ev1, C.byref(ms)))
    return ms.value / reps


def survey(size=1024, ns=2000, n=4):
    from radiosim_b200.survey.survey import pack_survey, JET_RECT
    np.random.seed(1337)
    b = pack_survey(n, size, ns, (0, 1, 2), True)   # gaussians + points (sub-jets are the jet kernel)
    total = b.src_params.shape[0]
    need = C.c_size_t()
    L.check(lib.rsb_survey_workspace_bytes(n, size, total, C.byref(need)))
    d_src, d_offs = L.dev_from(b.src_params), L.dev_from(b.src_offs)
    d_rects, d_joffs = L.dev_from(np.zeros(1, JET_RECT)), L.dev_from(np.zeros(n + 1, np.int32))
    d_pix = L.DeviceBuffer(16); d_w = L.DeviceBuffer(need.value)
    d_x = L.DeviceBuffer(n * size * size * 4); d_y = L.DeviceBuffer(n * 4 * size * size * 4)
    ms = timed(lambda: L.check(lib.rsb_survey_render(d_src.ptr, d_offs.ptr, total, d_rects.ptr, d_joffs.ptr,
                                                     d_pix.ptr, n, size, 0, d_x.ptr, d_y.ptr, d_w.ptr, d_w.nbytes, None)))
    return dict(config="C4 survey 1024x1024, 2000 gaussian/point sources per image", images=n,
                ms_per_image=ms / n, images_per_s=n / ms * 1e3,
                algorithmic_ex2_per_s=total * size * size / (ms * 1e-3),
                hbm_bytes_per_image=size * size * 4 * 11, sources_per_image=total / n)


def mojave(size=256, n=1000):
    from radiosim_b200.mojave import mojave as m
    rng = np.random.default_rng(1337)
    kids = rng.spawn(n)
    fns = [m._recipe_compact, m._recipe_one_jet, m._recipe_two_jet]
    t = time.perf_counter()
    src = [fns[i % 3](kids[i], size) for i in range(n)]
    pack_s = time.perf_counter() - t
    t = time.perf_counter()
    out = m.render_sources(src, size, f64=False)
    first = time.perf_counter() - t
    t = time.perf_counter()
    out = m.render_sources(src, size, f64=False)
    wall = time.perf_counter() - t
    return dict(config="C2 mojave 256x256 (1000-image sample of the 10k batch, 1/3 per class)", images=n,
                images_per_s_incl_copies=n / wall, host_recipe_images_per_s=n / pack_s)


def ppdisk(size=512, n=20):
    from radiosim_b200.ppdisks import ppdisks as pp
    np.random.seed(1337)
    packed = []
    for _ in range(n):
        a = np.random.randint(0, 180); r = np.random.randint(3, 15) / np.random.randint(3, 15)
        packed.append(pp.pack_proto(size, a, r, np.random.uniform(0.1, 1)))
    pp.render_protos(size, packed)
    t = time.perf_counter()
    pp.render_protos(size, packed)
    wall = time.perf_counter() - t
    return dict(config="C3 ppdisk 512x512, default ppdisk TOML ranges", images=n, images_per_s_incl_copies=n / wall)


if __name__ == "__main__":
    L.require_device()
    for f in (survey, mojave, ppdisk):
        print(json.dumps(f()))
