"""Plain pinned device-to-host copy rate with 1, 2, 4, ... GPUs copying at the same time (one
process per GPU, 1 GiB blocks, cudaMemcpyAsync + events), and the host's float32 -> float64
widening rate.  The e2e figures of bench.py are bounded by the first; the second says whether
sending float32 and widening on the host could beat sending float64.

    python tools/d2h_ceiling.py [max_gpus]
"""
import ctypes as C, json, multiprocessing as mp, os, sys, time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
GIB = 1 << 30


def worker(dev, start, q, reps=6):
    from radiosim_b200 import _lib as L

    lib = L.lib()
    L.set_device(dev)
    d = L.DeviceBuffer(GIB)
    h = L.pinned_empty((GIB,), np.uint8)
    ev = [C.c_void_p(), C.c_void_p()]
    for e in ev:
        L.check(lib.rsb_event_create(C.byref(e)))
    L.check(lib.rsb_memcpy_d2h(h.ctypes.data, d.ptr, GIB, None))
    L.device_sync()
    start.wait()
    L.check(lib.rsb_event_record(ev[0], None))
    for _ in range(reps):
        L.check(lib.rsb_memcpy_d2h(h.ctypes.data, d.ptr, GIB, None))
    L.check(lib.rsb_event_record(ev[1], None))
    ms = C.c_float()
    L.check(lib.rsb_event_elapsed_ms(ev[0], ev[1], C.byref(ms)))
    q.put((dev, reps * GIB / (ms.value * 1e-3) / 1e9))


def widen_rate(threads):
    from concurrent.futures import ThreadPoolExecutor

    src = np.ones(256 << 20, dtype=np.float32)
    dst = np.empty(256 << 20, dtype=np.float64)
    step = -(-src.size // threads)
    parts = [(dst[i : i + step], src[i : i + step]) for i in range(0, src.size, step)]
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda p: np.copyto(p[0], p[1]), parts))
        t = time.perf_counter()
        list(ex.map(lambda p: np.copyto(p[0], p[1]), parts))
        dt = time.perf_counter() - t
    return dst.nbytes / dt / 1e9


if __name__ == "__main__":
    from radiosim_b200 import _lib as L

    ndev = L.device_count()
    top = min(int(sys.argv[1]) if len(sys.argv) > 1 else ndev, ndev)
    ctx = mp.get_context("spawn")
    n = 1
    while n <= top:
        start, q = ctx.Barrier(n), ctx.Queue()
        ps = [ctx.Process(target=worker, args=(d, start, q)) for d in range(n)]
        for p in ps:
            p.start()
        rates = sorted(q.get() for _ in ps)
        for p in ps:
            p.join()
        print(json.dumps({"gpus_copying": n, "d2h_gb_per_s_per_gpu": [round(r, 1) for _, r in rates],
                          "total_gb_per_s": round(sum(r for _, r in rates), 1)}))
        n *= 2
    for t in (1, 4, 8, 16):
        if t <= (os.cpu_count() or 1):
            print(json.dumps({"host_widen_f32_to_f64_threads": t, "output_gb_per_s": round(widen_rate(t), 1)}))
