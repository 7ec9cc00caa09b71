"""ctypes binding of ``librsb200.so`` (declared in ``include/radiosim_b200.h``).

There is no CPU fallback: importing this module without the built library, or
calling a compute entry point without a CUDA device, raises.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librsb200.so")

JET_NORMALIZE = 1
JET_OUT_F64 = 2


class RsbError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m radiosim_b200.build` "
            "(needs nvcc; there is no CPU fallback)"
        )
    return C.CDLL(LIB_PATH)


_lib = _load()

_vp, _i32, _u32, _sz = C.c_void_p, C.c_int, C.c_uint, C.c_size_t
_SIGS = {
    "rsb_abi_version": ([], _i32),
    "rsb_last_error": ([], C.c_char_p),
    "rsb_device_count": ([C.POINTER(_i32)], _i32),
    "rsb_set_device": ([_i32], _i32),
    "rsb_get_device": ([C.POINTER(_i32)], _i32),
    "rsb_debug_scrub_shared": ([C.c_uint, _vp], _i32),
    "rsb_device_info": ([_i32, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_sz)], _i32),
    "rsb_malloc": ([C.POINTER(_vp), _sz], _i32),
    "rsb_free": ([_vp], _i32),
    "rsb_malloc_host": ([C.POINTER(_vp), _sz], _i32),
    "rsb_free_host": ([_vp], _i32),
    "rsb_memset": ([_vp, _i32, _sz, _vp], _i32),
    "rsb_memcpy_h2d": ([_vp, _vp, _sz, _vp], _i32),
    "rsb_memcpy_d2h": ([_vp, _vp, _sz, _vp], _i32),
    "rsb_memcpy2d_d2h": ([_vp, _sz, _vp, _sz, _sz, _sz, _vp], _i32),
    "rsb_stream_create": ([C.POINTER(_vp)], _i32),
    "rsb_stream_destroy": ([_vp], _i32),
    "rsb_stream_sync": ([_vp], _i32),
    "rsb_device_sync": ([], _i32),
    "rsb_event_create": ([C.POINTER(_vp)], _i32),
    "rsb_event_destroy": ([_vp], _i32),
    "rsb_event_record": ([_vp, _vp], _i32),
    "rsb_stream_wait_event": ([_vp, _vp], _i32),
    "rsb_event_elapsed_ms": ([_vp, _vp, C.POINTER(C.c_float)], _i32),
    "rsb_launch_count": ([C.POINTER(C.c_uint64)], _i32),
    "rsb_bench_sfu": ([_i32, C.POINTER(C.c_double)], _i32),
    "rsb_jet_workspace_bytes": ([_i32, _i32, _i32, C.POINTER(_sz)], _i32),
    "rsb_jet_render": ([_vp, _vp, _i32, _i32, _i32, _u32, _vp, _vp, _vp, _vp, _sz, _vp], _i32),
    "rsb_jet_render_var": ([_vp, _vp, _vp, _vp, _i32, _i32, _i32, _u32, _vp, _vp, _vp, _vp], _i32),
    "rsb_jet_planes": ([_vp, _vp, _vp, _vp, _i32, _i32, _i32, _u32, _vp, _vp], _i32),
    "rsb_jet_render_static": ([_vp, _vp, _i32, _i32, _i32, _u32, _vp, _vp, _vp, _vp], _i32),
    "rsb_survey_workspace_bytes": ([_i32, _i32, _i32, C.POINTER(_sz)], _i32),
    "rsb_survey_render": ([_vp, _vp, _i32, _vp, _vp, _vp, _i32, _i32, _u32, _vp, _vp, _vp, _sz, _vp], _i32),
    "rsb_ppdisk_workspace_bytes": ([_i32, _i32, C.POINTER(_sz)], _i32),
    "rsb_ppdisk_masks": ([_vp, _i32, _vp, _vp], _i32),
    "rsb_ppdisk_render": ([_vp, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _sz, _vp], _i32),
    "rsb_mojave_workspace_bytes": ([_i32, _i32, C.POINTER(_sz)], _i32),
    "rsb_mojave_render": ([_vp, _i32, _vp, _vp, _i32, _u32, _vp, _vp, _sz, _vp], _i32),
    "rsb_mojave_pack": ([_vp, _i32, _vp, _i32, C.c_longlong, _i32, _vp, _i32, _vp, _vp, _i32, C.POINTER(_i32)], _i32),
    "rsb_mojave_blobs": ([_vp, _i32, _vp, _i32, _i32, _vp, _vp, _vp], _i32),
    "rsb_swirl": ([_vp, _i32, _i32, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, _vp, _vp], _i32),
    "rsb_pcg64_spawn_states": ([_vp, _i32, _vp, _i32, C.c_longlong, _i32, _vp], _i32),
    "rsb_pcg64_draws": ([_vp, _i32, C.c_double, C.c_double, _i32, _vp], _i32),
    "rsb_source_margin": ([_vp, _i32, C.c_double, _vp, _vp], _i32),
    "rsb_resize_linear": ([_vp, _i32, _i32, _sz, _i32, _i32, _vp, _i32, _vp], _i32),
    "rsb_jet_render_host": ([_vp, _vp, _i32, _i32, _i32, _u32, _vp, _vp, _vp], _i32),
    "rsb_noise_workspace_bytes": ([_i32, _i32, C.POINTER(_sz)], _i32),
    "rsb_add_noise": ([_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _sz, _vp], _i32),
    "rsb_legacy_normal_fill": ([_vp, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(C.c_double), C.c_longlong, _vp], _i32),
    "rsb_ppdisk_pack": ([_vp, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(C.c_double), _i32, _i32, _i32, _i32, _i32,
                         _i32, C.c_double, C.c_double, _vp, _vp, _vp, _i32, _vp, C.POINTER(_i32)], _i32),
    "rsb_survey_pack": ([_vp, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(C.c_double), _i32, _i32, _i32, _vp, _i32,
                         _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.POINTER(_i32)], _i32),
    "rsb_jet_pack": ([_vp, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(C.c_double), _i32, _i32, _i32, _i32,
                      _vp, _vp, _vp, _vp, _vp], _i32),
}
for _name, (_args, _res) in _SIGS.items():
    if hasattr(_lib, _name):
        _f = getattr(_lib, _name)
        _f.argtypes = _args
        _f.restype = _res

EXPORTS = tuple(_SIGS)


def lib():
    return _lib


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = _lib.rsb_last_error().decode(errors="replace")
        raise RsbError(f"{what or 'librsb200'} failed ({code}): {msg}")


def device_count() -> int:
    n = _i32(0)
    if _lib.rsb_device_count(C.byref(n)) != 0:
        return 0
    return n.value


def require_device() -> None:
    if device_count() < 1:
        raise RsbError(
            "radiosim_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback: "
            + _lib.rsb_last_error().decode(errors="replace")
        )


def set_device(dev: int) -> None:
    check(_lib.rsb_set_device(int(dev)), "rsb_set_device")
    _CUR_DEV[0] = int(dev)


def launch_count() -> int:
    n = C.c_uint64(0)
    check(_lib.rsb_launch_count(C.byref(n)))
    return n.value


def device_sync() -> None:
    check(_lib.rsb_device_sync(), "rsb_device_sync")


# cudaMalloc / cudaFree cost ~0.1 ms each and synchronise the device; the drop-in functions need a
# dozen small buffers per call, so freed blocks are kept (per device, by size class) for the next
# call -- up to RSB_DEVICE_POOL_MB (default 8192) in total.
_DPOOL: dict = {}         # (device, size class) -> [ptr]
_DPOOL_BYTES = [0]
_DPOOL_CAP = int(os.environ.get("RSB_DEVICE_POOL_MB", "8192")) << 20
_CUR_DEV = [0]
_POISON = os.environ.get("RSB_POISON_DEVICE_BLOCKS", "0") == "1"


def current_device() -> int:
    """The calling thread's current CUDA device (blocks are pooled per device)."""
    d = _i32(0)
    if _lib.rsb_get_device(C.byref(d)) != 0:
        return _CUR_DEV[0]
    return d.value


def _size_class(nbytes: int) -> int:
    """Allocation size for a request: powers of two up to 1 MiB, then multiples of 1/8 of the
    power of two below (at most 12.5 % over-allocation)."""
    n = max(int(nbytes), 256)
    if n <= (1 << 20):
        return 1 << (n - 1).bit_length()
    step = 1 << ((n - 1).bit_length() - 4)
    return -(-n // step) * step


def device_pool_clear() -> None:
    """Return every cached device block to the driver."""
    for ptrs in _DPOOL.values():
        for ptr in ptrs:
            _lib.rsb_free(ptr)
    _DPOOL.clear()
    _DPOOL_BYTES[0] = 0


class DeviceBuffer:
    """Owning handle of a device allocation (rsb_malloc / rsb_free, pooled)."""

    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        self._cls = _size_class(nbytes)
        self._dev = current_device()
        free = _DPOOL.get((self._dev, self._cls))
        if free:
            self.ptr = free.pop()
            _DPOOL_BYTES[0] -= self._cls
        else:
            p = _vp()
            rc = _lib.rsb_malloc(C.byref(p), self._cls)
            if rc != 0 and _DPOOL:      # out of memory with blocks cached: give them back and retry
                device_pool_clear()
                rc = _lib.rsb_malloc(C.byref(p), self._cls)
            check(rc, "rsb_malloc")
            self.ptr = p.value
        if _POISON:  # test aid: whoever reads a block before writing it gets NaN bit patterns
            check(_lib.rsb_memset(self.ptr, 0xFF, self._cls, None), "rsb_memset")

    def free(self):
        if getattr(self, "ptr", None):
            if _DPOOL_BYTES[0] + self._cls <= _DPOOL_CAP:
                _DPOOL.setdefault((self._dev, self._cls), []).append(self.ptr)
                _DPOOL_BYTES[0] += self._cls
            else:
                _lib.rsb_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def upload(self, arr: np.ndarray, stream=None):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        check(_lib.rsb_memcpy_h2d(self.ptr, arr.ctypes.data, arr.nbytes, stream), "h2d")
        if stream is None:
            device_sync()
        return self

    def download(self, shape, dtype, stream=None, pinned=False) -> np.ndarray:
        """Copy to a fresh host array; ``pinned=True`` puts it in a (pooled) page-locked block,
        which the copy engine fills several times faster than pageable memory."""
        out = pinned_empty(shape, dtype) if pinned else np.empty(shape, dtype=dtype)
        assert out.nbytes <= self.nbytes
        check(_lib.rsb_memcpy_d2h(out.ctypes.data, self.ptr, out.nbytes, stream), "d2h")
        device_sync()
        return out


# Page-locked blocks are expensive to create (the driver pins every page: ~0.3 s per GB) and the
# drop-in functions return arrays backed by them, so released blocks are kept for the next call
# instead of being handed back to the driver -- up to RSB_PINNED_POOL_MB (default 4096) in total.
_POOL: list = []          # [(nbytes, ptr)]
_POOL_CAP = int(os.environ.get("RSB_PINNED_POOL_MB", "4096")) << 20


def _pool_take(nbytes: int):
    best = None
    for i, (sz, _) in enumerate(_POOL):
        if nbytes <= sz <= max(2 * nbytes, nbytes + (1 << 20)) and (best is None or sz < _POOL[best][0]):
            best = i
    return _POOL.pop(best) if best is not None else None


def pinned_pool_clear() -> None:
    """Return every cached page-locked block to the driver."""
    while _POOL:
        _, ptr = _POOL.pop()
        _lib.rsb_free_host(ptr)


class _PinnedOwner:
    def __init__(self, nbytes):
        nbytes = max(int(nbytes), 1)
        got = _pool_take(nbytes)
        if got is not None:
            self.nbytes, self.ptr = got
            return
        p = _vp()
        check(_lib.rsb_malloc_host(C.byref(p), nbytes), "rsb_malloc_host")
        self.ptr = p.value
        self.nbytes = nbytes

    def __del__(self):
        try:
            if self.ptr:
                if sum(sz for sz, _ in _POOL) + self.nbytes <= _POOL_CAP:
                    _POOL.append((self.nbytes, self.ptr))
                else:
                    _lib.rsb_free_host(self.ptr)
                self.ptr = None
        except Exception:
            pass


def pinned_empty(shape, dtype) -> np.ndarray:
    """numpy array backed by page-locked host memory.

    The allocation is owned by the ctypes buffer at the ROOT of the array's base
    chain, so it outlives every view numpy derives from the array.
    """
    dtype = np.dtype(dtype)
    count = int(np.prod(shape))
    nbytes = count * dtype.itemsize
    owner = _PinnedOwner(nbytes)
    buf = (C.c_char * max(nbytes, 1)).from_address(owner.ptr)
    buf._rsb_owner = owner
    return np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)


class CopyStream:
    """A second stream for device-to-host copies that overlap the default stream's kernels:
    ``after_default()`` makes its next operations wait for everything enqueued on the default stream
    so far."""

    _cache: dict = {}

    @classmethod
    def get(cls):
        dev = current_device()
        if dev not in cls._cache:
            cls._cache[dev] = cls()
        return cls._cache[dev]

    def __init__(self):
        s, e = _vp(), _vp()
        check(_lib.rsb_stream_create(C.byref(s)), "rsb_stream_create")
        check(_lib.rsb_event_create(C.byref(e)), "rsb_event_create")
        self.stream, self.event = s, e

    def after_default(self):
        check(_lib.rsb_event_record(self.event, None), "rsb_event_record")
        check(_lib.rsb_stream_wait_event(self.stream, self.event), "rsb_stream_wait_event")

    def d2h(self, dst: np.ndarray, dev_ptr: int):
        check(_lib.rsb_memcpy_d2h(dst.ctypes.data, dev_ptr, dst.nbytes, self.stream), "d2h")

    def sync(self):
        check(_lib.rsb_stream_sync(self.stream), "rsb_stream_sync")


def upload_blob(arrays):
    """Several small host arrays -> ONE page-locked staging block, one device block and one
    asynchronous copy on the default stream.  Returns ``(device buffer, [device pointers], staging
    array)``; keep the staging array alive until the stream has passed the copy."""
    arrays = [np.ascontiguousarray(a) for a in arrays]
    offs, total = [], 0
    for a in arrays:
        offs.append(total)
        total += (a.nbytes + 255) & ~255
    total = max(total, 256)
    stage = pinned_empty((total,), np.uint8)
    for a, o in zip(arrays, offs):
        if a.nbytes:
            stage[o : o + a.nbytes] = a.reshape(-1).view(np.uint8)
    buf = DeviceBuffer(total)
    check(_lib.rsb_memcpy_h2d(buf.ptr, stage.ctypes.data, total, None), "h2d")
    return buf, [buf.ptr + o for o in offs], stage


def dev_from(arr: np.ndarray) -> DeviceBuffer:
    arr = np.ascontiguousarray(arr)
    return DeviceBuffer(max(arr.nbytes, 16)).upload(arr)


def parallel_copy(dst: np.ndarray, src: np.ndarray, threads: int = 8) -> None:
    """``dst[...] = src`` for large contiguous arrays, split over a few threads (numpy releases
    the GIL inside the copy; one thread moves ~10-15 GB/s, far less than the memory system can)."""
    if dst.nbytes < (64 << 20) or not (dst.flags.c_contiguous and src.flags.c_contiguous):
        np.copyto(dst, src)
        return
    from concurrent.futures import ThreadPoolExecutor

    d, s_ = dst.reshape(-1), src.reshape(-1)
    step = -(-d.size // threads)
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda i: np.copyto(d[i : i + step], s_[i : i + step]), range(0, d.size, step)))
