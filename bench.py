#!/usr/bin/env python
"""Headline benchmark: images/s for 256x256 jet batches (BASELINE.json config 5).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one pass of the jet hot path (splat -> zoom -> normalise) over one
batch of pre-packed jets per GPU.  Parameters come from the reference's own RNG
chain (radiosim_b200.jets.packer, per-rank seed 1337+rank), so the live-component
count, widths and the ~70 % zoom rate are the reference's.

* ``value``  -- device-resident parameters, float32 images written to an HBM ring
  buffer (batch x 256 KiB >> L2), CUDA events on the launching stream, max over ranks.
* ``e2e``    -- same metric through the host-buffer C-ABI call
  (``rsb_jet_render_host``): pinned host parameters in, pinned host images out,
  copies inside the timed region.
* ``roofline`` -- HBM view (algorithmic bytes / launch time vs the measured copy
  bandwidth) plus the governing SFU view (``sfu_*``: algorithmic exp count vs the
  MUFU.EX2 rate measured on this very device).
* ``cpu_baseline`` / ``--impl reference`` -- the reference's CPU algorithm (numpy
  restatement in ``oracle/``; /root/reference does not exist on the GPU box) on the
  host cores, one process per core.
"""

from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SIZE = 256
COMPS = (3, 10)
SLOTS = 2 * COMPS[1] - 1
METRIC = "images/sec (whole box) for 256x256 jet batches"
UNIT = "images/s"
NCU_DRAM_BYTES_PER_IMAGE = 248657  # measured, see profiles/README.md
WORKLOAD = "C5: jet 256x256, default jet params (num_jet_components [3,10], scaling normalize)"


# ----------------------------------------------------------------------------- utils
def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled during the timed region."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None
        self.t = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.t = threading.Thread(target=self._pump, daemon=True)
        self.t.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        inwin = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.15] or [r for (_, r) in self.rows]
        sm, smax, reasons = [], [], set()
        for r in inwin:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                  "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)),
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------- CPU reference arm
def _cpu_worker(args):
    seed, n = args
    from oracle import jet as ojet

    np.random.seed(seed)
    t = time.perf_counter()
    ojet.create_jet((n, 3, SIZE, SIZE), COMPS, "normalize", "list")
    return time.perf_counter() - t


def cpu_reference_rate(images_per_proc: int, procs: int, seed0: int = 4242):
    """images/s of the reference algorithm (oracle port), one process per core."""
    import multiprocessing as mp

    ctx = mp.get_context("fork")
    t = time.perf_counter()
    with ctx.Pool(procs) as pool:
        pool.map(_cpu_worker, [(seed0 + i, images_per_proc) for i in range(procs)])
    dt = time.perf_counter() - t
    return images_per_proc * procs / dt, dt


def run_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    # bounded sample: ~28 ms per 256x256 jet per core; keep the whole run near 2 minutes
    per = max(4, min(200, int(120.0 / max(args.steps, 1) / 0.030)))
    for _ in range(min(args.warmup, 1)):
        cpu_reference_rate(4, cores)
    t_tot, n_tot = 0.0, 0
    for _ in range(args.steps):
        rate, dt = cpu_reference_rate(per, cores)
        t_tot += dt
        n_tot += per * cores
    value = n_tot / t_tot
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_tot / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_step": per * cores,
                   "note": "reference CPU algorithm (numpy port in oracle/), host cores only"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{per} images x {cores} processes per step"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------- CUDA arm
def run_cuda(args):
    rank, world = env_int("RANK", 0), env_int("WORLD_SIZE", 1)
    local = env_int("LOCAL_RANK", 0)
    dist = None
    near = []
    if world > 1:
        from radiosim_b200.sharder import bind_near_gpu

        near = bind_near_gpu(local)  # host buffers on the GPU's own socket (matters for the e2e leg)
        import torch
        import torch.distributed as dist_mod

        torch.cuda.set_device(local)
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod

    from radiosim_b200 import _lib as L
    from radiosim_b200.jets.packer import pack_jets

    L.require_device()
    L.set_device(local)
    lib = L.lib()

    batch = args.batch
    np.random.seed(1337 + rank)
    t_pack = time.perf_counter()
    jb = pack_jets(batch, SIZE, COMPS, "normalize")
    pack_rate = batch / (time.perf_counter() - t_pack)
    live = int((jb.params[:, :, 0] != 0).sum())
    zoom_frac = float((jb.zoom_pad >= 0).mean())

    flags = L.JET_NORMALIZE
    d_params = L.dev_from(jb.params)
    d_pad = L.dev_from(jb.zoom_pad)
    d_img = L.DeviceBuffer(batch * SIZE * SIZE * 4)
    d_margin = L.DeviceBuffer(batch * 4)
    d_peak = L.DeviceBuffer(batch * 8)
    wsz = C.c_size_t(0)
    L.check(lib.rsb_jet_workspace_bytes(batch, SLOTS, SIZE, C.byref(wsz)))
    d_work = L.DeviceBuffer(max(wsz.value, 16))
    stream = C.c_void_p()
    L.check(lib.rsb_stream_create(C.byref(stream)))
    ev = [C.c_void_p() for _ in range(2)]
    for e in ev:
        L.check(lib.rsb_event_create(C.byref(e)))

    def step():
        L.check(lib.rsb_jet_render(d_params.ptr, d_pad.ptr, batch, SLOTS, SIZE, flags, d_img.ptr,
                                   d_margin.ptr, d_peak.ptr, d_work.ptr, d_work.nbytes, stream),
                "rsb_jet_render")

    def barrier():
        L.check(lib.rsb_stream_sync(stream))
        L.device_sync()
        if dist is not None:
            dist.barrier()

    sfu = C.c_double(0.0)
    L.check(lib.rsb_bench_sfu(4096, C.byref(sfu)), "rsb_bench_sfu")

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    launches0 = L.launch_count()
    t0 = time.perf_counter()
    L.check(lib.rsb_event_record(ev[0], stream))
    for _ in range(args.steps):
        step()
    L.check(lib.rsb_event_record(ev[1], stream))
    ms = C.c_float(0.0)
    L.check(lib.rsb_event_elapsed_ms(ev[0], ev[1], C.byref(ms)))
    barrier()
    t1 = time.perf_counter()
    launches = L.launch_count() - launches0
    clocks = sampler.summary(t0, t1)
    elapsed_ms = float(ms.value)

    # ---- e2e: host-buffer C-ABI call, pinned buffers, copies inside the timed region
    eb = min(args.e2e_batch, batch)
    h_params = L.pinned_empty((eb, SLOTS, 6), np.float64)
    h_params[:] = jb.params[:eb]
    h_pad = L.pinned_empty((eb,), np.int32)
    h_pad[:] = jb.zoom_pad[:eb]
    h_img = L.pinned_empty((eb, SIZE, SIZE), np.float32)
    h_margin = L.pinned_empty((eb,), np.int32)
    h_peak = L.pinned_empty((eb,), np.float64)

    def e2e_step():
        L.check(lib.rsb_jet_render_host(h_params.ctypes.data, h_pad.ctypes.data, eb, SLOTS, SIZE,
                                        flags, h_img.ctypes.data, h_margin.ctypes.data,
                                        h_peak.ctypes.data), "rsb_jet_render_host")

    for _ in range(3):
        e2e_step()
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    te = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    L.device_sync()
    e2e_s = time.perf_counter() - te
    sampler.stop()

    # sanity of the result actually read back
    assert abs(float(h_img[0].max()) - 1.0) < 1e-5 and np.isfinite(h_peak).all()

    if dist is not None:
        import torch

        t = torch.tensor([elapsed_ms, e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms, e2e_s = float(t[0]), float(t[1])
        cnt = torch.tensor([float(live), float(launches)], device="cuda", dtype=torch.float64)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        live_all, launches_all = float(cnt[0]), int(cnt[1])
    else:
        live_all, launches_all = float(live), launches

    if rank == 0:
        value = world * batch * args.steps / (elapsed_ms * 1e-3)
        e2e_value = world * eb * e2e_steps / e2e_s
        hbm_peak, peak_src = measured_peaks()
        # algorithmic traffic per image: one float32 image out + parameters in
        bytes_per_img = 4 * SIZE * SIZE + SLOTS * 6 * 8 + 4 + 4 + 8
        per_gpu_rate = batch * args.steps / (elapsed_ms * 1e-3)
        achieved_gbs = per_gpu_rate * bytes_per_img / 1e9
        exps_per_img = SIZE * SIZE * (live_all / (world * batch))
        sfu_achieved = per_gpu_rate * exps_per_img
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {
                "workload": WORKLOAD, "batch_per_gpu": batch, "img_size": SIZE,
                "slots": SLOTS, "mean_live_components": live_all / (world * batch),
                "zoom_fraction": zoom_frac, "parallelism": f"dp{world} (independent shards, no collective)",
                "cpu_binding": f"rank 0 bound to {len(near)} CPUs near its GPU" if near else "none",
                "l2_policy": f"inputs larger than L2: {batch * SIZE * SIZE * 4 / 2**20:.0f} MiB written per step",
                "seed": "1337+rank", "packer_images_per_s_per_core": pack_rate,
            },
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT,
                    "h2d_bytes_per_step": eb * (SLOTS * 6 * 8 + 4),
                    "d2h_bytes_per_step": eb * (SIZE * SIZE * 4 + 12),
                    "batch_per_gpu": eb, "steps": e2e_steps,
                    "call": "rsb_jet_render_host (pinned host buffers, float32 images)"},
            "gpu_launches": launches_all,
            "roofline": {
                "bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved_gbs / hbm_peak,
                # dram__bytes_read+write of one launch from the ncu --set full capture, scaled to
                # this launch's batch (profiles/r01_jet2_kernel_ncu_full_raw.csv: 1.0185 GB / 4096 jets)
                "traffic": NCU_DRAM_BYTES_PER_IMAGE * batch, "traffic_unit": "bytes per launch",
                "traffic_source": "ncu --set full, 4096-jet launch, scaled by batch",
                "peak_source": peak_src,
                "kernel": "rsb::jet::v2::jet2_kernel<2,2>", "algorithmic_bytes_per_image": bytes_per_img,
                "algorithmic_bytes_per_launch": bytes_per_img * batch,
                "governing": "sfu",
                "sfu_achieved_ex2_per_s": sfu_achieved, "sfu_peak_ex2_per_s": sfu.value,
                "sfu_frac": sfu_achieved / sfu.value if sfu.value else None,
                "sfu_peak_source": "rsb_bench_sfu on this device (MUFU.EX2, best of 5)",
                "algorithmic_ex2_per_image": exps_per_img,
            },
        }
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            per = 12
            rate, dt = cpu_reference_rate(per, cores)
            line["cpu_baseline"] = {
                "value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": f"{per * cores} jets 256x256 ({per} per process x {cores} processes, "
                          f"{dt:.1f} s wall), oracle/jet.py create_jet"}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--batch", type=int, default=16384, help="jets per GPU per step")
    ap.add_argument("--e2e-batch", type=int, default=4096)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
