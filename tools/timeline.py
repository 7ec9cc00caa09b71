"""Debug: per-phase clock64 timeline of jet2_kernel (build with RSB_NVCC_EXTRA=-DRSB_TIMELINE)."""
import ctypes as C, sys
import numpy as np
sys.path.insert(0, ".")
from radiosim_b200 import _lib as L
from radiosim_b200.jets.packer import pack_jets
lib = L.lib()
n, size = 148 * 8, 256
np.random.seed(1)
b = pack_jets(n, size, (3, 10), "normalize")
d_p, d_z = L.dev_from(b.params), L.dev_from(b.zoom_pad)
d_img = L.DeviceBuffer(n * size * size * 4); d_m = L.DeviceBuffer(n * 4); d_k = L.DeviceBuffer(n * 8)
tl = L.DeviceBuffer(n * 2 * 96 * 8)
L.check(lib.rsb_memset(tl.ptr, 0, tl.nbytes, None))
lib.rsb_debug_set_timeline.argtypes = [C.c_void_p]
for rep in range(2):
    L.check(lib.rsb_debug_set_timeline(tl.ptr))
    L.check(lib.rsb_jet_render(d_p.ptr, d_z.ptr, n, 19, size, 1, d_img.ptr, d_m.ptr, d_k.ptr, None, 0, None))
    L.device_sync()
t = tl.download((n, 2, 96), np.int64)
zoom = b.zoom_pad >= 0
names = ["A->tiles done (this warp)", "->C (cluster sync / exch)", "->table built", "->H pass done (warp)", "->F barrier", "->halo", "->V pass done", "->best exch", "->stores done"]
for label, sel in [("zoom", zoom), ("plain", ~zoom)]:
    tt = t[sel]
    print(label, "images", sel.sum())
    if label == "zoom":
        idx = [0,1,2,3,4,5,6,7,8,9]
    else:
        idx = [0,1,2,9]
    for a_, b_ in zip(idx[:-1], idx[1:]):
        d = (tt[:, :, b_] - tt[:, :, a_]).astype(np.float64)
        print(f"  stamp {a_}->{b_}: mean {d.mean():9.0f}  p90 {np.percentile(d,90):9.0f}  rank0 {d[:,0].mean():9.0f} rank1 {d[:,1].mean():9.0f}")
    d = (tt[:, :, 9] - tt[:, :, 0]).astype(np.float64)
    print(f"  total per image: mean {d.mean():.0f}")
tz=t[zoom]
d=(tz[:,:,10]-tz[:,:,3]).astype(float); print("H: taps load", d.mean())
d=(tz[:,:,4]-tz[:,:,10]).astype(float); print("H: row loop", d.mean(), "rows by this warp", tz[:,:,11].mean(), "margin mean", b.zoom_pad[zoom].mean())
ok=(t[:,:,12]>0)&(t[:,:,13]>0)
print("warp0 coef build", (t[:,:,12]-t[:,:,0])[ok].mean(), "warp1 label job", (t[:,:,13]-t[:,:,0])[ok[:,0],0].mean())
# gap between images on the same CTA: next image's stamp0 - this image's stamp9
ncl = 74
gaps = []
for i in range(n - ncl):
    gaps.append(t[i + ncl, 0, 0] - t[i, 0, 9])
print("gap stamp9 -> next stamp0 (barrier A wait etc):", np.mean(gaps))
print("phase 1: slot-tile hits per CTA: zoom", t[zoom][:,:,14].mean(), "plain", t[~zoom][:,:,14].mean(), " tiles computed per CTA: zoom", t[zoom][:,:,15].mean(), "plain", t[~zoom][:,:,15].mean())
print("live slots: ", "see packer")
for label, sel in [("zoom", zoom), ("plain", ~zoom)]:
    tt = t[sel].astype(float)
    print(label, "warp2: fetch+cull", tt[:,:,16].mean(), "hit loop", tt[:,:,17].mean(), "epilogue", tt[:,:,18].mean(), "tiles", tt[:,:,19].mean(), "hits", tt[:,:,20].mean(), "| stores", tt[:,:,21].mean(), "shfl-wait", tt[:,:,22].mean(), "order-lds+skip", tt[:,:,23].mean())
tz = t[zoom].astype(float)
w = tz[:, :, 32:96].reshape(len(tz), 2, 16, 4)
s0 = tz[:, :, 0][:, :, None]
print("zoom per warp: start-after-A / finish-after-A / tiles / hits")
for k in range(16):
    print(f"  warp {k:2d}: start {np.mean(w[:,:,k,3]-s0[:,:,0]):8.0f} finish {np.mean(w[:,:,k,0]-s0[:,:,0]):8.0f} tiles {w[:,:,k,1].mean():5.2f} hits {w[:,:,k,2].mean():6.2f}")
print("  last finisher mean:", np.mean(np.max(w[:,:,:,0]-s0, axis=2)), " first finisher mean:", np.mean(np.min(w[:,:,:,0]-s0, axis=2)))
bt = t[:, :, 24:29].astype(float)
okb = bt[:, :, 4] > 0
print("builder sections (slots / spans / dlo / lists):", [float(np.mean((bt[:,:,k+1]-bt[:,:,k])[okb])) for k in range(4)])
print("hits per CTA by rank (zoom):", w[:, 0, :, 2].sum(axis=1).mean(), w[:, 1, :, 2].sum(axis=1).mean(), " tiles:", w[:, 0, :, 1].sum(axis=1).mean(), w[:, 1, :, 1].sum(axis=1).mean())
fin = w[:, :, 2:, 0] - s0
print("worker finish by rank (zoom): mean", fin[:, 0].mean(), fin[:, 1].mean(), " max", fin[:, 0].max(axis=1).mean(), fin[:, 1].max(axis=1).mean())
for rk in (0, 1):
    fw = 2 if rk == 0 else 1
    cost = w[:, rk, fw:, 1] + w[:, rk, fw:, 2]
    tm = (w[:, rk, fw:, 0] - w[:, rk, fw:, 3])
    print(f"rank {rk}: dealt cost max/mean {np.mean(cost.max(axis=1) / cost.mean(axis=1)):.3f}  time max/mean {np.mean(tm.max(axis=1) / tm.mean(axis=1)):.3f}  corr(cost,time) {np.corrcoef(cost.ravel(), tm.ravel())[0,1]:.3f}")
    # least squares time ~ a*tiles + b*hits
    A = np.stack([w[:, rk, fw:, 1].ravel(), w[:, rk, fw:, 2].ravel()], axis=1)
    coef, *_ = np.linalg.lstsq(A, tm.ravel(), rcond=None)
    print("   time ~ %.0f * tiles + %.0f * hits" % tuple(coef))
nc = t[zoom][:, :, 29]
print("borderline candidates per CTA (zoom): mean %.3f, fraction of CTAs with any %.3f, max %d" % (nc.mean(), (nc > 0).mean(), nc.max()))
sel = (nc > 0)
d12 = (t[zoom][:, :, 2] - t[zoom][:, :, 1]).astype(float)
print("stamp 1->2 with candidates %.0f, without %.0f" % (d12[sel].mean() if sel.any() else 0, d12[~sel].mean()))
tz3 = t[zoom].astype(float)
print("table: 2->sync %.0f, sync->loop done %.0f, loop done->3 %.0f" % ((tz3[:,:,30]-tz3[:,:,2]).mean(), (tz3[:,:,31]-tz3[:,:,30]).mean(), (tz3[:,:,3]-tz3[:,:,31]).mean()))
for rk in (0, 1):
    fw = 2 if rk == 0 else 1
    fin_w = w[:, rk, fw:, 0] - s0[:, rk, :]
    last = fin_w.argmax(axis=1) + fw
    print(f"rank {rk}: last finisher warp histogram", np.bincount(last, minlength=16).tolist())
    hits_w = w[:, rk, fw:, 2]; tiles_w = w[:, rk, fw:, 1]
    idx = np.arange(len(last))
    print(f"   last finisher: hits {hits_w[idx, last - fw].mean():.1f} tiles {tiles_w[idx, last - fw].mean():.2f}  vs mean hits {hits_w.mean():.1f} tiles {tiles_w.mean():.2f};  finish spread by scheduler (warp%4):",
          [round(float(fin_w[:, [k for k in range(16 - fw) if (k + fw) % 4 == m]].mean())) for m in range(4)])
