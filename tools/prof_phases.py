"""Development aid: per-phase clock64 stamps of the fused stage kernel (CFD3D_PROF=1), cfg2 by default."""
import ctypes as C
import os
import sys

os.environ["CFD3D_PROF"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# the instrumented build (python -m centerfusiondetect3d_b200.build --prof); the product library has no phase clocks
os.environ.setdefault("CFD3D_LIB", os.path.join(ROOT, "centerfusiondetect3d_b200", "libcfd3d_b200_prof.so"))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import bench
from centerfusiondetect3d_b200 import _lib, ops, synthetic as syn

wl = syn.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "cfg2"]
B = int(sys.argv[2]) if len(sys.argv) > 2 else min(wl.B, 32)
CS = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dev = "cuda:0"
keys = bench.STAGE_KEYS + ("calib", "pc_dep")
if wl.with_pillars:
    keys += ("radar_points", "radar_counts", "cam", "calib64", "A_out")
# like bench.py: back-to-back launches over rotating input sets (> 2 x L2 together), no other kernel in between --
# the SM's instruction cache keeps the kernel from launch to launch, the data is cold.  COLD=1: one input set and
# an L2-evicting read between the launches instead (that also evicts the code: every once-per-frame phase gets slower).
COLD = os.environ.get("COLD") == "1"
NSETS = 1 if COLD else 3
sets = []
for i in range(NSETS):
    host = bench.make_host_inputs(wl, B, seed=1234 + i)
    sets.append({k: torch.from_numpy(np.ascontiguousarray(host[k])).to(dev) for k in keys})
cfg = bench.Cfg(wl.K, wl.max_pc_dist)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for w in range(4):
    bench.run_stage(ops, wl, sets[w % NSETS], cfg)
L = _lib.lib()
L.cf_debug_read_prof.argtypes = [C.c_void_p, C.c_int]
names = ["start->sync0", "passA", "sync1", "tau", "passB", "sync2", "select", "sync3", "tail", "assoc", "sync4", "paint"]
order = [0, 1, 2, 3, 4, 5, 6, 7, 8, 24, 9, 10, 11]
for rep in range(6):
    d = sets[rep % NSETS]
    if COLD:
        _ = flush.view(torch.int32).sum()  # evict by READING: leaves clean lines, like the heads of a real pipeline
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    bench.run_stage(ops, wl, d, cfg)
    e1.record()
    torch.cuda.synchronize()
    plan = ops.last_plan()
    CS = plan["cluster_size"]
    n_cta = plan["n_clusters"] * CS
    buf = np.zeros((n_cta, 48), dtype=np.int64)
    L.cf_debug_read_prof(buf.ctypes.data, n_cta)
    dt = np.diff(buf[:, order], axis=1) / 1.965e3  # us at 1965 MHz
    print("rep %d  call %.1f us   per-phase us over %d CTAs (min / mean / max); total mean %.1f" % (
        rep, e0.elapsed_time(e1) * 1e3, n_cta, (buf[:, 11] - buf[:, 0]).mean() / 1.965e3))
    for i, nm in enumerate(names):
        print("   %-14s %7.2f %7.2f %7.2f" % (nm, dt[:, i].min(), dt[:, i].mean(), dt[:, i].max()))
    az = np.diff(buf[:, [24, 25, 26, 9]], axis=1) / 1.965e3
    print("   assoc finalize (thread 0): record ready %.2f  broadcast %.2f" % tuple((np.diff(buf[:, [26, 34, 9]], axis=1) / 1.965e3).mean(axis=0)))
    print("   assoc scans (warp 0): wait for the depth plane %.2f  first ROI %.2f  rest %.2f" % tuple((np.diff(buf[:, [24, 32, 33, 25]], axis=1) / 1.965e3).mean(axis=0)))
    print("   assoc: scans (warp 0) %.2f  sync %.2f  finalize+broadcast %.2f" % tuple(az.mean(axis=0)))
    ex = buf[:, [0, 15, 1]], buf[:, [4, 12, 13, 14, 5]]
    mg = np.diff(buf[:, 16:19], axis=1) / 1.965e3
    tl = np.diff(buf[:, [8, 19, 20]], axis=1) / 1.965e3
    for r in range(min(CS, 4)):
        print("   rank %d: select load %.2f  rank %.2f | tail gather %.2f  epilogue %.2f" % ((r,) + tuple(mg[r::CS].mean(axis=0)) + tuple(tl[r::CS].mean(axis=0))))
    print("   deferred decode epilogue (first thread of the decode warps): start after the tail %.2f  duration %.2f" % tuple((np.diff(buf[:, [24, 35, 36]], axis=1) / 1.965e3).mean(axis=0)))
    print("   passB: loop %.2f  sync %.2f  atomics+sync %.2f  copy %.2f" % tuple((np.diff(ex[1], axis=1) / 1.965e3).mean(axis=0)))
    print("   passB loop per CTA (warp 0):", " ".join("%.0f" % x for x in (buf[:, 12] - buf[:, 4]) / 1.965e3))
    print("   passB: barrier waits %.2f flags %.2f sync %.2f" % tuple((np.diff(buf[:, [4, 30, 31, 29]], axis=1) / 1.965e3).mean(axis=0)))
    print("   epilogue (thread 0): operands %.2f  trig %.2f  rest %.2f  wait for the other half %.2f" % tuple((np.diff(buf[:, [19, 21, 22, 23, 20]], axis=1) / 1.965e3).mean(axis=0)))
    print("   passB flag+sync %.2f; flagged groups per CTA:" % ((buf[:, 29] - buf[:, 4]).mean() / 1.965e3), " ".join(str(x) for x in buf[:, 27]), "\n   candidates per CTA:", " ".join(str(x) for x in buf[:, 28]))
    for r in range(min(CS, 4)):
        print("   rank %d mean:" % r, " ".join("%6.2f" % x for x in dt[r::CS].mean(axis=0)))
