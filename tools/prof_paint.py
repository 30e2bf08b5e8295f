"""Development aid: per-warp phase clocks of paint_kernel on the dense-radar workload (the -DCFD3D_PROFILE build)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("CFD3D_LIB", os.path.join(ROOT, "centerfusiondetect3d_b200", "libcfd3d_b200_prof.so"))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bench
from centerfusiondetect3d_b200 import _lib, ops, synthetic as syn
wl = syn.WORKLOADS["cfg4"]; B = 64; dev = "cuda:0"
host = bench.make_host_inputs(wl, B, seed=1234)
d = {k: torch.from_numpy(np.ascontiguousarray(host[k])).to(dev) for k in bench.PILLAR_KEYS}
for r in range(3):
    ops.radar_to_pc_dep(*[d[k] for k in bench.PILLAR_KEYS], (wl.H, wl.W), (wl.img_w, wl.img_h), wl.pillar_dims, wl.max_pc_dist)
torch.cuda.synchronize()
L = _lib.lib(); n = 64 * 7 * 4
buf = np.zeros((n, 16), np.int64)
L.cf_debug_read_paint_prof.argtypes = [C.c_void_p, C.c_int]
L.cf_debug_read_paint_prof(buf.ctypes.data, n)
us = lambda a: a / 1.965e3
tot = us(buf[:, 3] - buf[:, 0])
print("per warp us: total min/mean/max %.1f %.1f %.1f" % (tot.min(), tot.mean(), tot.max()))
for nm, a in (("staging", us(buf[:, 4])), ("walk", us(buf[:, 5])), ("own build", us(buf[:, 2] - buf[:, 1])), ("write", us(buf[:, 3] - buf[:, 2]))):
    print("  %-10s min/mean/max %.2f %.2f %.2f" % (nm, a.min(), a.mean(), a.max()))
print("  relevant records per warp min/mean/max", buf[:, 6].min(), buf[:, 6].mean(), buf[:, 6].max())
w = np.argmax(tot); print("slowest warp", w, [round(float(x), 2) for x in (us(buf[w, 4]), us(buf[w, 5]), us(buf[w, 2] - buf[w, 1]), us(buf[w, 3] - buf[w, 2]))], buf[w, 6])

g0, g1 = buf[:, 8], buf[:, 9]
print("globaltimer: first start -> last end %.1f us; starts spread %.1f us; ends spread %.1f us" % ((g1.max() - g0.min()) / 1e3, (g0.max() - g0.min()) / 1e3, (g1.max() - g1.min()) / 1e3))
order = np.argsort(g0)
print("start offsets (us) of warps sorted, every 200th:", [round(float(g0[o] - g0.min()) / 1e3, 1) for o in order[::200]])
