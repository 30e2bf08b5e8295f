import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch, bench
from centerfusiondetect3d_b200 import ops, synthetic as syn
wl = syn.WORKLOADS["cfg2"]; B = 32; dev = "cuda:0"
host = bench.make_host_inputs(wl, B, seed=1)
keys = bench.STAGE_KEYS + ("calib", "pc_dep")
pin = {k: torch.from_numpy(np.ascontiguousarray(host[k])).pin_memory() for k in keys}
d = {k: pin[k].to(dev) for k in keys}
cfg = bench.Cfg(wl.K, wl.max_pc_dist)
def run(mode):
    if mode == "dev": inp = dict(d)
    elif mode == "heads": inp = {k: (d[k] if k in ("heatmap", "pc_dep", "calib") else pin[k]) for k in keys}
    else:
        inp = {k: (d[k] if k in ("heatmap", "pc_dep", "calib") else pin[k]) for k in keys}; inp["pc_vel"] = pin["pc_dep"]
    for _ in range(3): bench.run_stage(ops, wl, inp, cfg)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): bench.run_stage(ops, wl, inp, cfg)
    e1.record(); torch.cuda.synchronize()
    print(mode, "%.1f us per call" % (e0.elapsed_time(e1) * 1e3 / 20))
for m in ("dev", "heads", "heads+vel"): run(m)
