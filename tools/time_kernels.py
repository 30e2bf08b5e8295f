"""Development aid: CUDA-event time of every kernel of one workload's step, one launch at a time (the ABI's measurement
hook, cf_set_kernel_events).  python tools/time_kernels.py cfg4"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import bench
from centerfusiondetect3d_b200 import _lib, ops, synthetic as syn

wl = syn.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "cfg4"]
B = int(sys.argv[2]) if len(sys.argv) > 2 else wl.B if wl.name.startswith(("cfg2", "cfg4")) else min(wl.B, 8)
dev = "cuda:0"
keys = bench.STAGE_KEYS + ("calib", "pc_dep") + (bench.PILLAR_KEYS if wl.with_pillars else ())
sets = []
for i in range(3):
    host = bench.make_host_inputs(wl, B, seed=1234 + i)
    sets.append({k: torch.from_numpy(np.ascontiguousarray(host[k])).to(dev) for k in keys})
cfg = bench.Cfg(wl.K, wl.max_pc_dist)
L = _lib.lib()
for kname, kid in (("pillar_rects", 4), ("paint", 3), ("fusion_stage", 6)):
    ts = []
    for rep in range(30):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); b.record()
        L.cf_set_kernel_events(kid, a.cuda_event, b.cuda_event)
        bench.run_stage(ops, wl, sets[rep % 3], cfg)
        L.cf_set_kernel_events(-1, None, None)
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = np.array(ts[5:])
    print("%-14s min %.1f  median %.1f us (event pair around one launch: ~4-5 us of event overhead included)" % (kname, ts.min(), np.median(ts)))
