"""Development aid: survivor counts per (frame, list) left in the workspace by one stage call."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import bench
from centerfusiondetect3d_b200 import ops, synthetic as syn

wl = syn.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "cfg2"]
B = int(sys.argv[2]) if len(sys.argv) > 2 else min(wl.B, 32)
dev = "cuda:0"
host = bench.make_host_inputs(wl, B, seed=1234)
keys = bench.STAGE_KEYS + ("calib", "pc_dep")
d = {k: torch.from_numpy(np.ascontiguousarray(host[k])).to(dev) for k in keys}
cfg = bench.Cfg(wl.K, wl.max_pc_dist)
bench.run_stage(ops, wl, d, cfg)
torch.cuda.synchronize()
ws, _ = ops._stage_ws(torch.device(dev), B, wl.C, wl.H, wl.W, wl.K)
ctr = ws[0:B * 2 * 4 * 4].view(torch.int32).cpu().numpy().reshape(B, 2, 4)
print("cluster", os.environ.get("CFD3D_CLUSTER"), "chunk", os.environ.get("CFD3D_CHUNK_BYTES"))
print("raw survivors per frame:", ctr[:, 0, 0])
print("nms survivors per frame:", ctr[:, 1, 0])
