import sys, torch
sys.path.insert(0, '.')
from centerfusiondetect3d_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
g = torch.Generator(device="cpu").manual_seed(11)
heat = torch.rand((B, 10, 112, 200), generator=g).sigmoid().cuda()
s, i, c, y, x = ops.topk(heat, 100)
torch.cuda.synchronize()
v, idx = heat.view(B, -1).topk(100, dim=1)
print("ok", bool((v == s).all()))
