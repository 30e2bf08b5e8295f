"""cf_frustum_boxes_static: the allocation-free variant of cf_frustum_boxes (scratch in a file-scope `static __device__`
array of the library instead of a caller-provided workspace; no counterpart in the reference).
Called through the C ABI, checked against the existing entry point on the same inputs -- bit for
bit, the bar of every pc_hm / association test here -- on two calls in a row: the second with fewer, different boxes, so
anything a call wrongly kept from the array's previous contents would show.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

from centerfusiondetect3d_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CF_STATIC_MAX_BOXES = 32768  # include/cfd3d_b200.h


def _case(seed, B, N, H, W):
    rng = np.random.default_rng(seed)
    pc_dep = np.zeros((B, 3, H, W), np.float32)
    for b in range(B):  # pillar-like rectangles of radar depth / velocity
        for _ in range(60):
            y, x = int(rng.integers(0, H - 6)), int(rng.integers(0, W - 3))
            pc_dep[b, :, y:y + int(rng.integers(2, 6)), x:x + int(rng.integers(1, 3))] = \
                np.array([rng.uniform(2, 58), rng.normal(0, 4), rng.normal(0, 4)], np.float32)[:, None, None]
    cx, cy = rng.uniform(0, W, (B, N)), rng.uniform(0, H, (B, N))
    w, h = rng.uniform(1, 30, (B, N)), rng.uniform(1, 20, (B, N))
    bboxes = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], -1).astype(np.float32)
    depth = rng.uniform(2, 58, (B, N)).astype(np.float32)
    thr = rng.uniform(0.5, 8, (B, N)).astype(np.float32)
    counts = rng.integers(N // 2, N + 1, (B,)).astype(np.int32)
    return pc_dep, bboxes, depth, thr, counts


def _run(L, name, arrs, B, N, H, W, with_ws):
    t = [torch.from_numpy(a).to(DEV) for a in arrs]
    pc_hm = torch.full((B, 3, H, W), float("nan"), device=DEV)  # the call must write every pixel
    assoc = torch.full((B, N), -7, dtype=torch.int32, device=DEV)
    args = [x.data_ptr() for x in t] + [B, N, H, W, C.c_float(60.0), pc_hm.data_ptr(), assoc.data_ptr()]
    ws = None
    if with_ws:
        ws = torch.empty(B * N * 32 + 256, dtype=torch.uint8, device=DEV)
        args += [ws.data_ptr(), ws.numel()]
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info()[0]
    rc = getattr(L, name)(*args, stream)
    torch.cuda.synchronize()
    after = torch.cuda.mem_get_info()[0]
    assert rc == 0, L.cf_last_error().decode()
    return pc_hm.cpu().numpy(), assoc.cpu().numpy(), before - after


def test_static_scratch_variant_matches_the_workspace_one():
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device")
    L = _lib.lib()
    H, W = 112, 200
    freed = []
    for call, (seed, B, N) in enumerate(((1, 4, 128), (2, 3, 50))):  # second call: fewer boxes over the same array
        arrs = _case(seed, B, N, H, W)
        hm_w, as_w, _ = _run(L, "cf_frustum_boxes", arrs, B, N, H, W, with_ws=True)
        hm_s, as_s, used = _run(L, "cf_frustum_boxes_static", arrs, B, N, H, W, with_ws=False)
        freed.append(used)
        assert not np.isnan(hm_s).any()
        assert np.array_equal(hm_s.view(np.uint32), hm_w.view(np.uint32)), "pc_hm differs on call %d" % call
        assert np.array_equal(as_s, as_w), "association indices differ on call %d" % call
        assert (as_w >= 0).sum() >= B and (hm_w != 0).any()  # the case exercises real matches
    # the steady-state call takes no device memory (the first may: lazy module loading, the one-time set-up)
    assert freed[1] <= 0, "device memory shrank by %d bytes across a cf_frustum_boxes_static call" % freed[1]
    # beyond the static array: refused with a clear error, nothing launched
    B, N = 2, CF_STATIC_MAX_BOXES // 2 + 1
    z = torch.zeros(1, device=DEV)
    rc = L.cf_frustum_boxes_static(z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), None, B, N, H, W, C.c_float(60.0),
                                   z.data_ptr(), None, torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and b"exceed the static scratch" in L.cf_last_error()
    assert L.cf_last_launch_count() == 0
