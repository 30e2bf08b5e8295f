"""Shared test helpers (test infrastructure; may use the oracle and, when present, the reference)."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False))


def tie_free(heat: np.ndarray, depth: int) -> bool:
    """True when, per frame, the top-`depth` values of both the raw and the 3x3-nms'd heat-map are pairwise distinct
    (torch.topk's tie order is unspecified, so parity inputs must be tie-free in the top-K region)."""
    from oracle import binding as ob

    B = heat.shape[0]
    depth = min(depth, heat.shape[2] * heat.shape[3])
    for apply_nms in (False, True):
        s = ob.topk(heat, depth, apply_nms=apply_nms)[0]
        for b in range(B):
            if len(np.unique(s[b])) != depth:
                return False
    return True


def ensure_margins(inp: dict, pc_dep: np.ndarray, K: int, max_pc_dist: float, tol: float = 5e-5) -> int:
    """Nudge `inp['depth']` (in place) at top-K raw peaks until every strict compare of the association
    (d < depth+thr, d > max(0, depth-thr); reference utils/pointcloud.py:452-455) clears each distinct radar depth of
    its frame by more than tol*max(1,|bound|).  thr comes through non-correctly-rounded libm in the reference, so
    borderline cases could legitimately flip (SURVEY.md 'Hard parts'); parity inputs exclude them by construction.
    Returns how many values were nudged."""
    from oracle import binding as ob

    nudged = 0
    B = pc_dep.shape[0]
    HW = pc_dep.shape[2] * pc_dep.shape[3]
    for _ in range(50):
        _, _, thr, inds = ob.frustum_pc_hm(inp["heatmap"], inp["widthHeight"], inp["depth"], inp["dimension"],
                                           inp["rotation"], pc_dep, inp["calib"], K, max_pc_dist)
        bad = []
        dflat = inp["depth"].reshape(B, HW)
        for b in range(B):
            dv = np.unique(pc_dep[b, 0][pc_dep[b, 0] != 0]).astype(np.float64)
            if dv.size == 0:
                continue
            dep = dflat[b, inds[b]].astype(np.float64)
            t = thr[b].astype(np.float64)
            for bound in (dep + t, np.maximum(0.0, dep - t)):
                gap = np.abs(dv[None, :] - bound[:, None]).min(axis=1)
                for k in np.nonzero(gap <= tol * np.maximum(1.0, np.abs(bound)))[0]:
                    bad.append((b, int(inds[b, k])))
        if not bad:
            return nudged
        for b, ind in set(bad):
            dflat[b, ind] = np.float32(dflat[b, ind] + np.float32(3e-3))
            nudged += 1
    raise RuntimeError("could not establish association margins")


# ----------------------------------------------------------------------------------------------
# reference-side helpers (only usable where /root/reference exists)
# ----------------------------------------------------------------------------------------------
def ref_pc_dep_from_points(ns, pts, calib, A_out, img_wh, cfg):
    """pc_dep (3,H,W) by the reference's own code path: the body of loadRadarPointCloud
    (dataset/datasets/nuscenes.py:172-205) minus file I/O, then processPointCloud."""
    import refimport

    n = pts.shape[0]
    radar_pc = np.zeros((18, n), dtype=np.float64)
    radar_pc[0:3] = pts[:, 0:3].T
    radar_pc[8] = pts[:, 3]
    radar_pc[9] = pts[:, 4]
    depth = radar_pc[2, :]
    maxd = cfg.DATASET.MAX_PC_DIST
    if maxd > 0:
        radar_pc = radar_pc[:, depth <= maxd]
    pc_2d, mask = ns.map_pointcloud_to_image(radar_pc, np.asarray(calib)[:, :3], img_shape=img_wh)
    pc_3d = radar_pc[:, mask]
    index = np.argsort(pc_2d[2, :])
    pc_2d = pc_2d[:, index]
    pc_3d = pc_3d[:, index]
    _, _, pc_dep = refimport.ref_process_point_cloud(ns, pc_2d, pc_3d, np.asarray(A_out), calib, cfg)
    return pc_dep


def ref_assoc_indices(ns, inp, pc_dep, K, max_pc_dist):
    """Association indices from the reference itself (SURVEY.md 8c trick): channel 1 of pc_dep carries the flat pixel
    index, each detection is run through the reference's cvtPcDepthToHeatmap on a fresh map.
    Returns assoc (B,K) int32: matched flat pixel, -1 no match, -2 undecidable (paint window empty); thr (B,K)."""
    import torch

    t = lambda a: torch.from_numpy(np.ascontiguousarray(a))  # noqa: E731
    heat = t(inp["heatmap"])
    B, _, H, W = heat.shape
    _, inds, _, ys, xs = ns.topk(heat, K=K)
    xs = xs.view(B, K, 1) + 0.5
    ys = ys.view(B, K, 1) + 0.5
    depth = ns.transposeAndGetFeature(t(inp["depth"]), inds)
    wh = ns.transposeAndGetFeature(t(inp["widthHeight"]), inds).view(B, K, 2)
    wh[wh < 0] = 0
    bboxes = torch.cat([xs - wh[..., 0:1] / 2, ys - wh[..., 1:2] / 2, xs + wh[..., 0:1] / 2, ys + wh[..., 1:2] / 2], dim=2)
    dim = ns.transposeAndGetFeature(t(inp["dimension"]), inds).view(B, K, -1)
    rot = ns.transposeAndGetFeature(t(inp["rotation"]), inds).view(B, K, -1)
    alpha = ns.get_alpha(rot)
    center = torch.stack([(bboxes[..., 0] + bboxes[..., 2]) / 2, (bboxes[..., 1] + bboxes[..., 3]) / 2], dim=2)
    thr = ns.getDistanceThresh(t(inp["calib"]).view(B, 3, 4), center, dim, alpha)
    assoc = np.full((B, K), -1, np.int32)
    idxmap = pc_dep.copy()
    idxmap[:, 1] = np.arange(H * W, dtype=np.float32).reshape(1, H, W)
    idxmap_t = t(idxmap)
    for b in range(B):
        for k in range(K):
            hm = torch.zeros(3, H, W)
            ns.cvtPcDepthToHeatmap(hm, idxmap_t[b], depth[b, k], bboxes[b, k], thr[b, k], max_pc_dist)
            w = hm[0] != 0
            if w.any():
                assoc[b, k] = int(hm[1][w][0].item())
            else:
                bb = bboxes[b, k].numpy()
                cx = np.float32(np.float32(bb[0] + bb[2]) / np.float32(2))
                cy = np.float32(np.float32(bb[1] + bb[3]) / np.float32(2))
                wi = np.float32(np.float32(0.3) * np.float32(bb[2] - bb[0])) / np.float32(2)
                hi = np.float32(np.float32(0.3) * np.float32(bb[3] - bb[1])) / np.float32(2)
                rs = slice(int(cy - hi), int(cy + hi) + 1).indices(H)
                cs = slice(int(cx - wi), int(cx + wi) + 2).indices(W)
                if rs[0] >= rs[1] or cs[0] >= cs[1]:
                    assoc[b, k] = -2
    return assoc, thr.numpy()


# ----------------------------------------------------------------------------------------------
# seeded parity cases shared by the GPU tests, smoke() and bench.py
# ----------------------------------------------------------------------------------------------
FWD_KEYS = ("heatmap", "widthHeight", "depth", "dimension", "rotation")
DEC_KEYS = ("heatmap", "reg", "widthHeight", "depth", "depth2", "rotation", "rotation2", "dimension",
            "amodal_offset", "nuscenes_att", "velocity")


def make_case(wl, B: int, seed: int = 1234, margins: bool = True) -> dict:
    """Synthetic stage inputs for workload `wl` at batch B: tie-free heat-map in the top-(K+32) region, pc_dep painted
    by the oracle's pillar code from the radar points, association margins established (see ensure_margins)."""
    from centerfusiondetect3d_b200 import synthetic as syn
    from oracle import binding as ob

    while True:
        inp = syn.make_inputs(wl, seed=seed, B=B)
        if tie_free(inp["heatmap"], wl.K + 32):
            break
        seed += 1
    pc_dep, owner = ob.radar_pillars(inp["radar_points"], inp["radar_counts"], inp["cam"], inp["calib64"],
                                     inp["A_out"], (wl.img_w, wl.img_h), (wl.H, wl.W), wl.pillar_dims,
                                     wl.max_pc_dist)
    inp["pc_dep"] = pc_dep
    inp["owner"] = owner
    inp["seed"] = seed
    inp["nudged"] = ensure_margins(inp, pc_dep, wl.K, wl.max_pc_dist) if margins else 0
    return inp


def to_cuda(inp: dict, keys, device="cuda:0") -> dict:
    import torch

    return {k: torch.from_numpy(np.ascontiguousarray(inp[k])).to(device) for k in keys}


def bits(a) -> np.ndarray:
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a
