"""Host side of the B200-native fusion stage: the reference's Python call signatures on top of the C ABI.

Every function here validates its tensors, allocates outputs with torch (device memory + current stream
only -- torch does no arithmetic on this path) and calls libcfd3d_b200.so through ctypes.  There is no
CPU / PyTorch fallback: non-CUDA tensors raise, a missing library raises.

Reference callables mirrored (paths relative to the reference tree):
    nms, topk, getFeature, transposeAndGetFeature          src/lib/model/utils.py:6-128
    get_alpha, getDistanceThresh                           src/lib/utils/pointcloud.py:195-328
    getPcFrustumHeatmap, cvtPcDepthToHeatmap               src/lib/utils/pointcloud.py:331-481
    fusionDecode                                           src/lib/model/decode.py:10-174
    processPointCloud (pillars) -> radar_to_pc_dep         src/lib/dataset/generic_dataset.py:738-942
Original CenterFusion names (_nms, _topk, _tranpose_and_gather_feat, generate_pc_hm, get_pc_hm,
pc_dep_to_hm_torch, fusion_decode) are exported as aliases.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import DecodeHeads, DecodeOut, PostFuse, PostIn, PostOut

__all__ = [
    "nms", "topk", "getFeature", "transposeAndGetFeature", "get_alpha", "getDistanceThresh",
    "getPcFrustumHeatmap", "cvtPcDepthToHeatmap", "fusionDecode", "radar_to_pc_dep", "fusion_stage",
    "frustum_association", "last_launch_count", "last_plan", "debug_set_plan", "release_buffers", "postProcess", "inverse_output_affine", "frustum_targets", "radar_sort",
    "_nms", "_topk", "_gather_feat", "_tranpose_and_gather_feat", "generate_pc_hm", "get_pc_hm",
    "pc_dep_to_hm_torch", "fusion_decode", "get_dist_thresh",
]


# ------------------------------------------------------------------------------------------------------
# plumbing
# ------------------------------------------------------------------------------------------------------
def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _ptr(t) -> int | None:
    return None if t is None else t.data_ptr()


def _map(t, name: str, ndim: int = 4, dtype=torch.float32, pinned_ok: bool = False) -> torch.Tensor:
    """Validates a map.  `pinned_ok`: the tensor is only GATHERED at the K peak pixels (every head but the heat-map),
    so a contiguous page-locked host tensor is accepted as it is -- under unified addressing the kernels read those few
    sectors straight out of host memory over PCIe instead of the caller copying the whole head to the device."""
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor, got %s" % (name, type(t).__name__))
    if not t.is_cuda:
        if pinned_ok and t.is_pinned() and t.is_contiguous():
            if t.dtype != dtype:
                raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
            if ndim is not None and t.dim() != ndim:
                raise ValueError("%s must have %d dimensions, got shape %s" % (name, ndim, tuple(t.shape)))
            return t
        raise ValueError("%s must be a CUDA tensor%s: the fusion stage has no CPU path" % (
            name, " or a contiguous pinned host tensor" if pinned_ok else ""))
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if ndim is not None and t.dim() != ndim:
        raise ValueError("%s must have %d dimensions, got shape %s" % (name, ndim, tuple(t.shape)))
    return t.contiguous()  # the reference accepts any strides


class _Lru(dict):
    """Small bounded cache of device buffers keyed by (device, stream, ...): streams come and go (their handles can
    even be reused), so the least recently used buffers are dropped instead of kept for the life of the process."""

    def __init__(self, cap: int):
        super().__init__()
        self.cap = cap

    def lookup(self, key):
        v = self.pop(key, None)
        if v is not None:
            self[key] = v  # most recently used last
        return v

    def store(self, key, v):
        self.pop(key, None)
        self[key] = v
        while len(self) > self.cap:
            self.pop(next(iter(self)))


_ws_cache = _Lru(16)
_staging_cache = _Lru(16)


def release_buffers() -> None:
    """Drops the cached workspaces / staging buffers (they are re-created on demand)."""
    _ws_cache.clear()
    _staging_cache.clear()


def _staging(dev: torch.device, name: str, shape: tuple) -> torch.Tensor:
    """Per-(device, stream, name, shape) device buffer for inputs a host-buffer caller hands over in pinned memory."""
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream, name, tuple(shape))
    buf = _staging_cache.lookup(key)
    if buf is None:
        buf = torch.zeros(shape, dtype=torch.float32, device=dev)
        _staging_cache.store(key, buf)
    return buf


def _workspace(dev: torch.device, nbytes: int) -> torch.Tensor:
    """Grow-only per-(device, stream) scratch buffer: no allocation on the steady-state path."""
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    buf = _ws_cache.lookup(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=dev)
        _ws_cache.store(key, buf)
    return buf


def _call(dev: torch.device, name: str, *args) -> None:
    """One ABI call with `dev` current: the kernels, their function attributes and the occupancy-based launch plans all
    act on the current device, which in a single-process multi-GPU program need not be the tensors' device."""
    fn = getattr(_lib.lib(), name)
    if dev.index is None or dev.index == torch.cuda.current_device():
        _lib.check(fn(*args))
    else:
        with torch.cuda.device(dev):
            _lib.check(fn(*args))


def _same_device(*tensors) -> None:
    """Every CUDA tensor of a call lives on one device (pinned host tensors are readable from any)."""
    devs = {t.device for t in tensors if isinstance(t, torch.Tensor) and t.is_cuda}
    if len(devs) > 1:
        raise ValueError("all CUDA tensors of a call must be on one device, got %s" % sorted(str(d) for d in devs))


def _stage_ws(dev, B, Cc, H, W, K):
    n = _lib.lib().cf_workspace_bytes(B, Cc, H, W, K)
    if n == 0:
        # let the ABI produce the precise error (K range, geometry)
        n = 256
    ws = _workspace(dev, n)
    return ws, ws.numel()


def last_launch_count() -> int:
    """Kernels launched by the last ABI call on this thread."""
    return int(_lib.lib().cf_last_launch_count())


def last_plan() -> dict:
    """Execution plan of the last fused-stage launch on this thread (cf_last_plan): cluster_size, n_clusters, rounds,
    resident, depth_plane_staged, chunk_rows, slots, smem_bytes."""
    info = _lib.PlanInfo()
    _lib.check(_lib.lib().cf_last_plan(C.byref(info)))
    return {n: int(getattr(info, n)) for n, _ in _lib.PlanInfo._fields_}


def debug_set_plan(force_cluster_size: int = 0, max_clusters: int = 0) -> None:
    """Test hook (cf_debug_set_plan): pin the cluster size / cap the co-resident clusters of later stage launches;
    (0, 0) restores the planner's own choice."""
    _lib.check(_lib.lib().cf_debug_set_plan(int(force_cluster_size), int(max_clusters)))


# ------------------------------------------------------------------------------------------------------
# a1-a3  heat-map peak decode primitives                                   src/lib/model/utils.py
# ------------------------------------------------------------------------------------------------------
def nms(heatmap: torch.Tensor, kernel: int = 3) -> torch.Tensor:
    """utils.py:112-128 -- heat * (maxpool3x3(heat) == heat)."""
    if kernel != 3:
        raise NotImplementedError("nms: only kernel=3 is on the fusion path (the reference never passes another)")
    heat = _map(heatmap, "heatmap")
    B, Cc, H, W = heat.shape
    out = torch.empty_like(heat)
    _call(heat.device, "cf_nms", heat.data_ptr(), B, Cc, H, W, out.data_ptr(), _stream(heat))
    return out


def topk(heatmap: torch.Tensor, K: int = 100, *, apply_nms: bool = False):
    """utils.py:6-38 -- (scores f32, inds i64, classes i32, ys i64, xs i64), each (B,K), scores descending.
    Ties are ordered (score desc, class asc, index asc); torch.topk leaves them unspecified.
    apply_nms=True selects on nms(heatmap) without materialising it."""
    heat = _map(heatmap, "heatmap")
    B, Cc, H, W = heat.shape
    K = int(K)
    dev = heat.device
    scores = torch.empty((B, K), dtype=torch.float32, device=dev)
    inds = torch.empty((B, K), dtype=torch.int64, device=dev)
    classes = torch.empty((B, K), dtype=torch.int32, device=dev)
    ys = torch.empty((B, K), dtype=torch.int64, device=dev)
    xs = torch.empty((B, K), dtype=torch.int64, device=dev)
    ws, wsn = _stage_ws(dev, B, Cc, H, W, K)
    _call(dev, "cf_peak_topk", heat.data_ptr(), B, Cc, H, W, K, int(bool(apply_nms)), scores.data_ptr(),
                                       inds.data_ptr(), classes.data_ptr(), ys.data_ptr(), xs.data_ptr(),
                                       ws.data_ptr(), wsn, _stream(heat))
    return scores, inds, classes, ys, xs


def transposeAndGetFeature(feature: torch.Tensor, indices: torch.Tensor) -> torch.Tensor:
    """utils.py:58-72 -- out[b,m,c] = feature[b,c,indices[b,m]] read straight from NCHW."""
    feat = _map(feature, "feature")
    ind = _map(indices, "indices", ndim=2, dtype=torch.int64)
    B, ch, H, W = feat.shape
    if ind.shape[0] != B:
        raise ValueError("indices batch %d != feature batch %d" % (ind.shape[0], B))
    M = ind.shape[1]
    out = torch.empty((B, M, ch), dtype=torch.float32, device=feat.device)
    _call(feat.device, "cf_gather_nchw", feat.data_ptr(), B, ch, H * W, ind.data_ptr(), M, out.data_ptr(),
                                         _stream(feat))
    return out


def getFeature(feature: torch.Tensor, indices: torch.Tensor) -> torch.Tensor:
    """utils.py:41-55 -- feature (B,N,ch), indices (B,M) -> (B,M,ch).  Same kernel as the NCHW gather with the
    roles of the strides swapped: served by viewing feature as (B*N... ) is not possible without a copy, so this
    one is a transposed view into the NCHW gather: out[b,m,c] = feature[b,indices[b,m],c]."""
    feat = _map(feature, "feature", ndim=3)
    # (B,N,ch) -> NCHW-style (B,ch,N,1); only the small (B,N,ch) candidate tensors ever take this path
    return transposeAndGetFeature(feat.permute(0, 2, 1).unsqueeze(-1), indices)


# ------------------------------------------------------------------------------------------------------
# a4  observation angle and frustum distance threshold                     src/lib/utils/pointcloud.py
# ------------------------------------------------------------------------------------------------------
def get_alpha(rotation: torch.Tensor) -> torch.Tensor:
    """pointcloud.py:195-211 -- rotation (...,8) -> alpha (...)."""
    rot = _map(rotation, "rotation", ndim=None)
    if rot.shape[-1] != 8:
        raise ValueError("rotation must end in 8 bins, got %s" % (tuple(rot.shape),))
    out = torch.empty(rot.shape[:-1], dtype=torch.float32, device=rot.device)
    _call(rot.device, "cf_get_alpha", rot.data_ptr(), out.numel(), out.data_ptr(), _stream(rot))
    return out


def getDistanceThresh(calib: torch.Tensor, center: torch.Tensor, dim: torch.Tensor, alpha: torch.Tensor):
    """pointcloud.py:299-328 -- calib (B,3,4), center (B,K,2), dim (B,K,3), alpha (B,K) -> (B,K)."""
    center = _map(center, "center", ndim=3)
    B, K, _ = center.shape
    calib = _map(calib, "calib", ndim=None).view(-1, 3, 4)
    if calib.shape[0] == 1 and B > 1:
        calib = calib.expand(B, 3, 4).contiguous()
    if calib.shape[0] != B:
        raise ValueError("calib must be view-able to (%d,3,4)" % B)
    dim = _map(dim, "dim", ndim=3)
    alpha = _map(alpha, "alpha", ndim=None).reshape(B, K)
    out = torch.empty((B, K), dtype=torch.float32, device=center.device)
    _call(center.device, "cf_dist_thresh", calib.data_ptr(), center.data_ptr(), dim.data_ptr(), alpha.data_ptr(), B, K,
                                         out.data_ptr(), _stream(center))
    return out


# ------------------------------------------------------------------------------------------------------
# a7 / a8  frustum association -> pc_hm                                     src/lib/utils/pointcloud.py
# ------------------------------------------------------------------------------------------------------
_FWD_ALIASES = {"heatmap": "hm", "widthHeight": "wh", "depth": "dep", "dimension": "dim", "rotation": "rot"}


def _head(output: dict, name: str) -> torch.Tensor:
    if name in output:
        return output[name]
    alt = _FWD_ALIASES.get(name)
    if alt is not None and alt in output:  # original CenterFusion head names
        return output[alt]
    raise KeyError(name)


def _cfg_k_dist(config):
    if hasattr(config, "MODEL"):
        return int(config.MODEL.K), float(config.DATASET.MAX_PC_DIST)
    return int(config.K), float(config.max_pc_dist)  # original CenterFusion `opt`


def frustum_association(output: dict, pc_dep: torch.Tensor, calib: torch.Tensor, K: int, max_pc_dist: float):
    """getPcFrustumHeatmap plus the by-products the reference computes but does not return:
    (pc_hm (B,3,H,W), assoc (B,K) i32 matched pc_dep pixel or -1, thr (B,K), inds (B,K) i64)."""
    heat = _map(_head(output, "heatmap"), "heatmap")
    B, Cc, H, W = heat.shape
    wh = _map(_head(output, "widthHeight"), "widthHeight", pinned_ok=True)
    dep = _map(_head(output, "depth"), "depth", pinned_ok=True)
    dim = _map(_head(output, "dimension"), "dimension", pinned_ok=True)
    rot = _map(_head(output, "rotation"), "rotation", pinned_ok=True)
    pcd = _map(pc_dep, "pc_dep")
    if pcd.shape != (B, 3, H, W):
        raise NotImplementedError("pc_dep must be (B,3,H,W) [depth, vx, vz]; ONE_HOT_PC is forbidden together with "
                                  "FRUSTUM by the reference (detectHeads.py:48)")
    for t, ch, nm in ((wh, 2, "widthHeight"), (dep, 1, "depth"), (dim, 3, "dimension"), (rot, 8, "rotation")):
        if t.shape != (B, ch, H, W):
            raise ValueError("%s must be (%d,%d,%d,%d), got %s" % (nm, B, ch, H, W, tuple(t.shape)))
    cal = _map(calib, "calib", ndim=None).view(B, 3, 4)
    _same_device(heat, wh, dep, dim, rot, pcd, cal)
    dev = heat.device
    pc_hm = torch.empty_like(pcd)
    assoc = torch.empty((B, K), dtype=torch.int32, device=dev)
    thr = torch.empty((B, K), dtype=torch.float32, device=dev)
    inds = torch.empty((B, K), dtype=torch.int64, device=dev)
    ws, wsn = _stage_ws(dev, B, Cc, H, W, K)
    _call(dev, "cf_frustum_pc_hm", heat.data_ptr(), wh.data_ptr(), dep.data_ptr(), dim.data_ptr(),
                                           rot.data_ptr(), pcd.data_ptr(), cal.data_ptr(), B, Cc, H, W, int(K),
                                           float(max_pc_dist), pc_hm.data_ptr(), assoc.data_ptr(), thr.data_ptr(),
                                           inds.data_ptr(), ws.data_ptr(), wsn, _stream(heat))
    return pc_hm, assoc, thr, inds


def getPcFrustumHeatmap(output: dict, pc_dep: torch.Tensor, calib: torch.Tensor, config) -> torch.Tensor:
    """pointcloud.py:331-394 -- drop-in for the call at detectHeads.py:176-177.  Returns a new pc_hm tensor."""
    K, max_pc_dist = _cfg_k_dist(config)
    return frustum_association(output, pc_dep, calib, K, max_pc_dist)[0]


def cvtPcDepthToHeatmap(pc_hm, pc_dep, depth, bbox, distanceThreshold, max_pc_dist) -> None:
    """pointcloud.py:397-481 -- one detection, IN PLACE on pc_hm (3,H,W).  depth / distanceThreshold may be 0-d or
    1-element tensors (as the reference passes them) or Python numbers."""
    hm = pc_hm
    if not (isinstance(hm, torch.Tensor) and hm.is_cuda and hm.dtype == torch.float32 and hm.dim() == 3
            and hm.is_contiguous()):
        raise ValueError("pc_hm must be a contiguous CUDA float32 (3,H,W) tensor (it is written in place)")
    pcd = _map(pc_dep, "pc_dep", ndim=3)
    if pcd.shape != hm.shape or hm.shape[0] != 3:
        raise ValueError("pc_hm and pc_dep must both be (3,H,W)")
    _, H, W = hm.shape
    dev = hm.device

    def scalar(v, name):
        if isinstance(v, torch.Tensor):
            return _map(v.reshape(-1)[:1], name, ndim=1)
        return torch.tensor([float(v)], dtype=torch.float32, device=dev)

    d = scalar(depth, "depth")
    t = scalar(distanceThreshold, "distanceThreshold")
    bb = _map(bbox, "bbox", ndim=1) if isinstance(bbox, torch.Tensor) else torch.tensor(
        [float(v) for v in bbox], dtype=torch.float32, device=dev)
    if bb.numel() != 4:
        raise ValueError("bbox must have 4 elements")
    _call(dev, "cf_pc_dep_to_hm_single", hm.data_ptr(), pcd.data_ptr(), H, W, d.data_ptr(), bb.data_ptr(),
                                                 t.data_ptr(), float(max_pc_dist), None, _stream(hm))


def frustum_targets(pc_dep: torch.Tensor, bboxes: torch.Tensor, depth: torch.Tensor, distanceThreshold: torch.Tensor,
                    counts: torch.Tensor | None = None, max_pc_dist: float = 60.0, return_assoc: bool = False):
    """Row f4 -- the training-target form of the frustum association (dataset/generic_dataset.py:673-687): for every
    frame, cvtPcDepthToHeatmap of annotation 0, 1, ... in list order on a zero pc_hm.  pc_dep (B,3,H,W); bboxes (B,N,4)
    in output-map pixels; depth, distanceThreshold (B,N); counts (B) int32 annotations per frame (None = N).
    Returns pc_hm (B,3,H,W) [, assoc (B,N) int32].

    All-float32 arguments follow the torch branch of cvtPcDepthToHeatmap (cf_frustum_boxes).  If any of bboxes / depth /
    distanceThreshold is float64 the call follows the NumPy branch the reference's DataLoader runs (cf_frustum_boxes_f64,
    NumPy >= 2 promotion): `depth` stands for the annotation's Python float (pass it as float64), and the dtypes of
    `bboxes` and `distanceThreshold` select float32 or float64 arithmetic exactly as they do in NumPy."""
    pcd = _map(pc_dep, "pc_dep")
    B, _, H, W = pcd.shape
    for t, nm in ((bboxes, "bboxes"), (depth, "depth"), (distanceThreshold, "distanceThreshold")):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype in (torch.float32, torch.float64)):
            raise ValueError("%s must be a float32 / float64 CUDA tensor: the fusion stage has no CPU path" % nm)
    np_branch = any(t.dtype == torch.float64 for t in (bboxes, depth, distanceThreshold))
    dt = torch.float64 if np_branch else torch.float32
    if bboxes.dim() != 3 or bboxes.shape[0] != B or bboxes.shape[2] != 4:
        raise ValueError("bboxes must be (B,N,4)")
    N = bboxes.shape[1]
    flags = (1 if bboxes.dtype == torch.float32 else 0) | (2 if distanceThreshold.dtype == torch.float32 else 0)
    bb = bboxes.to(dt).contiguous()
    dp = depth.to(dt).reshape(B, N).contiguous()
    th = distanceThreshold.to(dt).reshape(B, N).contiguous()
    cn = _map(counts, "counts", ndim=1, dtype=torch.int32) if counts is not None else None
    _same_device(pcd, bb, dp, th, cn)
    pc_hm = torch.empty_like(pcd)
    assoc = torch.empty((B, N), dtype=torch.int32, device=pcd.device) if return_assoc else None
    ws = _workspace(pcd.device, B * N * 32 + 256)
    if np_branch:
        _call(pcd.device, "cf_frustum_boxes_f64", pcd.data_ptr(), bb.data_ptr(), dp.data_ptr(), th.data_ptr(), _ptr(cn), B, N,
              H, W, float(max_pc_dist), flags, pc_hm.data_ptr(), _ptr(assoc), ws.data_ptr(), ws.numel(), _stream(pcd))
    else:
        _call(pcd.device, "cf_frustum_boxes", pcd.data_ptr(), bb.data_ptr(), dp.data_ptr(), th.data_ptr(), _ptr(cn), B, N,
              H, W, float(max_pc_dist), pc_hm.data_ptr(), _ptr(assoc), ws.data_ptr(), ws.numel(), _stream(pcd))
    return (pc_hm, assoc) if return_assoc else pc_hm


# ------------------------------------------------------------------------------------------------------
# a9  fusionDecode                                                          src/lib/model/decode.py:10-174
# ------------------------------------------------------------------------------------------------------
_DEC_CH = {"reg": 2, "widthHeight": 2, "depth": 1, "rotation": 8, "dimension": 3, "amodal_offset": 2,
           "nuscenes_att": 8, "velocity": 3, "uncertainty": 1}
_DEC_HEAD_FIELD = {"reg": "reg", "widthHeight": "wh", "depth": "depth", "rotation": "rot", "dimension": "dim",
                   "amodal_offset": "amodal", "nuscenes_att": "nuscenes_att", "velocity": "velocity",
                   "uncertainty": "uncertainty"}
_DEC_OUT = (("widthHeight", "bboxes", 4, "bboxes"), ("depth", "depth", 1, "depth"), ("rotation", "rot", 8, "rotation"),
            ("dimension", "dim", 3, "dimension"), ("amodal_offset", "amodal", 2, "amodal_offset"),
            ("nuscenes_att", "nuscenes_att", 8, "nuscenes_att"), ("velocity", "velocity", 3, "velocity"))


def _decode_structs(output: dict, K: int, keep: list):
    """Builds the ABI structs for one layer dict; applies the reference's rotation2 -> rotation rename
    (decode.py:120-121) to the caller's dict."""
    heat = _map(output["heatmap"], "heatmap")
    B, Cc, H, W = heat.shape
    dev = heat.device
    src = {}
    for name in _DEC_CH:
        key = name
        if name == "depth" and "depth2" in output:  # decode.py:101-110
            key = "depth2"
        if name == "rotation" and "rotation2" in output:
            key = "rotation2"
        if key in output:
            t = _map(output[key], key, pinned_ok=True)
            if t.shape != (B, _DEC_CH[name], H, W):
                raise ValueError("%s must be (%d,%d,%d,%d), got %s" % (key, B, _DEC_CH[name], H, W, tuple(t.shape)))
            src[name] = t
    if "rotation2" in output:
        output["rotation"] = output.pop("rotation2")
    heads = DecodeHeads()
    heads.heatmap = heat.data_ptr()
    for name, t in src.items():
        setattr(heads, _DEC_HEAD_FIELD[name], t.data_ptr())
    ret = {
        "scores": torch.empty((B, K), dtype=torch.float32, device=dev),
        "classIds": torch.empty((B, K), dtype=torch.float32, device=dev),
        "centers": torch.empty((B, K, 2), dtype=torch.float32, device=dev),
    }
    out = DecodeOut()
    out.scores, out.class_ids, out.centers = ret["scores"].data_ptr(), ret["classIds"].data_ptr(), ret["centers"].data_ptr()
    for head, field, ch, key in _DEC_OUT:
        if head in src:
            ret[key] = torch.empty((B, K, ch), dtype=torch.float32, device=dev)
            setattr(out, field, ret[key].data_ptr())
    keep.extend(src.values())
    keep.append(heat)
    return heads, out, ret, (B, Cc, H, W), heat


def _post_fuse(ret: dict, post: dict, keep: list):
    """cf_post_fuse_t for a decode whose result dict is `ret`: `post` = postProcess's remaining arguments
    {center, scale, height, width, calibs[, isGt]} (postProcess.py:13).  Returns (struct, finish) -- finish(ret) turns
    the decode's dict into what postProcess returns (centers / bboxes replaced, rotation -> alpha, new 3-D keys)."""
    B, K = ret["scores"].shape
    dev = ret["scores"].device
    height, width = int(post["height"]), int(post["width"])
    trans = inverse_output_affine(post["center"], post["scale"], width, height)
    has3d = all(k in ret for k in ("rotation", "depth", "dimension"))
    cal = None
    if has3d:
        cal = _map(post["calibs"], "calibs", ndim=None).reshape(-1, 3, 4)
        if cal.shape[0] != B:
            cal = cal.expand(B, 3, 4).contiguous()
    new = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)  # noqa: E731
    o_cen = new(B, K, 2)
    o_bb = new(B, K, 4) if "bboxes" in ret else None
    o_alpha = new(B, K) if "rotation" in ret else None
    o_loc, o_yaw, o_b3 = (new(B, K, 3), new(B, K), new(B, K, 8, 3)) if has3d else (None, None, None)
    tarr = (C.c_float * 6)(*[float(v) for v in trans.reshape(-1)])
    pf = PostFuse(C.cast(tarr, C.POINTER(C.c_float)), _ptr(cal), height, width, int(bool(post.get("isGt", False))),
                  PostOut(_ptr(o_cen), _ptr(o_bb), _ptr(o_alpha), _ptr(o_loc), _ptr(o_yaw), _ptr(o_b3)))
    keep.extend([tarr, cal])

    def finish(r: dict) -> dict:
        r["centers"] = o_cen
        if o_bb is not None:
            r["bboxes"] = o_bb
        if "depth" in r:
            r["depth"] = r["depth"].reshape(B, K)
        if o_alpha is not None:
            r.pop("rotation")
            r["alpha"] = o_alpha
        if has3d:
            r["locations"], r["yaws"], r["bboxes3d"] = o_loc, o_yaw, o_b3
        return r

    return pf, finish


def fusionDecode(outputs, outputSize=(112, 200), K=100, norm2d=False, post: dict | None = None) -> dict:
    """decode.py:10-174 -- drop-in.  Single pyramid layer (all the reference's networks emit one, dla.py:579).

    post = {center, scale, height, width, calibs[, isGt]}: postProcess (postProcess.py:13-85) runs inside the decode's
    own launch (row f1, cf_fusion_decode_post) and the returned dict is what `postProcess(fusionDecode(...), **post)`
    returns -- bit for bit, one launch fewer."""
    assert isinstance(outputs, list), "output must be a list of dictionaries"
    layers = [o for o in outputs if "heatmap" in o]
    if not layers:
        return {}
    if len(layers) != 1 or len(outputs) != 1:
        raise NotImplementedError("fusionDecode: multi-layer pyramid outputs are not produced by any reference "
                                  "network (dla.py:579) and are not implemented")
    K = int(K)
    keep: list = []
    heads, out, ret, (B, Cc, H, W), heat = _decode_structs(layers[0], K, keep)
    _same_device(*keep)
    ws, wsn = _stage_ws(heat.device, B, Cc, H, W, K)
    pf, finish = _post_fuse(ret, post, keep) if post is not None else (None, None)
    _call(heat.device, "cf_fusion_decode_post", C.byref(heads), B, Cc, H, W, K, int(outputSize[0]), int(outputSize[1]),
                                                int(bool(norm2d)), C.byref(out), C.byref(pf) if pf else None,
                                                ws.data_ptr(), wsn, _stream(heat))
    return finish(ret) if finish else ret


def stage_depth_plane(pc_dep_dev: torch.Tensor, pc_dep_host: torch.Tensor) -> None:
    """Copies plane 0 (depth) of every frame of a pinned host (B,3,H,W) pc_dep into the CUDA tensor `pc_dep_dev`
    (same shape) with one strided asynchronous copy on the current stream (cf_stage_depth_plane)."""
    dev_t = _map(pc_dep_dev, "pc_dep_dev")
    host_t = _map(pc_dep_host, "pc_dep_host", pinned_ok=True)
    if dev_t is not pc_dep_dev or host_t.is_cuda or dev_t.shape != host_t.shape or dev_t.shape[1] != 3:
        raise ValueError("stage_depth_plane: contiguous (B,3,H,W) CUDA destination and pinned host source required")
    B, _, H, W = dev_t.shape
    _call(dev_t.device, "cf_stage_depth_plane", dev_t.data_ptr(), host_t.data_ptr(), B, H, W, _stream(dev_t))


def fusion_stage(output: dict, pc_dep: torch.Tensor, calib: torch.Tensor, config, outputSize=None, norm2d=False,
                 decode_output: dict | None = None, sec_feats: torch.Tensor | None = None,
                 pc_vel: torch.Tensor | None = None, post: dict | None = None):
    """getPcFrustumHeatmap + fusionDecode with ONE pass over the heat-map (cf_fusion_stage).

    post = {center, scale, height, width, calibs[, isGt]}: postProcess runs in the decode epilogue of the same launch
    (row f1); `ret` is then what postProcess returns (see fusionDecode).

    `output` holds the first-stage heads (heatmap, widthHeight, depth, dimension, rotation); `decode_output` the dict
    fusionDecode would receive (defaults to `output`, which then typically also carries reg, depth2, rotation2, ...).
    Returns (pc_hm, ret) exactly as the two reference calls would.

    Host-buffer callers: every head except the heat-map, `calib`, and `pc_dep` may be contiguous PINNED host tensors.
    Heads and the velocity planes of pc_dep are then read in place at the K peak / matched pixels (a few sectors over
    PCIe instead of a copy of the whole tensor); a pinned `pc_dep` has its depth plane staged to the device here.
    `pc_vel`: the caller staged plane 0 into the CUDA tensor `pc_dep` itself and passes the pinned original here.

    `sec_feats` (row f3): a pre-allocated contiguous (B, C_img + 3, H, W) fp32 buffer, the input of the secondary heads
    (ConcateCombiner, fusionModules.py:18-35: cat([imgFeatures, pc_hm], 1)).  pc_hm is then painted straight into its
    last three channels and the returned pc_hm is that view: no torch.cat copy of the image features."""
    K, max_pc_dist = _cfg_k_dist(config)
    heat = _map(_head(output, "heatmap"), "heatmap")
    B, Cc, H, W = heat.shape
    wh = _map(_head(output, "widthHeight"), "widthHeight", pinned_ok=True)
    dep = _map(_head(output, "depth"), "depth", pinned_ok=True)
    dim = _map(_head(output, "dimension"), "dimension", pinned_ok=True)
    rot = _map(_head(output, "rotation"), "rotation", pinned_ok=True)
    pcd_vel = None
    if isinstance(pc_dep, torch.Tensor) and not pc_dep.is_cuda:
        # host-buffer caller: only the depth plane is read in full -> staged to the device here (stream-ordered copy of
        # plane 0 into a cached buffer); the velocity planes are read in place at the matched pixels
        pcd_vel = _map(pc_dep, "pc_dep", pinned_ok=True)
        if pcd_vel.shape != (B, 3, H, W):
            raise ValueError("pc_dep must be (B,3,H,W)")
        pcd = _staging(heat.device, "pc_dep", (B, 3, H, W))
        stage_depth_plane(pcd, pcd_vel)
    else:
        pcd = _map(pc_dep, "pc_dep")
    if pcd.shape != (B, 3, H, W):
        raise ValueError("pc_dep must be (B,3,H,W)")
    if pc_vel is not None:  # explicit split: pc_dep on the device with plane 0 filled, the whole map in pinned memory
        pcd_vel = _map(pc_vel, "pc_vel", pinned_ok=True)
        if pcd_vel.shape != (B, 3, H, W):
            raise ValueError("pc_vel must be (B,3,H,W)")
    cal = _map(calib, "calib", ndim=None, pinned_ok=True).view(B, 3, 4)
    dec = output if decode_output is None else decode_output
    dec_in = dict(dec)
    dec_in["heatmap"] = heat  # same storage: the ABI requires one heat-map
    keep: list = []
    heads, out, ret, _, _ = _decode_structs(dec_in, K, keep)
    _same_device(heat, wh, dep, dim, rot, pcd, cal, sec_feats, *keep)
    if "rotation2" in dec:
        dec["rotation"] = dec.pop("rotation2")
    if outputSize is None:
        outputSize = (H, W)
    if sec_feats is None:
        pc_hm = torch.empty_like(pcd)
        stride = 3 * H * W
    else:
        if not (sec_feats.is_cuda and sec_feats.dtype == torch.float32 and sec_feats.is_contiguous() and
                sec_feats.dim() == 4 and sec_feats.shape[0] == B and sec_feats.shape[1] >= 3 and
                tuple(sec_feats.shape[2:]) == (H, W)):
            raise ValueError("sec_feats must be a contiguous fp32 CUDA tensor (B, C_img + 3, H, W)")
        pc_hm = sec_feats[:, sec_feats.shape[1] - 3:]
        stride = sec_feats.shape[1] * H * W
    ws, wsn = _stage_ws(heat.device, B, Cc, H, W, K)
    pf, finish = _post_fuse(ret, post, keep) if post is not None else (None, None)
    _call(heat.device, "cf_fusion_stage_post", heat.data_ptr(), wh.data_ptr(), dep.data_ptr(), dim.data_ptr(),
                                               rot.data_ptr(), pcd.data_ptr(), _ptr(pcd_vel), cal.data_ptr(), B, Cc, H, W,
                                               K, max_pc_dist, pc_hm.data_ptr(), stride, None, None, None,
                                               C.byref(heads), int(outputSize[0]), int(outputSize[1]),
                                               int(bool(norm2d)), C.byref(out), C.byref(pf) if pf else None,
                                               ws.data_ptr(), wsn, _stream(heat))
    return pc_hm, (finish(ret) if finish else ret)


# ------------------------------------------------------------------------------------------------------
# f1  postProcess                                     src/lib/utils/postProcess.py:13-85
# ------------------------------------------------------------------------------------------------------
def inverse_output_affine(center, scale, width: int, height: int) -> np.ndarray:
    """getAffineTransform(center, scale, 0, (width, height), inverse=True) (utils/image.py:43-83) as a 2x3 fp32
    matrix: the map from output-map pixels back to image pixels.  With no rotation the three point pairs the
    reference hands to cv2.getAffineTransform describe a pure scale + shift, solved here in float64 like cv2 does."""
    c = np.asarray(center, dtype=np.float32).astype(np.float64)
    sc = np.asarray(scale, dtype=np.float32).reshape(-1).astype(np.float64)
    src_w = float(sc[0])
    src = np.zeros((3, 2)); dst = np.zeros((3, 2))
    src[0] = c
    src[1] = c + np.array([0.0, src_w * -0.5])
    dst[0] = [width * 0.5, height * 0.5]
    dst[1] = dst[0] + np.array([0.0, width * -0.5])
    for pts in (src, dst):  # get3rdPoint: b + (a - b) rotated by 90 degrees
        d = pts[0] - pts[1]
        pts[2] = pts[1] + np.array([-d[1], d[0]])
    A = np.hstack([dst.astype(np.float32).astype(np.float64), np.ones((3, 1))])  # inverse: dst -> src
    M = np.linalg.solve(A, src.astype(np.float32).astype(np.float64)).T
    return np.ascontiguousarray(M, dtype=np.float32)


def postProcess(y: dict, center, scale, height, width, calibs, isGt=False) -> dict:
    """postProcess.py:13-85 -- drop-in for the calls at detector.py:341 and progressBar.py:104,196.  Mutates and
    returns `y` like the reference: classIds and velocity are updated in place, rotation is replaced by alpha."""
    scores = _map(y["scores"], "scores", ndim=2)
    B, K = scores.shape
    dev = scores.device
    trans = inverse_output_affine(center, scale, int(width), int(height))
    cls = y["classIds"]
    if not (isinstance(cls, torch.Tensor) and cls.is_cuda and cls.dtype == torch.float32 and cls.is_contiguous()):
        raise ValueError("classIds must be a contiguous fp32 CUDA tensor (updated in place)")
    cen = _map(y["centers"], "centers", ndim=3)
    opt = lambda k, nd: (_map(y[k], k, ndim=nd) if k in y else None)  # noqa: E731
    bb, rot, dim, amo = opt("bboxes", 3), opt("rotation", 3), opt("dimension", 3), opt("amodal_offset", 3)
    dep = _map(y["depth"], "depth", ndim=None).reshape(B, K) if "depth" in y else None
    vel = y.get("velocity")
    if vel is not None and not (vel.is_cuda and vel.dtype == torch.float32 and vel.is_contiguous()):
        raise ValueError("velocity must be a contiguous fp32 CUDA tensor (updated in place)")
    has3d = rot is not None and dep is not None and dim is not None
    cal = _map(calibs, "calibs", ndim=None).reshape(-1, 3, 4) if has3d else None
    if cal is not None and cal.shape[0] != B:
        cal = cal.expand(B, 3, 4).contiguous()
    new = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)  # noqa: E731
    o_cen = new(B, K, 2)
    o_bb = new(B, K, 4) if bb is not None else None
    o_alpha = new(B, K) if rot is not None else None
    o_loc, o_yaw, o_b3 = (new(B, K, 3), new(B, K), new(B, K, 8, 3)) if has3d else (None, None, None)
    pin = PostIn(_ptr(cls), _ptr(cen), _ptr(bb), _ptr(dep), _ptr(rot), _ptr(dim), _ptr(amo), _ptr(vel))
    pout = PostOut(_ptr(o_cen), _ptr(o_bb), _ptr(o_alpha), _ptr(o_loc), _ptr(o_yaw), _ptr(o_b3))
    tarr = (C.c_float * 6)(*[float(v) for v in trans.reshape(-1)])
    _call(dev, "cf_post_process", C.byref(pin), B, K, int(height), int(width), tarr, _ptr(cal),
                                          int(bool(isGt)), C.byref(pout), _stream(scores))
    y["centers"] = o_cen
    if bb is not None:
        y["bboxes"] = o_bb
    if dep is not None:
        y["depth"] = dep
    if rot is not None:
        y.pop("rotation")
        y["alpha"] = o_alpha
    if has3d:
        y["locations"], y["yaws"], y["bboxes3d"] = o_loc, o_yaw, o_b3
    return y


# ------------------------------------------------------------------------------------------------------
# a5 + a6  radar pillar expansion -> pc_dep          src/lib/dataset/generic_dataset.py:738-942 (processPointCloud)
# ------------------------------------------------------------------------------------------------------
def radar_sort(points: torch.Tensor, counts: torch.Tensor, return_order: bool = False):
    """Row f2 -- stable ascending depth sort of every frame's first counts[b] points on the device
    (== points[b][np.argsort(points[b,:n,2], kind="stable")]; nuscenes.py:195-197 / detector.py:281 sort on the host)."""
    pts = _map(points, "points", ndim=3, dtype=torch.float64)
    B, Nmax, _ = pts.shape
    cnt = _map(counts, "counts", ndim=1, dtype=torch.int32)
    out = torch.empty_like(pts)
    order = torch.empty((B, Nmax), dtype=torch.int32, device=pts.device) if return_order else None
    _call(pts.device, "cf_radar_sort", _ptr(pts) if Nmax else None, cnt.data_ptr(), B, Nmax, out.data_ptr(), _ptr(order),
                                        _stream(pts))
    return (out, order) if return_order else out


def radar_ingest(raw: torch.Tensor, counts: torch.Tensor, flip=None, descending: bool = False,
                 return_order: bool = False, packed: bool = False):
    """Row f2, complete -- the host-side steps of loadRadarPointCloud between the pickle and processPointCloud
    (nuscenes.py:159-205; detector.py:255-283) on the device (cf_radar_ingest).

    raw (B, rows >= 10, Nmax) f64: per frame the stored radar array (rows 0-2 x, y, z; rows 8, 9 vx, vz), padded to Nmax
    columns -- or, with packed=True, (B, Nmax, 5) [X,Y,Z,vx,vz] rows; counts (B,) i32.  flip (B,) bool / i32 or None: the frame's image
    is flipped (X, vx change sign; hand the same flags to radar_to_pc_dep).  descending = not DATASET.PC_REVERSE
    (nuscenes.py:196-197); detector.py:281 sorts ascending.  Returns (B, Nmax, 5) f64 sorted [, order (B, Nmax) i32]."""
    r = _map(raw, "raw", ndim=3, dtype=torch.float64)
    B = r.shape[0]
    if packed:
        if r.shape[2] != 5:
            raise ValueError("packed points must be (B, Nmax, 5) [X,Y,Z,vx,vz], got %s" % (tuple(r.shape),))
        rows, Nmax = 0, r.shape[1]
    else:
        if r.shape[1] < 10:
            raise ValueError("raw must be (B, rows >= 10, Nmax), got %s" % (tuple(r.shape),))
        rows, Nmax = r.shape[1], r.shape[2]
    cnt = _map(counts, "counts", ndim=1, dtype=torch.int32)
    fl = None if flip is None else _map(torch.as_tensor(flip, device=r.device).to(torch.int32), "flip", ndim=1,
                                        dtype=torch.int32)
    _same_device(r, cnt, fl)
    out = torch.empty((B, Nmax, 5), dtype=torch.float64, device=r.device)
    order = torch.empty((B, Nmax), dtype=torch.int32, device=r.device) if return_order else None
    _call(r.device, "cf_radar_ingest", _ptr(r) if Nmax else None, cnt.data_ptr(), B, rows, Nmax, _ptr(fl),
                                          int(bool(descending)), out.data_ptr(), _ptr(order), _stream(r))
    return (out, order) if return_order else out


def radar_to_pc_dep(points: torch.Tensor, counts: torch.Tensor, cam: torch.Tensor, calib: torch.Tensor,
                    A_out: torch.Tensor, out_hw, img_wh=(1600, 900), pillar_dims=(1.5, 0.2, 0.2), max_dist=60.0,
                    z_offset=0.0, return_owner=False, sort_points=False, flip=None):
    """points (B,Nmax,5) f64 [X,Y,Z,vx,vz] camera frame, each frame sorted ascending by Z (or sort_points=True: sorted
    on the device first, row f2; owner indices then refer to the SORTED order); counts (B,) i32;
    cam (B,3,3) f64; calib (B,3,4) f64; A_out (B,2,3) f64 -> pc_dep (B,3,H,W) f32 [, owner (B,H,W) i32].
    flip (B,) or None: frames whose image is flipped (nuscenes.py:201-205); their points carry X and vx negated, as
    radar_ingest(..., flip=flip) leaves them."""
    pts = _map(points, "points", ndim=3, dtype=torch.float64)
    B, Nmax, five = pts.shape
    if five != 5:
        raise ValueError("points must be (B,Nmax,5) [X,Y,Z,vx,vz]")
    cnt = _map(counts, "counts", ndim=1, dtype=torch.int32)
    if sort_points and Nmax:
        pts = radar_sort(pts, cnt)
    fl = None if flip is None else _map(torch.as_tensor(flip, device=pts.device).to(torch.int32), "flip", ndim=1,
                                        dtype=torch.int32)
    cam = _map(cam, "cam", ndim=None, dtype=torch.float64).view(B, 3, 3)
    cal = _map(calib, "calib", ndim=None, dtype=torch.float64).view(B, 3, 4)
    A = _map(A_out, "A_out", ndim=None, dtype=torch.float64).view(B, 2, 3)
    if cnt.shape[0] != B:
        raise ValueError("counts must be (B,)")
    H, W = int(out_hw[0]), int(out_hw[1])
    dev = pts.device
    pc_dep = torch.empty((B, 3, H, W), dtype=torch.float32, device=dev)
    owner = torch.empty((B, H, W), dtype=torch.int32, device=dev) if return_owner else None
    n = _lib.lib().cf_pillars_workspace_bytes(B, Nmax)
    ws = _workspace(dev, n)
    pd = (C.c_double * 3)(*[float(v) for v in pillar_dims])
    _call(dev, "cf_radar_pillars_flip", _ptr(pts) if Nmax else None, cnt.data_ptr(), _ptr(fl), B, Nmax, cam.data_ptr(),
                                                cal.data_ptr(), A.data_ptr(), int(img_wh[0]), int(img_wh[1]), H, W, pd,
                                                float(max_dist), float(z_offset), pc_dep.data_ptr(), _ptr(owner),
                                                ws.data_ptr(), ws.numel(), _stream(pts))
    return (pc_dep, owner) if return_owner else pc_dep


# ------------------------------------------------------------------------------------------------------
# original CenterFusion (mrnabati) names
# ------------------------------------------------------------------------------------------------------
_nms = nms
_topk = topk
_gather_feat = getFeature
_tranpose_and_gather_feat = transposeAndGetFeature
generate_pc_hm = getPcFrustumHeatmap
get_pc_hm = cvtPcDepthToHeatmap
pc_dep_to_hm_torch = cvtPcDepthToHeatmap
fusion_decode = fusionDecode
get_dist_thresh = getDistanceThresh
