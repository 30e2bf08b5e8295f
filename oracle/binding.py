"""ctypes/numpy binding of the CPU oracle (oracle/cfd3d_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of cfd3d_oracle.c.  Imported by tests/, by
__graft_entry__.smoke() (as the checker) and by bench.py's cpu_baseline / --impl reference legs.
The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libcfd3d_oracle.so")
_lib = None

_f32p = C.POINTER(C.c_float)
_f64p = C.POINTER(C.c_double)
_i64p = C.POINTER(C.c_int64)
_i32p = C.POINTER(C.c_int32)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "cfd3d_oracle.c")
    if force or not os.path.isfile(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libcfd3d_oracle.so"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.isfile(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.cfo_get_alpha.restype = C.c_float
        _lib.cfo_dist_thresh_one.restype = C.c_float
        _lib.cfo_pc_dep_to_hm.restype = C.c_int32
        _lib.cfo_pc_dep_to_hm_np.restype = C.c_int32
    return _lib


def set_threads(n: int) -> None:
    lib().cfo_set_threads(int(n))


def max_threads() -> int:
    return int(lib().cfo_max_threads())


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a, t):
    return None if a is None else a.ctypes.data_as(t)


def nms(heat):
    heat = _f32(heat)
    B, Cc, H, W = heat.shape
    out = np.empty_like(heat)
    lib().cfo_nms(_p(heat, _f32p), B, Cc, H, W, _p(out, _f32p))
    return out


def topk(heat, K=100, apply_nms=False):
    heat = _f32(heat)
    B, Cc, H, W = heat.shape
    scores = np.empty((B, K), np.float32)
    inds = np.empty((B, K), np.int64)
    classes = np.empty((B, K), np.int32)
    ys = np.empty((B, K), np.int64)
    xs = np.empty((B, K), np.int64)
    rc = lib().cfo_topk(_p(heat, _f32p), B, Cc, H, W, K, int(bool(apply_nms)), _p(scores, _f32p), _p(inds, _i64p),
                        _p(classes, _i32p), _p(ys, _i64p), _p(xs, _i64p))
    if rc != 0:
        raise ValueError("K out of range")
    return scores, inds, classes, ys, xs


def gather_nchw(feat, inds):
    feat = _f32(feat)
    inds = np.ascontiguousarray(inds, dtype=np.int64)
    B, ch = feat.shape[:2]
    HW = int(np.prod(feat.shape[2:]))
    M = inds.shape[1]
    out = np.empty((B, M, ch), np.float32)
    lib().cfo_gather_nchw(_p(feat, _f32p), B, ch, C.c_long(HW), _p(inds, _i64p), M, _p(out, _f32p))
    return out


def get_alpha(rot):
    rot = _f32(rot)
    n = rot.size // 8
    out = np.empty(rot.shape[:-1], np.float32)
    lib().cfo_get_alpha_batch(_p(rot, _f32p), C.c_long(n), _p(out, _f32p))
    return out


def dist_thresh(calib, center, dim, alpha):
    calib = _f32(calib).reshape(-1, 3, 4)
    center = _f32(center)
    dim = _f32(dim)
    alpha = _f32(alpha)
    B, K = center.shape[:2]
    out = np.empty((B, K), np.float32)
    lib().cfo_dist_thresh(_p(calib, _f32p), _p(center, _f32p), _p(dim, _f32p), _p(alpha, _f32p), B, K, _p(out, _f32p))
    return out


def pc_dep_to_hm(pc_hm, pc_dep, depth, bbox, thr, max_pc_dist):
    """In-place on pc_hm (3,H,W) float32 C-contiguous; returns matched flat pixel or -1."""
    assert pc_hm.dtype == np.float32 and pc_hm.flags.c_contiguous
    pc_dep = _f32(pc_dep)
    bbox = _f32(bbox)
    _, H, W = pc_hm.shape
    return int(lib().cfo_pc_dep_to_hm(_p(pc_hm, _f32p), _p(pc_dep, _f32p), H, W, C.c_float(float(depth)),
                                      _p(bbox, _f32p), C.c_float(float(thr)), C.c_float(float(max_pc_dist))))


def pc_dep_to_hm_np(pc_hm, pc_dep, depth, bbox, thr, max_pc_dist):
    """The NumPy branch as the DataLoader calls it (generic_dataset.py:673-687): depth a Python float, thr np.float32 or
    np.float64, bbox float32 or float64 (the dtypes select the arithmetic, NumPy >= 2 promotion).  In place on pc_hm; returns the matched pixel or -1."""
    assert pc_hm.dtype == np.float32 and pc_hm.flags.c_contiguous
    pc_dep = _f32(pc_dep)
    bbox = np.asarray(bbox)
    f32 = 1 if bbox.dtype == np.float32 else 0
    bb = _f64(bbox)
    t32 = 1 if isinstance(thr, np.float32) or getattr(thr, "dtype", None) == np.float32 else 0
    _, H, W = pc_hm.shape
    return int(lib().cfo_pc_dep_to_hm_np(_p(pc_hm, _f32p), _p(pc_dep, _f32p), H, W, C.c_double(float(depth)),
                                         _p(bb, _f64p), C.c_double(float(thr)), C.c_float(float(max_pc_dist)), f32, t32))


def frustum_pc_hm(heat, wh, depth, dim, rot, pc_dep, calib, K, max_pc_dist):
    heat, wh, depth, dim, rot, pc_dep = map(_f32, (heat, wh, depth, dim, rot, pc_dep))
    calib = _f32(calib).reshape(-1, 3, 4)
    B, Cc, H, W = heat.shape
    pc_hm = np.empty((B, 3, H, W), np.float32)
    assoc = np.empty((B, K), np.int32)
    thr = np.empty((B, K), np.float32)
    inds = np.empty((B, K), np.int64)
    rc = lib().cfo_frustum_pc_hm(_p(heat, _f32p), _p(wh, _f32p), _p(depth, _f32p), _p(dim, _f32p), _p(rot, _f32p),
                                 _p(pc_dep, _f32p), _p(calib, _f32p), B, Cc, H, W, K, C.c_float(float(max_pc_dist)),
                                 _p(pc_hm, _f32p), _p(assoc, _i32p), _p(thr, _f32p), _p(inds, _i64p))
    if rc != 0:
        raise ValueError("K out of range")
    return pc_hm, assoc, thr, inds


_DEC_HEADS = ("reg", "widthHeight", "depth", "rotation", "dimension", "amodal_offset", "nuscenes_att", "velocity",
              "uncertainty")
_DEC_CH = {"reg": 2, "widthHeight": 2, "depth": 1, "rotation": 8, "dimension": 3, "amodal_offset": 2,
           "nuscenes_att": 8, "velocity": 3, "uncertainty": 1}


def fusion_decode(output: dict, out_hw=(112, 200), K=100, norm2d=False):
    """`output`: one layer's dict of numpy maps using the reference's key names (depth2 / rotation2 honoured)."""
    heat = _f32(output["heatmap"])
    B, Cc, H, W = heat.shape
    heads = {}
    for name in _DEC_HEADS:
        src = name
        if name == "depth" and "depth2" in output:
            src = "depth2"
        if name == "rotation" and "rotation2" in output:
            src = "rotation2"
        heads[name] = _f32(output[src]) if src in output else None
    ret = {
        "scores": np.empty((B, K), np.float32),
        "classIds": np.empty((B, K), np.float32),
        "centers": np.empty((B, K, 2), np.float32),
    }
    outs = {}
    if heads["widthHeight"] is not None:
        outs["bboxes"] = np.empty((B, K, 4), np.float32)
    for name in ("depth", "rotation", "dimension", "amodal_offset", "nuscenes_att", "velocity"):
        if heads[name] is not None:
            outs[name] = np.empty((B, K, _DEC_CH[name]), np.float32)
    inds = np.empty((B, K), np.int64)
    rc = lib().cfo_fusion_decode(
        _p(heat, _f32p), _p(heads["reg"], _f32p), _p(heads["widthHeight"], _f32p), _p(heads["depth"], _f32p),
        _p(heads["rotation"], _f32p), _p(heads["dimension"], _f32p), _p(heads["amodal_offset"], _f32p),
        _p(heads["nuscenes_att"], _f32p), _p(heads["velocity"], _f32p), _p(heads["uncertainty"], _f32p),
        B, Cc, H, W, K, int(out_hw[0]), int(out_hw[1]), int(bool(norm2d)),
        _p(ret["scores"], _f32p), _p(ret["classIds"], _f32p), _p(ret["centers"], _f32p), _p(outs.get("bboxes"), _f32p),
        _p(outs.get("depth"), _f32p), _p(outs.get("rotation"), _f32p), _p(outs.get("dimension"), _f32p),
        _p(outs.get("amodal_offset"), _f32p), _p(outs.get("nuscenes_att"), _f32p), _p(outs.get("velocity"), _f32p),
        _p(inds, _i64p))
    if rc != 0:
        raise ValueError("K out of range")
    ret.update(outs)
    ret["_inds"] = inds
    return ret


def radar_ingest(raw, counts, flip=None, descending=False):
    """nuscenes.py:159-205 / detector.py:255-283 between the pickle and processPointCloud, restated in NumPy: rows
    0,1,2,8,9 of the stored (rows, N) array, stable depth sort (descending = the reversed stable order, index[::-1] of
    nuscenes.py:196-197), flip of X and vx (:203-204).  raw (B, rows, Nmax) f64, counts (B,) -> (B, Nmax, 5) f64 and
    the order (B, Nmax) i32 (source column of every sorted row).  The filters the reference applies before its sort drop
    points without changing the order of the others: they live in radar_pillars."""
    raw = _f64(raw)
    B, rows, Nmax = raw.shape
    out = np.empty((B, Nmax, 5), np.float64)
    order = np.empty((B, Nmax), np.int32)
    for b in range(B):
        n = min(Nmax, max(0, int(counts[b])))
        pts = raw[b][[0, 1, 2, 8, 9], :].T.copy()  # (Nmax, 5)
        idx = np.argsort(pts[:n, 2], kind="stable")
        if descending:
            idx = idx[::-1]
        idx = np.concatenate([idx, np.arange(n, Nmax)]).astype(np.int64)
        srt = pts[idx]
        if flip is not None and flip[b]:
            srt[:n, 0] *= -1
            srt[:n, 3] *= -1
        out[b] = srt
        order[b] = idx
    return out, order


def radar_pillars(points, counts, cam, calib, A, img_wh, out_hw, pillar_dims=(1.5, 0.2, 0.2), max_dist=60.0,
                  z_offset=0.0, want_owner=True, flip=None):
    points = _f64(points)
    B, Nmax, five = points.shape
    assert five == 5
    counts = np.ascontiguousarray(counts, dtype=np.int32)
    cam = _f64(cam).reshape(B, 3, 3)
    calib = _f64(calib).reshape(B, 3, 4)
    A = _f64(A).reshape(B, 2, 3)
    H, W = out_hw
    pd = _f64(pillar_dims)
    pc_dep = np.empty((B, 3, H, W), np.float32)
    owner = np.empty((B, H, W), np.int32) if want_owner else None
    flip = None if flip is None else np.ascontiguousarray(flip, dtype=np.int32)
    lib().cfo_radar_pillars(_p(points, _f64p), _p(counts, _i32p), _p(flip, _i32p), B, Nmax, _p(cam, _f64p), _p(calib, _f64p),
                            _p(A, _f64p), int(img_wh[0]), int(img_wh[1]), H, W, _p(pd, _f64p),
                            C.c_double(float(max_dist)), C.c_double(float(z_offset)), _p(pc_dep, _f32p),
                            _p(owner, _i32p))
    return pc_dep, owner


def post_process(y: dict, trans, out_hw, calib, is_gt=False):
    """postProcess (utils/postProcess.py:13-85) on NumPy copies of a decode dict; returns a new dict with the
    reference's keys.  `trans` = the 2x3 fp32 inverse output affine (getAffineTransform(..., inverse=True))."""
    B, K = y["scores"].shape
    opt = lambda k: (_f32(y[k]).copy() if k in y else None)  # noqa: E731
    cls, cen = _f32(y["classIds"]).copy(), _f32(y["centers"])
    bb, dep, rot, dim, amo, vel = (opt(k) for k in ("bboxes", "depth", "rotation", "dimension", "amodal_offset", "velocity"))
    has3d = rot is not None and dep is not None and dim is not None
    o_cen = np.empty((B, K, 2), np.float32)
    o_bb = np.empty((B, K, 4), np.float32) if bb is not None else None
    o_alpha = np.empty((B, K), np.float32) if rot is not None else None
    o_loc = np.empty((B, K, 3), np.float32) if has3d else None
    o_yaw = np.empty((B, K), np.float32) if has3d else None
    o_b3 = np.empty((B, K, 8, 3), np.float32) if has3d else None
    pp = lambda a: (_p(a, _f32p) if a is not None else None)  # noqa: E731
    lib().cfo_post_process(pp(cls), pp(cen), pp(bb), pp(dep), pp(rot), pp(dim), pp(amo), pp(vel), B, K,
                           int(out_hw[0]), int(out_hw[1]), pp(_f32(trans)), pp(_f32(calib).reshape(-1, 3, 4)),
                           int(bool(is_gt)), pp(o_cen), pp(o_bb), pp(o_alpha), pp(o_loc), pp(o_yaw), pp(o_b3))
    out = {k: v for k, v in y.items() if k != "rotation"}
    out["classIds"], out["centers"] = cls, o_cen
    if bb is not None:
        out["bboxes"] = o_bb
    if dep is not None:
        out["depth"] = dep.reshape(B, K)
    if rot is not None:
        out["alpha"] = o_alpha
    if vel is not None:
        out["velocity"] = vel
    if has3d:
        out["locations"], out["yaws"], out["bboxes3d"] = o_loc, o_yaw, o_b3
    return out
