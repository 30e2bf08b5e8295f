"""ctypes binding of libcfd3d_b200.so (include/cfd3d_b200.h).

The library is the product: there is NO CPU or PyTorch fallback.  If the shared object is missing or a
symbol is absent, importing this module's `lib()` raises; the build recipe is
`python -m centerfusiondetect3d_b200.build` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CFD3D_LIB: development override (the -DCFD3D_PROFILE build of tools/prof_phases.py, A/B builds)
LIB_PATH = os.environ.get("CFD3D_LIB") or os.path.join(_HERE, "libcfd3d_b200.so")

_vp = C.c_void_p


class DecodeHeads(C.Structure):
    """cf_decode_heads_t"""
    _fields_ = [(n, _vp) for n in ("heatmap", "reg", "wh", "depth", "rot", "dim", "amodal", "nuscenes_att",
                                   "velocity", "uncertainty")]


class DecodeOut(C.Structure):
    """cf_decode_out_t"""
    _fields_ = [(n, _vp) for n in ("scores", "class_ids", "centers", "bboxes", "depth", "rot", "dim", "amodal",
                                   "nuscenes_att", "velocity", "inds")]


class PostIn(C.Structure):
    """cf_post_in_t"""
    _fields_ = [(n, _vp) for n in ("class_ids", "centers", "bboxes", "depth", "rot", "dim", "amodal", "velocity")]


class PlanInfo(C.Structure):
    """cf_plan_info_t"""
    _fields_ = [(n, C.c_int) for n in ("cluster_size", "n_clusters", "rounds", "resident", "depth_plane_staged",
                                       "chunk_rows", "slots", "smem_bytes")]


class PostOut(C.Structure):
    """cf_post_out_t"""
    _fields_ = [(n, _vp) for n in ("centers", "bboxes", "alpha", "locations", "yaws", "bboxes3d")]


# every symbol include/cfd3d_b200.h declares: name -> (restype, argtypes)
_I, _F, _D, _Z, _L = C.c_int, C.c_float, C.c_double, C.c_size_t, C.c_int64
class PostFuse(C.Structure):  # cf_post_fuse_t
    _fields_ = [("trans_host", C.POINTER(C.c_float)), ("calib", C.c_void_p), ("height", C.c_int), ("width", C.c_int),
                ("is_gt", C.c_int), ("out", PostOut)]


SYMBOLS = {
    "cf_version": (_I, []),
    "cf_last_error": (C.c_char_p, []),
    "cf_last_launch_count": (_I, []),
    "cf_set_kernel_events": (_I, [_I, _vp, _vp]),
    "cf_last_plan": (_I, [C.POINTER(PlanInfo)]),
    "cf_debug_set_plan": (_I, [_I, _I]),
    "cf_workspace_bytes": (_Z, [_I] * 5),
    "cf_pillars_workspace_bytes": (_Z, [_I, _I]),
    "cf_nms": (_I, [_vp, _I, _I, _I, _I, _vp, _vp]),
    "cf_peak_topk": (_I, [_vp, _I, _I, _I, _I, _I, _I, _vp, _vp, _vp, _vp, _vp, _vp, _Z, _vp]),
    "cf_gather_nchw": (_I, [_vp, _I, _I, _L, _vp, _I, _vp, _vp]),
    "cf_get_alpha": (_I, [_vp, _L, _vp, _vp]),
    "cf_dist_thresh": (_I, [_vp, _vp, _vp, _vp, _I, _I, _vp, _vp]),
    "cf_frustum_pc_hm": (_I, [_vp] * 7 + [_I] * 5 + [_F] + [_vp] * 4 + [_vp, _Z, _vp]),
    "cf_pc_dep_to_hm_single": (_I, [_vp, _vp, _I, _I, _vp, _vp, _vp, _F, _vp, _vp]),
    "cf_fusion_decode": (_I, [C.POINTER(DecodeHeads), _I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(DecodeOut), _vp, _Z,
                              _vp]),
    "cf_fusion_stage": (_I, [_vp] * 7 + [_I] * 5 + [_F] + [_vp] * 4 + [C.POINTER(DecodeHeads), _I, _I, _I,
                                                                       C.POINTER(DecodeOut), _vp, _Z, _vp]),
    "cf_fusion_stage_strided": (_I, [_vp] * 7 + [_I] * 5 + [_F] + [_vp, _L] + [_vp] * 3 + [C.POINTER(DecodeHeads), _I, _I, _I,
                                                                                        C.POINTER(DecodeOut), _vp, _Z, _vp]),
    "cf_fusion_stage_split": (_I, [_vp] * 8 + [_I] * 5 + [_F] + [_vp, _L] + [_vp] * 3 + [C.POINTER(DecodeHeads), _I, _I, _I,
                                                                                      C.POINTER(DecodeOut), _vp, _Z, _vp]),
    "cf_fusion_decode_post": (_I, [C.POINTER(DecodeHeads), _I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(DecodeOut),
                                   C.POINTER(PostFuse), _vp, _Z, _vp]),
    "cf_fusion_stage_post": (_I, [_vp] * 8 + [_I] * 5 + [_F] + [_vp, _L] + [_vp] * 3 + [C.POINTER(DecodeHeads), _I, _I, _I,
                                                                                     C.POINTER(DecodeOut),
                                                                                     C.POINTER(PostFuse), _vp, _Z, _vp]),
    "cf_stage_depth_plane": (_I, [_vp, _vp, _I, _I, _I, _vp]),
    "cf_radar_sort": (_I, [_vp, _vp, _I, _I, _vp, _vp, _vp]),
    "cf_radar_ingest": (_I, [_vp, _vp, _I, _I, _I, _vp, _I, _vp, _vp, _vp]),
    "cf_frustum_boxes": (_I, [_vp] * 5 + [_I] * 4 + [_F, _vp, _vp, _vp, _Z, _vp]),
    "cf_frustum_boxes_static": (_I, [_vp] * 5 + [_I] * 4 + [_F, _vp, _vp, _vp]),
    "cf_frustum_boxes_f64": (_I, [_vp] * 5 + [_I] * 4 + [_F, _I, _vp, _vp, _vp, _Z, _vp]),
    "cf_post_process": (_I, [C.POINTER(PostIn), _I, _I, _I, _I, C.POINTER(C.c_float), _vp, _I, C.POINTER(PostOut), _vp]),
    "cf_radar_pillars": (_I, [_vp, _vp, _I, _I, _vp, _vp, _vp, _I, _I, _I, _I, C.POINTER(_D), _D, _D, _vp, _vp, _vp,
                              _Z, _vp]),
    "cf_radar_pillars_flip": (_I, [_vp, _vp, _vp, _I, _I, _vp, _vp, _vp, _I, _I, _I, _I, C.POINTER(_D), _D, _D, _vp, _vp,
                                   _vp, _Z, _vp]),
}

_lib = None


class CudaLibraryMissing(ImportError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise CudaLibraryMissing(
                "%s not found: build it with `python -m centerfusiondetect3d_b200.build` "
                "(there is no CPU fallback for the fusion stage)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the .so does not export what the header declares
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class CfError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__("libcfd3d_b200: %s (code %d)" % (msg, code))
        self.code = code


def check(rc: int) -> None:
    """Maps the ABI's return code to the exception the reference's torch code would raise."""
    if rc == 0:
        return
    msg = lib().cf_last_error().decode("utf-8", "replace")
    if rc == -3:  # CF_E_K: torch.topk raises RuntimeError("selected index k out of range")
        raise RuntimeError(msg)
    if rc < 0:
        raise ValueError("libcfd3d_b200: %s (code %d)" % (msg, rc))
    raise CfError(rc, msg)
