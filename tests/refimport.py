"""Load the reference's own fusion-stage modules *by file path* (read-only, from /root/reference).

TEST INFRASTRUCTURE ONLY.  Used (a) by tests/golden/make_golden.py to generate the committed
golden vectors and (b) by differential tests that are skipped when /root/reference is absent
(it does not exist on the GPU box).  Nothing in the product package imports this file.

Follows the recipes of SURVEY.md section 8(c): the hot path never calls into nuscenes-devkit,
pyquaternion, lightning, pycocotools, yacs or matplotlib, so those absent packages are replaced by
inert stubs purely to let the reference's `import` lines succeed.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np

REF_ROOT = os.environ.get("CFD3D_REFERENCE_ROOT", "/root/reference")
REF_LIB = os.path.join(REF_ROOT, "src", "lib")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_LIB, "utils", "pointcloud.py"))


def _view_points(points, view, normalize):
    # Same published math as nuscenes-devkit geometry_utils.view_points
    # (src/nuscenes-devkit/python-sdk/nuscenes/utils/geometry_utils.py:18-54 @583dff7):
    # homogeneous multiply by the padded 4x4 view, optional divide by the third row.
    viewpad = np.eye(4)
    viewpad[: view.shape[0], : view.shape[1]] = view
    n = points.shape[1]
    pts = np.concatenate((points, np.ones((1, n))))
    pts = np.dot(viewpad, pts)[:3, :]
    if normalize:
        pts = pts / pts[2:3, :].repeat(3, 0).reshape(3, n)
    return pts


class _Anything:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        return _Anything()


def _stub(name, **attrs):
    mod = sys.modules.get(name)
    if mod is None:
        mod = types.ModuleType(name)
        mod.__path__ = []  # behave as a package so sub-imports resolve
        sys.modules[name] = mod
    for k, v in attrs.items():
        setattr(mod, k, v)
    return mod


class AttrDict(dict):
    """Minimal stand-in for yacs CfgNode: attribute access on a dict tree."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:  # pragma: no cover
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def defrost(self):
        pass

    def freeze(self):
        pass

    def clone(self):
        return self


def make_config(K=100, out_hw=(112, 200), max_pc_dist=60.0, pillar_dims=(1.5, 0.2, 0.2)):
    return AttrDict(
        DEBUG=0,
        MODEL=AttrDict(K=K, OUTPUT_SIZE=list(out_hw), INPUT_SIZE=[out_hw[0] * 4, out_hw[1] * 4]),
        DATASET=AttrDict(
            MAX_PC_DIST=max_pc_dist,
            ONE_HOT_PC=False,
            PC_ROI_METHOD="pillars",
            PILLAR_DIMS=list(pillar_dims),
            PC_Z_OFFSET=0,
            PC_REVERSE=True,
            MAX_PC=1000,
        ),
    )


_loaded = {}


def _load(modname, relpath):
    if modname in _loaded:
        return _loaded[modname]
    path = os.path.join(REF_LIB, relpath)
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    _loaded[modname] = mod
    return mod


def _install_stubs():
    ident = lambda f: f  # noqa: E731
    _stub("nuscenes")
    _stub("nuscenes.nuscenes", NuScenes=_Anything)
    _stub("nuscenes.utils")
    _stub("nuscenes.utils.data_classes", RadarPointCloud=type("RadarPointCloud", (), {}), Box=_Anything)
    _stub(
        "nuscenes.utils.geometry_utils",
        view_points=_view_points,
        transform_matrix=_Anything(),
        BoxVisibility=_Anything,
    )
    _stub("nuscenes.eval")
    _stub("nuscenes.eval.common")
    _stub("nuscenes.eval.common.data_classes", EvalBoxes=_Anything)
    _stub("nuscenes.eval.common.loaders", load_gt=_Anything(), add_center_dist=_Anything(), filter_eval_boxes=_Anything())
    _stub("nuscenes.eval.common.config", config_factory=_Anything())
    _stub("nuscenes.eval.detection")
    _stub("nuscenes.eval.detection.data_classes", DetectionBox=_Anything)
    _stub("nuscenes.eval.detection.render", visualize_sample=_Anything())
    _stub("nuscenes.eval.detection.utils", category_to_detection_name=_Anything())
    _stub("pyquaternion", Quaternion=_Anything)
    _stub("pycocotools")
    _stub("pycocotools.coco", COCO=_Anything)
    _stub("lightning")
    _stub("lightning.pytorch")
    _stub("lightning.pytorch.utilities", rank_zero_only=ident)
    _stub("matplotlib")
    _stub("matplotlib.pyplot")
    # reference-internal packages: give them a path so `from .utils import` style imports work
    model_pkg = _stub("model")
    model_pkg.__path__ = [os.path.join(REF_LIB, "model")]
    utils_pkg = _stub("utils")
    utils_pkg.__path__ = [os.path.join(REF_LIB, "utils")]
    _stub("utils.utils", safe_run=ident, createFolder=_Anything(), AverageMeter=_Anything)
    _stub("config", config=make_config(), updateDatasetAndModelConfig=_Anything())
    ds_pkg = _stub("dataset")
    ds_pkg.__path__ = [os.path.join(REF_LIB, "dataset")]
    dsd_pkg = _stub("dataset.datasets")
    dsd_pkg.__path__ = [os.path.join(REF_LIB, "dataset", "datasets")]


def load_reference():
    """Returns a namespace with the reference's hot-path callables (unmodified code)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    if "ns" in _loaded:
        return _loaded["ns"]
    _install_stubs()
    mu = _load("model.utils", "model/utils.py")
    pc = _load("utils.pointcloud", "utils/pointcloud.py")
    dec = _load("model.decode", "model/decode.py")
    img = _load("utils.image", "utils/image.py")
    ddd = _load("utils.ddd", "utils/ddd.py")
    gd = _load("dataset.generic_dataset", "dataset/generic_dataset.py")
    nus = _load("dataset.datasets.nuscenes", "dataset/datasets/nuscenes.py")
    pp = _load("utils.postProcess", "utils/postProcess.py")
    ns = types.SimpleNamespace(
        model_utils=mu, pointcloud=pc, decode=dec, image=img, ddd=ddd, generic_dataset=gd, nuscenes=nus,
        nms=mu.nms, topk=mu.topk, transposeAndGetFeature=mu.transposeAndGetFeature,
        getPcFrustumHeatmap=pc.getPcFrustumHeatmap, cvtPcDepthToHeatmap=pc.cvtPcDepthToHeatmap,
        getDistanceThresh=pc.getDistanceThresh, get_alpha=pc.get_alpha,
        map_pointcloud_to_image=pc.map_pointcloud_to_image, fusionDecode=dec.fusionDecode,
        getAffineTransform=img.getAffineTransform, postProcess=pp.postProcess,
    )
    _loaded["ns"] = ns
    return ns


def ref_process_point_cloud(ns, pc_2d, pc_3d, A_out, calib, cfg):
    """Run the reference's processPointCloud (generic_dataset.py:738-828) on a bare namespace object."""
    GD = ns.generic_dataset.GenericDataset
    NU = ns.nuscenes.nuScenes
    self = types.SimpleNamespace(config=cfg)
    for name in ("transformPointCloud", "getPcPillarsSize", "processPointCloud"):
        setattr(self, name, types.MethodType(getattr(GD, name), self))
    for name in ("getDepthMap", "drawPcHeat"):
        setattr(self, name, types.MethodType(getattr(NU, name), self))
    img_info = {"calib": np.asarray(calib, dtype=np.float64).tolist()}
    return self.processPointCloud(pc_2d, pc_3d, None, A_out, A_out, img_info)


# ----------------------------------------------------------------------------------------------
# Recipe B: the reference's whole model, random init, on CPU
# ----------------------------------------------------------------------------------------------
def make_model_config(K=100, in_hw=(448, 800), num_classes=10):
    """The keys config/utils.py:69-166 derives for configs/centerfusion_debug.yaml (MODEL.FRUSTUM true,
    FUSION_STRATEGY middle, DLA-34), as an attribute dict.  LOAD_DIR is non-empty so that dla.py:596-598 does not try
    to download ImageNet weights; loadModel is never called (random init)."""
    out_hw = [in_hw[0] // 4, in_hw[1] // 4]
    heads = AttrDict(heatmap=num_classes, reg=2, widthHeight=2, depth=1, rotation=8, dimension=3, amodal_offset=2,
                     nuscenes_att=8, velocity=3, depth2=1, rotation2=8)
    head_conv = AttrDict({h: [256] for h in heads})
    for h in ("depth2", "rotation2", "velocity", "nuscenes_att"):
        head_conv[h] = [256, 256, 256]
    return AttrDict(
        DEBUG=0,
        heads=heads,
        head_conv=head_conv,
        MODEL=AttrDict(ARCH="dla_34", FUSION_STRATEGY="middle", FRUSTUM=True, K=K, INPUT_SIZE=list(in_hw),
                       OUTPUT_SIZE=out_hw, LOAD_DIR="random-init", FREEZE_BACKBONE=True, NORM_EVAL=False, NORM_2D=False,
                       DLA=AttrDict(NODE="DeformConv")),
        DATASET=AttrDict(DATASET="nuscenes", RADAR_PC=True, ONE_HOT_PC=False, MAX_PC_DIST=60.0, MAX_PC=1000,
                         PC_ROI_METHOD="pillars", PILLAR_DIMS=[1.5, 0.2, 0.2], PC_Z_OFFSET=0, PC_REVERSE=True,
                         NUM_CLASSES=num_classes),
        TRAIN=AttrDict(UNCERTAINTY_LOSS=False),
    )


def load_model_modules():
    """The reference's model-side modules (unmodified): model.networks.{fusionModules,detectHeads,base_model,dla},
    model.model, model.losses.  Returns a namespace; load_reference() has run first."""
    ns = load_reference()
    if "model_ns" in _loaded:
        return _loaded["model_ns"]
    if "tune_mode_convbn" not in sys.modules:
        spec = importlib.util.spec_from_file_location("tune_mode_convbn", os.path.join(REF_ROOT, "src", "tune_mode_convbn.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["tune_mode_convbn"] = mod
        spec.loader.exec_module(mod)
    nets = _stub("model.networks")
    nets.__path__ = [os.path.join(REF_LIB, "model", "networks")]
    fm = _load("model.networks.fusionModules", "model/networks/fusionModules.py")
    dh = _load("model.networks.detectHeads", "model/networks/detectHeads.py")
    bm = _load("model.networks.base_model", "model/networks/base_model.py")
    dla = _load("model.networks.dla", "model/networks/dla.py")
    mm = _load("model.model", "model/model.py")
    losses = _load("model.losses", "model/losses.py")
    model_pkg = sys.modules["model"]
    # what model/__init__.py:1-2 exports (the package object here is a stub with the real __path__)
    model_pkg.getModel, model_pkg.loadModel, model_pkg.fusionDecode = mm.getModel, getattr(mm, "loadModel", None), ns.decode.fusionDecode
    mns = types.SimpleNamespace(ref=ns, fusionModules=fm, detectHeads=dh, base_model=bm, dla=dla, model=mm, losses=losses,
                                model_pkg=model_pkg)
    _loaded["model_ns"] = mns
    return mns


def build_model(cfg, seed=0):
    """Random-init DLASeg (21.4 M parameters) in eval mode."""
    import torch

    mns = load_model_modules()
    torch.manual_seed(seed)
    m = mns.model.getModel(cfg)
    m.eval()
    return m
