"""Drop-in switch: swaps the B200 ops into an imported reference tree (module-level replacement of exactly the
callables SURVEY.md 8(b) lists -- the reference has no plugin registry, its boundary is plain Python names).

    import utils.pointcloud, model.utils, model.decode          # the reference's own modules
    from centerfusiondetect3d_b200 import patch
    undo = patch.apply()        # BaseModel.forward / Detector.process / ProgressBar now run the CUDA ops
    ...
    undo()

What is replaced (reference src/lib/...):
  * utils.pointcloud.getPcFrustumHeatmap -- reached through the module attribute `pc.getPcFrustumHeatmap` by
    CenterFusionHead.forward (model/networks/detectHeads.py:8,176-177);
  * fusionDecode where it was imported BY NAME: model/__init__.py:2, detector.py:13, model/progressBar.py:12, and
    model.decode itself;
  * nms / topk / transposeAndGetFeature / getFeature in model.utils and in the modules that imported them for the
    fusion path: model/decode.py:7, utils/pointcloud.py:14;
  * get_alpha / getDistanceThresh / cvtPcDepthToHeatmap in utils.pointcloud; postProcess in utils.postProcess,
    detector.py:11, model/progressBar.py:11.

What is deliberately NOT replaced: model.losses (model/losses.py:15,18 imports transposeAndGetFeature / get_alpha for
the training losses, whose inputs need autograd), the DataLoader (dataset/generic_dataset.py:673-687, NumPy), and
everything else of the reference.

Every replacement is a DISPATCHER, not the bare op: CUDA fp32 tensors that do not take part in autograd go to the
kernels; anything else -- NumPy arrays (the DataLoader's target generation), CPU tensors, other dtypes (AMP), tensors
that require grad while grad mode is on (a loss that picked the name up after apply()) -- keeps the reference's own
code.  That is the reference's unchanged CPU / training path, not a fallback of the CUDA ops: ops.* themselves still
refuse anything but CUDA tensors.
"""
from __future__ import annotations

import functools
import sys

import torch

from . import ops

# reference module -> attributes replaced on the module itself (definitions and module-attribute call sites)
_TARGETS = {
    "utils.pointcloud": ("getPcFrustumHeatmap", "cvtPcDepthToHeatmap", "getDistanceThresh", "get_alpha"),
    "model.utils": ("nms", "topk", "getFeature", "transposeAndGetFeature"),
    "model.decode": ("fusionDecode",),
    "utils.postProcess": ("postProcess",),
}
# modules that hold the names through `from ... import ...` and sit on the fusion / decode path
_BY_NAME = {
    "model": ("fusionDecode",),
    "model.decode": ("nms", "topk", "transposeAndGetFeature"),
    "utils.pointcloud": ("topk", "transposeAndGetFeature"),
    "detector": ("fusionDecode", "postProcess"),
    "model.progressBar": ("fusionDecode", "postProcess"),
    "lib.detector": ("fusionDecode", "postProcess"),
    "lib.model": ("fusionDecode",),
    "lib.model.progressBar": ("fusionDecode", "postProcess"),
}
# never touched, whatever they import (checked by tests/test_patch_reference.py)
NOT_PATCHED = ("model.losses", "dataset.generic_dataset", "dataset.datasets.nuscenes")


def _tensors(obj):
    """The tensors / arrays an argument carries (dicts and lists of dicts are the heads' containers)."""
    if isinstance(obj, dict):
        for v in obj.values():
            yield from _tensors(v)
    elif isinstance(obj, (list, tuple)):
        for v in obj:
            yield from _tensors(v)
    elif isinstance(obj, torch.Tensor) or type(obj).__module__ == "numpy":
        yield obj


def takes_cuda_path(*args, check_grad: bool = True) -> bool:
    """True when a call with these arguments belongs to the kernels: at least one tensor, every tensor on a CUDA device,
    every floating tensor fp32, and (check_grad) none of them part of an autograd graph being recorded."""
    seen = False
    grad_on = check_grad and torch.is_grad_enabled()
    for a in args:
        for t in _tensors(a):
            if not isinstance(t, torch.Tensor) or not t.is_cuda:
                return False
            if t.is_floating_point() and t.dtype != torch.float32:
                return False
            if grad_on and t.requires_grad:
                return False
            seen = True
    return seen


# name -> (positions of the arguments that decide the path, whether autograd participation keeps the reference's code).
# The stage-level callables are post-processing the reference itself never differentiates through (top-K indices, a
# Python loop of slice writes; Detector.process and the Lightning validation loop run them under no_grad) -- for the
# primitives the losses share (model/losses.py:102-353) a tensor that requires grad must keep its grad_fn.
_RULES = {
    "getPcFrustumHeatmap": ((0, 1, 2), False),
    "fusionDecode": ((0,), False),
    "postProcess": ((0, 5), False),
    "cvtPcDepthToHeatmap": ((0, 1), True),
    "getDistanceThresh": ((0, 1, 2, 3), True),
    "get_alpha": ((0,), True),
    "nms": ((0,), True),
    "topk": ((0,), True),
    "getFeature": ((0, 1), True),
    "transposeAndGetFeature": ((0, 1), True),
}
_KWNAMES = {
    "getPcFrustumHeatmap": ("output", "pc_dep", "calib", "config"),
    "fusionDecode": ("outputs", "outputSize", "K", "norm2d"),
    "postProcess": ("y", "center", "scale", "height", "width", "calibs", "isGt"),
    "cvtPcDepthToHeatmap": ("pc_hm", "pc_dep", "depth", "bbox", "distanceThreshold", "max_pc_dist"),
    "getDistanceThresh": ("calib", "center", "dim", "alpha"),
    "get_alpha": ("rotation",),
    "nms": ("heatmap", "kernel"),
    "topk": ("heatmap", "K"),
    "getFeature": ("feature", "indices"),
    "transposeAndGetFeature": ("feature", "indices"),
}


def _dispatch(name, orig, new):
    pos, check_grad = _RULES[name]
    kwn = _KWNAMES[name]

    @functools.wraps(orig)
    def call(*args, **kwargs):
        deciding = [args[i] if i < len(args) else kwargs.get(kwn[i]) for i in pos]
        if takes_cuda_path(*deciding, check_grad=check_grad):
            return new(*args, **kwargs)
        return orig(*args, **kwargs)

    call.__cfd3d_original__ = orig
    call.__cfd3d_op__ = new
    return call


def apply(modules=None):
    """Patches every already-imported reference module; returns a function that restores the originals.
    `modules`: optional {name: module} to patch instead of sys.modules (tests load the reference by file path)."""
    mods = dict(sys.modules) if modules is None else dict(modules)
    saved = []
    made = {}  # one dispatcher per original callable, shared by every module that holds the name

    def swap(mod, attr):
        if mod is None or not hasattr(mod, attr):
            return
        orig = getattr(mod, attr)
        if hasattr(orig, "__cfd3d_original__"):
            return  # already patched
        key = id(orig)
        if key not in made:
            made[key] = _dispatch(attr, orig, getattr(ops, attr))
        saved.append((mod, attr, orig))
        setattr(mod, attr, made[key])

    for mname, attrs in _TARGETS.items():
        for attr in attrs:
            swap(mods.get(mname), attr)
    for mname, attrs in _BY_NAME.items():
        for attr in attrs:
            swap(mods.get(mname), attr)

    def undo():
        for mod, attr, orig in reversed(saved):
            setattr(mod, attr, orig)
        saved.clear()

    return undo
