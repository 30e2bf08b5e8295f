"""Drop-in switch: swaps the B200 ops into an imported reference tree (module-level replacement of exactly the
callables SURVEY.md 8(b) lists -- the reference has no plugin registry, its boundary is plain Python names).

    import utils.pointcloud, model.utils, model.decode          # the reference's own modules
    from centerfusiondetect3d_b200 import patch
    undo = patch.apply()        # BaseModel.forward / Detector.process / ProgressBar now run the CUDA ops
    ...
    undo()

Call sites covered: CenterFusionHead.forward calls `pc.getPcFrustumHeatmap` through the module attribute
(src/lib/model/networks/detectHeads.py:8,177); fusionDecode / nms / topk / transposeAndGetFeature are imported BY NAME
into model/__init__.py:2, detector.py:13, model/progressBar.py:12, model/decode.py:7, utils/pointcloud.py:14 and
model/losses.py:15, so every module that already holds the name is patched too.
"""
from __future__ import annotations

import sys

from . import ops

# reference module name -> {attribute: replacement}
_TARGETS = {
    "utils.pointcloud": ("getPcFrustumHeatmap", "cvtPcDepthToHeatmap", "getDistanceThresh", "get_alpha"),
    "model.utils": ("nms", "topk", "getFeature", "transposeAndGetFeature"),
    "model.decode": ("fusionDecode",),
    "utils.postProcess": ("postProcess",),
}
# modules that imported the names with `from ... import ...`
_BY_NAME = ("model", "detector", "model.progressBar", "model.decode", "utils.pointcloud", "model.losses",
            "lib.detector", "lib.model", "lib.model.progressBar")
_NAMES = ("nms", "topk", "getFeature", "transposeAndGetFeature", "fusionDecode", "getPcFrustumHeatmap",
          "cvtPcDepthToHeatmap", "postProcess")


def apply(modules=None, torch_branch_only: bool = True):
    """Patches every already-imported reference module; returns a function that restores the originals.
    `modules`: optional {name: module} to patch instead of sys.modules (tests load the reference by file path)."""
    mods = dict(sys.modules) if modules is None else dict(modules)
    saved = []

    def swap(mod, attr):
        if mod is None or not hasattr(mod, attr):
            return
        orig = getattr(mod, attr)
        new = getattr(ops, attr)
        if attr == "cvtPcDepthToHeatmap" and torch_branch_only:
            new = _dual_cvt(orig)
        if orig is new:
            return
        saved.append((mod, attr, orig))
        setattr(mod, attr, new)

    for mname, attrs in _TARGETS.items():
        for attr in attrs:
            swap(mods.get(mname), attr)
    for mname in _BY_NAME:
        for attr in _NAMES:
            swap(mods.get(mname), attr)

    def undo():
        for mod, attr, orig in reversed(saved):
            setattr(mod, attr, orig)
        saved.clear()

    return undo


def _dual_cvt(orig):
    """cvtPcDepthToHeatmap is dual NumPy / torch in the reference (the NumPy form builds training targets in the
    DataLoader, src/lib/dataset/generic_dataset.py:680-687): CUDA tensors go to the kernel, anything else keeps
    the reference's own code -- that is the reference's CPU data path, not a fallback of the CUDA op."""
    import torch

    def cvtPcDepthToHeatmap(pc_hm, pc_dep, depth, bbox, distanceThreshold, max_pc_dist):
        if isinstance(pc_hm, torch.Tensor) and pc_hm.is_cuda:
            return ops.cvtPcDepthToHeatmap(pc_hm, pc_dep, depth, bbox, distanceThreshold, max_pc_dist)
        return orig(pc_hm, pc_dep, depth, bbox, distanceThreshold, max_pc_dist)

    return cvtPcDepthToHeatmap
