"""centerfusiondetect3d_b200 -- B200-native (sm_100a CUDA) implementation of CenterFusion's frustum-based
radar-camera fusion stage, behind the reference's own Python call signatures.

    from centerfusiondetect3d_b200 import ops          # nms, topk, getPcFrustumHeatmap, fusionDecode, ...
    from centerfusiondetect3d_b200 import patch        # swap the ops into an imported reference tree

The compute lives in libcfd3d_b200.so (csrc/, C ABI in include/cfd3d_b200.h); build it in-tree with
`python -m centerfusiondetect3d_b200.build`.  There is no CPU fallback.
"""
__version__ = "0.1.0"
