"""Seeded synthetic inputs for the fusion stage.

Pure NumPy, deterministic for a given seed on any machine.  These are the first-stage head maps a
CenterFusion network would hand to the fusion stage (reference call site:
src/lib/model/networks/detectHeads.py:165-191) plus camera-frame radar points in the layout of
src/convert_nuScenes.py's pickles (rows x,y,z and vx,vz), already depth-sorted as
src/lib/dataset/datasets/nuscenes.py:193-199 leaves them.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

NUSC_CALIB = np.array(
    [[1266.417203046554, 0.0, 816.2670197447984, 0.0],
     [0.0, 1266.417203046554, 491.50706579294757, 0.0],
     [0.0, 0.0, 1.0, 0.0]], dtype=np.float64)

# channel counts of the heads the stage touches (reference: src/lib/config/utils.py:69-106)
FORWARD_HEADS = {"widthHeight": 2, "depth": 1, "dimension": 3, "rotation": 8}
DECODE_HEADS = {"reg": 2, "widthHeight": 2, "depth2": 1, "rotation2": 8, "dimension": 3, "amodal_offset": 2,
                "nuscenes_att": 8, "velocity": 3}


@dataclass
class Workload:
    """One named configuration of BASELINE.json:configs."""
    name: str
    B: int
    H: int
    W: int
    K: int
    n_radar: int
    C: int = 10
    blobs: int = 40
    img_w: int = 1600
    img_h: int = 900
    with_pillars: bool = False
    max_pc_dist: float = 60.0
    pillar_dims: tuple = (1.5, 0.2, 0.2)
    extra: dict = field(default_factory=dict)


WORKLOADS = {
    "cfg1": Workload("cfg1_debug_1x448x800_K100_N100", 1, 112, 200, 100, 100),
    "cfg2": Workload("cfg2_middle_B32_448x800_K100_N300", 32, 112, 200, 100, 300),
    "cfg3": Workload("cfg3_B256_448x800_K100_N300", 256, 112, 200, 100, 300),
    "cfg4": Workload("cfg4_dense_B64_448x800_K100_N2000_pillars", 64, 112, 200, 100, 2000, with_pillars=True),
    "cfg5": Workload("cfg5_fullres_B64_900x1600_K500_N300", 64, 225, 400, 500, 300, blobs=200),
}


def output_affine(H: int, W: int, img_w: int = 1600, img_h: int = 900) -> np.ndarray:
    """2x3 affine image px -> output-map px, what utils/image.py:43-83 getAffineTransform returns for
    center=(img_w/2, img_h/2), scale=max(img_w,img_h), rot=0, outputSize=(W,H)."""
    s = W / float(max(img_w, img_h))
    return np.array([[s, 0.0, W * 0.5 - s * img_w * 0.5], [0.0, s, H * 0.5 - s * img_h * 0.5]], dtype=np.float64)


def make_heatmap(rng: np.random.Generator, B, C, H, W, blobs) -> np.ndarray:
    bg = rng.normal(-4.6, 1.0, size=(B, C, H, W))
    heat = 1.0 / (1.0 + np.exp(-bg))
    yy = np.arange(H, dtype=np.float64)[:, None]
    xx = np.arange(W, dtype=np.float64)[None, :]
    for b in range(B):
        cls = rng.integers(0, C, size=blobs)
        cy = rng.uniform(0, H, size=blobs)   # sub-pixel centres: symmetric neighbours would tie exactly
        cx = rng.uniform(0, W, size=blobs)
        sig = rng.uniform(1.0, 4.0, size=blobs)
        peak = rng.uniform(0.1, 0.95, size=blobs)
        for j in range(blobs):
            r = int(np.ceil(4 * sig[j]))
            y0, y1 = max(0, int(cy[j]) - r), min(H, int(cy[j]) + r + 1)
            x0, x1 = max(0, int(cx[j]) - r), min(W, int(cx[j]) + r + 1)
            g = peak[j] * np.exp(-((yy[y0:y1] - cy[j]) ** 2 + (xx[:, x0:x1] - cx[j]) ** 2) / (2 * sig[j] ** 2))
            np.maximum(heat[b, cls[j], y0:y1, x0:x1], g, out=heat[b, cls[j], y0:y1, x0:x1])
    return np.clip(heat, 1e-4, 1 - 1e-4).astype(np.float32)


def make_heads(rng: np.random.Generator, B, H, W, scale_wh=1.0) -> dict:
    u = lambda lo, hi, ch: rng.uniform(lo, hi, size=(B, ch, H, W)).astype(np.float32)  # noqa: E731
    n = lambda sd, ch: rng.normal(0.0, sd, size=(B, ch, H, W)).astype(np.float32)  # noqa: E731
    out = {
        "widthHeight": u(-2.0, 40.0 * scale_wh, 2),
        "depth": u(1.0, 70.0, 1),
        "dimension": np.concatenate([u(0.5, 4.0, 1), u(0.5, 3.0, 1), u(0.5, 12.0, 1)], axis=1),
        "rotation": n(1.0, 8),
        "reg": u(0.0, 1.0, 2),
        "amodal_offset": n(2.0, 2),
        "velocity": n(5.0, 3),
        "nuscenes_att": n(1.0, 8),
        "depth2": u(1.0, 70.0, 1),
        "rotation2": n(1.0, 8),
    }
    return out


def make_calib(rng: np.random.Generator, B) -> np.ndarray:
    calib = np.repeat(NUSC_CALIB[None], B, axis=0).copy()
    jit = 1.0 + rng.uniform(-0.01, 0.01, size=(B, 3))
    calib[:, 0, 0] *= jit[:, 0]
    calib[:, 1, 1] *= jit[:, 0]
    calib[:, 0, 2] *= jit[:, 1]
    calib[:, 1, 2] *= jit[:, 2]
    return calib


def make_radar(rng: np.random.Generator, B, n) -> tuple:
    """(B, n, 5) float64 [x, y, z, vx, vz], each frame stably sorted ascending by z; counts (B,)."""
    z = rng.uniform(1.0, 80.0, size=(B, n))
    x = z * np.tan(rng.uniform(-0.6, 0.6, size=(B, n)))
    y = rng.uniform(-1.0, 2.0, size=(B, n))
    vx = rng.normal(0.0, 5.0, size=(B, n))
    vz = rng.normal(0.0, 5.0, size=(B, n))
    pts = np.stack([x, y, z, vx, vz], axis=-1)
    order = np.argsort(pts[:, :, 2], axis=1, kind="stable")
    pts = np.take_along_axis(pts, order[:, :, None], axis=1)
    return np.ascontiguousarray(pts), np.full((B,), n, dtype=np.int32)


def make_inputs(wl: Workload, seed: int = 1234, B: int | None = None) -> dict:
    """All stage inputs for `wl` (optionally with a different batch size), as NumPy arrays."""
    B = wl.B if B is None else B
    rng = np.random.default_rng(seed)
    scale_wh = wl.W / 200.0
    out = {"heatmap": make_heatmap(rng, B, wl.C, wl.H, wl.W, wl.blobs)}
    out.update(make_heads(rng, B, wl.H, wl.W, scale_wh))
    calib = make_calib(rng, B)
    out["calib"] = calib.astype(np.float32)
    out["calib64"] = calib
    out["cam"] = np.ascontiguousarray(calib[:, :, :3])
    pts, counts = make_radar(rng, B, wl.n_radar)
    out["radar_points"] = pts
    out["radar_counts"] = counts
    out["A_out"] = np.repeat(output_affine(wl.H, wl.W, wl.img_w, wl.img_h)[None], B, axis=0)
    return out
