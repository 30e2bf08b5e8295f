"""Generate the committed golden vectors by running the UNMODIFIED reference on CPU.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
Writes tests/golden/*.npz.  The reference has no tests or fixtures of its own for this path
(SURVEY.md section 4), so these outputs of the reference itself are what pins the oracle
(oracle/cfd3d_oracle.c) and, through it and directly, the CUDA kernels.

Every array stored under a key starting with "in_" is an input, "out_" an output of reference code.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import refimport  # noqa: E402
from centerfusiondetect3d_b200 import synthetic as syn  # noqa: E402
from helpers import ensure_margins, ref_assoc_indices, ref_pc_dep_from_points, tie_free  # noqa: E402

torch.set_num_threads(1)
ns = refimport.load_reference()


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def save(name, **arrs):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **arrs)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


def gen_topk_nms():
    wl = syn.Workload("g_topk", 2, 24, 40, 20, 0, C=4, blobs=12)
    for seed in range(100, 200):
        rng = np.random.default_rng(seed)
        heat = syn.make_heatmap(rng, wl.B, wl.C, wl.H, wl.W, wl.blobs)
        if tie_free(heat, wl.K + 8):
            break
    nmsd = ns.nms(t(heat))
    raw = ns.topk(t(heat), K=wl.K)
    sup = ns.topk(nmsd, K=wl.K)
    feat = rng.normal(size=(2, 5, 24, 40)).astype(np.float32)
    g = ns.transposeAndGetFeature(t(feat), raw[1])
    save("topk_nms", in_heat=heat, in_K=np.int64(wl.K), out_nms=nmsd.numpy(),
         out_raw_scores=raw[0].numpy(), out_raw_inds=raw[1].numpy(), out_raw_classes=raw[2].numpy(),
         out_raw_ys=raw[3].numpy(), out_raw_xs=raw[4].numpy(),
         out_nms_scores=sup[0].numpy(), out_nms_inds=sup[1].numpy(), out_nms_classes=sup[2].numpy(),
         out_nms_ys=sup[3].numpy(), out_nms_xs=sup[4].numpy(),
         in_feat=feat, out_gather=g.numpy())


def gen_alpha_thr():
    rng = np.random.default_rng(7)
    B, K = 3, 64
    rot = rng.normal(size=(B, K, 8)).astype(np.float32)
    rot[0, 0, 1] = rot[0, 0, 5]  # tie in the bin classifier -> second branch
    dim = np.stack([rng.uniform(0.5, 4, (B, K)), rng.uniform(0.5, 3, (B, K)), rng.uniform(0.5, 12, (B, K))], -1).astype(np.float32)
    center = np.stack([rng.uniform(0, 200, (B, K)), rng.uniform(0, 112, (B, K))], -1).astype(np.float32)
    calib = syn.make_calib(rng, B).astype(np.float32)
    alpha = ns.get_alpha(t(rot))
    thr = ns.getDistanceThresh(t(calib), t(center), t(dim), alpha)
    save("alpha_thr", in_rot=rot, in_dim=dim, in_center=center, in_calib=calib, out_alpha=alpha.numpy(), out_thr=thr.numpy())


def gen_pillars():
    cfg_a = refimport.make_config(K=100, out_hw=(112, 200))
    cfg_b = refimport.make_config(K=100, out_hw=(225, 400))
    arrs = {}
    for tag, cfg, n, seed in (("a", cfg_a, 100, 11), ("b", cfg_b, 300, 12), ("c", cfg_a, 2000, 13)):
        H, W = cfg.MODEL.OUTPUT_SIZE
        rng = np.random.default_rng(seed)
        pts, counts = syn.make_radar(rng, 1, n)
        calib = syn.make_calib(rng, 1)
        A = syn.output_affine(H, W)
        pc_dep = ref_pc_dep_from_points(ns, pts[0], calib[0], A, (1600, 900), cfg)
        arrs.update({f"in_points_{tag}": pts, f"in_counts_{tag}": counts, f"in_calib_{tag}": calib,
                     f"in_A_{tag}": A[None], f"in_hw_{tag}": np.array([H, W]), f"out_pc_dep_{tag}": pc_dep[None]})
    save("pillars", **arrs)


def _small_pc_dep(H, W, n, seed, cfg):
    rng = np.random.default_rng(seed)
    pts, _ = syn.make_radar(rng, 1, n)
    calib = syn.make_calib(rng, 1)
    A = syn.output_affine(H, W)
    return ref_pc_dep_from_points(ns, pts[0], calib[0], A, (1600, 900), cfg)


def gen_pc2hm_quirks():
    H, W = 24, 40
    cfg = refimport.make_config(K=10, out_hw=(H, W))
    pc_dep = _small_pc_dep(H, W, 60, 21, cfg)
    # two pillars of EQUAL depth but different velocity in one ROI (argmin tie -> first in row-major order)
    pc_dep[:, 3:6, 30:32] = np.array([33.0, 1.0, 2.0], np.float32)[:, None, None]
    pc_dep[:, 4:7, 34:36] = np.array([33.0, -3.0, 4.0], np.float32)[:, None, None]
    cases = [
        # bbox x1,y1,x2,y2, depth, thr
        ([10.2, 5.1, 20.7, 15.9], 30.0, 40.0),    # ordinary
        ([-3.5, 4.0, 8.5, 12.0], 30.0, 40.0),     # x1 < 0: negative slice start wraps -> ROI empty or shifted
        ([-45.0, 4.0, 8.5, 12.0], 30.0, 40.0),    # x1 <= -W: start clamps to 0
        ([2.0, -2.5, 12.0, 6.0], 30.0, 40.0),     # y1 < 0
        ([30.0, 15.0, 47.0, 30.0], 30.0, 40.0),   # beyond right/bottom edge
        ([12.5, 8.5, 12.5, 8.5], 30.0, 40.0),     # w = h = 0 box
        ([28.5, 2.0, 37.5, 8.0], 33.0, 5.0),      # equal-depth tie between the two painted pillars
        ([10.2, 5.1, 20.7, 15.9], 3.0, 40.0),     # depth - thr < 0 -> lower bound 0
        ([10.2, 5.1, 20.7, 15.9], 200.0, 1.0),    # nothing within threshold
        ([0.2, 0.3, 39.9, 23.8], 30.0, 100.0),    # whole map
        ([1.0, 1.0, 3.0, 3.0], 30.0, 40.0),       # small box near the origin (paint w_min may be 1)
        ([0.5, 0.5, 1.9, 1.7], 30.0, 40.0),       # paint window starts at 0 after trunc
        ([36.0, 20.0, 39.5, 23.5], 30.0, 40.0),   # bottom-right corner
    ]
    # pillar depth exactly at depth +- thr (strict compares): take a depth present in the map
    dvals = np.unique(pc_dep[0][pc_dep[0] != 0])
    d0 = float(dvals[len(dvals) // 2])
    cases.append(([0.2, 0.3, 39.9, 23.8], d0 - 2.0, 2.0))   # d0 == depth + thr -> excluded
    cases.append(([0.2, 0.3, 39.9, 23.8], d0 + 2.0, 2.0))   # d0 == depth - thr -> excluded
    outs, assoc = [], []
    idxmap = pc_dep.copy()
    idxmap[1] = np.arange(H * W, dtype=np.float32).reshape(H, W)
    for bbox, depth, thr in cases:
        hm = torch.zeros(3, H, W)
        ns.cvtPcDepthToHeatmap(hm, t(pc_dep), torch.tensor([depth]), torch.tensor(bbox), torch.tensor(thr), 60.0)
        outs.append(hm.numpy())
        hm2 = torch.zeros(3, H, W)
        ns.cvtPcDepthToHeatmap(hm2, t(idxmap), torch.tensor([depth]), torch.tensor(bbox), torch.tensor(thr), 60.0)
        w = hm2[0] != 0
        assoc.append(int(hm2[1][w][0].item()) if w.any() else -1)
    save("pc2hm_quirks", in_pc_dep=pc_dep, in_bbox=np.array([c[0] for c in cases], np.float32),
         in_depth=np.array([c[1] for c in cases], np.float32), in_thr=np.array([c[2] for c in cases], np.float32),
         in_max_pc_dist=np.float32(60.0), out_pc_hm=np.stack(outs), out_assoc=np.array(assoc, np.int32))


def gen_frustum_and_decode():
    wl = syn.Workload("g_frustum", 2, 48, 80, 40, 300, C=4, blobs=25)
    # PILLAR_DIMS is a config knob (configs/*.yaml DATASET.PILLAR_DIMS); larger pillars give the small map real coverage
    cfg = refimport.make_config(K=wl.K, out_hw=(wl.H, wl.W), pillar_dims=(3.0, 0.8, 0.8))
    for seed in range(300, 400):
        inp = syn.make_inputs(wl, seed=seed)
        if tie_free(inp["heatmap"], wl.K + 8):
            break
    inp["widthHeight"] *= 0.5  # boxes sized for the small map
    pc_dep = np.stack([ref_pc_dep_from_points(ns, inp["radar_points"][b], inp["calib64"][b], inp["A_out"][b],
                                              (1600, 900), cfg) for b in range(wl.B)])
    nudged = ensure_margins(inp, pc_dep, wl.K, wl.max_pc_dist)
    print("frustum golden: nudged", nudged, "depth values for margin")
    output = {k: t(inp[k]) for k in ("heatmap", "widthHeight", "depth", "dimension", "rotation")}
    pc_hm = ns.getPcFrustumHeatmap(output, t(pc_dep), t(inp["calib"]), cfg)
    assoc, thr = ref_assoc_indices(ns, inp, pc_dep, wl.K, wl.max_pc_dist)
    print("frustum golden: matched", int((assoc >= 0).sum()), "of", assoc.size)
    arrs = {"in_" + k: inp[k] for k in ("heatmap", "widthHeight", "depth", "dimension", "rotation", "calib", "reg",
                                         "amodal_offset", "velocity", "nuscenes_att", "depth2", "rotation2")}
    arrs.update(in_pc_dep=pc_dep, in_K=np.int64(wl.K), in_max_pc_dist=np.float32(wl.max_pc_dist),
                out_pc_hm=pc_hm.numpy(), out_assoc=assoc, out_thr=thr)
    # decode variants on the same maps
    unc = np.random.default_rng(5).normal(-1.0, 0.5, size=(wl.B, 1, wl.H, wl.W)).astype(np.float32)
    arrs["in_uncertainty"] = unc
    variants = {
        "full": dict(keys=("heatmap", "reg", "widthHeight", "depth", "depth2", "rotation", "rotation2", "dimension",
                           "amodal_offset", "nuscenes_att", "velocity"), norm2d=False),
        "noreg_norm2d": dict(keys=("heatmap", "widthHeight", "depth", "rotation", "dimension", "amodal_offset"), norm2d=True),
        "unc": dict(keys=("heatmap", "reg", "widthHeight", "depth2", "rotation2", "dimension", "velocity", "uncertainty"), norm2d=False),
    }
    for vname, v in variants.items():
        out = {k: t(unc if k == "uncertainty" else inp[k]).clone() for k in v["keys"]}
        ret = ns.fusionDecode([out], outputSize=(wl.H, wl.W), K=wl.K, norm2d=v["norm2d"])
        for k, val in ret.items():
            arrs[f"out_dec_{vname}_{k}"] = val.numpy()
        arrs[f"in_dec_{vname}_keys"] = np.array(",".join(v["keys"]))
        arrs[f"in_dec_{vname}_norm2d"] = np.bool_(v["norm2d"])
    save("frustum_decode", **arrs)


def gen_post_process():
    """postProcess (utils/postProcess.py:13-85) on the decode outputs stored in frustum_decode.npz: every branch of
    the centre choice (amodal offset / 2-D box centre / isGt) and a detection with a non-positive dimension."""
    g = np.load(os.path.join(HERE, "frustum_decode.npz"))
    H, W = g["in_heatmap"].shape[2:]
    center, scale = np.array([800.0, 450.0], dtype=np.float32), 1600.0
    calib = g["in_calib"]
    arrs = dict(in_center=center, in_scale=np.float32(scale), in_H=np.int64(H), in_W=np.int64(W), in_calib=calib,
                in_trans=ns.getAffineTransform(center, scale, 0, (W, H), inverse=True).astype(np.float32))
    variants = {"amodal": ("full", (), False), "bboxcenter": ("full", ("amodal_offset",), False),
                "gt": ("full", (), True), "no3d": ("full", ("rotation", "velocity"), False)}
    for vname, (dec, drop, is_gt) in variants.items():
        y = {k[len("out_dec_%s_" % dec):]: g[k].copy() for k in g.files if k.startswith("out_dec_%s_" % dec)}
        for k in drop:
            y.pop(k)
        if "dimension" in y:
            y["dimension"][0, 3, 1] = -0.25  # any(dim <= 0) -> the 3-D box is zeroed
            y["dimension"][1, 5, 0] = 0.0
        for k, v in y.items():
            arrs["in_%s_%s" % (vname, k)] = v.copy()
        arrs["in_%s_isgt" % vname] = np.bool_(is_gt)
        out = ns.postProcess({k: t(v).clone() for k, v in y.items()}, center, scale, H, W, t(calib), isGt=is_gt)
        for k, v in out.items():
            arrs["out_%s_%s" % (vname, k)] = v.numpy()
    save("post_process", **arrs)


def gen_cfg1_model():
    """BASELINE.json configs[0] / SURVEY.md 8(c) Recipe B: the reference's OWN network (random-init DLA-34 `DLASeg`,
    seed 0, one 448x800 frame, 100 radar points through the reference pillar code) run end to end on CPU.  Stored: the
    first- and second-stage head maps the model produced, pc_dep, calib, and the reference's pc_hm (from
    CenterFusionHead.forward and from getPcFrustumHeatmap on those maps), association indices, fusionDecode and
    postProcess outputs.  The -m gpu test replays the stored head maps through the ops.

    The reference initialises its head convolutions with std 0.001 (model/utils.py initConv2dWeights), which leaves
    every untrained head within 1e-5 of its bias: no boxes, no associations.  The fixture therefore rescales the LAST
    convolution of every head (weight *= target_std / measured_std, bias = target mean) so that the heads span the
    ranges of a trained model; the maps are still outputs of the reference's network code (spatially correlated conv
    features, post-sigmoid heat-map, depth = 1/(sigmoid+1e-6)-1), which is what distinguishes them from synthetic.py."""
    import refimport as R

    K, H, W = 100, 112, 200
    cfg = R.make_model_config(K=K)
    model = R.build_model(cfg, seed=0)
    # An untrained DLA-34 in eval mode (BatchNorm running statistics 0 / 1) maps any image to spatially constant
    # features (relative variation 1e-4: a few hundred distinct fp32 levels, i.e. a heat-map full of exact ties).  The
    # backbone's BatchNorm layers therefore normalise with the statistics of the frame itself (their train-mode
    # arithmetic), which is what gives an untrained network feature maps with structure.  The model and its heads stay
    # in eval mode: CenterFusionHead.forward takes its frustum branch (detectHeads.py:176).
    for m_ in model.modules():
        if isinstance(m_, torch.nn.BatchNorm2d):
            m_.train()
    assert not model.training and not model.detectHead_0.training
    head = model.detectHead_0
    targets = {  # head -> (mean per channel or scalar, std) of the pre-activation output
        "heatmap": (-4.6, 0.6), "reg": (0.5, 0.2), "widthHeight": (12.0, 9.0), "depth": (-3.0, 0.45),
        "depth2": (-3.0, 0.45), "rotation": (0.0, 1.0), "rotation2": (0.0, 1.0), "dimension": ([1.6, 1.8, 4.5], 0.3),
        "amodal_offset": (0.0, 2.0), "nuscenes_att": (0.0, 1.0), "velocity": (0.0, 3.0)}
    x = None  # drawn below: the first image seed whose heat-map is tie-free in the top-K region
    rng = np.random.default_rng(2024)
    pts, _ = syn.make_radar(rng, 1, 100)
    calib64 = syn.make_calib(rng, 1)
    A = syn.output_affine(H, W)
    pc_dep = ref_pc_dep_from_points(ns, pts[0], calib64[0], A, (1600, 900), R.make_config(K=K, out_hw=(H, W)))[None]
    calib = calib64.astype(np.float32)

    def last_conv(seq):
        return [m for m in seq.modules() if isinstance(m, torch.nn.Conv2d)][-1]

    def forward():
        with torch.no_grad():
            return model(x, pc_dep=t(pc_dep).clone(), calib=t(calib))[0]

    def calibrate():
        pre = {}
        hooks = [last_conv(getattr(head, h)).register_forward_hook(lambda m, i, o, h=h: pre.__setitem__(h, o.detach().clone()))
                 for h in cfg.heads]
        forward()
        for hk in hooks:
            hk.remove()
        for h, (mean, std) in targets.items():
            conv = last_conv(getattr(head, h))
            ch_mean = pre[h].mean(dim=(0, 2, 3))
            f = std / max(float((pre[h] - ch_mean.view(1, -1, 1, 1)).std()), 1e-12)
            with torch.no_grad():  # new output = f * (old - old bias) + new bias: per-channel mean = target
                want = torch.as_tensor(mean, dtype=torch.float32).expand_as(conv.bias)
                new_bias = want - f * (ch_mean - conv.bias)
                conv.weight.mul_(f)
                conv.bias.copy_(new_bias)

    for x_seed in range(20):  # the first image seed whose heat-map is tie-free in the top-K region
        x = torch.randn(1, 3, 448, 800, generator=torch.Generator().manual_seed(x_seed))
        calibrate()
        calibrate()  # second round: the secondary heads see the pc_hm of calibrated first-stage heads
        y = forward()
        heads = {k: y[k].numpy().copy() for k in cfg.heads}
        if tie_free(heads["heatmap"], K + 32):
            print("cfg1 model: image seed", x_seed)
            break
    else:
        raise RuntimeError("no tie-free heat-map in 20 image seeds")
    probe = {k: heads[k].copy() for k in ("heatmap", "widthHeight", "depth", "dimension", "rotation")}
    probe["calib"] = calib
    assert ensure_margins(probe, pc_dep, K, 60.0) == 0, "a borderline association: change the seed"
    fcfg = R.make_config(K=K, out_hw=(H, W))
    out5 = {k: t(heads[k]) for k in ("heatmap", "widthHeight", "depth", "dimension", "rotation")}
    pc_hm = ns.getPcFrustumHeatmap(out5, t(pc_dep), t(calib), fcfg).numpy()
    assert np.array_equal(pc_hm[:, 0:1], y["pc_hm"].numpy()), "forward's pc_hm != getPcFrustumHeatmap on its heads"
    assoc, thr = ref_assoc_indices(ns, {**heads, "calib": calib}, pc_dep, K, 60.0)
    print("cfg1 model: matched", int((assoc >= 0).sum()), "of", assoc.size, "| heat max %.3f" % heads["heatmap"].max(),
          "| negative widthHeight at %.1f %% of the pixels" % (100.0 * (heads["widthHeight"] < 0).mean()))
    dec_in = {k: t(v).clone() for k, v in heads.items()}
    dec = ns.fusionDecode([dec_in], outputSize=(H, W), K=K)
    assert "rotation2" not in dec_in
    center, scale = np.array([800.0, 450.0], dtype=np.float32), 1600.0
    post = ns.postProcess({k: v.clone() for k, v in dec.items()}, center, scale, H, W, t(calib))
    arrs = {"in_" + k: v for k, v in heads.items()}
    arrs.update(in_pc_dep=pc_dep, in_calib=calib, in_K=np.int64(K), in_max_pc_dist=np.float32(60.0),
                in_center=center, in_scale=np.float32(scale),
                in_trans=ns.getAffineTransform(center, scale, 0, (W, H), inverse=True).astype(np.float32),
                out_pc_hm=pc_hm, out_assoc=assoc, out_thr=thr)
    for k, v in dec.items():
        arrs["out_dec_" + k] = v.numpy()
    for k, v in post.items():
        arrs["out_post_" + k] = v.numpy()
    save("cfg1_model", **arrs)


def gen_gt_assoc():
    """Row f4 on the reference's REAL branch: the DataLoader's per-annotation loop (dataset/generic_dataset.py:673-687)
    calls getDistanceThresh and cvtPcDepthToHeatmap with NumPy arguments (the `lib = np` branch of
    utils/pointcloud.py:411-481): pc_dep / pc_hm float32 maps, depth a Python float, distanceThreshold np.float32 (what
    the NumPy branch of getDistanceThresh returns), bbox float32 (transformBbox, generic_dataset.py:507-526).  Two frames x 128 annotations in that arrangement (boxes
    clipped to the map like transformBbox does), plus one frame of float64 boxes crossing every border with float64 thresholds (what any
    other NumPy caller of the function may pass: float64 compares and paint-window arithmetic under NumPy >= 2).  Stored: inputs, the reference's thresholds, pc_hm after the whole loop and
    the matched pixel of every annotation (index trick of SURVEY.md 8c)."""
    H, W, N = 112, 200, 128
    cfg = refimport.make_config(K=100, out_hw=(H, W))
    rng = np.random.default_rng(77)
    frames = []
    for f in range(3):
        pts, _ = syn.make_radar(rng, 1, 300)
        calib64 = syn.make_calib(rng, 1)
        pc_dep = ref_pc_dep_from_points(ns, pts[0], calib64[0], syn.output_affine(H, W), (1600, 900), cfg)
        calib = calib64[0].astype(np.float32)
        cx, cy = rng.uniform(-8, W + 8, N), rng.uniform(-8, H + 8, N)
        w, h = rng.uniform(0.5, 70, N), rng.uniform(0.5, 45, N)
        bb = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], -1)
        f32 = f < 2
        if f32:  # transformBbox: float32, clipped into the map
            bb = bb.astype(np.float32)
            bb[:, [0, 2]] = np.clip(bb[:, [0, 2]], 0, W - 1)
            bb[:, [1, 3]] = np.clip(bb[:, [1, 3]], 0, H - 1)
        dvals = np.unique(pc_dep[0][pc_dep[0] != 0]).astype(np.float64)
        depth = dvals[rng.integers(0, len(dvals), N)] + rng.normal(0, 2.5, N)  # near real pillar depths: many matches
        depth = np.maximum(depth, 0.5)
        dim = np.stack([rng.uniform(1.2, 3.5, N), rng.uniform(0.5, 2.6, N), rng.uniform(0.6, 11, N)], -1)
        alpha = rng.uniform(-np.pi, np.pi, N)
        thr = np.zeros(N, np.float64)
        assoc = np.full(N, -1, np.int32)
        pc_hm = np.zeros((3, H, W), np.float32)
        idxmap = pc_dep.copy()
        idxmap[1] = np.arange(H * W, dtype=np.float32).reshape(H, W)
        for i in range(N):
            bbox = bb[i]
            center = np.array([(bbox[0] + bbox[2]) / 2, (bbox[1] + bbox[3]) / 2], dtype=np.float32)  # :545-547
            t_ = ns.getDistanceThresh(calib.reshape(1, 3, 4), center.reshape(1, 1, 2), np.array(list(dim[i])).reshape(1, 1, 3),
                                      np.array(float(alpha[i])).reshape(1, 1, 1))[0, 0]
            t_ = t_.reshape(-1)[0]
            assert isinstance(t_, np.float32)  # the NumPy branch of get3DCorners works in float32 arrays
            if not f32:
                t_ = np.float64(t_)  # the generic caller: a float64 threshold (float64 compares)
            thr[i] = t_
            # keep the strict compares clear of every pillar depth (thr is an INPUT of the op here, so only the
            # decimal -> binary of the stored values matters; a margin keeps the fixture robust all the same)
            for bound in (float(depth[i]) + float(t_), max(0.0, float(depth[i]) - float(t_))):
                if np.abs(dvals - bound).min() < 1e-4:
                    depth[i] += 3e-3
            ns.cvtPcDepthToHeatmap(pc_hm, pc_dep, float(depth[i]), bbox, t_, 60.0)
            probe = np.zeros((3, H, W), np.float32)
            ns.cvtPcDepthToHeatmap(probe, idxmap, float(depth[i]), bbox, t_, 60.0)
            wr = probe[0] != 0
            if wr.any():
                assoc[i] = int(probe[1][wr][0])
            else:
                assoc[i] = -2  # no match, or a match whose paint window is empty: undecidable from the outside
        print("gt_assoc frame", f, "float32 boxes" if f32 else "float64 boxes", "matched", int((assoc >= 0).sum()), "of", N)
        frames.append(dict(pc_dep=pc_dep, bbox=bb.astype(np.float64), bbox_f32=np.bool_(f32), depth=depth, thr=thr,  # (f32 frames: thr holds float32 values)
                           pc_hm=pc_hm, assoc=assoc))
    arrs = {}
    for f, fr in enumerate(frames):
        arrs.update({"in_pc_dep_%d" % f: fr["pc_dep"], "in_bbox_%d" % f: fr["bbox"], "in_bbox_f32_%d" % f: fr["bbox_f32"],
                     "in_depth_%d" % f: fr["depth"], "in_thr_%d" % f: fr["thr"], "out_pc_hm_%d" % f: fr["pc_hm"],
                     "out_assoc_%d" % f: fr["assoc"]})
    save("gt_assoc", in_max_pc_dist=np.float32(60.0), in_frames=np.int64(len(frames)), **arrs)


def gen_radar_ingest():
    """Row f2, complete: the reference's OWN loadRadarPointCloud (dataset/datasets/nuscenes.py:131-218) run on pickled
    (18, N) radar arrays written where it looks for them -- distance filter, z offset, projection + masks, depth sort
    (+ reversal when PC_REVERSE is false), flip, processPointCloud.  Outputs: pc_dep and the padded, sorted pc_3d."""
    import pickle
    import tempfile
    import types

    GD, NU = ns.generic_dataset.GenericDataset, ns.nuscenes.nuScenes
    cases = [  # tag, out_hw, N, flipped, PC_REVERSE, PC_Z_OFFSET, seed
        ("a", (112, 200), 150, False, False, 0, 31),
        ("b", (112, 200), 150, True, False, 0, 32),
        ("c", (112, 200), 300, True, True, 0.5, 33),
        ("d", (225, 400), 300, False, True, 0, 34),
    ]
    arrs = {"in_cases": np.array([c[0] for c in cases])}
    with tempfile.TemporaryDirectory() as tmp:
        for tag, (H, W), n, flipped, reverse, zoff, seed in cases:
            cfg = refimport.make_config(K=100, out_hw=(H, W))
            cfg.DATASET.PC_REVERSE = reverse
            cfg.DATASET.PC_Z_OFFSET = zoff
            cfg.DATASET.RADAR_PC = True
            rng = np.random.default_rng(seed)
            pts, _ = syn.make_radar(rng, 1, n)
            pts = pts[0][rng.permutation(n)]  # as stored: NOT sorted
            raw = rng.normal(size=(18, n))     # the other rows (rcs, raw velocities, ...) are never read on this path
            raw[0:3] = pts[:, 0:3].T
            raw[8], raw[9] = pts[:, 3], pts[:, 4]
            calib = syn.make_calib(rng, 1)[0]
            A = syn.output_affine(H, W)
            d = os.path.join(tmp, "annotations", "radar_pc", "CAM_FRONT")
            os.makedirs(d, exist_ok=True)
            with open(os.path.join(d, "tok_%s.bin" % tag), "wb") as f:
                pickle.dump(raw.tolist(), f)
            self = types.SimpleNamespace(config=cfg, img_dir=tmp, SENSOR_NAME={1: "CAM_FRONT"})
            for name in ("transformPointCloud", "getPcPillarsSize", "processPointCloud"):
                setattr(self, name, types.MethodType(getattr(GD, name), self))
            for name in ("getDepthMap", "drawPcHeat", "loadRadarPointCloud"):
                setattr(self, name, types.MethodType(getattr(NU, name), self))
            img_info = {"sensor_id": 1, "sample_token": "tok_%s" % tag, "camera_intrinsic": calib[:, :3].tolist(),
                        "width": 1600, "height": 900, "calib": calib.tolist()}
            img = np.zeros((900, 1600, 3), np.uint8)
            pc_z, pc_N, pc_dep, pc_3dz = self.loadRadarPointCloud(img, img_info, A, A, flipped)
            print("radar_ingest", tag, "points in", n, "kept", int(pc_N), "painted px", int((pc_dep[0] != 0).sum()))
            arrs.update({"in_raw_" + tag: raw, "in_calib_" + tag: calib, "in_A_" + tag: A, "in_hw_" + tag: np.array([H, W]),
                         "in_flip_" + tag: np.bool_(flipped), "in_reverse_" + tag: np.bool_(reverse),
                         "in_zoff_" + tag: np.float64(zoff), "out_pc_dep_" + tag: pc_dep,
                         "out_pc_N_" + tag: np.int64(pc_N), "out_pc_3d_" + tag: pc_3dz, "out_pc_2d_" + tag: pc_z})
    save("radar_ingest", **arrs)


if __name__ == "__main__":
    if len(sys.argv) > 1:  # python tests/golden/make_golden.py gen_cfg1_model
        for name in sys.argv[1:]:
            globals()[name]()
        sys.exit(0)
    gen_topk_nms()
    gen_alpha_thr()
    gen_pillars()
    gen_pc2hm_quirks()
    gen_frustum_and_decode()
    gen_post_process()
    gen_cfg1_model()
    gen_gt_assoc()
    gen_radar_ingest()
