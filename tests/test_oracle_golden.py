"""Pin the CPU oracle (oracle/cfd3d_oracle.c) against golden vectors produced by the reference itself
(tests/golden/make_golden.py).  CPU-only; runs everywhere."""
import numpy as np

from helpers import load_golden

RTOL_LIBM = 2e-6  # reference trig is 1-ulp SLEEF fp32, oracle is fp64-rounded: a few ulp on derived values


def test_nms_bit_exact(oracle):
    g = load_golden("topk_nms")
    out = oracle.nms(g["in_heat"])
    assert np.array_equal(out.view(np.uint32), g["out_nms"].view(np.uint32))


def test_topk_raw_and_nms_bit_exact(oracle):
    g = load_golden("topk_nms")
    K = int(g["in_K"])
    for tag, apply_nms in (("raw", False), ("nms", True)):
        s, i, c, y, x = oracle.topk(g["in_heat"], K, apply_nms=apply_nms)
        assert np.array_equal(s.view(np.uint32), g[f"out_{tag}_scores"].view(np.uint32))
        assert np.array_equal(i, g[f"out_{tag}_inds"])
        assert c.dtype == g[f"out_{tag}_classes"].dtype == np.int32
        assert np.array_equal(c, g[f"out_{tag}_classes"])
        assert np.array_equal(y, g[f"out_{tag}_ys"])
        assert np.array_equal(x, g[f"out_{tag}_xs"])


def test_gather_bit_exact(oracle):
    g = load_golden("topk_nms")
    out = oracle.gather_nchw(g["in_feat"], g["out_raw_inds"])
    assert np.array_equal(out, g["out_gather"])


def test_alpha_and_dist_thresh(oracle):
    g = load_golden("alpha_thr")
    alpha = oracle.get_alpha(g["in_rot"])
    np.testing.assert_allclose(alpha, g["out_alpha"], rtol=RTOL_LIBM, atol=1e-6)
    # feed the reference's alpha so only yaw/corners/threshold arithmetic is compared
    thr = oracle.dist_thresh(g["in_calib"], g["in_center"], g["in_dim"], g["out_alpha"])
    np.testing.assert_allclose(thr, g["out_thr"], rtol=RTOL_LIBM, atol=1e-6)


def test_pc_dep_to_hm_quirks_bit_exact(oracle):
    g = load_golden("pc2hm_quirks")
    pc_dep = g["in_pc_dep"]
    n = len(g["in_depth"])
    matched = 0
    for j in range(n):
        hm = np.zeros_like(pc_dep)
        a = oracle.pc_dep_to_hm(hm, pc_dep, g["in_depth"][j], g["in_bbox"][j], g["in_thr"][j], float(g["in_max_pc_dist"]))
        assert np.array_equal(hm.view(np.uint32), g["out_pc_hm"][j].view(np.uint32)), f"case {j}"
        if g["out_assoc"][j] >= 0:
            assert a == g["out_assoc"][j], f"case {j}"
            matched += 1
        elif a >= 0:  # reference wrote nothing: only legal when the paint window is empty
            assert not hm.any()
    assert matched >= 5


def test_pillars_bit_exact(oracle):
    g = load_golden("pillars")
    for tag in "abc":
        H, W = (int(v) for v in g[f"in_hw_{tag}"])
        calib = g[f"in_calib_{tag}"]
        pc_dep, owner = oracle.radar_pillars(g[f"in_points_{tag}"], g[f"in_counts_{tag}"], calib[:, :, :3], calib,
                                             g[f"in_A_{tag}"], (1600, 900), (H, W))
        ref = g[f"out_pc_dep_{tag}"]
        assert np.array_equal(pc_dep.view(np.uint32), ref.view(np.uint32)), tag
        assert ((owner >= 0) == (ref[:, 0] != 0)).all()
        assert (ref[:, 0] != 0).mean() > 0.05


def radar_ingest_cases(g):
    """(tag, raw (1,18,N), counts, calib (1,3,4), A (1,2,3), (H,W), flip (1,), descending, z_offset) per fixture case."""
    for tag in (str(t) for t in g["in_cases"]):
        raw = g["in_raw_" + tag][None]
        yield (tag, raw, np.array([raw.shape[2]], np.int32), g["in_calib_" + tag][None], g["in_A_" + tag][None],
               tuple(int(v) for v in g["in_hw_" + tag]), np.array([bool(g["in_flip_" + tag])]),
               not bool(g["in_reverse_" + tag]), float(g["in_zoff_" + tag]))


def check_sorted_points_against_reference(g, tag, pts, zoff):
    """The reference's padded pc_3d (18, MAX_PC) holds the points it kept, in its order: they must appear in the
    ingested (sorted, flipped) points in the same order, values bit-equal (Y after the z offset)."""
    n_kept = min(int(g["out_pc_N_" + tag]), g["out_pc_3d_" + tag].shape[1])
    ref = g["out_pc_3d_" + tag][[0, 1, 2, 8, 9], :n_kept].T
    mine = pts.copy()
    mine[:, 1] -= zoff
    pos = 0
    for row in ref:
        while pos < len(mine) and not np.array_equal(mine[pos], row):
            pos += 1
        assert pos < len(mine), "%s: a kept point is missing or out of order" % tag
        pos += 1


def test_radar_ingest_and_flip(oracle):
    """Row f2, complete: stored (18, N) array -> pc_dep, against the reference's own loadRadarPointCloud
    (nuscenes.py:131-218; fixture gen_radar_ingest): layout, sort order incl. PC_REVERSE, flip, z offset -- bit-exact."""
    g = load_golden("radar_ingest")
    flips = 0
    for tag, raw, cnt, calib, A, (H, W), flip, desc, zoff in radar_ingest_cases(g):
        pts, order = oracle.radar_ingest(raw, cnt, flip, desc)
        z = pts[0, :, 2]
        assert (np.diff(z) <= 0).all() if desc else (np.diff(z) >= 0).all()
        assert np.array_equal(np.sort(order[0]), np.arange(raw.shape[2]))
        check_sorted_points_against_reference(g, tag, pts[0], zoff)
        pc_dep, _ = oracle.radar_pillars(pts, cnt, calib[:, :, :3], calib, A, (1600, 900), (H, W), z_offset=zoff,
                                         flip=flip)
        ref = g["out_pc_dep_" + tag]
        assert np.array_equal(pc_dep[0].view(np.uint32), ref.view(np.uint32)), tag
        assert (ref[0] != 0).mean() > 0.05
        flips += int(flip[0])
    assert flips >= 2


def test_frustum_pc_hm(oracle):
    g = load_golden("frustum_decode")
    K = int(g["in_K"])
    pc_hm, assoc, thr, _ = oracle.frustum_pc_hm(g["in_heatmap"], g["in_widthHeight"], g["in_depth"], g["in_dimension"],
                                                g["in_rotation"], g["in_pc_dep"], g["in_calib"], K,
                                                float(g["in_max_pc_dist"]))
    # end to end thr passes through alpha -> yaw -> sin/cos times up to l/2 = 6 m: 1-ulp libm differences in the
    # reference (SLEEF fp32) amplify to a few 1e-6 absolute
    np.testing.assert_allclose(thr, g["out_thr"], rtol=2e-5, atol=5e-6)
    known = g["out_assoc"] != -2
    assert np.array_equal(assoc[known], g["out_assoc"][known])  # association indices bit-exact
    assert (g["out_assoc"] >= 0).sum() >= 10
    # every painted value is a copy of a pc_dep value (or d/60): bit-exact
    assert np.array_equal(pc_hm.view(np.uint32), g["out_pc_hm"].view(np.uint32))


def _decode_inputs(g, vname):
    keys = str(g[f"in_dec_{vname}_keys"]).split(",")
    return {k: g["in_" + k] for k in keys}, bool(g[f"in_dec_{vname}_norm2d"])


def test_fusion_decode_variants(oracle):
    g = load_golden("frustum_decode")
    K = int(g["in_K"])
    H, W = g["in_heatmap"].shape[2:]
    for vname in ("full", "noreg_norm2d", "unc"):
        out, norm2d = _decode_inputs(g, vname)
        ret = oracle.fusion_decode(out, (H, W), K, norm2d)
        ref_keys = sorted(k[len(f"out_dec_{vname}_"):] for k in g if k.startswith(f"out_dec_{vname}_"))
        assert sorted(k for k in ret if not k.startswith("_")) == ref_keys, vname
        for k in ref_keys:
            refv = g[f"out_dec_{vname}_{k}"]
            assert ret[k].shape == refv.shape and ret[k].dtype == refv.dtype, (vname, k)
            if k == "scores" and vname == "unc":
                np.testing.assert_allclose(ret[k], refv, rtol=1e-6)  # expf differs by an ulp
            else:
                assert np.array_equal(ret[k].view(np.uint32), refv.view(np.uint32)), (vname, k)


POST_VARIANTS = ("amodal", "bboxcenter", "gt", "no3d")


def post_case(g, v):
    y = {k[len("in_%s_" % v):]: g[k] for k in g if k.startswith("in_%s_" % v) and not k.endswith("_isgt")}
    ref = {k[len("out_%s_" % v):]: g[k] for k in g if k.startswith("out_%s_" % v)}
    return y, bool(g["in_%s_isgt" % v]), ref


def test_post_process_variants(oracle):
    """f1: postProcess (utils/postProcess.py:13-85) against outputs of the reference itself.  Tolerance, not bits: the
    reference leaves the summation order of its 2x3 / 3x3 products to the BLAS and uses SLEEF fp32 trigonometry;
    1e-5 relative is the bar north_star states for decoded boxes (observed: 1.2e-6)."""
    g = load_golden("post_process")
    H, W = int(g["in_H"]), int(g["in_W"])
    for v in POST_VARIANTS:
        y, is_gt, ref = post_case(g, v)
        out = oracle.post_process(y, g["in_trans"], (H, W), g["in_calib"], is_gt)
        assert set(out) == set(ref), (v, set(out) ^ set(ref))
        for k, r in ref.items():
            assert out[k].shape == r.shape and out[k].dtype == r.dtype, (v, k)
            np.testing.assert_allclose(out[k], r, rtol=1e-5, atol=1e-5, err_msg="%s %s" % (v, k))
    # a detection with a non-positive dimension has its 3-D box zeroed (postProcess.py:81)
    _, _, ref = post_case(g, "amodal")
    assert not ref["bboxes3d"][0, 3].any() and not ref["bboxes3d"][1, 5].any()


CFG1_HEADS = ("heatmap", "reg", "widthHeight", "depth", "depth2", "rotation", "rotation2", "dimension",
              "amodal_offset", "nuscenes_att", "velocity")


def test_cfg1_whole_model_fixture(oracle):
    """BASELINE.json configs[0]: head maps produced by the reference's own random-init DLA-34 forward (Recipe B,
    tests/golden/make_golden.py gen_cfg1_model) -- the oracle reproduces the reference's pc_hm (also what
    CenterFusionHead.forward itself emitted), association indices, fusionDecode and postProcess outputs on them."""
    g = load_golden("cfg1_model")
    K = int(g["in_K"])
    H, W = g["in_heatmap"].shape[2:]
    pc_hm, assoc, thr, _ = oracle.frustum_pc_hm(g["in_heatmap"], g["in_widthHeight"], g["in_depth"], g["in_dimension"],
                                                g["in_rotation"], g["in_pc_dep"], g["in_calib"], K,
                                                float(g["in_max_pc_dist"]))
    np.testing.assert_allclose(thr, g["out_thr"], rtol=2e-5, atol=5e-6)
    known = g["out_assoc"] != -2
    assert np.array_equal(assoc[known], g["out_assoc"][known])
    assert (g["out_assoc"] >= 0).sum() >= 10 and (g["in_widthHeight"] < 0).any()
    assert np.array_equal(pc_hm.view(np.uint32), g["out_pc_hm"].view(np.uint32))
    ret = oracle.fusion_decode({k: g["in_" + k] for k in CFG1_HEADS}, (H, W), K)
    ref_keys = sorted(k[len("out_dec_"):] for k in g if k.startswith("out_dec_"))
    assert sorted(k for k in ret if not k.startswith("_")) == ref_keys
    for k in ref_keys:
        assert np.array_equal(ret[k].view(np.uint32), g["out_dec_" + k].view(np.uint32)), k
    post = oracle.post_process({k: g["out_dec_" + k].copy() for k in ref_keys}, g["in_trans"], (H, W), g["in_calib"], False)
    for k in (k[len("out_post_"):] for k in g if k.startswith("out_post_")):
        np.testing.assert_allclose(post[k], g["out_post_" + k], rtol=1e-5, atol=1e-5, err_msg=k)


def gt_assoc_frames():
    g = load_golden("gt_assoc")
    for f in range(int(g["in_frames"])):
        f32 = bool(g["in_bbox_f32_%d" % f])
        bb = g["in_bbox_%d" % f].astype(np.float32) if f32 else g["in_bbox_%d" % f]
        thr = g["in_thr_%d" % f].astype(np.float32) if f32 else g["in_thr_%d" % f]
        yield f, f32, g["in_pc_dep_%d" % f], bb, g["in_depth_%d" % f], thr, g["out_pc_hm_%d" % f], g["out_assoc_%d" % f], \
            float(g["in_max_pc_dist"])


def test_gt_association_numpy_branch(oracle):
    """f4: the DataLoader's NumPy branch of cvtPcDepthToHeatmap (generic_dataset.py:673-687) -- fixture generated by
    the reference with float32 boxes / thresholds (its own arrangement) and with float64 ones (generic caller)."""
    total = 0
    for f, f32, pc_dep, bb, depth, thr, hm_ref, assoc_ref, mx in gt_assoc_frames():
        hm = np.zeros_like(pc_dep)
        assoc = np.array([oracle.pc_dep_to_hm_np(hm, pc_dep, float(depth[i]), bb[i], thr[i], mx) for i in range(len(bb))])
        known = assoc_ref != -2
        assert np.array_equal(assoc[known], assoc_ref[known]), f
        assert (assoc[~known] == -1).sum() + (assoc[~known] >= 0).sum() == (~known).sum()
        assert np.array_equal(hm.view(np.uint32), hm_ref.view(np.uint32)), f
        total += int((assoc_ref >= 0).sum())
        if f32:  # the all-float32 arrangement is the torch branch's arithmetic: the fp32 restatement agrees
            hm2 = np.zeros_like(pc_dep)
            for i in range(len(bb)):
                oracle.pc_dep_to_hm(hm2, pc_dep, np.float32(depth[i]), bb[i], thr[i], mx)
            assert np.array_equal(hm2.view(np.uint32), hm_ref.view(np.uint32))
    assert total >= 80
