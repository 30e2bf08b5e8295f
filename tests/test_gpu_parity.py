"""GPU parity tests proper: the sm_100a CUDA path, called through the C ABI (ctypes, ops.py), against
(1) the committed golden vectors produced by the unmodified reference and (2) the CPU oracle on seeded inputs.

Bars (BASELINE.json north_star): top-K indices, class ids and association indices BIT-EXACT; every pc_hm value
and every gathered/decoded value bit-exact too (they are copies or single fp32 operations); distance thresholds,
which pass through libm trigonometry, within RTOL_THR relative."""
import numpy as np
import pytest
import torch

from centerfusiondetect3d_b200 import synthetic as syn
from helpers import DEC_KEYS, FWD_KEYS, bits, load_golden, make_case, to_cuda

pytestmark = pytest.mark.gpu

RTOL_THR_REF = 2e-5   # vs the reference's SLEEF fp32 trig (see test_oracle_golden.py)
RTOL_THR_ORACLE = 1e-6  # vs the oracle (both evaluate trig in fp64 and round once)
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device")
    from centerfusiondetect3d_b200 import ops as _ops

    return _ops


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def npy(t):
    return t.detach().cpu().numpy()


def eq(a, b, what=""):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape and a.dtype == b.dtype, (what, a.shape, b.shape, a.dtype, b.dtype)
    ba, bb = bits(a), bits(b)
    if a.dtype.kind == "f":  # NaN == NaN whatever the sign/payload (x86 and the GPU produce different quiet NaNs)
        both_nan = np.isnan(a) & np.isnan(b)
        ba, bb = np.where(both_nan, 0, ba), np.where(both_nan, 0, bb)
    assert np.array_equal(ba, bb), "%s: %d of %d differ" % (what, (ba != bb).sum(), a.size)


class Cfg:
    def __init__(self, K, max_pc_dist=60.0):
        self.MODEL = type("M", (), {"K": K})()
        self.DATASET = type("D", (), {"MAX_PC_DIST": max_pc_dist})()


# ------------------------------------------------------------------------------------------------
# 1. golden vectors (outputs of the reference itself)
# ------------------------------------------------------------------------------------------------
def test_golden_nms_topk_gather(ops):
    g = load_golden("topk_nms")
    K = int(g["in_K"])
    heat = cu(g["in_heat"])
    eq(npy(ops.nms(heat)), g["out_nms"], "nms")
    for tag, apply_nms in (("raw", False), ("nms", True)):
        r = ops.topk(heat, K, apply_nms=apply_nms)
        for name, t in zip(("scores", "inds", "classes", "ys", "xs"), r):
            eq(npy(t), g[f"out_{tag}_{name}"], f"{tag} {name}")
    # the reference composition nms -> topk on the materialised map
    r2 = ops.topk(ops.nms(heat), K)
    for name, t in zip(("scores", "inds", "classes", "ys", "xs"), r2):
        eq(npy(t), g[f"out_nms_{name}"], f"nms-then-topk {name}")
    eq(npy(ops.transposeAndGetFeature(cu(g["in_feat"]), cu(g["out_raw_inds"]))), g["out_gather"], "gather")


def test_golden_alpha_thr(ops):
    g = load_golden("alpha_thr")
    alpha = npy(ops.get_alpha(cu(g["in_rot"])))
    np.testing.assert_allclose(alpha, g["out_alpha"], rtol=2e-6, atol=1e-6)
    thr = npy(ops.getDistanceThresh(cu(g["in_calib"]), cu(g["in_center"]), cu(g["in_dim"]), cu(g["out_alpha"])))
    np.testing.assert_allclose(thr, g["out_thr"], rtol=2e-6, atol=1e-6)


def test_golden_pc2hm_quirks(ops):
    g = load_golden("pc2hm_quirks")
    pc_dep = cu(g["in_pc_dep"])
    for j in range(len(g["in_depth"])):
        hm = torch.zeros_like(pc_dep)
        ops.cvtPcDepthToHeatmap(hm, pc_dep, cu(g["in_depth"][j:j + 1])[0], cu(g["in_bbox"][j]), cu(g["in_thr"][j:j + 1])[0],
                                float(g["in_max_pc_dist"]))
        eq(npy(hm), g["out_pc_hm"][j], f"quirk case {j}")


def test_golden_pillars(ops):
    g = load_golden("pillars")
    for tag in "abc":
        H, W = (int(v) for v in g[f"in_hw_{tag}"])
        calib = g[f"in_calib_{tag}"]
        pc_dep, owner = ops.radar_to_pc_dep(cu(g[f"in_points_{tag}"]), cu(g[f"in_counts_{tag}"].astype(np.int32)),
                                            cu(calib[:, :, :3]), cu(calib), cu(g[f"in_A_{tag}"]), (H, W),
                                            return_owner=True)
        ref = g[f"out_pc_dep_{tag}"]
        eq(npy(pc_dep), ref, f"pillars {tag}")
        assert ((npy(owner) >= 0) == (ref[:, 0] != 0)).all()


def test_pillars_with_a_records_only_workspace(ops):
    """cf_radar_pillars sizes its workspace with cf_pillars_workspace_bytes (records + the dense rectangle array the
    paint stages from); a caller that still passes the records-only size of earlier builds (B * Nmax * 32 + 256) gets the
    same pc_dep -- the paint then reads the rectangles out of the records."""
    import ctypes as C
    from centerfusiondetect3d_b200 import _lib
    g = load_golden("pillars")
    L = _lib.lib()
    for tag in "ac":
        H, W = (int(v) for v in g[f"in_hw_{tag}"])
        calib = g[f"in_calib_{tag}"]
        pts, cnt = cu(g[f"in_points_{tag}"]), cu(g[f"in_counts_{tag}"].astype(np.int32))
        cam, cal, A = cu(np.ascontiguousarray(calib[:, :, :3])), cu(calib), cu(g[f"in_A_{tag}"])
        B, Nmax, _ = pts.shape
        small = B * Nmax * 32 + 256
        assert small < L.cf_pillars_workspace_bytes(B, Nmax)
        ws = torch.empty(small, dtype=torch.uint8, device=DEV)
        out = torch.full((B, 3, H, W), float("nan"), device=DEV)
        pd = (C.c_double * 3)(1.5, 0.2, 0.2)
        rc = L.cf_radar_pillars(pts.data_ptr(), cnt.data_ptr(), B, Nmax, cam.data_ptr(), cal.data_ptr(), A.data_ptr(), 1600, 900,
                                H, W, pd, 60.0, 0.0, out.data_ptr(), None, ws.data_ptr(), small,
                                torch.cuda.current_stream().cuda_stream)
        assert rc == 0, L.cf_last_error().decode()
        eq(npy(out), g[f"out_pc_dep_{tag}"], f"pillars {tag}, records-only workspace")


def test_golden_frustum(ops):
    g = load_golden("frustum_decode")
    K = int(g["in_K"])
    out = {k: cu(g["in_" + k]) for k in FWD_KEYS}
    pc_hm, assoc, thr, _ = ops.frustum_association(out, cu(g["in_pc_dep"]), cu(g["in_calib"]), K,
                                                   float(g["in_max_pc_dist"]))
    np.testing.assert_allclose(npy(thr), g["out_thr"], rtol=RTOL_THR_REF, atol=5e-6)
    known = g["out_assoc"] != -2
    assert np.array_equal(npy(assoc)[known], g["out_assoc"][known])
    eq(npy(pc_hm), g["out_pc_hm"], "pc_hm")
    # reference signature
    eq(npy(ops.getPcFrustumHeatmap(out, cu(g["in_pc_dep"]), cu(g["in_calib"]), Cfg(K, float(g["in_max_pc_dist"])))),
       g["out_pc_hm"], "getPcFrustumHeatmap")


def test_golden_decode_variants(ops):
    g = load_golden("frustum_decode")
    K = int(g["in_K"])
    H, W = g["in_heatmap"].shape[2:]
    for vname in ("full", "noreg_norm2d", "unc"):
        keys = str(g[f"in_dec_{vname}_keys"]).split(",")
        out = {k: cu(g["in_" + k]) for k in keys}
        had_rot2 = "rotation2" in out
        ret = ops.fusionDecode([out], (H, W), K, bool(g[f"in_dec_{vname}_norm2d"]))
        if had_rot2:
            assert "rotation2" not in out and "rotation" in out  # decode.py:120-121 side effect
        ref_keys = sorted(k[len(f"out_dec_{vname}_"):] for k in g if k.startswith(f"out_dec_{vname}_"))
        assert sorted(ret) == ref_keys, vname
        for k in ref_keys:
            refv = g[f"out_dec_{vname}_{k}"]
            if k == "scores" and vname == "unc":
                np.testing.assert_allclose(npy(ret[k]), refv, rtol=1e-6)
            else:
                eq(npy(ret[k]), refv, f"{vname} {k}")


def test_golden_cfg1_whole_model(ops):
    """BASELINE.json configs[0] / SURVEY.md 8(c) item (3): the head maps the reference's own random-init DLA-34 forward
    produced (CenterFusionHead.forward, detectHeads.py:165-191; fixture from tests/golden/make_golden.py
    gen_cfg1_model) replayed through the ops that replace its calls -- getPcFrustumHeatmap, fusionDecode,
    postProcess -- against the reference's own outputs: indices and associations bit-exact, values 1e-5."""
    from test_oracle_golden import CFG1_HEADS
    g = load_golden("cfg1_model")
    K = int(g["in_K"])
    H, W = g["in_heatmap"].shape[2:]
    cfg = Cfg(K, float(g["in_max_pc_dist"]))
    y = {k: cu(g["in_" + k]) for k in CFG1_HEADS}
    pc_dep, calib = cu(g["in_pc_dep"]), cu(g["in_calib"])
    # the call of detectHeads.py:176-177
    eq(npy(ops.getPcFrustumHeatmap(y, pc_dep, calib, cfg)), g["out_pc_hm"], "pc_hm")
    _, assoc, thr, _ = ops.frustum_association(y, pc_dep, calib, K, cfg.DATASET.MAX_PC_DIST)
    known = g["out_assoc"] != -2
    assert np.array_equal(npy(assoc)[known], g["out_assoc"][known])
    np.testing.assert_allclose(npy(thr), g["out_thr"], rtol=RTOL_THR_REF, atol=5e-6)
    # detector.py:414-424 / progressBar.py:96-104: fusionDecode then postProcess
    ret = ops.fusionDecode([dict(y)], (H, W), K)
    ref_keys = sorted(k[len("out_dec_"):] for k in g if k.startswith("out_dec_"))
    assert sorted(ret) == ref_keys
    for k in ref_keys:
        eq(npy(ret[k]), g["out_dec_" + k], "decode " + k)
    out = ops.postProcess(ret, g["in_center"], float(g["in_scale"]), H, W, calib)
    post_keys = sorted(k[len("out_post_"):] for k in g if k.startswith("out_post_"))
    assert sorted(out) == post_keys
    for k in post_keys:
        np.testing.assert_allclose(npy(out[k]), g["out_post_" + k], rtol=1e-5, atol=1e-5, err_msg="post " + k)
    # and the fused call on the same maps
    pc_hm2, ret2 = ops.fusion_stage(dict(y), pc_dep, calib, cfg)
    eq(npy(pc_hm2), g["out_pc_hm"], "fused pc_hm")
    for k in ref_keys:
        eq(npy(ret2[k]), g["out_dec_" + k], "fused decode " + k)
    # ... and with postProcess folded into the same launch (row f1)
    _, out3 = ops.fusion_stage(dict(y), pc_dep, calib, cfg,
                               post=dict(center=g["in_center"], scale=float(g["in_scale"]), height=H, width=W, calibs=calib))
    assert ops.last_launch_count() == 1 and sorted(out3) == post_keys
    for k in post_keys:
        np.testing.assert_allclose(npy(out3[k]), g["out_post_" + k], rtol=1e-5, atol=1e-5, err_msg="fused post " + k)
        eq(npy(out3[k]), npy(out[k]), "fused post == postProcess launch, " + k)


def test_get_feature_bnc_form(ops):
    """a3, the (B,N,ch) form (model/utils.py:41-55): out[b,m,c] = feature[b,indices[b,m],c]."""
    rng = np.random.default_rng(12)
    feat = rng.normal(size=(3, 50, 7)).astype(np.float32)
    ind = rng.integers(0, 50, size=(3, 11)).astype(np.int64)
    want = np.take_along_axis(feat, ind[:, :, None].repeat(7, axis=2), axis=1)
    eq(npy(ops.getFeature(cu(feat), cu(ind))), want, "getFeature")
    # an index outside the tensor reads nothing and yields NaN (torch.gather: device-side assert)
    ind[1, 4], ind[2, 0] = 50, -1
    got = npy(ops.getFeature(cu(feat), cu(ind)))
    assert np.isnan(got[1, 4]).all() and np.isnan(got[2, 0]).all()
    ok = np.ones((3, 11), bool)
    ok[1, 4] = ok[2, 0] = False
    eq(got[ok], want[ok], "getFeature, in-range rows")


# ------------------------------------------------------------------------------------------------
# 2. oracle differential at BASELINE.json shapes
# ------------------------------------------------------------------------------------------------
def _check_stage(ops, oracle, wl, inp, B, need_matches=True):
    fwd = to_cuda(inp, FWD_KEYS, DEV)
    pc_dep = cu(inp["pc_dep"])
    calib = cu(inp["calib"])
    # F
    hm_o, assoc_o, thr_o, inds_o = oracle.frustum_pc_hm(inp["heatmap"], inp["widthHeight"], inp["depth"],
                                                        inp["dimension"], inp["rotation"], inp["pc_dep"],
                                                        inp["calib"], wl.K, wl.max_pc_dist)
    pc_hm, assoc, thr, inds = ops.frustum_association(fwd, pc_dep, calib, wl.K, wl.max_pc_dist)
    eq(npy(inds), inds_o, "raw top-K inds")
    np.testing.assert_allclose(npy(thr), thr_o, rtol=RTOL_THR_ORACLE, atol=0)
    eq(npy(assoc), assoc_o, "association indices")
    eq(npy(pc_hm), hm_o, "pc_hm")
    assert not need_matches or (assoc_o >= 0).sum() >= B  # the case exercises real matches
    # D
    dec_np = {k: inp[k] for k in DEC_KEYS}
    ret_o = oracle.fusion_decode(dec_np, (wl.H, wl.W), wl.K)
    dec = to_cuda(inp, DEC_KEYS, DEV)
    ret = ops.fusionDecode([dec], (wl.H, wl.W), wl.K)
    for k in ret:
        eq(npy(ret[k]), ret_o[k], "decode " + k)
    # F + D fused: one pass over the heat-map
    dec2 = to_cuda(inp, DEC_KEYS, DEV)
    pc_hm2, ret2 = ops.fusion_stage(dec2, pc_dep, calib, Cfg(wl.K, wl.max_pc_dist))
    plan = ops.last_plan()  # of the fused call: the one bench.py measures
    eq(npy(pc_hm2), hm_o, "fused pc_hm")
    for k in ret2:
        eq(npy(ret2[k]), ret_o[k], "fused decode " + k)
    # topk both flavours incl. classes / ys / xs
    for apply_nms in (False, True):
        r_o = oracle.topk(inp["heatmap"], wl.K, apply_nms=apply_nms)
        r = ops.topk(fwd["heatmap"], wl.K, apply_nms=apply_nms)
        for a, b, nm in zip(r, r_o, ("scores", "inds", "classes", "ys", "xs")):
            eq(npy(a), b, "topk %s nms=%s" % (nm, apply_nms))
    return plan


@pytest.mark.parametrize("cfg_name,B", [("cfg1", 1), ("cfg2", 4), ("cfg5", 2)])
def test_oracle_stage(ops, oracle, cfg_name, B):
    wl = syn.WORKLOADS[cfg_name]
    inp = make_case(wl, B)
    _check_stage(ops, oracle, wl, inp, B)


@pytest.mark.parametrize("cs", [1, 2, 4, 8, 16])
@pytest.mark.parametrize("cfg_name,B", [("cfg1", 3), ("cfg5", 3)])
def test_oracle_stage_forced_plans(ops, oracle, cfg_name, B, cs):
    """Every execution plan of the fused stage kernel gives the oracle's result: cluster sizes 1..16 (heat-map share
    resident in shared memory, or streamed through the slot ring and re-fetched by the filter pass) and clusters that
    process several frames one after the other.  The host picks the plan per call; the ABI's test hook
    (cf_debug_set_plan) pins it."""
    wl = syn.WORKLOADS[cfg_name]
    inp = make_case(wl, B)
    ops.debug_set_plan(cs, 2)
    try:
        plan = _check_stage(ops, oracle, wl, inp, B)
        assert plan["cluster_size"] == cs and plan["n_clusters"] == 2 and plan["rounds"] == 2, plan
    finally:
        ops.debug_set_plan(0, 0)


# the plans bench.py runs (no forcing): workload, frames per GPU, expected (cluster size, resident share, staged plane)
BENCH_PLANS = [("cfg2", 32, dict(cluster_size=4, n_clusters=32, rounds=1, resident=1, depth_plane_staged=1)),
               ("cfg3", 256, None),
               ("cfg4", 64, dict(cluster_size=2, n_clusters=64, resident=0, rounds=1)),
               ("cfg5", 8, dict(cluster_size=8, n_clusters=8, rounds=1))]


@pytest.mark.parametrize("cfg_name,B,expect", BENCH_PLANS)
def test_oracle_stage_bench_shapes(ops, oracle, cfg_name, B, expect):
    """Oracle differential at the EXACT batch sizes and execution plans bench.py measures (BASELINE.json configs 2-5:
    the planner chooses from B, so small-batch tests do not reach these plans).  The chosen plan is asserted through
    cf_last_plan so that a planner change cannot silently move this test off the benchmarked path."""
    wl = syn.WORKLOADS[cfg_name]
    inp = make_case(wl, B, seed=4242)
    ops.debug_set_plan(0, 0)
    plan = _check_stage(ops, oracle, wl, inp, B)
    for k, v in (expect or {}).items():
        assert plan[k] == v, (cfg_name, k, plan)
    n_sm = torch.cuda.get_device_properties(0).multi_processor_count
    assert plan["n_clusters"] * plan["cluster_size"] <= n_sm
    if expect is None:  # cfg3 on one GPU: more frames than clusters -> every cluster loops over several frames
        assert plan["rounds"] >= 2 and plan["n_clusters"] * plan["cluster_size"] > n_sm // 2, plan


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6, 7, 8])
def test_oracle_stage_seeds_and_plans(ops, oracle, seed):
    """Seeded stress (tools/stress_parity.py folded into the suite): a fresh cfg2-shaped case per seed, run under the
    planner's plan and under a forced plan that changes with the seed (cluster size x frames per cluster)."""
    wl = syn.WORKLOADS["cfg2"]
    B = 2 + seed % 4
    inp = make_case(wl, B, seed=9000 + 17 * seed)
    _check_stage(ops, oracle, wl, inp, B)
    cs = (1, 2, 4, 8, 16)[seed % 5]
    ops.debug_set_plan(cs, 1 + seed % 3)
    try:
        _check_stage(ops, oracle, wl, inp, B)
    finally:
        ops.debug_set_plan(0, 0)


def test_paint_into_combiner_input(ops, oracle):
    """f3: pc_hm painted straight into channels 64..66 of the (B,67,H,W) secondary-head input equals the stand-alone
    pc_hm, and the 64 image-feature channels are left untouched (ConcateCombiner, fusionModules.py:18-35)."""
    wl = syn.WORKLOADS["cfg2"]
    B = 5
    inp = make_case(wl, B)
    cfg = type("Cfg", (), {"MODEL": type("M", (), {"K": wl.K})(), "DATASET": type("D", (), {"MAX_PC_DIST": wl.max_pc_dist})()})()
    pc_hm, ret = ops.fusion_stage(to_cuda(inp, DEC_KEYS, DEV), cu(inp["pc_dep"]), cu(inp["calib"]), cfg)
    sec = torch.full((B, 67, wl.H, wl.W), 7.0, dtype=torch.float32, device=DEV)
    pc_hm2, ret2 = ops.fusion_stage(to_cuda(inp, DEC_KEYS, DEV), cu(inp["pc_dep"]), cu(inp["calib"]), cfg, sec_feats=sec)
    assert pc_hm2.data_ptr() == sec[:, 64:].data_ptr()
    eq(npy(sec[:, 64:]), npy(pc_hm), "pc_hm inside the combiner input")
    assert bool((sec[:, :64] == 7.0).all())
    for k in ret:
        eq(npy(ret2[k]), npy(ret[k]), k)
    hm_o = oracle.frustum_pc_hm(inp["heatmap"], inp["widthHeight"], inp["depth"], inp["dimension"], inp["rotation"],
                                inp["pc_dep"], inp["calib"], wl.K, wl.max_pc_dist)[0]
    eq(npy(sec[:, 64:]), hm_o, "pc_hm vs oracle")


def test_pinned_host_heads_read_in_place(ops, oracle):
    """Host-buffer callers: the gather-only heads may stay in pinned host memory (read in place at the K peak pixels);
    results are bit-identical to the all-device call and to the oracle."""
    wl = syn.WORKLOADS["cfg2"]
    B = 3
    inp = make_case(wl, B)
    cfg = type("Cfg", (), {"MODEL": type("M", (), {"K": wl.K})(), "DATASET": type("D", (), {"MAX_PC_DIST": wl.max_pc_dist})()})()
    pc_hm, ret = ops.fusion_stage(to_cuda(inp, DEC_KEYS, DEV), cu(inp["pc_dep"]), cu(inp["calib"]), cfg)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()  # noqa: E731
    mk = lambda: {k: (cu(inp[k]) if k == "heatmap" else pin(inp[k])) for k in DEC_KEYS}  # noqa: E731 (calls rename keys)
    mixed = mk()
    pc_hm2, ret2 = ops.fusion_stage(mk(), pin(inp["pc_dep"]), pin(inp["calib"]), cfg)  # depth plane staged by ops
    eq(npy(pc_hm2), npy(pc_hm), "pc_hm")
    staged = torch.zeros_like(cu(inp["pc_dep"]))
    staged[:, 0].copy_(cu(inp["pc_dep"])[:, 0])
    pc_hm3, _ = ops.fusion_stage(mk(), staged, cu(inp["calib"]), cfg, pc_vel=pin(inp["pc_dep"]))
    eq(npy(pc_hm3), npy(pc_hm), "pc_hm, caller-staged depth plane")
    assert set(ret2) == set(ret)
    for k in ret:
        eq(npy(ret2[k]), npy(ret[k]), k)
    hm_o = oracle.frustum_pc_hm(inp["heatmap"], inp["widthHeight"], inp["depth"], inp["dimension"], inp["rotation"],
                                inp["pc_dep"], inp["calib"], wl.K, wl.max_pc_dist)[0]
    eq(npy(pc_hm2), hm_o, "pc_hm vs oracle")
    with pytest.raises(ValueError):  # pageable host memory is not readable by the device
        bad = dict(mixed)
        bad["depth"] = torch.from_numpy(np.ascontiguousarray(inp["depth"]))
        ops.fusion_stage(bad, cu(inp["pc_dep"]), cu(inp["calib"]), cfg)


def test_radar_sort_on_device(ops, oracle):
    """f2: the device sort is np.argsort(kind="stable") on depth (ties keep source order, padding rows stay), and the
    pillar expansion of device-sorted points equals the one of host-sorted points."""
    wl = syn.WORKLOADS["cfg4"]
    B = 3
    inp = make_case(wl, B)
    pts, cnt = inp["radar_points"].copy(), inp["radar_counts"].copy()
    rng = np.random.default_rng(3)
    shuf = pts.copy()
    for b in range(B):
        n = int(cnt[b])
        shuf[b, :n] = pts[b, rng.permutation(n)]
        shuf[b, 5:25, 2] = shuf[b, 4, 2]  # a run of equal depths: ties must keep their source order
    cnt[1] = max(0, cnt[1] - 17)  # ragged
    want = shuf.copy()
    order_w = np.tile(np.arange(shuf.shape[1], dtype=np.int32), (B, 1))
    for b in range(B):
        n = int(cnt[b])
        o = np.argsort(shuf[b, :n, 2], kind="stable")
        want[b, :n] = shuf[b, :n][o]
        order_w[b, :n] = o
    got, order = ops.radar_sort(cu(shuf), cu(cnt), return_order=True)
    eq(npy(got), want, "sorted points")
    eq(npy(order), order_w, "order")
    args = (cu(inp["cam"]), cu(inp["calib64"]), cu(inp["A_out"]), (wl.H, wl.W), (wl.img_w, wl.img_h), wl.pillar_dims,
            wl.max_pc_dist)
    a = ops.radar_to_pc_dep(cu(want), cu(cnt), *args)
    b_ = ops.radar_to_pc_dep(cu(shuf), cu(cnt), *args, sort_points=True)
    eq(npy(a), npy(b_), "pc_dep of device-sorted points")
    pc_o, _ = oracle.radar_pillars(want, cnt, inp["cam"], inp["calib64"], inp["A_out"], (wl.img_w, wl.img_h), (wl.H, wl.W),
                                   wl.pillar_dims, wl.max_pc_dist, want_owner=False)
    eq(npy(b_), pc_o, "pc_dep vs oracle")


def test_golden_radar_ingest_and_flip(ops, oracle):
    """Row f2, complete: the stored (18, N) radar array through ops.radar_ingest + ops.radar_to_pc_dep(flip=...) against
    the reference's own loadRadarPointCloud (fixture gen_radar_ingest) -- pc_dep bit-exact -- and against the oracle's
    restatement of the ingestion; then all cases as ONE ragged batch (Nmax = the largest, per-frame flip flags)."""
    from test_oracle_golden import check_sorted_points_against_reference, radar_ingest_cases
    g = load_golden("radar_ingest")
    cases = list(radar_ingest_cases(g))
    for tag, raw, cnt, calib, A, (H, W), flip, desc, zoff in cases:
        pts, order = ops.radar_ingest(cu(raw), cu(cnt), flip=cu(flip), descending=desc, return_order=True)
        pts_o, order_o = oracle.radar_ingest(raw, cnt, flip, desc)
        eq(npy(pts), pts_o, "ingested points " + tag)
        eq(npy(order), order_o, "order " + tag)
        check_sorted_points_against_reference(g, tag, npy(pts)[0], zoff)
        pc_dep = ops.radar_to_pc_dep(pts, cu(cnt), cu(calib[:, :, :3]), cu(calib), cu(A), (H, W), z_offset=zoff,
                                     flip=cu(flip))
        eq(npy(pc_dep)[0], g["out_pc_dep_" + tag], "pc_dep " + tag)
    # one batch of the three 112x200 cases (they differ in N, flip and order direction: direction is per call)
    for desc in (False, True):
        sel = [c for c in cases if c[5] == (112, 200) and c[7] == desc and c[8] == 0.0]
        if not sel:
            continue
        Nmax = max(c[1].shape[2] for c in sel)
        raw = np.zeros((len(sel), 18, Nmax))
        for i, c in enumerate(sel):
            raw[i, :, :c[1].shape[2]] = c[1][0]
        cnt = np.array([c[1].shape[2] for c in sel], np.int32)
        calib = np.concatenate([c[3] for c in sel])
        A = np.concatenate([c[4] for c in sel])
        flip = np.array([bool(c[6][0]) for c in sel])
        pts = ops.radar_ingest(cu(raw), cu(cnt), flip=cu(flip), descending=desc)
        pc_dep = ops.radar_to_pc_dep(pts, cu(cnt), cu(calib[:, :, :3]), cu(calib), cu(A), (112, 200), flip=cu(flip))
        for i, c in enumerate(sel):
            eq(npy(pc_dep)[i], g["out_pc_dep_" + c[0]], "batched pc_dep " + c[0])
    # packed [X,Y,Z,vx,vz] input, ties: descending == the reversed stable order
    rng = np.random.default_rng(5)
    p5 = rng.normal(size=(2, 40, 5))
    p5[:, 10:20, 2] = p5[:, 3:4, 2]
    cnt = np.array([40, 33], np.int32)
    got = npy(ops.radar_ingest(cu(p5), cu(cnt), descending=True, packed=True))
    for b in range(2):
        n = int(cnt[b])
        eq(got[b, :n], p5[b, :n][np.argsort(p5[b, :n, 2], kind="stable")[::-1]], "descending stable")
        eq(got[b, n:], p5[b, n:], "padding rows stay")


def test_frustum_targets(ops, oracle):
    """f4: association of given (ground-truth style) boxes in list order == the oracle's cvtPcDepthToHeatmap loop
    (pinned on the reference by the quirk fixtures), ragged counts included."""
    wl = syn.WORKLOADS["cfg2"]
    B, N = 3, 37
    inp = make_case(wl, B)
    rng = np.random.default_rng(7)
    cx, cy = rng.uniform(-5, wl.W + 5, (B, N)), rng.uniform(-5, wl.H + 5, (B, N))
    w, h = rng.uniform(0, 60, (B, N)), rng.uniform(0, 40, (B, N))
    bb = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], -1).astype(np.float32)
    depth = rng.uniform(1, 70, (B, N)).astype(np.float32)
    thr = rng.uniform(0.3, 6.0, (B, N)).astype(np.float32)
    counts = np.array([N, 0, 11], dtype=np.int32)
    hm_o = np.zeros_like(inp["pc_dep"])
    assoc_o = np.full((B, N), -1, np.int32)
    for b in range(B):
        for n in range(counts[b]):
            assoc_o[b, n] = oracle.pc_dep_to_hm(hm_o[b], inp["pc_dep"][b], depth[b, n], bb[b, n], thr[b, n], wl.max_pc_dist)
    hm, assoc = ops.frustum_targets(cu(inp["pc_dep"]), cu(bb), cu(depth), cu(thr), cu(counts), wl.max_pc_dist,
                                    return_assoc=True)
    eq(npy(assoc), assoc_o, "assoc")
    eq(npy(hm), hm_o, "pc_hm")
    assert (assoc_o >= 0).sum() >= 5


def test_golden_gt_association_numpy_branch(ops):
    """f4 on the reference's real branch: fixture generated by the NumPy form of cvtPcDepthToHeatmap under the
    DataLoader's loop (generic_dataset.py:673-687) -- float32 boxes + np.float32 thresholds + Python-float depths (its
    own arrangement) and float64 boxes / thresholds crossing every border (generic NumPy caller)."""
    from test_oracle_golden import gt_assoc_frames
    for f, f32, pc_dep, bb, depth, thr, hm_ref, assoc_ref, mx in gt_assoc_frames():
        hm, assoc = ops.frustum_targets(cu(pc_dep[None]), cu(bb[None]), cu(depth[None]), cu(thr[None]), None, mx,
                                        return_assoc=True)
        known = assoc_ref != -2
        assert np.array_equal(npy(assoc)[0][known], assoc_ref[known]), f
        eq(npy(hm)[0], hm_ref, "pc_hm frame %d" % f)
        if f32:  # all-float32 arguments (depth rounded like NumPy's weak Python float): the torch-branch kernel agrees
            hm2 = ops.frustum_targets(cu(pc_dep[None]), cu(bb[None]), cu(depth.astype(np.float32)[None]), cu(thr[None]), None, mx)
            eq(npy(hm2)[0], hm_ref, "pc_hm frame %d, fp32 entry point" % f)


def test_post_process(ops, oracle):
    """f1 postProcess: the kernel against the oracle (bit-exact apart from the fp64-rounded trigonometry, 1e-6) and
    against the reference's own outputs (golden fixture, 1e-5 relative: north_star's bar for decoded boxes)."""
    from test_oracle_golden import POST_VARIANTS, post_case
    g = load_golden("post_process")
    H, W = int(g["in_H"]), int(g["in_W"])
    trans = ops.inverse_output_affine(g["in_center"], float(g["in_scale"]), W, H)
    eq(trans, g["in_trans"], "inverse output affine")
    for v in POST_VARIANTS:
        y, is_gt, ref = post_case(g, v)
        out_o = oracle.post_process(y, g["in_trans"], (H, W), g["in_calib"], is_gt)
        yd = {k: cu(a) for k, a in y.items()}
        cls_t, vel_t = yd["classIds"], yd.get("velocity")
        out = ops.postProcess(yd, g["in_center"], float(g["in_scale"]), H, W, cu(g["in_calib"]), isGt=is_gt)
        assert out is yd and out["classIds"] is cls_t and (vel_t is None or out["velocity"] is vel_t)  # in place
        assert set(out) == set(ref), (v, set(out) ^ set(ref))
        for k, r in ref.items():
            a = npy(out[k])
            assert a.shape == r.shape and a.dtype == r.dtype, (v, k, a.shape, r.shape)
            np.testing.assert_allclose(a, r, rtol=1e-5, atol=1e-5, err_msg="vs reference %s %s" % (v, k))
            if k in ("alpha", "yaws", "velocity", "bboxes3d", "locations"):
                np.testing.assert_allclose(a, out_o[k], rtol=1e-6, atol=1e-6, err_msg="vs oracle %s %s" % (v, k))
            else:
                eq(a, out_o[k], "vs oracle %s %s" % (v, k))
    # full-size property: B=32, K=100 through fusionDecode -> postProcess; boxes of dead dimensions are zero,
    # everything finite, yaw wrapped into [-pi, pi]
    wl = syn.WORKLOADS["cfg2"]
    inp = make_case(wl, 4)
    ret = ops.fusionDecode([to_cuda(inp, DEC_KEYS, DEV)], (wl.H, wl.W), wl.K)
    out = ops.postProcess(ret, (800.0, 450.0), 1600.0, wl.H, wl.W, cu(inp["calib"]))
    assert out["bboxes3d"].shape == (4, wl.K, 8, 3) and bool(torch.isfinite(out["bboxes3d"]).all())
    assert float(out["yaws"].abs().max()) <= np.pi + 1e-6
    dead = (out["dimension"] <= 0).any(dim=2)
    assert not bool(out["bboxes3d"][dead].any())


def _same_dict(a: dict, b: dict, what: str):
    assert sorted(a) == sorted(b), (what, sorted(a), sorted(b))
    for k in a:
        eq(npy(a[k]), npy(b[k]), "%s %s" % (what, k))


@pytest.mark.parametrize("cfg_name,B,heads_dropped", [("cfg2", 5, ()), ("cfg2", 3, ("amodal_offset",)),
                                                     ("cfg2", 2, ("amodal_offset", "widthHeight", "reg")),
                                                     ("cfg2", 2, ("depth2", "depth")), ("cfg5", 2, ())])
def test_post_process_fused_into_decode(ops, cfg_name, B, heads_dropped):
    """Row f1, fused: fusionDecode(post=...) and fusion_stage(post=...) run postProcess in the decode epilogue of their
    one launch; the result is bit for bit what fusionDecode -> postProcess (two launches) returns, for every mix of
    heads postProcess branches on (postProcess.py:39-64: amodal centre / box centre / plain centre, 3-D keys or not)."""
    wl = syn.WORKLOADS[cfg_name]
    inp = make_case(wl, B)
    cfg = Cfg(wl.K, wl.max_pc_dist)
    post = dict(center=(800.0, 450.0), scale=1600.0, height=wl.H, width=wl.W, calibs=cu(inp["calib"]))
    dec_keys = [k for k in DEC_KEYS if k not in heads_dropped]
    # two launches
    ret = ops.fusionDecode([to_cuda(inp, dec_keys, DEV)], (wl.H, wl.W), wl.K)
    n_dec = ops.last_launch_count()
    want = ops.postProcess(ret, post["center"], post["scale"], wl.H, wl.W, post["calibs"])
    # one launch: decode alone ...
    got = ops.fusionDecode([to_cuda(inp, dec_keys, DEV)], (wl.H, wl.W), wl.K, post=post)
    assert ops.last_launch_count() == n_dec
    _same_dict(got, want, "decode+post")
    # ... and the whole stage (frustum heads always complete; the decode dict is the varied one)
    fwd = to_cuda(inp, FWD_KEYS, DEV)
    pc_dep, calib = cu(inp["pc_dep"]), cu(inp["calib"])
    pc_hm0, _ = ops.fusion_stage(dict(fwd), pc_dep, calib, cfg, decode_output=to_cuda(inp, dec_keys, DEV))
    pc_hm, got2 = ops.fusion_stage(dict(fwd), pc_dep, calib, cfg, decode_output=to_cuda(inp, dec_keys, DEV), post=post)
    assert ops.last_launch_count() == 1
    eq(npy(pc_hm), npy(pc_hm0), "pc_hm unchanged by the fused post")
    _same_dict(got2, want, "stage+post")
    # ground-truth flavour (isGt: centres stay the plain ones, velocity untouched; postProcess.py:54,71)
    want_gt = ops.postProcess(ops.fusionDecode([to_cuda(inp, dec_keys, DEV)], (wl.H, wl.W), wl.K), post["center"],
                              post["scale"], wl.H, wl.W, post["calibs"], isGt=True)
    got_gt = ops.fusionDecode([to_cuda(inp, dec_keys, DEV)], (wl.H, wl.W), wl.K, post=dict(post, isGt=True))
    _same_dict(got_gt, want_gt, "decode+post isGt")


def test_oracle_dense_radar_with_pillars(ops, oracle):
    """cfg4: 2000 points/frame, pillar expansion on the GPU feeding the association."""
    wl = syn.WORKLOADS["cfg4"]
    B = 3
    inp = make_case(wl, B)
    pc_dep, owner = ops.radar_to_pc_dep(cu(inp["radar_points"]), cu(inp["radar_counts"]), cu(inp["cam"]),
                                        cu(inp["calib64"]), cu(inp["A_out"]), (wl.H, wl.W), (wl.img_w, wl.img_h),
                                        wl.pillar_dims, wl.max_pc_dist, return_owner=True)
    eq(npy(pc_dep), inp["pc_dep"], "pc_dep")
    eq(npy(owner), inp["owner"], "owner")
    assert (inp["pc_dep"][:, 0] != 0).mean() > 0.5
    _check_stage(ops, oracle, wl, inp, B)


def test_ragged_and_empty_radar(ops, oracle):
    wl = syn.WORKLOADS["cfg2"]
    B = 4
    inp = syn.make_inputs(wl, seed=77, B=B)
    counts = np.array([0, 1, 137, wl.n_radar], np.int32)
    for dims in ((112, 200), (57, 101)):
        A = np.repeat(syn.output_affine(dims[0], dims[1])[None], B, axis=0)
        pd_o, ow_o = oracle.radar_pillars(inp["radar_points"], counts, inp["cam"], inp["calib64"], A, (1600, 900), dims)
        pd, ow = ops.radar_to_pc_dep(cu(inp["radar_points"]), cu(counts), cu(inp["cam"]), cu(inp["calib64"]), cu(A),
                                     dims, return_owner=True)
        eq(npy(pd), pd_o, "ragged pc_dep %s" % (dims,))
        eq(npy(ow), ow_o, "ragged owner")
        assert not pd_o[0].any()
    # Nmax = 0
    pd = ops.radar_to_pc_dep(torch.zeros((2, 0, 5), dtype=torch.float64, device=DEV), cu(np.zeros(2, np.int32)),
                             cu(inp["cam"][:2]), cu(inp["calib64"][:2]), cu(inp["A_out"][:2]), (112, 200))
    assert not npy(pd).any()
    # no radar at all: pc_hm all zeros, nothing associated
    fwd = to_cuda(inp, FWD_KEYS, DEV)
    hm, assoc, _, _ = ops.frustum_association(fwd, torch.zeros((B, 3, 112, 200), device=DEV), cu(inp["calib"]), wl.K, 60.0)
    assert not npy(hm).any() and (npy(assoc) == -1).all()


@pytest.mark.parametrize("shape", [(1, 1, 5, 7, 35), (2, 3, 9, 13, 20), (1, 2, 1, 64, 10), (2, 10, 33, 50, 1),
                                   (1, 4, 24, 40, 700), (1, 1, 300, 8, 64)])
def test_odd_geometries(ops, oracle, shape):
    """W not a multiple of 4 (no TMA path), K == H*W, K == 1, single row, K > 512, tall-thin maps."""
    B, Cc, H, W, K = shape
    rng = np.random.default_rng(5)
    heat = rng.permutation(B * Cc * H * W).astype(np.float32).reshape(B, Cc, H, W) / np.float32(B * Cc * H * W + 1)
    h = cu(heat)
    eq(npy(ops.nms(h)), oracle.nms(heat), "nms")
    for apply_nms in (False,) + ((True,) if K <= 8 else ()):
        r_o = oracle.topk(heat, K, apply_nms=apply_nms)
        r = ops.topk(h, K, apply_nms=apply_nms)
        for a, b, nm in zip(r, r_o, ("scores", "inds", "classes", "ys", "xs")):
            eq(npy(a), b, "%s nms=%s" % (nm, apply_nms))


def test_ties_follow_the_documented_total_order(ops, oracle):
    """After nms most pixels are exactly 0 and plateaus tie: order is (score desc, class asc, index asc)."""
    rng = np.random.default_rng(6)
    heat = (rng.integers(0, 6, size=(2, 3, 16, 20)) / 8.0).astype(np.float32)  # heavy ties, K beyond the positives
    heat[0, 1, 3, 4] = -0.0
    for K in (7, 150):
        for apply_nms in (False, True):
            r_o = oracle.topk(heat, K, apply_nms=apply_nms)
            r = ops.topk(cu(heat), K, apply_nms=apply_nms)
            for a, b, nm in zip(r, r_o, ("scores", "inds", "classes", "ys", "xs")):
                eq(npy(a), b, "K=%d %s nms=%s" % (K, nm, apply_nms))


@pytest.mark.parametrize("levels", [4, 64])
def test_dense_ties_at_reference_size(ops, oracle, levels):
    """A quantised heat-map at the reference map size: thousands of pixels share the K-th value, so the bound of the
    stage kernel admits far more candidates than its fast paths hold -- the flagged-group list, the work list and the
    survivor lists overflow and the exact fall-backs (block scan, in-place append, radix select over the list or the
    frame) must give the oracle's total order all the same.  The whole stage is compared, not only the top-K."""
    wl = syn.WORKLOADS["cfg2"]
    B = 2
    inp = make_case(wl, B)
    rng = np.random.default_rng(levels)
    inp["heatmap"] = (rng.integers(0, levels, size=inp["heatmap"].shape) / float(levels)).astype(np.float32)
    _check_stage(ops, oracle, wl, inp, B, need_matches=False)  # (the winners of a tie sit wherever the index order puts them)


def test_special_values(ops, oracle):
    """Negative scores, +-inf and plateaus through nms and topk (heat-maps are sigmoid outputs in the
    reference, but the op must stay exact for any non-NaN input)."""
    rng = np.random.default_rng(8)
    heat = rng.normal(size=(1, 2, 12, 16)).astype(np.float32)
    heat[0, 0, 0, 0] = np.inf
    heat[0, 1, 5, 5] = -np.inf
    heat[0, 1, 6:8, 6:8] = 2.5
    eq(npy(ops.nms(cu(heat))), oracle.nms(heat), "nms special")
    r_o = oracle.topk(heat, 30)
    r = ops.topk(cu(heat), 30)
    for a, b in zip(r, r_o):
        eq(npy(a), b, "topk special")


def test_errors_and_no_cpu_fallback(ops):
    heat = torch.rand(1, 2, 4, 5)
    with pytest.raises(ValueError):
        ops.topk(heat, 3)  # CPU tensor: no fallback
    with pytest.raises(TypeError):
        ops.topk(heat.to(DEV).double(), 3)
    with pytest.raises(RuntimeError):
        ops.topk(heat.to(DEV), 21)  # K > H*W, like torch.topk
    assert ops.fusionDecode([{"foo": 1}]) == {}
    with pytest.raises(AssertionError):
        ops.fusionDecode({"heatmap": heat})
    # B == 0
    r = ops.topk(torch.empty((0, 2, 4, 5), device=DEV), 3)
    assert r[0].shape == (0, 3)


def test_strided_inputs_and_stream(ops, oracle):
    wl = syn.WORKLOADS["cfg2"]
    inp = make_case(wl, 2, seed=4321)
    heat_nhwc = cu(inp["heatmap"]).permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)  # non-contiguous NCHW view
    assert not heat_nhwc.is_contiguous()
    r_o = oracle.topk(inp["heatmap"], wl.K)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        r = ops.topk(heat_nhwc, wl.K)
    s.synchronize()
    for a, b in zip(r, r_o):
        eq(npy(a), b, "strided / side stream")


# ------------------------------------------------------------------------------------------------
# 3. full-size properties (no oracle: sizes of BASELINE.json configs 2 and 5)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("cfg_name,B", [("cfg2", 32), ("cfg5", 8)])
def test_full_size_properties(ops, cfg_name, B):
    wl = syn.WORKLOADS[cfg_name]
    g = torch.Generator(device="cpu").manual_seed(11)
    heat = torch.rand((B, wl.C, wl.H, wl.W), generator=g).to(DEV)
    s, i, c, y, x = ops.topk(heat, wl.K)
    assert (s[:, :-1] >= s[:, 1:]).all()                                   # sortedness
    assert (heat[torch.arange(B, device=DEV)[:, None], c.long(), y, x] == s).all()   # scores are the map values
    assert (i == y * wl.W + x).all()
    kth = s[:, -1:]
    # the K-th score is the K-th largest value of the frame (torch.rand has ties: < K beat it, >= K reach it)
    assert ((heat.view(B, -1) > kth).sum(dim=1) <= wl.K - 1).all()
    assert ((heat.view(B, -1) >= kth).sum(dim=1) >= wl.K).all()
    n = ops.nms(heat)
    assert (ops.nms(n) == n).all() or True                                  # (nms of nms may differ on plateaus)
    s2, i2, c2, _, _ = ops.topk(heat, wl.K, apply_nms=True)
    s3, i3, c3, _, _ = ops.topk(n, wl.K)
    assert (s2 == s3).all() and (i2 == i3).all() and (c2 == c3).all()       # fused nms == materialised nms
    # sharding invariance: frames are independent
    s4 = ops.topk(heat[B // 2:], wl.K)[0]
    assert (s4 == s[B // 2:]).all()
    # pc_hm: painted values are d/60 of a real pc_dep pixel; association indices point at non-zero pixels
    pc_dep = torch.zeros((B, 3, wl.H, wl.W), device=DEV)
    pc_dep[:, 0, ::3, ::5] = torch.rand((B, (wl.H + 2) // 3, (wl.W + 4) // 5), generator=g).to(DEV) * 59 + 1
    pc_dep[:, 1] = 7.0 * (pc_dep[:, 0] != 0)
    out = {"heatmap": heat, "widthHeight": torch.rand((B, 2, wl.H, wl.W), generator=g).to(DEV) * 30,
           "depth": torch.rand((B, 1, wl.H, wl.W), generator=g).to(DEV) * 60 + 1,
           "dimension": torch.rand((B, 3, wl.H, wl.W), generator=g).to(DEV) * 4 + 0.5,
           "rotation": torch.randn((B, 8, wl.H, wl.W), generator=g).to(DEV)}
    calib = torch.tensor(syn.NUSC_CALIB, dtype=torch.float32).repeat(B, 1, 1).to(DEV)
    hm, assoc, thr, inds = ops.frustum_association(out, pc_dep, calib, wl.K, 60.0)
    assert (inds == i).all()
    m = assoc >= 0
    assert m.any()
    dsel = pc_dep[:, 0].reshape(B, -1).gather(1, assoc.clamp(min=0).long())
    assert (dsel[m] != 0).all()
    painted = hm[:, 0] != 0
    assert (hm[:, 1][painted] == 7.0).all() and (hm[:, 1][~painted] == 0).all()
    # true fp32 division on the CPU (torch's CUDA div-by-scalar multiplies by the reciprocal: 1 ulp off)
    allowed = torch.unique(dsel[m].cpu() / 60.0).to(DEV)
    assert torch.isin(hm[:, 0][painted], allowed).all()
    hm_again = ops.frustum_association(out, pc_dep, calib, wl.K, 60.0)[0]
    assert (hm_again == hm).all()                                           # deterministic
