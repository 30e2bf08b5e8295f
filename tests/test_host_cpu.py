"""CPU-side tests (no GPU, no compute calls): the C-ABI library loads and exports every symbol the header declares,
argument errors come back without touching a device, the host side refuses CPU tensors (no fallback), the
drop-in patch swaps exactly the reference's callables, and the N>1 host logic (frame sharding + gather of the
detections) works over a world_size-2 gloo group."""
import ctypes as C
import os
import re
import types

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "cfd3d_b200.h")


@pytest.fixture(scope="module")
def lib():
    from centerfusiondetect3d_b200 import _lib, build

    build.build()
    return _lib


def declared_symbols():
    txt = open(HEADER).read()
    txt = re.sub(r"#ifdef CFD3D_PROFILE.*?#endif", "", txt, flags=re.S)  # development build only
    return sorted(set(re.findall(r"CF_API\s+[\w\s\*]+?\b(cf_\w+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(lib):
    names = declared_symbols()
    assert len(names) >= 15, names
    L = C.CDLL(lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), "libcfd3d_b200.so does not export %s" % n
        assert n in lib.SYMBOLS, "ctypes table misses %s" % n
    assert sorted(lib.SYMBOLS) == names  # nothing bound that the header does not declare
    assert lib.lib().cf_version() >= 100


def test_abi_argument_errors_need_no_device(lib):
    L = lib.lib()
    assert L.cf_workspace_bytes(32, 10, 112, 200, 100) > 0
    assert L.cf_workspace_bytes(32, 10, 112, 200, 5000) == 0       # K > CF_MAX_K
    assert L.cf_pillars_workspace_bytes(4, 2000) >= 4 * 2000 * 32
    rc = L.cf_peak_topk(None, 1, 1, 4, 4, 2, 0, None, None, None, None, None, None, 0, None)
    assert rc == -1 and b"NULL" in L.cf_last_error()               # CF_E_NULL
    buf = (C.c_float * 64)()
    rc = L.cf_peak_topk(C.addressof(buf), 1, 1, 4, 4, 17, 0, None, None, None, None, None, None, 0, None)
    assert rc == -3                                                # CF_E_K: K > H*W like torch.topk
    with pytest.raises(RuntimeError):
        lib.check(rc)
    rc = L.cf_peak_topk(C.addressof(buf), 1, 1, 4, 4, 2, 0, None, None, None, None, None, None, 0, None)
    assert rc == -4                                                # CF_E_WORKSPACE
    with pytest.raises(ValueError):
        lib.check(rc)
    assert L.cf_peak_topk(None, 0, 1, 4, 4, 2, 0, None, None, None, None, None, None, 0, None) == 0  # empty batch


def test_ops_have_no_cpu_fallback():
    from centerfusiondetect3d_b200 import ops

    heat = torch.rand(1, 2, 4, 5)
    with pytest.raises(ValueError, match="CUDA"):
        ops.topk(heat, 3)
    with pytest.raises(ValueError, match="CUDA"):
        ops.nms(heat)
    with pytest.raises(ValueError, match="CUDA"):
        ops.getPcFrustumHeatmap({"heatmap": heat}, heat, heat, types.SimpleNamespace(K=1, max_pc_dist=60.0))
    assert ops.fusionDecode([{"not_a_heatmap": 1}]) == {}         # reference: returns {} (decode.py:25-27)
    with pytest.raises(AssertionError):
        ops.fusionDecode({"heatmap": heat})                        # reference: assert isinstance(list) (decode.py:23)
    for alias in ("_nms", "_topk", "_tranpose_and_gather_feat", "generate_pc_hm", "get_pc_hm", "fusion_decode"):
        assert callable(getattr(ops, alias))


def test_missing_library_fails_loudly(monkeypatch, lib):
    monkeypatch.setattr(lib, "_lib", None)
    monkeypatch.setattr(lib, "LIB_PATH", "/nonexistent/libcfd3d_b200.so")
    with pytest.raises(ImportError, match="no CPU fallback"):
        lib.lib()


def test_patch_swaps_and_restores_reference_names():
    """Mock modules (the real-tree version is tests/test_patch_reference.py): every swapped name is a dispatcher that
    sends CUDA fp32 no-grad tensors to the op and everything else to the original."""
    from centerfusiondetect3d_b200 import ops, patch

    def mk(**kw):
        m = types.ModuleType("m")
        for k, v in kw.items():
            setattr(m, k, v)
        return m

    o = {n: (lambda *a, _n=n, **k: _n) for n in ("getPcFrustumHeatmap", "cvtPcDepthToHeatmap", "nms", "topk",
                                                   "transposeAndGetFeature", "fusionDecode", "get_alpha")}
    mods = {"utils.pointcloud": mk(getPcFrustumHeatmap=o["getPcFrustumHeatmap"], cvtPcDepthToHeatmap=o["cvtPcDepthToHeatmap"],
                                   topk=o["topk"], transposeAndGetFeature=o["transposeAndGetFeature"], get_alpha=o["get_alpha"]),
            "model.utils": mk(nms=o["nms"], topk=o["topk"], transposeAndGetFeature=o["transposeAndGetFeature"]),
            "model.decode": mk(fusionDecode=o["fusionDecode"], nms=o["nms"], topk=o["topk"]),
            "model.losses": mk(transposeAndGetFeature=o["transposeAndGetFeature"], get_alpha=o["get_alpha"]),
            "detector": mk(fusionDecode=o["fusionDecode"])}
    undo = patch.apply(mods)
    assert mods["utils.pointcloud"].getPcFrustumHeatmap.__cfd3d_op__ is ops.getPcFrustumHeatmap
    assert mods["model.utils"].topk.__cfd3d_op__ is ops.topk and mods["model.decode"].nms.__cfd3d_op__ is ops.nms
    assert mods["detector"].fusionDecode.__cfd3d_op__ is ops.fusionDecode
    assert mods["model.utils"].topk is mods["model.decode"].topk       # one dispatcher per original
    # the training losses keep the reference's autograd-capable functions (model/losses.py:15,18)
    assert mods["model.losses"].transposeAndGetFeature is o["transposeAndGetFeature"]
    assert mods["model.losses"].get_alpha is o["get_alpha"]
    # NumPy, CPU tensors, non-fp32 and grad-requiring tensors keep the reference's own code
    assert mods["utils.pointcloud"].cvtPcDepthToHeatmap(np.zeros((3, 2, 2)), None, 0, None, 0, 60) == "cvtPcDepthToHeatmap"
    assert mods["model.utils"].topk(torch.rand(1, 1, 2, 2), K=2) == "topk"
    assert mods["utils.pointcloud"].get_alpha(rotation=torch.rand(2, 8)) == "get_alpha"
    assert not patch.takes_cuda_path(np.zeros(3)) and not patch.takes_cuda_path(torch.zeros(3))
    assert not patch.takes_cuda_path() and not patch.takes_cuda_path({"a": None})
    undo()
    assert mods["utils.pointcloud"].getPcFrustumHeatmap is o["getPcFrustumHeatmap"]
    assert mods["model.utils"].topk is o["topk"] and mods["detector"].fusionDecode is o["fusionDecode"]


def test_frame_shard_partitions():
    from centerfusiondetect3d_b200.shard import frame_shard

    for B in (0, 1, 7, 32, 64, 256):
        for world in (1, 2, 3, 4, 8):
            cuts = [frame_shard(B, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == B
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


def _gloo_worker(rank, world, port, B, tmp):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from centerfusiondetect3d_b200 import synthetic as syn
        from centerfusiondetect3d_b200.shard import gather_detections, shard_batch
        from helpers import DEC_KEYS, make_case
        from oracle import binding as ob

        wl = syn.WORKLOADS["cfg1"]
        inp = make_case(wl, B, seed=77)                      # identical on both ranks (seeded)
        mine = shard_batch({k: inp[k] for k in DEC_KEYS}, rank, world)
        # the oracle stands in for the per-rank compute (no GPU here): what is under test is shard + gather
        ret = ob.fusion_decode(mine, (wl.H, wl.W), wl.K)
        full = gather_detections({k: torch.from_numpy(v) for k, v in ret.items()})
        ref = ob.fusion_decode({k: inp[k] for k in DEC_KEYS}, (wl.H, wl.W), wl.K)
        ok = all(np.array_equal(full[k].numpy(), ref[k]) for k in ref)
        open(os.path.join(tmp, "ok%d" % rank), "w").write("1" if ok else "0")
    finally:
        dist.destroy_process_group()


def test_sharded_ranks_reproduce_the_full_batch_gloo(tmp_path):
    """world_size 2 over gloo: rank r decodes its frame shard; the gathered detections equal the unsharded decode
    (B = 3: uneven shards)."""
    import socket

    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_gloo_worker, args=(2, port, 3, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / ("ok%d" % r)).read() for r in range(2)] == ["1", "1"]
