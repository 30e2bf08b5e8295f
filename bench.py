#!/usr/bin/env python
"""bench.py -- fusion-stage frames/s on N B200s (BASELINE.json metric), with the roofline of the dominant kernel,
the end-to-end number through the public API from host buffers, and the host-CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg4|cfg5] [--impl native|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One "step" = one pass of the fusion stage (getPcFrustumHeatmap + fusionDecode, one fused C-ABI call; cfg4 adds the
radar pillar expansion) over one batch of B synthetic frames per GPU.  Frames are independent, so ranks own whole
frames and the data path has no collective ("scaling": "weak": B frames per GPU at every N).

Timing: CUDA events on the launching stream around exactly K steps, barrier + synchronize on both sides, max over
ranks.  Cold-cache discipline: the step rotates through NSETS distinct resident input sets whose fully-read maps
(heat-map + pc_dep) together exceed 2x the 126 MB L2, so no step finds its inputs in L2 ("l2": "rotating_inputs").
The timed region is ONE CUDA-graph launch whose nodes are the K steps (one fused kernel each; cfg4 adds the two pillar
kernels); nothing is allocated, launched from the host or synchronised inside it.

Both arms (--impl native / reference) print the same `config` dict; run-time details of the native arm go to `run`.
At N > 1 the line also carries per-rank step times (min / median / max), the NCCL gather of the detections for
reporting (shard.gather_detections: microseconds, bytes, checked on rank 0), per-rank H2D rates of the e2e leg and, as
`config3_strong`, BASELINE config 3's split (256 frames over the N GPUs) timed the same way.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "fusion_stage_frames_per_s"
UNIT = "frames/s"
L2_BYTES = 126 * 1024 * 1024


def algorithmic_bytes_per_frame(wl, with_pillars: bool) -> int:
    """heat read + pc_dep read + pc_hm write + gathers (14 fwd + 29 decode ch) + 33 decoded floats
    + calib; pillars add the pc_dep write and the fp64 points."""
    HW = wl.H * wl.W
    b = 4 * HW * (wl.C + 3 + 3) + 4 * wl.K * (14 + 29) + 4 * wl.K * 33 + 48
    if with_pillars:
        b += 4 * HW * 3 + 8 * 5 * wl.n_radar
    return b


# --------------------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------------------
def bind_to_gpu_numa(index: int) -> str:
    """Pins this process to the CPUs NVML reports as local to its GPU before any pinned buffer is allocated (first-touch
    puts the pages on that NUMA node), so that at N > 1 every rank's H2D traffic starts from its own socket.  Returns a
    note for the record; a host that reports one node for every GPU is left as it is."""
    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {w * 64 + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        allowed = set(os.sched_getaffinity(0))
        want = (cpus & allowed) or allowed
        if want != allowed:
            os.sched_setaffinity(0, want)
        return "gpu %d: %d local cpus of %d allowed%s" % (phys, len(cpus & allowed), len(allowed),
                                                          " (bound)" if want != allowed else " (all local: unchanged)")
    except Exception as e:  # pragma: no cover
        return "unavailable: %r" % (e,)


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.004):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self.period = period_s
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # pragma: no cover
            self.nv = None
            self.err = repr(e)

    _NAMES = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
              0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in self._NAMES.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def merge(self, other: "ClockSampler"):
        self.samples += other.samples
        self.reasons |= other.reasons

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------------------
# inputs
# --------------------------------------------------------------------------------------------------------
def make_host_inputs(wl, B: int, seed: int):
    """One batch of synthetic stage inputs (NumPy); pc_dep painted from the radar points by the oracle's pillar code
    (input generation, not the measured path)."""
    from centerfusiondetect3d_b200 import synthetic as syn
    from oracle import binding as ob

    gen_B = min(B, 8)  # generate 8 distinct frames, tile to B (values do not affect traffic; addresses do)
    inp = syn.make_inputs(wl, seed=seed, B=gen_B)
    ob.set_threads(min(ob.max_threads(), 16))
    pc_dep, _ = ob.radar_pillars(inp["radar_points"], inp["radar_counts"], inp["cam"], inp["calib64"], inp["A_out"],
                                 (wl.img_w, wl.img_h), (wl.H, wl.W), wl.pillar_dims, wl.max_pc_dist, want_owner=False)
    inp["pc_dep"] = pc_dep
    if gen_B < B:
        reps = (B + gen_B - 1) // gen_B
        for k, v in list(inp.items()):
            if isinstance(v, np.ndarray) and v.ndim >= 1 and v.shape[0] == gen_B:
                inp[k] = np.ascontiguousarray(np.concatenate([np.roll(v, r, axis=0) for r in range(reps)], axis=0)[:B])
    return inp


def config_dict(wl, B: int, world: int) -> dict:
    """What both arms measure (identical dict from --impl native and --impl reference)."""
    return {"workload": wl.name, "frames_per_gpu_per_step": B, "global_frames_per_step": B * world, "K": wl.K,
            "classes": wl.C, "map": [wl.H, wl.W], "radar_points": wl.n_radar,
            "stage": ("radar_pillars + " if wl.with_pillars else "") + "getPcFrustumHeatmap + fusionDecode"}


# the unmodified Python reference measured in the survey container (it cannot travel to the GPU box)
PY_REFERENCE_NOTE = ("the unmodified Python reference (getPcFrustumHeatmap alone, 8 torch threads, survey container, "
                     "BASELINE.md section 2) ran 11.3 frames/s at 112x200 / K=100 and 7.9 frames/s at 225x400 / K=500: "
                     "the C/OpenMP port timed here is ~500x faster than it, i.e. a conservative baseline")


STAGE_KEYS = ("heatmap", "reg", "widthHeight", "depth", "depth2", "rotation", "rotation2", "dimension",
              "amodal_offset", "nuscenes_att", "velocity")
PILLAR_KEYS = ("radar_points", "radar_counts", "cam", "calib64", "A_out")


class Cfg:
    def __init__(self, K, max_pc_dist):
        self.MODEL = type("M", (), {"K": K})()
        self.DATASET = type("D", (), {"MAX_PC_DIST": max_pc_dist})()


def run_stage(ops, wl, dev_in: dict, cfg):
    """The measured step through the public API.  Returns (pc_hm, ret, launches)."""
    n = 0
    if wl.with_pillars:
        pc_dep = ops.radar_to_pc_dep(dev_in["radar_points"], dev_in["radar_counts"], dev_in["cam"], dev_in["calib64"],
                                     dev_in["A_out"], (wl.H, wl.W), (wl.img_w, wl.img_h), wl.pillar_dims,
                                     wl.max_pc_dist)
        n += ops.last_launch_count()
    else:
        pc_dep = dev_in["pc_dep"]
    heads = {k: dev_in[k] for k in STAGE_KEYS}  # fresh dict: fusionDecode renames rotation2 -> rotation in it
    pc_hm, ret = ops.fusion_stage(heads, pc_dep, dev_in["calib"], cfg, pc_vel=dev_in.get("pc_vel"))
    n += ops.last_launch_count()
    return pc_hm, ret, n


# --------------------------------------------------------------------------------------------------------
# CPU baseline (oracle port of the reference's algorithm; the reference itself is Python and cannot travel)
# --------------------------------------------------------------------------------------------------------
def cpu_stage(ob, wl, inp):
    if wl.with_pillars:
        pc_dep, _ = ob.radar_pillars(inp["radar_points"], inp["radar_counts"], inp["cam"], inp["calib64"],
                                     inp["A_out"], (wl.img_w, wl.img_h), (wl.H, wl.W), wl.pillar_dims, wl.max_pc_dist,
                                     want_owner=False)
    else:
        pc_dep = inp["pc_dep"]
    ob.frustum_pc_hm(inp["heatmap"], inp["widthHeight"], inp["depth"], inp["dimension"], inp["rotation"], pc_dep,
                     inp["calib"], wl.K, wl.max_pc_dist)
    ob.fusion_decode({k: inp[k] for k in STAGE_KEYS}, (wl.H, wl.W), wl.K)


def host_threads() -> int:
    """Host threads the CPU legs use: every core this process may run on.  (Not omp_get_max_threads(): torchrun exports
    OMP_NUM_THREADS=1 to its workers, which would time the reference arm on one thread at N > 1; the oracle's parallel
    regions carry an explicit num_threads clause, so the environment variable does not bind them.)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        return max(1, os.cpu_count() or 1)


def cpu_baseline(wl, inp, B, budget_s=12.0, min_reps=2):
    from oracle import binding as ob

    threads = host_threads()
    ob.set_threads(threads)
    cpu_stage(ob, wl, inp)  # warm-up
    t0 = time.perf_counter()
    reps = 0
    while reps < min_reps or (time.perf_counter() - t0) < budget_s:
        cpu_stage(ob, wl, inp)
        reps += 1
        if reps >= 10000:
            break
    dt = time.perf_counter() - t0
    return {"value": B * reps / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "%d passes over one %d-frame batch of %s (oracle/cfd3d_oracle.c, OpenMP over frames x classes, "
                      "%.1f s); %s" % (reps, B, wl.name, dt, PY_REFERENCE_NOTE)}


def run_reference_arm(args, wl, B):
    """--impl reference: the reference's CPU algorithm (oracle port) on the host cores, same config / metric."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import binding as ob

    ob.build()
    inp = make_host_inputs(wl, B, seed=1234)
    threads = host_threads()
    ob.set_threads(threads)
    for _ in range(max(1, min(args.warmup, 3))):
        cpu_stage(ob, wl, inp)
    steps = args.steps
    # bounded: never let the reference arm run longer than ~3 minutes whatever --steps says
    t_probe = time.perf_counter()
    cpu_stage(ob, wl, inp)
    per = time.perf_counter() - t_probe
    steps_eff = max(1, min(steps, int(150.0 / max(per, 1e-6))))
    t0 = time.perf_counter()
    for _ in range(steps_eff):
        cpu_stage(ob, wl, inp)
    dt = time.perf_counter() - t0
    val = B * steps_eff / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps_eff,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / steps_eff, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(wl, B, int(os.environ.get("WORLD_SIZE", "1"))),
        "note": "reference algorithm on host CPU cores (C/OpenMP port in oracle/; the Python reference cannot travel to "
                "the GPU box); rank 0 only, one batch of frames_per_gpu_per_step frames per step",
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "%d passes over one %d-frame batch; %s" % (steps_eff, B, PY_REFERENCE_NOTE)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(line)
    return 0


# --------------------------------------------------------------------------------------------------------
# native arm
# --------------------------------------------------------------------------------------------------------
_JSON_FD = None


def own_stdout() -> None:
    """stdout carries exactly ONE line, the JSON result: anything a library prints there (NCCL's version banner under
    NCCL_DEBUG, for one) is sent to stderr instead -- file descriptor 1 is pointed at stderr for the rest of the run and
    the result goes out through a private duplicate of the original."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit_line(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main() -> int:
    own_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--in-flight", type=int, default=1,
                    help="independent batches in flight in the timed region (parallel branches of its graph)")
    ap.add_argument("--no-pipelined", action="store_true", help="skip the two-batches-in-flight leg")
    ap.add_argument("--batch", type=int, default=0, help="frames per GPU per step (default: the workload's)")
    ap.add_argument("--e2e-steps", type=int, default=0)
    ap.add_argument("--e2e-copy-all", action="store_true", help="e2e leg: copy every head to the device instead of reading the gather-only heads in place")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-graph", action="store_true", help="launch through ctypes every step instead of graph replay")
    ap.add_argument("--no-cycle-graph", action="store_true", help="one graph launch per step instead of one per cycle of input sets")
    args = ap.parse_args()

    from centerfusiondetect3d_b200 import synthetic as syn

    wl = syn.WORKLOADS[args.workload]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # frames per GPU per step: cfg2 = 32; cfg3 = 256 total on one GPU at N=1, 32/GPU at N=8; cfg5 = 8/GPU; cfg4 = 64
    B = args.batch or {"cfg2": 32, "cfg3": 256 // max(world, 1) if world > 1 else 256, "cfg4": 64, "cfg5": 8}[args.workload]

    if args.impl == "reference":
        return run_reference_arm(args, wl, B)

    import torch
    import torch.distributed as dist

    assert torch.cuda.is_available(), "bench.py (native arm) needs a CUDA device: there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # (N = 1 keeps the whole affinity mask: the cpu_baseline leg of this process uses every host core)
    affinity_note = bind_to_gpu_numa(local) if world > 1 else "not bound (single rank)"
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from centerfusiondetect3d_b200 import _lib, ops

    cfg = Cfg(wl.K, wl.max_pc_dist)
    host = make_host_inputs(wl, B, seed=1234 + rank)
    keys = STAGE_KEYS + ("calib", "pc_dep") + (PILLAR_KEYS if wl.with_pillars else ())
    fully_read = 4 * B * wl.H * wl.W * (wl.C + 3)
    nsets = max(2, -(-2 * L2_BYTES // fully_read) + 1)
    nsets = min(nsets, 24)
    sets = []
    for j in range(nsets):
        sets.append({k: torch.from_numpy(np.ascontiguousarray(np.roll(host[k], j, axis=0))).to(dev) for k in keys})
    torch.cuda.synchronize()

    # -- correctness guard: the measured call agrees with the oracle on the bench's own inputs, ALL B frames of input set 0,
    #    under the very execution plan the timed loop runs (same shapes, same batch) ------------------------------------
    plan = None
    if rank == 0 and not os.environ.get("CFD3D_STOP_AFTER"):
        from oracle import binding as ob

        hm, ret, _ = run_stage(ops, wl, sets[0], cfg)
        plan = ops.last_plan()
        ob.set_threads(host_threads())
        hm_o = ob.frustum_pc_hm(host["heatmap"], host["widthHeight"], host["depth"], host["dimension"], host["rotation"],
                                host["pc_dep"], host["calib"], wl.K, wl.max_pc_dist)[0]
        ret_o = ob.fusion_decode({k: host[k] for k in STAGE_KEYS}, (wl.H, wl.W), wl.K)
        torch.cuda.synchronize()
        check_ok = bool(np.array_equal(hm.cpu().numpy().view(np.uint32), hm_o.view(np.uint32))) and all(
            np.array_equal(v.cpu().numpy().view(np.uint32), ret_o[k].view(np.uint32)) for k, v in ret.items())
        check_frames = B
    else:
        check_ok, check_frames = True, 0

    # -- capture one CUDA graph per input set -----------------------------------------------------------------
    launches_per_step = 0
    stream = torch.cuda.Stream(device=dev)
    graphs, outs = [], []
    with torch.cuda.stream(stream):
        for s in sets:  # eager warm-up (workspace allocation, kernel attribute setup)
            _, _, launches_per_step = run_stage(ops, wl, s, cfg)
        stream.synchronize()
        if not args.no_graph:
            for s in sets:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    o = run_stage(ops, wl, s, cfg)
                graphs.append(g)
                outs.append(o)

    # ... and one graph of a whole cycle through the input sets: the steps of a cycle are consecutive kernel nodes of ONE
    # graph launch (a launch-bound inner loop belongs in a graph); the timed loop replays cycles, then single steps for
    # the remainder
    cycle = None
    if graphs and not args.no_cycle_graph:
        with torch.cuda.stream(stream):
            cycle = torch.cuda.CUDAGraph()
            with torch.cuda.graph(cycle, stream=stream):
                cycle_outs = [run_stage(ops, wl, s, cfg) for s in sets]

    # The timed region itself: ONE graph whose nodes are the K steps (set i % nsets at step i), so that no host-side
    # launch sits inside a sub-millisecond window (at N > 1 the ranks' Python processes contend for the host).  The
    # steps' outputs are dropped as they are captured: the graph's private pool reuses their memory step after step.
    side_streams = {}

    def capture_timed(S):
        """The K steps as ONE graph.  S > 1: S independent batches in flight -- step i runs on branch i % S of the graph
        (its own stream, hence its own workspace); a branch's steps stay ordered, the branches only share the GPU."""
        if S not in side_streams:
            side_streams[S] = [torch.cuda.Stream(device=dev) for _ in range(S)] if S > 1 else []
            for st_ in side_streams[S]:  # (the ops keep one workspace per stream: allocated here, outside the capture)
                with torch.cuda.stream(st_):
                    run_stage(ops, wl, sets[0], cfg)
                st_.synchronize()
        side = side_streams[S]
        with torch.cuda.stream(stream):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                if S == 1:
                    for i in range(args.steps):
                        run_stage(ops, wl, sets[i % nsets], cfg)
                else:
                    fork = torch.cuda.Event()
                    fork.record(stream)
                    for st_ in side:
                        st_.wait_event(fork)
                    for i in range(args.steps):
                        with torch.cuda.stream(side[i % S]):
                            run_stage(ops, wl, sets[i % nsets], cfg)
                    for st_ in side:
                        join = torch.cuda.Event()
                        join.record(st_)
                        stream.wait_event(join)
        return g

    timed = None
    in_flight = max(1, args.in_flight)
    if graphs and not args.no_cycle_graph and args.steps <= 8192:
        timed = capture_timed(in_flight)

    def step(i):
        if graphs:
            graphs[i % nsets].replay()
        else:
            run_stage(ops, wl, sets[i % nsets], cfg)

    def run_steps(n):  # steps 0 .. n-1, set i % nsets at step i
        i = 0
        if cycle is not None:
            while i + nsets <= n:
                cycle.replay()
                i += nsets
        while i < n:
            step(i)
            i += 1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        run_steps(max(args.warmup, 3))
        if timed is not None:
            timed.replay()  # (untimed: the graph's first launch uploads it)
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clocks = ClockSampler(local)
        with clocks:
            ev0.record(stream)
            if timed is not None:
                timed.replay()
            else:
                run_steps(args.steps)
            ev1.record(stream)
            barrier()
        ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    per_rank_ms = [ms]
    if world > 1:
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        per_rank_ms = [float(x.item()) for x in allt]
    ms_max = max(per_rank_ms)
    frames_total = B * world * args.steps
    value = frames_total / (ms_max * 1e-3)

    # -- pipelined leg: the same K steps with TWO independent batches in flight (a serving loop that double-buffers its
    #    batches).  The fused kernel is one cluster per frame and latency-bound per frame: a second launch fills the SMs
    #    the first leaves idle (its last round, its stragglers, the launch gap).  Reported beside `value`, never as it.
    pipelined = None
    if timed is not None and in_flight == 1 and not args.no_pipelined:
        g2 = capture_timed(2)
        with torch.cuda.stream(stream):
            g2.replay()
            barrier()
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record(stream)
            g2.replay()
            p1.record(stream)
            barrier()
        pt = torch.tensor([p0.elapsed_time(p1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(pt, op=dist.ReduceOp.MAX)
        pms = float(pt.item())
        pipelined = {"in_flight": 2, "ms_per_step": pms / args.steps, "value": frames_total / (pms * 1e-3),
                     "how": "step i on branch i % 2 of one graph: two streams, two workspaces, each branch ordered"}
        del g2

    # -- roofline leg: every kernel of the step timed alone with CUDA events on the launching stream (the ABI's
    #    measurement hook records the events right around the chosen kernel), inputs rotating through > 2 x L2 --------
    roof = None
    if rank == 0:
        n_roof = min(args.steps, 100)
        L = _lib.lib()
        kernel_ids = {"fusion_stage": 6}  # CF_KERNEL_STAGE
        kernel_us = {}
        for kname, kid in kernel_ids.items():
            pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_roof)]
            with torch.cuda.stream(stream):
                for i, (a, b) in enumerate(pairs):
                    # L2 holds the previous input sets, not this one (rotating inputs > 2 x L2); no write-flush: dirty
                    # lines would have to be written back under the kernel's reads
                    a.record(stream)  # events must exist before their handles are read
                    b.record(stream)
                    L.cf_set_kernel_events(kid, a.cuda_event, b.cuda_event)
                    if wl.with_pillars:  # the hook is per ABI call: keep the pillar call out of it
                        L.cf_set_kernel_events(-1, None, None)
                        pcd = ops.radar_to_pc_dep(*[sets[i % nsets][k] for k in PILLAR_KEYS], (wl.H, wl.W),
                                                  (wl.img_w, wl.img_h), wl.pillar_dims, wl.max_pc_dist)
                        L.cf_set_kernel_events(kid, a.cuda_event, b.cuda_event)
                        ops.fusion_stage({k: sets[i % nsets][k] for k in STAGE_KEYS}, pcd, sets[i % nsets]["calib"], cfg)
                    else:
                        run_stage(ops, wl, sets[i % nsets], cfg)
                    L.cf_set_kernel_events(-1, None, None)
                stream.synchronize()
            d = np.array([a.elapsed_time(b) for a, b in pairs]) * 1e3
            kernel_us[kname] = {"mean": float(d.mean()), "min": float(d.min()), "median": float(np.median(d))}
        dom = "fusion_stage"
        dur = kernel_us[dom]["mean"] * 1e-6
        timed_how = "CUDA events the ABI records around the kernel (cf_set_kernel_events), one launch at a time"
        if launches_per_step == 1:
            # the step IS this one kernel: its average duration over the timed region (graph replay, events on the
            # launching stream) is the step time itself; the event-pair figure above includes ~5 us of event overhead
            dur = ms_max * 1e-3 / args.steps
            timed_how = "the timed region itself: one kernel per step, CUDA events around one graph launch of %d steps" % args.steps
        peaks = {}
        pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak, peak_src = 6650.0, "fallback"
        if os.path.isfile(pk_path):
            try:
                peaks = json.load(open(pk_path))
                peak, peak_src = float(peaks["hbm_gbs"]), "measured"
            except Exception:
                pass
        # the fused kernel is the whole stage (SURVEY.md 8d bytes per frame, pillar bytes excluded: separate kernels)
        alg = algorithmic_bytes_per_frame(wl, False) * B
        ach = alg / dur / 1e9
        stage_alg = algorithmic_bytes_per_frame(wl, wl.with_pillars) * B
        stage_ach = stage_alg / (ms_max * 1e-3 / args.steps) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.isfile(tpath):
            try:
                traffic = json.load(open(tpath)).get(wl.name, {}).get("dominant_kernel_dram_bytes_per_launch")
            except Exception:
                traffic = None
        roof = {"bound": "hbm", "kernel": dom + "_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg, "kernel_us": dur * 1e6,
                "launches_timed": args.steps if launches_per_step == 1 else n_roof, "timed": timed_how,
                "l2": "rotating input sets larger than 2 x L2",
                "kernels_us": kernel_us,
                "stage": {"algorithmic_bytes_per_step": stage_alg, "achieved": stage_ach, "frac": stage_ach / peak}}

    # -- e2e: public API from pinned host buffers, H2D + compute + D2H inside the timed region --------------------
    e2e_steps = args.e2e_steps or max(10, min(args.steps, 100))
    pinned = {k: torch.from_numpy(np.ascontiguousarray(host[k])).pin_memory() for k in keys if not (
        wl.with_pillars and k == "pc_dep")}
    # The heat-map, pc_dep (or the radar points) and calib are read in full: copied to the device every step.  The other
    # heads are only gathered at the K peak pixels: ops accepts them as pinned host tensors and the kernel reads those
    # sectors in place over PCIe (--e2e-copy-all: copy them too, the way a caller without that option would).
    zero_keys = [] if args.e2e_copy_all else [k for k in pinned if k in STAGE_KEYS and k != "heatmap"]
    e_keys = [k for k in pinned if k not in zero_keys]
    h2d_copied = sum(pinned[k].numel() * pinned[k].element_size() for k in e_keys)
    # gathered in place: one 32-byte sector per (detection, head channel); the first-stage heads serve the frustum job
    # (14 channels) and, where fusionDecode reads the same head, the decode job as well
    gather_ch = sum(pinned[k].shape[1] for k in zero_keys) + (14 if zero_keys else 0)
    # pc_dep: the depth plane is read in full (copied), the two velocity planes only at the matched pixels (in place)
    split_dep = bool(zero_keys) and "pc_dep" in pinned
    if split_dep:
        h2d_copied -= pinned["pc_dep"].numel() * 4 * 2 // 3
        gather_ch += 2
    h2d = h2d_copied + B * wl.K * gather_ch * 32
    NBUF = 3
    dbuf = [{**{k: torch.zeros_like(pinned[k], device=dev) for k in e_keys}, **{k: pinned[k] for k in zero_keys}}
            for _ in range(NBUF)]
    if split_dep:
        for d_ in dbuf:
            d_["pc_vel"] = pinned["pc_dep"]

    def copy_in(dst, k):
        if split_dep and k == "pc_dep":
            ops.stage_depth_plane(dst[k], pinned[k])  # one strided copy of the depth planes
        else:
            dst[k].copy_(pinned[k], non_blocking=True)

    for k in e_keys:
        copy_in(dbuf[0], k)
    hm0, ret0, _ = run_stage(ops, wl, dbuf[0], cfg)
    res_names = ["pc_hm"] + list(ret0.keys())
    hout = [{n: torch.empty_like(tns, device="cpu").pin_memory() for n, tns in zip(res_names, [hm0] + list(ret0.values()))}
            for _ in range(NBUF)]
    d2h = sum(v.numel() * v.element_size() for v in hout[0].values())
    s_in, s_cmp, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    ev_in = [torch.cuda.Event() for _ in range(NBUF)]
    ev_cmp = [torch.cuda.Event() for _ in range(NBUF)]
    ev_out = [torch.cuda.Event() for _ in range(NBUF)]
    results = [None] * NBUF

    def e2e_step(i):
        j = i % NBUF
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev_cmp[j])  # buffer j free again once its previous compute finished
            for k in e_keys:
                copy_in(dbuf[j], k)
            ev_in[j].record(s_in)
        with torch.cuda.stream(s_cmp):
            s_cmp.wait_event(ev_in[j])
            s_cmp.wait_event(ev_out[j])
            hm, ret, _ = run_stage(ops, wl, dbuf[j], cfg)
            results[j] = (hm, ret)
            ev_cmp[j].record(s_cmp)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_cmp[j])
            hout[j]["pc_hm"].copy_(hm, non_blocking=True)
            for n, v in ret.items():
                hout[j][n].copy_(v, non_blocking=True)
            ev_out[j].record(s_out)

    for i in range(3):
        e2e_step(i)
    barrier()
    # the link's own rate on this host, measured the same minute: the heat-map buffer copied pinned -> device, alone
    # (every rank at the same time: at N > 1 they share the host's memory and PCIe root)
    big_k = max(e_keys, key=lambda k: pinned[k].numel())
    n_bw = 5
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s_in):
        dbuf[0][big_k].copy_(pinned[big_k], non_blocking=True)
        b0.record(s_in)
        for _ in range(n_bw):
            dbuf[0][big_k].copy_(pinned[big_k], non_blocking=True)
        b1.record(s_in)
    barrier()
    h2d_link_gbs = n_bw * pinned[big_k].numel() * pinned[big_k].element_size() / (b0.elapsed_time(b1) * 1e-3) / 1e9
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s_in)
    for i in range(e2e_steps):
        e2e_step(i)
    s_out.wait_stream(s_cmp)
    s_out.wait_stream(s_in)
    e1.record(s_out)
    barrier()
    e2e_ms = max(e0.elapsed_time(e1), 0.0)
    wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max(e2e_ms, 0.0)
    t = torch.tensor([e2e_ms, wall_ms], dtype=torch.float64, device=dev)
    e2e_rank_ms = [e2e_ms]
    if world > 1:
        alle = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(alle, t)
        e2e_rank_ms = [float(x[0].item()) for x in alle]
        t = torch.stack(alle).max(dim=0).values
    e2e_ms_max = float(t[0].item())
    e2e_val = B * world * e2e_steps / (e2e_ms_max * 1e-3)
    # achieved host -> device rate of every rank over its own e2e region (copied bytes only): at N > 1 the ranks share
    # the host's memory and PCIe root, which is what bounds this leg
    h2d_gbs_per_rank = [h2d_copied * e2e_steps / (m * 1e-3) / 1e9 for m in e2e_rank_ms]
    link = torch.tensor([h2d_link_gbs], dtype=torch.float64, device=dev)
    link_all = [h2d_link_gbs]
    if world > 1:
        ll = [torch.zeros_like(link) for _ in range(world)]
        dist.all_gather(ll, link)
        link_all = [float(x.item()) for x in ll]
    # time the step's host -> device bytes (copies + in-place gathers) would take on the link alone: the floor of this leg
    e2e_link_floor_ms = h2d / (min(link_all) * 1e9) * 1e3

    # -- reporting only: the detections of all ranks gathered over NCCL (shard.gather_detections replaces the reference's
    #    all_gather of the whole head-output dict, model/progressBar.py:85-91); untimed by the bench metric ------------
    gather_info = None
    if world > 1:
        from centerfusiondetect3d_b200.shard import gather_detections

        _, ret_g, _ = run_stage(ops, wl, sets[0], cfg)
        ret_g = {k: v.clone() for k, v in ret_g.items()}
        gather_detections(ret_g)  # warm-up (communicator set-up)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        full = gather_detections(ret_g)
        g1.record()
        torch.cuda.synchronize()
        g_us = torch.tensor([g0.elapsed_time(g1) * 1e3], dtype=torch.float64, device=dev)
        dist.all_reduce(g_us, op=dist.ReduceOp.MAX)
        if rank == 0:
            ok = all(tuple(v.shape[1:]) == tuple(ret_g[k].shape[1:]) and v.shape[0] == B * world for k, v in full.items())
            ok = ok and all(torch.equal(full[k][:B], ret_g[k]) for k in ret_g)  # rank 0's shard leads the gathered tensor
            # every rank decoded a rolled copy of a seeded batch: scores of every gathered frame are sorted and finite
            sc = full["scores"]
            ok = ok and bool(torch.isfinite(sc).all()) and bool((sc[:, :-1] >= sc[:, 1:]).all())
            gather_info = {"us": float(g_us.item()), "bytes_per_rank": int(sum(v.numel() * v.element_size() for v in ret_g.values())),
                           "gathered_frames": int(sc.shape[0]), "tensors": len(full), "checked": bool(ok),
                           "how": "shard.gather_detections: all_gather_into_tensor of the (B_local, K, .) detection tensors "
                                  "over NCCL, after the timed region (reporting only)"}

    # -- BASELINE config 3 as written (256 frames split over the N GPUs: strong scaling), timed like the main region --
    strong = None
    if world > 1 and wl.name.startswith("cfg2") and not os.environ.get("CFD3D_NO_STRONG"):
        Bs = max(1, 256 // world)
        reps = -(-Bs // B)
        big = {k: torch.cat([sets[r % nsets][k] for r in range(reps)], 0)[:Bs].contiguous() for k in keys}
        bigs = [big, {k: torch.roll(v, 1, 0) for k, v in big.items()}]
        with torch.cuda.stream(stream):
            for sset in bigs:
                run_stage(ops, wl, sset, cfg)
            stream.synchronize()
            n_st = 20
            gs = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gs, stream=stream):
                for i in range(n_st):
                    run_stage(ops, wl, bigs[i % 2], cfg)
            gs.replay()
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record(stream)
            gs.replay()
            s1.record(stream)
            barrier()
        st = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
        strong = {"workload": "cfg3: 256 frames split over the GPUs", "frames_per_gpu_per_step": Bs, "steps": n_st,
                  "ms_per_step": float(st.item()) / n_st, "value": Bs * world * n_st / (float(st.item()) * 1e-3), "unit": UNIT,
                  "scaling": "strong"}

    if rank == 0:
        cpu = None if args.no_cpu or world > 1 else cpu_baseline(wl, host, B)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(wl, B, world),
            "run": {"l2": "rotating_inputs", "input_sets": nsets, "cuda_graph": not args.no_graph,
                    "graph_steps_per_launch": (args.steps if timed is not None else (nsets if cycle is not None else 1)),
                    "parallelism": "frames sharded across ranks, no data-path collective", "plan": plan,
                    "cpu_affinity": affinity_note},
            "per_rank_ms_per_step": {"min": min(per_rank_ms) / args.steps, "median": float(np.median(per_rank_ms)) / args.steps,
                                     "max": max(per_rank_ms) / args.steps, "all": [m / args.steps for m in per_rank_ms]},
            "clocks": clocks.summary(),
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": e2e_ms_max / e2e_steps, "wall_ms_per_step": float(t[1].item()) / e2e_steps,
                    "h2d_copied_bytes_per_step": h2d_copied, "heads_read_in_place": zero_keys,
                    "per_rank_ms_per_step": [m / e2e_steps for m in e2e_rank_ms], "h2d_gbs_per_rank": h2d_gbs_per_rank,
                    "h2d_link_gbs_per_rank": link_all, "link_floor_ms_per_step": e2e_link_floor_ms,
                    "frac_of_link_floor": e2e_link_floor_ms / (e2e_ms_max / e2e_steps),
                    "how": "ops.fusion_stage from pinned host buffers; 3-deep pipeline over copy-in / compute / "
                           "copy-out streams; every step copies the fully-read inputs (heat-map, depth plane of pc_dep / radar "
                           "points, calib) H2D and pc_hm + detections D2H; the gather-only heads and the velocity planes "
                           "of pc_dep stay in pinned host memory and the kernel reads their K peak / matched pixels in "
                           "place over PCIe (counted at 32 B per detection and channel)" if zero_keys else
                           "ops.fusion_stage from pinned host buffers; 3-deep pipeline; every step copies ALL inputs "
                           "H2D and pc_hm + detections D2H"},
            "gpu_launches": launches_per_step * args.steps,
            "launches_per_step": launches_per_step,
            "roofline": roof,
            "cpu_baseline": cpu,
            "parity_check": check_ok,
            "parity_check_frames": check_frames,
            "detection_gather": gather_info,
            "config3_strong": strong,
            "pipelined": pipelined,
        }
        if pipelined is not None and roof is not None:
            stage_alg_step = roof["stage"]["algorithmic_bytes_per_step"]
            pipelined["frac_of_hbm_peak"] = stage_alg_step / (pipelined["ms_per_step"] * 1e-3) / 1e9 / roof["peak"]
        if in_flight > 1:
            line["run"]["batches_in_flight"] = in_flight
        emit_line(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0 if check_ok else 1


if __name__ == "__main__":
    sys.exit(main())
