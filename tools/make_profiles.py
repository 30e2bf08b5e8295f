"""Turns one GPU call's scratch files (tools/gpu_profiles.sh -> gpurun_out/<tag>_*) into the tracked evidence under
profiles/<round>_*:   python tools/make_profiles.py <tag> <round>      e.g.  r2 r2"""
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, rnd = sys.argv[1], sys.argv[2]
G = lambda n: os.path.join(ROOT, "gpurun_out", "%s_%s" % (tag, n))  # noqa: E731
P = lambda n: os.path.join(ROOT, "profiles", "%s_%s" % (rnd, n))  # noqa: E731
ALG = {"cfg2": 1464048 * 32, "cfg3": 1464048 * 256, "cfg4": 1464048 * 64, "cfg5": 5912048 * 8}


def line(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


tpath = os.path.join(ROOT, "profiles", "traffic.json")
traffic = json.load(open(tpath)) if os.path.isfile(tpath) else {}
for w in ("cfg2", "cfg4", "cfg5"):
    rep = G("stage_full_%s.ncu-rep" % w)
    if not os.path.isfile(rep):
        continue
    body = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep], capture_output=True,
                          text=True, check=True).stdout
    step = line(G("bench_%s.json" % w))
    us = step["ms_per_step"] * 1e3
    dur = float(re.search(r"gpu__time_duration.sum\s+([\d.]+) us", body).group(1))

    def mbytes(name):
        m = re.search(name + r"\s+([\d.]+) (\w+)", body)
        return float(m.group(1)) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[m.group(2)]

    rd, wr = mbytes("dram__bytes_read.sum"), mbytes("dram__bytes_write.sum")
    tot = int(rd + wr)
    plan = step["run"]["plan"]
    hdr = ("ncu --set full --clock-control none --import-source on -k regex:fusion_stage -s 6 -c 1  (python bench.py --workload %s "
           "--steps 5 --warmup 3 --no-cpu --no-graph --e2e-steps 2)\n%s: %d frames, %dx%d, K=%d; plan: %s\n"
           "numbers under a profiler are not bench values (ncu serialises launches and the instruction cache starts cold): %.1f us "
           "here, %.1f us per step in the bench's graph replay (%s_bench_%s.json)\n\n"
           % (w, w, step["config"]["frames_per_gpu_per_step"], step["config"]["map"][0], step["config"]["map"][1],
              step["config"]["K"], json.dumps(plan), dur, us, rnd, w))
    tail = ("\nDRAM traffic of the stage kernel per launch = %d B (read %.2f MB + write %.2f MB; pc_hm and the detections "
            "mostly sit in L2 write-back at capture time)\nalgorithmic bytes per launch (SURVEY.md 8d, stage only) = %d B  -> "
            "DRAM traffic / algorithmic = %.3f\n" % (tot, rd / 1e6, wr / 1e6, ALG[w], tot / ALG[w]))
    open(P("stage_full_summary_%s.txt" % w), "w").write(hdr + body + tail)
    traffic[step["config"]["workload"]] = {"dominant_kernel": "fusion_stage_kernel",
                                           "dominant_kernel_dram_bytes_per_launch": tot,
                                           "source": "profiles/%s_stage_full_summary_%s.txt" % (rnd, w)}
json.dump(traffic, open(tpath, "w"), indent=1)
for w in ("cfg2", "cfg4"):
    if os.path.isfile(G("launches_%s.csv" % w)):
        shutil.copy(G("launches_%s.csv" % w), P("launches_%s.csv" % w))
for w in ("cfg2", "cfg3", "cfg4", "cfg5"):
    shutil.copy(G("bench_%s.json" % w), P("bench_%s.json" % w))
for n, dst in (("bench_default.json", "bench_cfg2_default_e2e.json"), ("bench_driver.json", "bench_cfg2_driver_invocation.json"),
               ("bench_reference.json", "bench_reference_arm.json"), ("gather_floor.txt", "gather_floor.txt"),
               ("plan_sweep.txt", "plan_sweep.txt")):
    if os.path.isfile(G(n)):
        shutil.copy(G(n), P(dst))


def trim(x):
    return "\n".join(l[:200] for l in x.splitlines() if "per CTA" not in l)


warm = open(G("phases.txt")).read().split("rep 5")[1]
cold = open(G("phases_cold.txt")).read().split("rep 5")[1]
open(P("phase_clocks_cfg2.txt"), "w").write(
    """In-kernel phase clocks of the fused stage kernel (library built with -DCFD3D_PROFILE, clock64 per CTA at the phase
boundaries; tools/prof_phases.py cfg2 32 4).  cfg2, 32 clusters of 4 CTAs, one frame each.  Microseconds at 1965 MHz;
min / mean / max over the 128 CTAs.  "sync3" is the cluster barrier round 1 had between the ranking and the tail (gone).

A. like the bench: back-to-back launches over rotating input sets (data cold in L2, the kernel's code warm in L2)
rep 5""" + trim(warm) + """

B. COLD=1: an L2-evicting read kernel between the launches (evicts the stage kernel's instructions as well: every
   once-per-frame phase gets slower -- what a stage call in the middle of a network sees)
rep 5""" + trim(cold) + "\n")
sass = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_summary.py")], capture_output=True, text=True, check=True).stdout
open(P("sass_summary.txt"), "w").write(sass)
print("profiles written for", rnd)
