"""Turns one GPU cycle's scratch files (gpurun_out/<tag>_*) into the tracked evidence under profiles/<round>_*:
python tools/make_profiles.py <tag> <round>   e.g.  r1m r1d"""
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

body = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), G("stage_full.ncu-rep")],
                      capture_output=True, text=True, check=True).stdout
step = json.loads(open(G("bench_cfg2.json")).read().strip().splitlines()[-1])
us = step["ms_per_step"] * 1e3
dur = float(re.search(r"gpu__time_duration.sum\s+([\d.]+) us", body).group(1))
rd = float(re.search(r"dram__bytes_read.sum\s+([\d.]+) Mbyte", body).group(1))
wr = float(re.search(r"dram__bytes_write.sum\s+([\d.]+) Kbyte", body).group(1))
tot = int(rd * 1e6 + wr * 1e3)
alg = 1464048 * 32
hdr = """ncu --set full --clock-control none --import-source on -k regex:fusion_stage -s 6 -c 2  (python bench.py --steps 5 --warmup 3 --no-cpu --no-graph --e2e-steps 2)
cfg2: B=32, 112x200, C=10, K=100; plan: cluster size 4, 32 clusters (128 CTAs x 512 threads), heat-map share resident in shared memory
numbers under a profiler are not bench values (ncu serialises launches and the instruction cache starts cold: %.1f us here,
%.1f us per step in the bench's graph replay); the launch list (%s_launches_cfg2.csv) has the serialised per-launch times

""" % (dur, us, rnd)
tail = """
DRAM traffic per launch = %d B (read %.2f Mbyte + write %.2f Kbyte; the 8.6 MB of pc_hm + detections stay in L2 write-back at capture time)
algorithmic bytes per launch (SURVEY.md 8d: 1 464 048 B/frame x 32) = %d B  -> DRAM read / algorithmic = %.3f: the heat-map is read from HBM once
shared-memory bank conflicts: the pass-B rescan reads one float4 per 512 bytes (a group = every 32nd float4 of a chunk), 16 lanes on the same banks; ~130 such reads per CTA, not on any critical path
""" % (tot, rd, wr, alg, rd * 1e6 / alg)
open(P("stage_full_summary.txt"), "w").write(hdr + body + tail)
shutil.copy(G("launches_cfg2.csv"), P("launches_cfg2.csv"))
for w in ("cfg2", "cfg3", "cfg4", "cfg5"):
    shutil.copy(G("bench_%s.json" % w), P("bench_%s.json" % w))
if os.path.isfile(G("bench_default.json")):
    shutil.copy(G("bench_default.json"), P("bench_cfg2_default_e2e.json"))
tpath = os.path.join(ROOT, "profiles", "traffic.json")
t = json.load(open(tpath))
t[step["config"]["workload"]] = {"dominant_kernel": "fusion_stage_kernel", "dominant_kernel_dram_bytes_per_launch": tot,
                                 "source": "profiles/%s_stage_full_summary.txt" % rnd}
json.dump(t, open(tpath, "w"), indent=1)


def trim(x):
    return "\n".join(l[:200] for l in x.splitlines() if "per CTA" not in l)


warm = open(G("phases.txt")).read().split("rep 5")[1]
cold = open(G("phases_cold.txt")).read().split("rep 5")[1]
open(P("phase_clocks_cfg2.txt"), "w").write(
    """In-kernel phase clocks of the fused stage kernel (CFD3D_PROF=1, clock64 per CTA at the phase boundaries; tools/prof_phases.py cfg2 32 4).
cfg2, 32 clusters of 4 CTAs, one frame each.  Microseconds at 1965 MHz; min / mean / max over the 128 CTAs.

A. like the bench: back-to-back launches over rotating input sets (data cold in L2, the kernel's code warm in the SMs)
rep 5""" + trim(warm) + """

B. COLD=1: an L2-evicting read kernel between the launches (evicts the stage kernel's instructions from the SMs as well:
   every once-per-frame phase after pass B gets slower -- what a stage call in the middle of a network sees)
rep 5""" + trim(cold) + "\n")
print("step %.2f us, ncu %.2f us, traffic %d B" % (us, dur, tot))
