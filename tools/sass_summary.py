#!/usr/bin/env python
"""Blackwell evidence from the shipped library: per-kernel counts of the SASS mnemonics that show what the hardware is
asked to do (TMA-unit bulk copies, bulk L2 prefetch, mbarrier, cluster barriers, DSMEM traffic, warp match / redux,
128-bit shared loads and global stores) plus the registers / shared memory ptxas reported.

    python tools/sass_summary.py [path/to/libcfd3d_b200.so] > profiles/r2_sass_summary.txt

Needs only cuobjdump (no GPU).  Regenerate after every kernel change so that the file tracks the code."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "centerfusiondetect3d_b200", "libcfd3d_b200.so")
LOG = os.path.join(ROOT, "centerfusiondetect3d_b200", "build.log")

# mnemonic prefix -> what it is
WHAT = [
    ("UBLKCP", "cp.async.bulk global->shared: 1-D bulk copy issued to the TMA unit"),
    ("UBLKPF", "cp.async.bulk.prefetch.L2: bulk L2 prefetch"),
    ("UTMALDG", "tensor-map TMA load (none expected: rows are contiguous, 1-D bulk copies are the right form)"),
    ("SYNCS", "mbarrier arrive / expect_tx / try_wait"),
    ("UCGABAR_ARV", "barrier.cluster.arrive"),
    ("UCGABAR_WAIT", "barrier.cluster.wait"),
    ("MAPA", "mapa: shared::cluster address of a peer CTA"),
    ("ST.E.U8.STRONG.CTA", None),
    ("ATOMS", "shared-memory atomics (local and, through mapa, distributed shared memory)"),
    ("REDUX", "redux.sync warp reductions"),
    ("MATCH", "match.any: per-warp aggregation of equal histogram bins"),
    ("VOTE", "ballot / any"),
    ("SHFL", "warp shuffles"),
    ("LDS.128", "128-bit shared-memory loads"),
    ("LDS", "shared-memory loads (all widths)"),
    ("STG.E.128", "128-bit global stores"),
    ("LDG", "global loads"),
    ("CCTL", "cache control (prefetch.global.L1 / L2)"),
    ("BAR.SYNC", "CTA / named barriers"),
    ("ACQBULK", "bulk-copy proxy acquire"),
    ("FENCE", "proxy / memory fences"),
    ("DADD", "fp64 adds"), ("DMUL", "fp64 multiplies"), ("DFMA", "fp64 fused multiply-adds (libdevice atan2 / sin / cos only)"),
    ("HMMA", "tensor-core MMA (none expected: nothing on this path is a contraction)"),
    ("UTCHMMA", "tcgen05 MMA (none expected)"),
]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            kernels[cur]["_n"] = 0
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur:
            kernels[cur]["_n"] += 1
            kernels[cur][m.group(1)] += 1
    regs = {}
    if os.path.isfile(LOG):
        name = None
        for line in open(LOG):
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                name = m.group(1)
            m = re.search(r"Used (\d+) registers.*?(?:, (\d+) bytes smem)?$", line.strip())
            if m and name:
                regs[name] = line.strip().replace("ptxas info    : ", "")
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m and name:
                regs[name + ":spill"] = line.strip()
    print("SASS summary of", os.path.relpath(LIB, ROOT), "(cuobjdump -sass, sm_100a; tools/sass_summary.py)")
    print("nvcc flags: see centerfusiondetect3d_b200/build.py (-gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false)\n")
    for k, c in kernels.items():
        short = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip().split("(")[0]
        print("%s   [%d SASS instructions]" % (short, c["_n"]))
        if k in regs:
            print("   ptxas: %s; %s" % (regs[k], regs.get(k + ":spill", "")))
        for pref, what in WHAT:
            if what is None:
                continue
            n = sum(v for name, v in c.items() if name.startswith(pref))
            if n or pref in ("UBLKCP", "UTMALDG", "HMMA", "UTCHMMA") and "fusion_stage" in k:
                print("   %-13s %5d   %s" % (pref, n, what))
        print()


if __name__ == "__main__":
    main()
