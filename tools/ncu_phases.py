#!/usr/bin/env python
"""Per-phase instruction, sample and stall shares of contacts_fused_kernel<1> from an ncu report.

    python tools/ncu_phases.py gpurun_out/fused_full_r1y.ncu-rep

Joins the report's SASS page (`ncu -i rep --page source --csv`) with the line table of the
library that was profiled (`nvdisasm -g` on the cubin inside gaudi_b200/libgaudimm_b200.so; the
library must be the build that produced the report) and buckets source lines into the phases
named in the header of gaudi_b200/csrc/contacts_fused.cu.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gaudi_b200", "libgaudimm_b200.so")
SRC = os.path.join(ROOT, "gaudi_b200", "csrc", "contacts_fused.cu")
KERNEL = os.environ.get("GMM_PHASE_KERNEL", "_ZN3gmm21contacts_fused_kernelILi1ELb1ELb1")


def line_table():
    """[(offset, (file, line))] of the kernel's SASS instructions."""
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, check=True, capture_output=True)
        cubin = [f for f in os.listdir(tmp) if f.startswith("contacts_fused.")][0]
        dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], check=True, capture_output=True,
                             text=True).stdout.split("\n")
    start = next(i for i, l in enumerate(dis) if re.match(r"\s*\.text\." + KERNEL, l))
    out, cur = [], None
    for l in dis[start + 1:]:
        if l.strip().startswith(".text.") and out:
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            out.append((int(m.group(1), 16), cur))
    return out


def phases():
    src = open(SRC).read().split("\n")

    def find(pat):
        return next(i + 1 for i, l in enumerate(src) if pat in l)
    marks = [("drain_pairs", find("void fused_drain_pairs")), ("expand_candidates", find("void fused_drain_candidates")),
             ("kernel_head", find("contacts_fused_kernel(const FusedArgs f)")), ("stage", find("---- 0. stage")),
             ("cull", find("---- 1. surviving")), ("refine", find("Refine: a leaf stays only")),
             ("sweep_prep", find("---- 2. sweep")), ("sweep_loop", find("two mobile atoms per step")),
             ("post_sweep", find("bool in[4], band[4];")), ("static_pairs", find("---- 3. static-static")),
             ("intra_mobile", find("---- 4. pairs inside")), ("finish", find("---- 5. finish")), ("end", len(src) + 1)]
    return [(marks[i][1], marks[i + 1][1] - 1, marks[i][0]) for i in range(len(marks) - 1)]


def main():
    rep = sys.argv[1]
    page = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(page)))
    hdr = next(i for i, r in enumerate(rows) if "Address" in r)
    cols, data = rows[hdr], rows[hdr + 1:]
    ia, ie, isamp = cols.index("Address"), cols.index("Instructions Executed"), cols.index("# Samples")
    stall = [(i, h[6:]) for i, h in enumerate(cols) if h.startswith("stall_") and "Not Issued" not in h]
    table = dict(line_table())
    ph = phases()

    def phase(c):
        if c is None:
            return "unattributed"
        f, l = c
        if f == "common.cuh":
            return "score_pair"
        if f == "contacts_common.cuh":
            return "helpers (dist2f, loose_c2, ms_excluded)"
        if f != "contacts_fused.cu":
            return "intrinsics (shfl, ballot, atomics)"
        return next((name for lo, hi, name in ph if lo <= l <= hi), "other")
    base = int(data[0][ia], 16)
    inst, samp, why = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
    for r in data:
        if len(r) <= isamp:
            continue
        p = phase(table.get(int(r[ia], 16) - base))
        inst[p] += int(r[ie] or 0)
        samp[p] += int(r[isamp] or 0)
        for i, h in stall:
            why[p][h] += int(r[i] or 0)
    ti, ts = sum(inst.values()), sum(samp.values())
    total = collections.Counter()
    for c in why.values():
        total.update(c)
    print("static SASS instructions: %d; warp instructions executed: %d; samples: %d" % (len(table), ti, ts))
    print("stall reasons overall: " + ", ".join("%s %.1f%%" % (h, 100.0 * v / sum(total.values()))
                                                for h, v in total.most_common(8)))
    print("%-40s %7s %8s  %s" % ("phase", "inst %", "sample %", "top stall reasons of the phase"))
    for p, n in samp.most_common():
        s = sum(why[p].values()) or 1
        print("%-40s %7.2f %8.2f  %s" % (p, 100.0 * inst[p] / ti, 100.0 * n / ts,
                                         ", ".join("%s %.0f%%" % (h, 100.0 * v / s) for h, v in why[p].most_common(3))))


if __name__ == "__main__":
    main()
