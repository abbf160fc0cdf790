#!/usr/bin/env python
"""SASS census of libgaudimm_b200.so: per kernel, the instruction count and how often the opcodes
that identify a hardware path occur (tcgen05 MMA = UTCHMMA, TMEM loads = LDTM, tensor TMA =
UTMALDG, bulk TMA = UBLKCP, commit / mbarrier = UTCBAR / SYNCS, legacy tensor cores = HMMA, warp
reductions = REDUX, 3-input min/max = FMNMX3, fp64 = DFMA ...).
    python tools/sass_census.py > profiles/sass_census_rNN.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gaudi_b200", "libgaudimm_b200.so")
KEY = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "UTCATOM", "SYNCS", "HMMA", "LDSM",
       "LDGSTS", "REDUX", "FMNMX3", "FFMA", "FADD", "FMUL", "DFMA", "MUFU", "LDG", "STG", "LDS", "STS", "ATOMS", "ATOMG", "RED",
       "BAR", "SHFL", "VOTE", "MATCH"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], check=True, capture_output=True, text=True).stdout
kernels = collections.OrderedDict()
cur = None
for line in sass.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kernels[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        kernels[cur][m.group(1)] += 1

demangle = subprocess.run(["c++filt"] + list(kernels), capture_output=True, text=True).stdout.split("\n")
print("SASS census of gaudi_b200/libgaudimm_b200.so (sm_100a), %d kernels" % len(kernels))
print("columns: total instructions, then the occurrences of path-identifying opcodes (zero counts omitted)\n")
for (name, ops), pretty in zip(kernels.items(), demangle):
    total = sum(ops.values())
    short = re.sub(r"\(.*", "", pretty) or name
    marks = ["%s %d" % (k, sum(v for o, v in ops.items() if o == k or o.startswith(k + "_"))) for k in KEY]
    marks = [m for m in marks if not m.endswith(" 0")]
    print("%-64s %6d  %s" % (short[:64], total, ", ".join(marks)))
