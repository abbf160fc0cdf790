#!/usr/bin/env python
"""Static SASS instruction counts of the fused Contacts kernels, per kernel variant and, for one
variant, per phase of the source (same phase boundaries as tools/ncu_phases.py):

    python tools/sass_size.py [mangled-prefix, default _ZN3gmm21contacts_fused_kernelILi1ELb1ELb1]"""
import collections, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
want = sys.argv[1] if len(sys.argv) > 1 else "_ZN3gmm21contacts_fused_kernelILi1ELb1ELb1"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "gaudi_b200", "libgaudimm_b200.so")], cwd=tmp, check=True, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.startswith("contacts_fused.")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], check=True, capture_output=True, text=True).stdout.split("\n")
cur_k, cur_l = None, None
size = collections.Counter(); lines = collections.Counter()
for l in dis:
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m: cur_k = m.group(1); continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur_l = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if cur_k and re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l):
        size[cur_k] += 1
        if cur_k.startswith(want): lines[cur_l] += 1
for k, v in sorted(size.items()): print("%-70s %5d" % (k[:70], v))
src = open(os.path.join(ROOT, "gaudi_b200", "csrc", "contacts_fused.cu")).read().split("\n")
def find(p): return next(i + 1 for i, l in enumerate(src) if p in l)
marks = [("drain_pairs", find("void fused_drain_pairs")), ("expand_candidates", find("void fused_drain_candidates")),
         ("kernel_head", find("contacts_fused_kernel(const FusedArgs f)")), ("stage", find("---- 0. stage")),
         ("cull", find("---- 1. surviving")), ("refine", find("Refine: a leaf stays only")), ("sweep_prep", find("---- 2. sweep")),
         ("sweep_loop", find("two mobile atoms per step")), ("post_sweep", find("float ma[4]")), ("static_pairs", find("---- 3. static-static")),
         ("intra_mobile", find("---- 4. pairs inside")), ("finish", find("---- 5. finish")), ("end", len(src) + 1)]
ph = collections.Counter()
for (f, ln), n in ((k, v) for k, v in lines.items() if k):
    if f == "common.cuh": ph["score_pair (common.cuh)"] += n
    elif f != "contacts_fused.cu": ph[f] += n
    else: ph[next((nm for (nm, lo), (_, hi) in zip(marks, marks[1:]) if lo <= ln < hi), "other")] += n
print("--- %s by phase" % want)
for k, v in ph.most_common(): print("%-32s %5d" % (k, v))
