#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares of one kernel from an ncu report.

    python tools/ncu_lines.py REPORT.ncu-rep CUBIN_PREFIX MANGLED_KERNEL_PREFIX SOURCE.cu [inst|samp] [rows]

e.g. tools/ncu_lines.py gpurun_out/fin_fused_culled.ncu-rep contacts_fused \\
         _ZN3gmm21contacts_fused_kernelILi1ELb1 gaudi_b200/csrc/contacts_fused.cu samp 30
Joins the report's SASS page with the `nvdisasm -g` line table of the library that was profiled
(gaudi_b200/libgaudimm_b200.so must be that build)."""
import collections, csv, io, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, cubin_prefix, kernel, srcfile = sys.argv[1:5]
key = sys.argv[5] if len(sys.argv) > 5 else "inst"
rows_out = int(sys.argv[6]) if len(sys.argv) > 6 else 28
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "gaudi_b200", "libgaudimm_b200.so")], cwd=tmp,
               check=True, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.startswith(cubin_prefix + ".")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], check=True, capture_output=True,
                     text=True).stdout.split("\n")
start = next(i for i, l in enumerate(dis) if re.match(r"\s*\.text\." + kernel, l))
table, cur = {}, None
for l in dis[start + 1:]:
    if l.strip().startswith(".text.") and table:
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        table[int(m.group(1), 16)] = cur
page = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], check=True, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(page)))
h = next(i for i, r in enumerate(rows) if "Address" in r)
cols, data = rows[h], rows[h + 1:]
ia, ie, isamp = cols.index("Address"), cols.index("Instructions Executed"), cols.index("# Samples")
stall = [(i, c[6:]) for i, c in enumerate(cols) if c.startswith("stall_") and "Not Issued" not in c]
base = int(data[0][ia], 16)
inst, samp, why = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
for r in data:
    if len(r) <= isamp:
        continue
    c = table.get(int(r[ia], 16) - base)
    inst[c] += int(r[ie] or 0)
    samp[c] += int(r[isamp] or 0)
    for i, name in stall:
        why[c][name] += int(r[i] or 0)
ti, ts = sum(inst.values()), sum(samp.values())
print("static SASS %d, warp instructions %d, samples %d" % (len(table), ti, ts))
src = open(srcfile).read().split("\n")
order = (samp if key == "samp" else inst).most_common(rows_out)
for c, _ in order:
    txt = src[c[1] - 1].strip()[:64] if c and c[0] == os.path.basename(srcfile) else ""
    top = ", ".join("%s %.0f%%" % (k, 100.0 * v / max(1, sum(why[c].values()))) for k, v in why[c].most_common(2))
    print("%-26s inst %5.1f%% samp %5.1f%%  [%s]  %s" % (str(c), 100.0 * inst[c] / ti, 100.0 * samp[c] / ts, top, txt))
