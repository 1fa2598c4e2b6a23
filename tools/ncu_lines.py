"""Aggregates an ncu report's per-instruction samples by source line / function.
usage: ncu_lines.py report.ncu-rep object.o [top]"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

rep, obj = sys.argv[1], os.path.abspath(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
addr2 = {}
cur, fn = None, None
for l in dis.splitlines():
    m = re.match(r"^(\$?[_A-Za-z0-9\$]+):", l)
    if m and "_Z" in m.group(1):
        fn = m.group(1).split("$")[-1]
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
    if m:
        addr2[int(m.group(1), 16)] = (cur, fn)
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]
ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
sidx = {h: hdr.index(h) for h in stalls}
base = None
inst, samp = collections.Counter(), collections.Counter()
finst, fsamp = collections.Counter(), collections.Counter()
st_line = collections.defaultdict(collections.Counter)
st_tot = collections.Counter()
for r in rows[2:]:
    a = int(r[ia], 16)
    base = a if base is None else base
    line, f = addr2.get(a - base, (None, None))
    n, s = int(r[ii] or 0), int(r[isamp] or 0)
    inst[line] += n
    samp[line] += s
    finst[f] += n
    fsamp[f] += s
    for h in stalls:
        v = int(r[sidx[h]] or 0)
        st_line[line][h] += v
        st_tot[h] += v
ti, ts = sum(inst.values()), sum(samp.values())
print(f"total warp-instructions {ti}  samples {ts}")
print("stall totals:", ", ".join(f"{h[6:]} {100*v/ts:.1f}%" for h, v in st_tot.most_common(8)))
print("--- by function")
for f, v in fsamp.most_common(16):
    d = subprocess.run(["c++filt", str(f)], capture_output=True, text=True).stdout.strip()
    d = re.sub(r"\(anonymous namespace\)::|_INTERNAL_[0-9a-f_]+localize_kernels_cu_[0-9a-f]+::", "", d)
    print(f"{100*v/ts:6.2f}% samples {100*finst[f]/ti:6.2f}% inst  {d[:90]}")
print("--- by line")
for l, v in samp.most_common(top):
    tops = ", ".join(f"{h[6:]} {100*x/max(v,1):.0f}%" for h, x in st_line[l].most_common(3))
    print(f"{100*v/ts:6.2f}% samples {100*inst[l]/ti:6.2f}% inst  {l}  [{tops}]")
