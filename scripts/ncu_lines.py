#!/usr/bin/env python
"""Per-source-line view of an ncu report for code that lives in inlined headers.

ncu's `--page source --csv` lists SASS instructions with sampling counts but no file:line; nvdisasm
-g on the same cubin lists file:line per SASS address. This joins the two by address.
  python scripts/ncu_lines.py <report.ncu-rep> <kernel name> <cubin> [launch index] [mangled-name substring]
"""
import csv, re, subprocess, sys, collections, io

rep, kern, cubin = sys.argv[1:4]
skip = sys.argv[4] if len(sys.argv) > 4 else "0"
mangled = sys.argv[5] if len(sys.argv) > 5 else None  # substring of the cubin function (templated kernels), e.g. 7k_traceILi252E
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
line_of, cur, infn = {}, None, False
for ln in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        infn = re.search(r"\d+%s[A-Z]" % re.escape(kern), m.group(1)) is not None or m.group(1).endswith(kern)
        infn = (mangled in m.group(1)) if mangled else (kern in m.group(1))
        fn = m.group(1)
        cur = None
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m and infn and cur:
        line_of.setdefault((fn, int(m.group(1), 16)), cur)
# pick the function whose name matches exactly the kernel (first with most addresses)
fns = collections.Counter(f for f, _ in line_of)
best = None
for f in fns:
    if re.search(r"pb2\d+%sE" % re.escape(kern), f) or f.endswith(kern):
        best = f
if best is None:
    best = fns.most_common(1)[0][0]
base = None
agg = collections.defaultdict(lambda: [0, 0, 0, 0, collections.Counter()])
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = 0
for r in rows[2:]:
    try:
        addr = int(r[ix["Address"]], 16)
    except ValueError:
        continue
    if base is None:
        base = addr
    key = line_of.get((best, addr - base), ("?", 0))
    a = agg[key]
    s = int(r[ix["# Samples"]] or 0)
    a[0] += s
    tot += s
    a[1] += int(r[ix["Instructions Executed"]] or 0)
    a[2] += int(r[ix["Thread Instructions Executed"]] or 0)
    for c in stall_cols:
        v = int(r[ix[c]] or 0)
        if v:
            a[4][c] += v
print(f"kernel {kern} function {best} total samples {tot}")
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
    lanes = a[2] / a[1] if a[1] else 0
    top = ", ".join(f"{k[6:]}:{v}" for k, v in a[4].most_common(3))
    print(f"{key[0]:>22}:{key[1]:<4} samples {a[0]:7d} ({100*a[0]/max(tot,1):5.1f}%) inst {a[1]:10d} lanes {lanes:5.1f}  {top}")
