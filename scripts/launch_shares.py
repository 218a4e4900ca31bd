#!/usr/bin/env python
"""Per-kernel totals and shares of an ncu launch list (`--metrics gpu__time_duration.sum --csv`):
  python scripts/launch_shares.py <launches.csv> [first_launch [last_launch]]
The times are cold-cache and serialised (ncu replays each kernel alone), so only the SHARES are comparable
with bench.py's `kernel_time_shares`."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr, rows = rows[0], rows[1:]
ix = {h: i for i, h in enumerate(hdr)}
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 10 ** 9
tot, cnt = collections.Counter(), collections.Counter()
for r in rows:
    if not lo <= int(r[ix["ID"]]) < hi:
        continue
    name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "")
    ns = float(r[ix["Metric Value"]].replace(",", ""))
    tot[name] += ns
    cnt[name] += 1
total = sum(tot.values())
group = collections.Counter()
for k, v in tot.items():
    g = ("trace" if k.startswith("k_trace") else "shadow" if k.startswith("k_shadow") else
         "shade" if k.startswith(("k_shade", "k_classify")) else "other")
    group[g] += v
print(f"{len(rows)} launches in file, {sum(cnt.values())} in range, {total / 1e6:.3f} ms under ncu")
print("kernel | launches | total ms | share")
for k, v in tot.most_common():
    print(f"{k} | {cnt[k]} | {v / 1e6:.3f} | {v / total:.3f}")
print("groups (as bench.py kernel_time_shares):", {g: round(v / total, 3) for g, v in group.items()})
