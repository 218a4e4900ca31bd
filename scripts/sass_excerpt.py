#!/usr/bin/env python
"""SASS of one source range of one kernel, from the built library (cuobjdump -xelf + nvdisasm -g):
  python scripts/sass_excerpt.py <lib.so> <cubin stem> <kernel regex> <file> <first line> <last line>
Prints the instructions attributed to file:first..last in address order and an opcode histogram."""
import collections, os, re, subprocess, sys, tempfile
lib, stem, kern, fname, lo, hi = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4], int(sys.argv[5]), int(sys.argv[6])
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, check=True, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.startswith(stem) and f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
cur, infn, hist, n = None, False, collections.Counter(), 0
for line in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", line)
    if m:
        infn, cur = re.search(kern, m.group(1)) is not None, None
        if infn:
            print("// function", m.group(1))
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+((?:@!?U?P\d+\s+)?)([A-Z0-9_]+)(\S*)(.*?);", line)
    if m and cur and cur[0] == fname and lo <= cur[1] <= hi:
        print(f"/*{m.group(1)}*/ L{cur[1]:<4} {m.group(2)}{m.group(3)}{m.group(4)}{m.group(5)}")
        hist[m.group(3)] += 1
        n += 1
print(f"// {n} instructions attributed to {fname}:{lo}-{hi}")
print("// " + ", ".join(f"{k} {v}" for k, v in hist.most_common()))
