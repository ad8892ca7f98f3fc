#!/usr/bin/env python
"""Dynamic instruction mix and stall samples per opcode from an .ncu-rep source page."""
import csv, subprocess, sys, io, collections, re
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('"Address"')][0]
rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
h = rows[0]
ia, isrc, iex, ismp = h.index("Address"), h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
ex = collections.Counter(); smp = collections.Counter(); tot = 0; tots = 0
for r in rows[1:]:
    if len(r) <= iex or not r[iex].isdigit():
        continue
    m = re.match(r"\s*(@!?U?P\w+\s+)?([A-Z0-9_]+)", r[isrc])
    op = m.group(2) if m else "?"
    ex[op] += int(r[iex]); smp[op] += int(r[ismp]); tot += int(r[iex]); tots += int(r[ismp])
print("total warp instructions", tot, "samples", tots)
for op, n in ex.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 25):
    print(f"{op:10s} {n:12d} {100*n/tot:5.1f}%   stall samples {100*smp[op]/max(tots,1):5.1f}%")
