#!/usr/bin/env python
"""Hot spots of an .ncu-rep's SASS source page: python scripts/ncu_hot.py file.ncu-rep [N]
Prints total samples by opcode class and the top-N instructions by stall samples with their dominant stall reason."""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
isrc, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = 0; byop = collections.Counter(); bystall = collections.Counter(); exop = collections.Counter()
recs = []
for k, r in enumerate(rows[2:]):
    if len(r) < len(hdr): continue
    s = int(r[isamp] or 0); tot += s
    op = r[isrc].split()[0] if not r[isrc].strip().startswith("@") else r[isrc].split()[1]
    op = op.split(".")[0]
    byop[op] += s; exop[op] += int(r[iex] or 0)
    st = {h: int(r[i] or 0) for i, h in stall_cols}
    for h, v in st.items(): bystall[h] += v
    recs.append((s, k, r[isrc].strip(), max(st, key=st.get) if s else "", int(r[iex] or 0)))
print("total samples", tot)
print("by stall:", [(h, v) for h, v in bystall.most_common(10)])
print("by opcode (samples, executed):", [(o, v, exop[o]) for o, v in byop.most_common(16)])
for s, k, src, why, ex in sorted(recs, reverse=True)[:N]:
    print(f"{s:7d} #{k:5d} ex={ex:9d} {why:22s} {src[:90]}")
