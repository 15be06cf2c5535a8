#!/usr/bin/env python
"""Top stall locations of one kernel launch in an ncu report (SASS view).  usage: ncu_hot.py report.ncu-rep [launch=0] [top=40] [out.csv]"""
import csv, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
secs = []
for i, r in enumerate(rows):
    if r and r[0] == "Kernel Name":
        secs.append(i)
secs.append(len(rows))
s0, s1 = secs[which], secs[which + 1]
name = rows[s0][1]
hdr = rows[s0 + 1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[s0 + 2:s1] if len(r) == len(hdr)]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[ix["# Samples"]] or 0) for r in data) or 1
lines = [["address", "samples", "pct", "sass", "top_stall", "second_stall"]]
agg = {}
for r in data:
    for h in stalls:
        agg[h] = agg.get(h, 0) + int(r[ix[h]] or 0)
for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:top]:
    s = int(r[ix["# Samples"]] or 0)
    st = sorted(((h, int(r[ix[h]] or 0)) for h in stalls), key=lambda kv: -kv[1])[:2]
    lines.append([r[ix["Address"]][-5:], s, f"{100 * s / tot:.1f}", r[ix["Source"]].strip()[:90], f"{st[0][0]}={st[0][1]}", f"{st[1][0]}={st[1][1]}"])
print("#", name[:150], "samples", tot)
print("# stall totals:", sorted(((k, round(100 * v / tot, 1)) for k, v in agg.items()), key=lambda kv: -kv[1])[:8])
if len(sys.argv) > 4:
    with open(sys.argv[4], "w") as f:
        f.write(f"# {name}\n")
        csv.writer(f).writerows(lines)
else:
    for l in lines:
        print(*l, sep=" | ")
