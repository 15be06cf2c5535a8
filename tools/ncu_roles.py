#!/usr/bin/env python
"""Per-warp-role view of an ncu source page for the warp-specialised GEMM kernels: the code is split at the USETMAXREG
instructions (each role starts with one), and for every role the samples, the top stall reasons and the hottest
instructions are listed.  usage: ncu_roles.py report.ncu-rep [launch=0] [top=6]"""
import csv, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 6
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
s0, s1 = secs[which], secs[which + 1]
hdr = rows[s0 + 1]; ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[s0 + 2:s1] if len(r) == len(hdr) and r[0].startswith("0x")]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("#", rows[s0][1][:120])
tot = sum(int(r[ix["# Samples"]] or 0) for r in data) or 1
# split into regions at USETMAXREG and at EXIT
regions = []; cur = []
for r in data:
    src = r[ix["Source"]]
    if "USETMAXREG" in src and cur:
        regions.append(cur); cur = []
    cur.append(r)
regions.append(cur)
for k, reg in enumerate(regions):
    s = sum(int(r[ix["# Samples"]] or 0) for r in reg)
    agg = {}
    for r in reg:
        for h in stalls: agg[h] = agg.get(h, 0) + int(r[ix[h]] or 0)
    ex = max(int(r[ix["Instructions Executed"]] or 0) for r in reg)
    print(f"\n== region {k}: {reg[0][ix['Address']][-5:]}..{reg[-1][ix['Address']][-5:]} first='{reg[0][ix['Source']].strip()[:40]}' instrs={len(reg)} samples={s} ({100*s/tot:.1f}%) max_exec={ex}")
    print("   stalls:", sorted(((h[6:], v) for h, v in agg.items() if v), key=lambda kv: -kv[1])[:7])
    for r in sorted(reg, key=lambda r: -int(r[ix["# Samples"]] or 0))[:top]:
        st = sorted(((h[6:], int(r[ix[h]] or 0)) for h in stalls), key=lambda kv: -kv[1])[0]
        print("    ", r[ix["Address"]][-5:], r[ix["# Samples"]], r[ix["Instructions Executed"]], r[ix["Source"]].strip()[:80], st)
