#!/usr/bin/env python
"""Top instructions of one launch in an ncu report by stall samples / executed count (SASS page).
usage: python tools/ncu_hot.py rep.ncu-rep LAUNCH_INDEX [N]"""
import csv, subprocess, sys
rep, k = sys.argv[1], int(sys.argv[2])
n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-skip", str(k),
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
def f(r, name):
    try: return float(r[col[name]])
    except Exception: return 0.0
tot_s = sum(f(r, "# Samples") for r in body); tot_i = sum(f(r, "Instructions Executed") for r in body)
print(f"launch {k}: {len(body)} SASS instructions, samples {tot_s:.0f}, warp instructions executed {tot_i:.0f}")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(f(r, s) for r in body) for s in stalls}
print("stall totals:", ", ".join(f"{s[6:]}={v:.0f}" for s, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
print("-- by samples")
for r in sorted(body, key=lambda r: -f(r, "# Samples"))[:n]:
    top = max(stalls, key=lambda s: f(r, s))
    print(f"{f(r,'# Samples'):7.0f} {f(r,'Instructions Executed'):9.0f} {top[6:]:>14} {r[col['Source']][:90]}")
print("-- by executed")
for r in sorted(body, key=lambda r: -f(r, "Instructions Executed"))[:n]:
    print(f"{f(r,'# Samples'):7.0f} {f(r,'Instructions Executed'):9.0f} {r[col['Source']][:90]}")
