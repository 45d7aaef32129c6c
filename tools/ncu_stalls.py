"""Stall-reason summary of one kernel of an .ncu-rep (source page): python tools/ncu_stalls.py rep [launch_index]"""
import csv, subprocess, sys
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + (["--launch-skip", sys.argv[2], "--launch-count", "1"] if len(sys.argv) > 2 else ["--launch-count", "1"]),
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
print(rows[0][:2])
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr) and r[ix["# Samples"]].isdigit()]
tot = sum(int(r[ix['# Samples']] or 0) for r in data)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(int(r[ix[h]] or 0) for r in data) for h in stalls}
print("samples", tot, [(k[6:], round(100 * v / tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]])
inst = sum(int(r[ix['Instructions Executed']] or 0) for r in data)
print("warp instructions", inst)
for r in sorted(data, key=lambda r: -int(r[ix['# Samples']] or 0))[:18]:
    n = int(r[ix['# Samples']])
    print(f"{100*n/tot:5.1f}%  {r[ix['Source']][:70]:70s}", {h[6:]: r[ix[h]] for h in stalls if int(r[ix[h]] or 0) > n * 0.25})
