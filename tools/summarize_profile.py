"""Summarise gpurun_out/ ncu artefacts into profiles/ (tracked).

  python tools/summarize_profile.py <tag> [launches.csv] [prof.ncu-rep]
writes profiles/<tag>_launches.md (per-kernel share of the step from the --metrics gpu__time_duration.sum pass)
and    profiles/<tag>_ncu_full.md  (raw-page metrics of the --set full capture, one row per profiled launch).
"""
import collections
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RAW = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
       "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
       "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
       "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
       "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "sm__inst_executed_pipe_tensor.sum"]


def launches(path, out):
    rows = list(csv.reader(open(path)))
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[start]
    ix = {h: i for i, h in enumerate(hdr)}
    agg = collections.OrderedDict()
    for r in rows[start + 1:]:
        if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = r[ix["Kernel Name"]].split("(")[0]
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        v = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write(f"# ncu launch list: {os.path.basename(path)}\n\n")
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares, "
                "not absolutes).\n\n| kernel | launches | total us | share |\n|---|---:|---:|---:|\n")
        for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write(f"| `{k[:90]}` | {n} | {t:.1f} | {100 * t / tot:.1f}% |\n")
        f.write(f"\ntotal {tot:.1f} us over {sum(a[0] for a in agg.values())} launches\n")


def full(path, out):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    cols = [c for c in RAW if c in ix]
    with open(out, "w") as f:
        f.write(f"# ncu --set full: {os.path.basename(path)}\n\n| kernel | " +
                " | ".join(f"{c} [{units[ix[c]]}]" for c in cols) + " |\n|---|" + "---:|" * len(cols) + "\n")
        for r in rows[2:]:
            f.write(f"| `{r[ix['Kernel Name']].split('(')[0][:60]}` | " + " | ".join(r[ix[c]] for c in cols) + " |\n")


if __name__ == "__main__":
    tag = sys.argv[1]
    lc = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "launches.csv")
    rep = sys.argv[3] if len(sys.argv) > 3 else None
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    if os.path.exists(lc):
        launches(lc, os.path.join(ROOT, "profiles", f"{tag}_launches.md"))
    if rep and os.path.exists(rep):
        full(rep, os.path.join(ROOT, "profiles", f"{tag}_ncu_full.md"))
