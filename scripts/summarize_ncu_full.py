"""Transpose `ncu -i X.ncu-rep --page raw --csv` of one or more reports into metric rows
(one column per captured kernel launch), keeping the metrics a roofline review needs.
usage: python scripts/summarize_ncu_full.py a.ncu-rep [b.ncu-rep ...] > profiles/rNN_ncu_full_summary.csv"""
import csv
import subprocess
import sys

KEEP = ("gpu__time_duration", "dram__bytes", "dram__throughput", "lts__t_sector_hit_rate", "lts__t_bytes",
        "sm__throughput", "sm__warps_active", "smsp__issue_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe", "sm__pipe_", "launch__", "smsp__average_warp", "smsp__warp_issue_stalled",
        "l1tex__t_bytes", "smsp__cycles_active", "sm__cycles_elapsed", "smsp__pcsamp_warps_issue_stalled")


def main():
    cols = []          # (kernel name, {metric: (unit, value)})
    for rep in sys.argv[1:]:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader([ln for ln in out.splitlines() if ln.startswith('"')]))
        hdr, units = rows[0], rows[1]
        for r in rows[2:]:
            rec = dict(zip(hdr, zip(units, r)))
            cols.append((rep.split("/")[-1] + ": " + rec["Kernel Name"][1][:70], rec))
    names = []
    for _, rec in cols:
        for k in rec:
            if k not in names and any(k.startswith(p) or ("." + p) in k for p in KEEP):
                names.append(k)
    w = csv.writer(sys.stdout)
    w.writerow(["metric", "unit"] + [c for c, _ in cols])
    for k in names:
        unit = next((rec[k][0] for _, rec in cols if k in rec), "")
        w.writerow([k, unit] + [rec.get(k, ("", ""))[1] for _, rec in cols])


if __name__ == "__main__":
    main()
