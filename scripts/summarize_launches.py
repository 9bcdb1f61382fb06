"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel.
usage: python scripts/summarize_launches.py gpurun_out/launches.csv [header line ...] > profiles/rNN_launches_summary.txt"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    rows = collections.OrderedDict()
    with open(path, newline="") as fh:
        lines = [ln for ln in fh if ln.startswith('"')]
    for rec in csv.DictReader(lines):
        if rec["Metric Name"] != "gpu__time_duration.sum":
            continue
        k = rec["Kernel Name"]
        n, t = rows.get(k, (0, 0.0))
        rows[k] = (n + 1, t + float(rec["Metric Value"].replace(",", "")))
    total = sum(t for _, t in rows.values())
    for extra in sys.argv[2:]:
        print("# " + extra)
    print(f"# {sum(n for n, _ in rows.values())} launches, {total / 1e6:.3f} ms of kernel time (cold-cache, serialised)")
    print("# launches  total_ns  share  mean_us  kernel")
    for k, (n, t) in sorted(rows.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:6d} {int(t):14d} {100 * t / total:5.1f}% {t / n / 1e3:9.1f}  {k[:150]}")


if __name__ == "__main__":
    main()
