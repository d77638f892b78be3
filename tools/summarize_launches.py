"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into a markdown table.
usage: python tools/summarize_launches.py launches.csv "title" > profiles/xxx.md"""
import csv
import collections
import sys


def main():
    path, title = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
    rows = []
    with open(path, newline="") as f:
        lines = f.readlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    rd = csv.DictReader(lines[start:])
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
        rows.append((r["Kernel Name"], us))
    agg = collections.defaultdict(lambda: [0.0, 0])
    for k, us in rows:
        agg[k][0] += us
        agg[k][1] += 1
    total = sum(v[0] for v in agg.values())
    print(f"# {title}\n")
    print(f"total {total:.0f} us over {len(rows)} launches\n")
    print("| us | share | launches | avg us | kernel |\n|---|---|---|---|---|")
    for k, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"| {us:.1f} | {100 * us / total:.1f}% | {n} | {us / n:.2f} | `{k[:110]}` |")


if __name__ == "__main__":
    main()
