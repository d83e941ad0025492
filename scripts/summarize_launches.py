"""Turn an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table.
usage: python scripts/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches.md"""
import collections
import csv
import sys


def main(path: str) -> None:
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg: dict[str, list[float]] = collections.OrderedDict()
    tot = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = {"ns": v / 1e3, "ms": v * 1e3}.get(row["Metric Unit"], v)
        name = row["Kernel Name"].split("(")[0].replace("void ", "")[:80]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    print(f"source: {path} (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised launches:")
    print("compare SHARES, not absolutes)\n")
    print("| kernel | launches | total us | avg us | share |")
    print("|---|---:|---:|---:|---:|")
    for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {int(n)} | {v:.1f} | {v / n:.1f} | {100 * v / tot:.1f}% |")


if __name__ == "__main__":
    main(sys.argv[1])
