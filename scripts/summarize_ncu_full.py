"""Summarise an `ncu --set full` report (.ncu-rep) as a markdown table, one row per captured launch,
and emit the per-launch DRAM traffic of each kernel as JSON (bench.py's roofline.traffic reads it).

usage: python scripts/summarize_ncu_full.py gpurun_out/prof.ncu-rep profiles/rNN_ncu_full.md profiles/traffic.json
"""
import csv
import io
import json
import re
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "us"),
    ("sm__cycles_elapsed.max", "cycles"),
    ("dram__bytes_read.sum", "dram rd MB"),
    ("dram__bytes_write.sum", "dram wr MB"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "SFU %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
]


def short(name: str) -> str:
    name = re.sub(r"^void ", "", name)
    name = name.replace("b200::", "")
    name = re.sub(r"\(CUtensorMap.*", "", name)
    name = re.sub(r"\(const .*", "", name)
    return name.replace("(int)", "")


def main():
    rep, out_md, out_json = sys.argv[1:4]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", ",".join(m for m, _ in METRICS)],
                         capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    lines = ["source: `%s` (ncu --set full --clock-control none; one row per captured launch)" % rep, "",
             "| kernel | " + " | ".join(lbl for _, lbl in METRICS) + " |", "|---|" + "---:|" * len(METRICS)]
    traffic = {}
    for r in data:
        name = short(r[col["Kernel Name"]])
        cells = []
        for m, lbl in METRICS:
            v = r[col[m]] if m in col else ""
            try:
                f = float(v.replace(",", ""))
                u = units[col[m]]
                if "MB" in lbl:
                    f = f * {"Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "byte": 1e-6}.get(u, 1.0)
                if lbl == "us":      # ncu picks one unit per column (ms when a long kernel is in the report)
                    f = f * {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}.get(u, 1.0)
                cells.append("%.1f" % f if abs(f) < 1e5 and f != int(f) else "%d" % f)
            except ValueError:
                cells.append(v)
        lines.append("| `%s` | " % name + " | ".join(cells) + " |")
        rd = float(r[col["dram__bytes_read.sum"]].replace(",", "")) * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}[units[col["dram__bytes_read.sum"]]]
        wr = float(r[col["dram__bytes_write.sum"]].replace(",", "")) * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}[units[col["dram__bytes_write.sum"]]]
        traffic.setdefault(name, []).append(rd + wr)
    open(out_md, "w").write("\n".join(lines) + "\n")
    json.dump({"source": rep, "dram_bytes_per_launch": {k: sum(v) / len(v) for k, v in traffic.items()},
               "launches_captured": {k: len(v) for k, v in traffic.items()}}, open(out_json, "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
