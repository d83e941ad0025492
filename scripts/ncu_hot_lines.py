"""Top stall-sample lines (SASS + CUDA source line) of a kernel in an .ncu-rep captured with --import-source on.
usage: python scripts/ncu_hot_lines.py prof.ncu-rep [kernel-regex] > hot.txt"""
import csv, io, subprocess, sys
rep = sys.argv[1]
kern = sys.argv[2] if len(sys.argv) > 2 else ".*"
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr_i = [i for i, r in enumerate(rows) if "Source" in r and any("Sampling (All" in c for c in r)]
if not hdr_i:
    print("no source table", raw[:500]); sys.exit(0)
h = rows[hdr_i[0]]
end = hdr_i[1] if len(hdr_i) > 1 else len(rows)
si = h.index("Source"); col = [i for i, c in enumerate(h) if "Warp Stall Sampling (All" in c][0]
data, last_src = [], ""
for r in rows[hdr_i[0] + 1:end]:
    if len(r) <= col: continue
    try: v = float(r[col])
    except ValueError: continue
    data.append((v, r[0], r[si][:120]))
tot = sum(d[0] for d in data) or 1.0
print("columns:", h[:6])
for v, a, s in sorted(data, reverse=True)[:40]:
    print("%5.1f%%  %s  %s" % (100 * v / tot, a[-6:], s))
