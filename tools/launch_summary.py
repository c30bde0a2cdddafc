"""ncu launch list (--metrics gpu__time_duration.sum --csv) -> per-kernel totals and shares
   python tools/launch_summary.py gpurun_out/x_launches.csv profiles/x_launch_summary.csv"""
import csv, collections, re, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value"); ui = h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki]); name = re.sub(r"<.*", "", name) if len(name) > 60 else name
    v = float(r[vi].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[ui], 1.0)
    agg[name][0] += 1; agg[name][1] += v
tot = sum(v[1] for v in agg.values())
with open(sys.argv[2], "w") as f:
    f.write("kernel,launches,total_ms,share\n")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write('"%s",%d,%.4f,%.4f\n' % (k, v[0], v[1], v[1] / tot))
