"""Per-kernel table of ONE iteration from an ncu launch list (gpu__time_duration.sum, --csv --log-file):
usage: python profiles/launch_table.py launches.csv launches_per_iteration  -> markdown table on stdout.
ncu times are cold-cache and serialised: the shares are what should agree with bench.py, not the absolute times."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
last = rows[1:][-int(sys.argv[2]):]
agg = collections.OrderedDict()
for r in last:
    name = r[ki]
    name = name[:name.find(">(") + 1] if ">(" in name else name.split("(")[0]
    name = name.replace("void ", "").replace("<unnamed>::", "").replace("(int)", "")
    v = float(r[vi].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}[r[ui]]
    agg.setdefault(name, [0, 0.0])
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print("| kernel | launches | ms (ncu, serialised) | share |\n|---|---|---|---|")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("| `%s` | %d | %.3f | %.1f %% |" % (k, c, t, 100 * t / tot))
print("| total | %d | %.3f | |" % (sum(v[0] for v in agg.values()), tot))
