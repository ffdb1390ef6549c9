"""Per-kernel launch count, total time and share of a `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections
import csv
import sys

all_rows = list(csv.reader(open(sys.argv[1])))
head = next(r for r in all_rows if "Kernel Name" in r)          # (an --nvtx capture has two extra range columns)
i_name, i_val = head.index("Kernel Name"), head.index("Metric Value")
rows = [r for r in all_rows if len(r) > i_val and r[0].isdigit()]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name = r[i_name].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    agg[name[:90]][0] += 1
    agg[name[:90]][1] += float(r[i_val])
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot / 1e6:.3f} ms of kernel time (under ncu: serialised, cold caches)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[0]:5d} x  mean {v[1] / v[0] / 1e3:9.1f} us  total {v[1] / 1e6:8.3f} ms  {100 * v[1] / tot:5.1f} %  {k}")
