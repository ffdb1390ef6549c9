"""Per-source-line instruction / stall-sample table from an .ncu-rep (needs -lineinfo):
   python tools/ncu_lines.py report.ncu-rep units  (units = rays, samples, ... the kernel processed)"""
import csv
import subprocess
import sys

rep, units = sys.argv[1], float(sys.argv[2])
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fn, h, agg = None, None, []
for r in csv.reader(txt.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        fn = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        h = r
        ix, isamp = h.index("Instructions Executed"), h.index("# Samples")
    elif h and len(r) > ix and r[0].isdigit() and r[2] == "-":
        try:
            agg.append((fn, int(r[0]), int(r[ix]), int(r[isamp] or 0), r[1][:105]))
        except ValueError:
            pass
tot, ts = sum(a[2] for a in agg), sum(a[3] for a in agg)
print(f"total instructions / unit: {tot / units:.1f}   stall samples: {ts}")
for fn, ln, n, smp, src in agg:
    if n > tot * 0.004 or smp > ts * 0.006:
        print(f"{n / units:7.1f} {100 * smp / max(ts, 1):5.1f}% {fn[:16]:16}:{ln:<4} {src}")
