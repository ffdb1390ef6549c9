"""dram__bytes_read.sum + dram__bytes_write.sum of the (last) kernel in an .ncu-rep, as the JSON bench.py reads for
`roofline.traffic`:   python tools/ncu_traffic.py report.ncu-rep profiles/rNN_k1_traffic.json "<how it was captured>" """
import csv
import json
import subprocess
import sys

rep, out, how = sys.argv[1], sys.argv[2], sys.argv[3]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h, units, v = rows[0], rows[1], rows[-1]


def val(key):
    i = h.index(key)
    x = float(v[i].replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]


d = {"kernel": v[h.index("Kernel Name")], "dram_bytes_read": val("dram__bytes_read.sum"), "dram_bytes_write": val("dram__bytes_write.sum"),
     "gpu_time_ms_under_ncu": float(v[h.index("gpu__time_duration.sum")].replace(",", "")) *
     {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[units[h.index("gpu__time_duration.sum")]], "how": how}
d["dram_bytes_per_launch"] = d["dram_bytes_read"] + d["dram_bytes_write"]
json.dump(d, open(out, "w"), indent=1)
print(json.dumps(d))
