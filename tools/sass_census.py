"""Opcode census of the shipped library per kernel (cuobjdump -sass): the tensor-core / TMA / TMEM mnemonics that show
which hardware path each kernel uses.   python tools/sass_census.py > profiles/rNN_sass_census.txt"""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nerf_tex_b200", "libnerftex_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "IMMA", "MUFU", "ATOMS", "ATOMG", "RED",
        "SHFL", "LDG", "STG", "LDS", "STS", "BAR", "F2FP", "FFMA", "DFMA"]
txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
kern, counts, total = None, collections.OrderedDict(), collections.Counter()
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(anonymous namespace\)::|<unnamed>::|^void ", "", kern).split("(")[0]
        counts[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        total[kern] += 1
        for k in KEYS:
            if op == k or op.startswith(k + "."):
                counts[kern][k] += 1
                if k in ("UTCHMMA", "UTCBAR", "UBLKCP") and ("2CTA" in op or "MULTICAST" in op):
                    counts[kern][op] += 1
print(f"SASS opcode census of {os.path.basename(LIB)} (sm_100a), instructions per kernel\n")
for kern, c in counts.items():
    items = ", ".join(f"{k} {v}" for k, v in c.items())
    print(f"{kern[:100]:100s} total {total[kern]:6d} | {items}")
