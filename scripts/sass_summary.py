"""Count the SASS mnemonics that prove tcgen05 / TMEM / TMA use, per kernel of the built library.
    python scripts/sass_summary.py > profiles/sass_summary.txt"""
import collections
import re
import subprocess
import sys

LIB = "mink_octtree_stablediffusion_b200/lib/libmink_b200.so"
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMALDG.GATHER4", "UBLKCP", "LDGSTS", "SYNCS", "FFMA", "ATOM", "RED"]
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()[:90]
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        for k in KEYS:
            if k == "UTMALDG.GATHER4":
                if "UTMALDG" in op and "GATHER4" in op:
                    counts[cur][k] += 1
            elif op.split(".")[0] == k:
                counts[cur][k] += 1
print("SASS mnemonic counts per kernel of", LIB, "(cuobjdump -sass; sm_100a)")
print(f"{'kernel':92s}" + "".join(f"{k[:9]:>10s}" for k in KEYS))
tot = collections.Counter()
for name, c in counts.items():
    if any(c[k] for k in KEYS[:7]) or "conv" in name:
        print(f"{name:92s}" + "".join(f"{c[k]:10d}" for k in KEYS))
    tot.update(c)
print(f"{'TOTAL (all kernels)':92s}" + "".join(f"{tot[k]:10d}" for k in KEYS))
