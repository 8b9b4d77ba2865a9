"""Counts the SASS mnemonics that prove the tcgen05 / TMEM / bulk-copy path per kernel of libdba_b200.so.
    python tools/sass_check.py [out.txt]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "dba_mod_b200", "ops", "libdba_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
WANT = ["UTCHMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UBLKCP", "UTMALDG", "SYNCS", "REDG", "RED.E", "ATOMG", "HMMA",
        "MEMBAR", "LDG.E.128", "STG.E.128", "MULTIMEM"]
per = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::", "", name).split("(")[0]
        per[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        per[cur]["_total"] += 1
        for w in WANT:
            if op.startswith(w):
                per[cur][w] += 1
lines = [f"{'kernel':58s} {'instrs':>7s}  tensor/TMEM/bulk-copy mnemonics"]
for k, c in per.items():
    tags = "  ".join(f"{w}={c[w]}" for w in WANT if c[w])
    lines.append(f"{k[:58]:58s} {c['_total']:7d}  {tags}")
out = "\n".join(lines) + "\n"
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write(out)
print(out)
