"""SASS evidence for libdba_b200.so: per kernel, the count of the mnemonics that prove the tcgen05 / TMEM / bulk-copy /
NVLS paths, plus real listing excerpts (the MMA issue loops of conv_tc / wgrad_tc, the switch-reduced load loop and the
release / acquire flag protocol of the fused collectives).
    python tools/sass_check.py [counts.txt [excerpts.txt]]

What to look for (PTX -> SASS): tcgen05.mma -> UTCHMMA (kind::f16 and kind::tf32), tcgen05.ld -> LDTM, tcgen05.commit
-> UTCBAR, cp.async.bulk -> UBLKCP, multimem.ld_reduce -> LDGMC.E.ADD.F32 (multimem.st is an ordinary STG whose address
lies in the multicast mapping), st.release.sys / ld.acquire.sys -> STG / LDG .E.STRONG.SYS, griddepcontrol -> ACQBULK /
PREEXIT-class control, red.global.add.v4.f32 -> RED / REDG."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "dba_mod_b200", "ops", "libdba_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
WANT = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "REDG", "RED.E",
        "ATOMG", "HMMA", "MEMBAR", "LDG.E.128", "STG.E.128", "LDGMC", "LDG.E.STRONG.SYS", "STG.E.STRONG.SYS",
        "STG.E.128.STRONG.SYS", "MULTIMEM"]
EXCERPTS = [("conv_tc_kernel<4, true>", "UTCHMMA", 14), ("conv_tc_kernel<4, false>", "UTCHMMA", 14),
            ("wgrad_tc_kernel<true>", "UTCHMMA", 14), ("fedavg_p2p_kernel", "LDGMC", 16),
            ("fedavg_p2p_kernel", "STG.E.STRONG.SYS", 8), ("weiszfeld_p2p_kernel", "LDG.E.STRONG.SYS", 8)]
body = collections.OrderedDict()
per = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::", "", name).split("(")[0]
        per[cur] = collections.Counter()
        body[cur] = []
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        body[cur].append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line.rstrip()))
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
ex = []
for kern, needle, ctx in EXCERPTS:
    for name, lines_ in body.items():
        if kern not in name:
            continue
        hit = next((i for i, l in enumerate(lines_) if needle in l), None)
        if hit is None:
            continue
        ex.append(f"==== {name}: first {needle} (+/- {ctx} instructions)")
        ex += lines_[max(0, hit - ctx // 2): hit + ctx]
        ex.append("")
        break
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write("\n".join(ex) + "\n")
