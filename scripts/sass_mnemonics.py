"""Per-kernel count of the SASS mnemonics that show which Blackwell units a kernel uses (tcgen05 MMA = UTCHMMA / UTCQMMA,
TMA = UTMALDG / UTMASTG, TMEM loads = LDTM, mbarriers = SYNCS, 256-bit global accesses, packed fp32 = FFMA2 / FADD2),
from `cuobjdump -sass` of the built library.  Runs without a GPU:
    python scripts/sass_mnemonics.py > profiles/r02_sass_mnemonics.txt"""
import collections
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "video-query-in-the-wild_b200", "libvqwild.so")
KEYS = [("tcgen05.mma", r"\bUTC[HQIO]?MMA\b(?!\.2CTA)"), ("mma 2CTA", r"\bUTC[HQIO]?MMA\.2CTA"), ("tcgen05.commit", r"\bUTCBAR"),
        ("TMEM alloc", r"\bUTCATOMSWS"), ("tcgen05.ld", r"\bLDTM"), ("TMA load", r"\bUTMALDG"), ("TMA store", r"\bUTMASTG"),
        ("mbarrier", r"\bSYNCS"), ("cluster bar", r"\bUCGABAR"), ("LDG.256", r"\bLDG\.[A-Z0-9_.]*256"),
        ("STG.256", r"\bSTG\.[A-Z0-9_.]*256"), ("FFMA2", r"\bFFMA2\b"), ("FADD2", r"\bFADD2\b"), ("FFMA", r"\bFFMA\b"),
        ("LDS", r"\bLDS\b"), ("SHFL", r"\bSHFL\b"), ("ATOMG/RED", r"\b(ATOMG|REDG|RED)\b")]
KEYS = [(n, re.compile(p)) for n, p in KEYS]
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
counts, order, cur, lines = {}, [], None, collections.Counter()
it = iter(names)
for ln in sass.split("\n"):
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = next(it)
        cur = re.sub(r"\(.*", "", cur).replace("vq::", "")
        cur = re.sub(r"^void ", "", cur)
        while cur in counts:
            cur += "'"
        counts[cur] = collections.Counter()
        order.append(cur)
        continue
    if cur is None or "/*" not in ln:
        continue
    body = ln.split("*/", 1)[-1]
    if not re.search(r"\b[A-Z][A-Z0-9_]+", body):
        continue
    lines[cur] += 1
    for n, p in KEYS:
        if p.search(body):
            counts[cur][n] += 1
print("SASS mnemonics per kernel of %s (sm_100a; `cuobjdump -sass`, scripts/sass_mnemonics.py)" % os.path.basename(so))
print("columns: " + " | ".join(n for n, _ in KEYS))
w = max(len(k) for k in order)
print("%-*s %7s " % (w, "kernel", "instr") + " ".join("%*s" % (max(len(n), 5), n) for n, _ in KEYS))
for k in sorted(order, key=lambda k: -counts[k]["tcgen05.mma"] - counts[k]["mma 2CTA"] - 0.001 * counts[k]["TMA load"]):
    print("%-*s %7d " % (w, k, lines[k]) + " ".join("%*s" % (max(len(n), 5), counts[k][n] or ".") for n, _ in KEYS))
