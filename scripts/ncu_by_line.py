"""Aggregates an `ncu --page source --csv` SASS listing per CUDA source line, using `nvdisasm --print-line-info` of the same
cubin.  Usage: python scripts/ncu_by_line.py <source.csv> <nvdisasm.txt> <kernel substring> [top N]"""
import collections
import csv
import re
import sys

src_csv, dis, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
# nvdisasm: address offset -> (file, line) inside the kernel's text section
line_of, cur, inside = {}, None, False
for ln in open(dis, errors="replace"):
    if ln.startswith(".text.") or ".section\t.text." in ln:
        inside = kern in ln
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur
rows = [r for r in csv.reader(open(src_csv, errors="replace"))]
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ci = {k: hdr.index(k) for k in ("Address", "Source", "# Samples", "Instructions Executed", "L1 Wavefronts Shared")}
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
base = None
agg = collections.defaultdict(lambda: [0, 0, 0, collections.Counter()])
tot = [0, 0, 0]
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    a = int(r[ci["Address"]], 16) if r[ci["Address"]].startswith("0x") else int(r[ci["Address"]])
    if base is None:
        base = a
    key = line_of.get(a - base, ("?", 0))
    f = lambda x: int(float(x.replace(",", ""))) if x not in ("", "-") else 0
    s, n, w = f(r[ci["# Samples"]]), f(r[ci["Instructions Executed"]]), f(r[ci["L1 Wavefronts Shared"]])
    e = agg[key]
    e[0] += s; e[1] += n; e[2] += w
    for i, h in stall_cols:
        e[3][h] += f(r[i])
    tot[0] += s; tot[1] += n; tot[2] += w
print("total samples %d  warp instructions %d  shared wavefronts %d" % tuple(tot))
print("%-22s %8s %6s %10s %6s %10s %6s  top stalls" % ("file:line", "samples", "%", "instr", "%", "smem wf", "%"))
for key, e in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    st = ", ".join("%s %d" % (h[6:], v) for h, v in e[3].most_common(3))
    print("%-22s %8d %6.2f %10d %6.2f %10d %6.2f  %s" % ("%s:%d" % key, e[0], 100.0 * e[0] / max(tot[0], 1), e[1], 100.0 * e[1] / max(tot[1], 1),
                                                   e[2], 100.0 * e[2] / max(tot[2], 1), st))
if len(sys.argv) > 5:  # phase ranges "name:lo-hi,..." over the main file
    print()
    for spec in sys.argv[5].split(","):
        name, rng = spec.split(":")
        lo, hi_ = map(int, rng.split("-"))
        s = n = w = 0
        st = collections.Counter()
        for (f_, l), e in agg.items():
            if f_.endswith(".cu") and lo <= l <= hi_:
                s += e[0]; n += e[1]; w += e[2]; st.update(e[3])
        print("%-10s lines %4d-%4d  samples %6d (%5.2f%%)  instr %11d (%5.2f%%)  smem wf %10d (%5.2f%%)  %s" % (
            name, lo, hi_, s, 100.0 * s / tot[0], n, 100.0 * n / tot[1], w, 100.0 * w / max(tot[2], 1),
            ", ".join("%s %d" % (h[6:], v) for h, v in st.most_common(4))))
