"""Turns the raw ncu outputs in gpurun_out/ into the text summaries committed under profiles/.
Usage: python scripts/summarize_profiles.py <tag>      (e.g. v7)"""
import collections
import csv
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "vX"
KEEP = ("Kernel Name", "Block Size", "Grid Size", "pipe_tensor", "dram__bytes", "dram__cycles_elapsed.avg.per_second",
        "gpc__cycles_elapsed.avg.per_second", "gpu__dram_throughput", "gpu__time_duration.sum", "launch__cluster_dim_x",
        "launch__grid_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "lts__t_sector_hit_rate.pct",
        "sm__cycles_active.avg", "sm__cycles_elapsed", "sm__inst_executed.sum", "smsp__issue_active.avg.pct",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__warp_issue_stalled", "sm__throughput.avg.pct",
        "l1tex__throughput.avg.pct", "lts__throughput.avg.pct", "smsp__inst_executed_op_shared")


def launches(path, out, header):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if not r[0].isdigit():
            continue
        a = agg.setdefault(r[ki], [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", "")) / 1e6
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write(header)
        for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write("%-100s n=%4d total_ms=%10.3f share=%.4f\n" % (k[:100], a[0], a[1], a[1] / tot))


def full(rep, out, header):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw"], capture_output=True, text=True).stdout
    with open(out, "w") as f:
        f.write(header)
        for line in txt.splitlines():
            s = line.strip()
            if any(k in s for k in KEEP) or s.startswith("void ") or s.startswith("vq::"):
                f.write(s + "\n")


if __name__ == "__main__":
    import os
    g = "gpurun_out/"
    if os.path.exists(g + "launches.csv"):
        launches(g + "launches.csv", "profiles/r01_ncu_launches_bench_%s.txt" % tag,
                 "# ncu --metrics gpu__time_duration.sum --clock-control none -c 400: python bench.py --steps 2 --warmup 3 --no-aux --no-cpu-baseline\n"
                 "# (cold-cache, serialised; compare SHARES). at::* = torch's synthetic-data generation, before the timed region.\n")
    for name, cmd in (("prof_gemm_topk", "scripts/prof_gemm.py 18944 1048576 2"), ("prof_triplet_tc", "scripts/time_triplet.py"),
                      ("prof_moment", "scripts/prof_moment.py")):
        if os.path.exists(g + name + ".ncu-rep"):
            full(g + name + ".ncu-rep", "profiles/r01_ncu_%s_full_%s.txt" % (name[5:], tag),
                 "# ncu --set full --clock-control none --import-source on: python %s\n" % cmd)
