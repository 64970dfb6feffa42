#!/usr/bin/env python3
"""Summarise ncu output read back on the CPU box.
   python tools/ncu_summary.py launches <launches.csv>      per-kernel totals of a launch list
   python tools/ncu_summary.py raw <file.ncu-rep>            key metrics of the first captured launch
   python tools/ncu_summary.py sass <file.ncu-rep>           opcode histogram + hot address ranges
"""
import collections
import csv
import re
import subprocess
import sys


def launches(path):
    rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(list)
    for r in rows[1:]:
        if len(r) > vi:
            agg[r[ki][:72]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("%-74s n=%4d total=%9.3f ms (%5.1f%%) avg=%9.1f us" % (k, len(v), sum(v) / 1e6, 100 * sum(v) / tot, sum(v) / len(v) / 1e3))


KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "launch__shared_mem_per_block_dynamic",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "sass__inst_executed_shared_loads",
        "sass__inst_executed_global_loads", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.avg"]


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = dict(zip(hdr, zip(units, vals)))
    print("kernel:", d.get("Kernel Name", ("", ""))[1][:100])
    for k in KEYS:
        if k in d:
            print("%-80s %-16s %s" % (k, d[k][0], d[k][1]))
    stalls = [(float(v[1].replace(",", "")), k) for k, v in d.items() if "issue_stalled" in k and k.endswith("per_issue_active.ratio")]
    for v, k in sorted(stalls, reverse=True)[:7]:
        print("   stall %-62s %.2f" % (k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))


def sass(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[1]
    ci, si, ti = h.index("Instructions Executed"), h.index("Source"), h.index("Avg. Threads Executed")
    cur = []
    for r in rows[2:]:
        try:
            cur.append((r[si].strip(), float(r[ci]), float(r[ti] or 0)))
        except (ValueError, IndexError):
            pass
    tot = sum(c[1] for c in cur)
    op = collections.Counter()
    for s, v, _ in cur:
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", s)
        op[m.group(2).split(".")[0] if m else s[:8]] += v
    print("total warp instructions %.0f over %d SASS lines" % (tot, len(cur)))
    print("  ".join("%s %.1f%%" % (k, 100 * v / tot) for k, v in op.most_common(16)))
    for i in range(0, len(cur), 100):
        ch = cur[i:i + 100]
        sh = sum(c[1] for c in ch)
        if sh / tot > 0.01:
            print("lines %4d-%4d  %5.1f%%  avg threads %.1f" % (i, i + 99, 100 * sh / tot, sum(c[1] * c[2] for c in ch) / max(sh, 1)))


if __name__ == "__main__":
    {"launches": launches, "raw": raw, "sass": sass}[sys.argv[1]](sys.argv[2])
