#!/usr/bin/env python3
"""Brief of one `ncu --set full` report: the metrics the round's write-ups quote.  usage: ncu_brief.py report.ncu-rep [out.csv]
(the raw page is exported with `ncu -i report --page raw --csv`; with a second argument the CSV is kept)"""
import csv
import io
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__bytes.sum.per_second',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'lts__t_sector_hit_rate.pct', 'smsp__inst_executed_op_local_ld.sum', 'smsp__inst_executed_op_local_st.sum',
        'sm__cycles_elapsed.max', 'smsp__cycles_active.avg']


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(raw)
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = dict(zip(hdr, vals))
        print("==", d.get("Kernel Name", "?")[:90])
        for w in WANT:
            if w in d:
                print("  %-62s %s %s" % (w, d[w], units[hdr.index(w)]))
        stalls = sorted(((float(d[h]), h) for h in hdr if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and d[h]), reverse=True)
        for v, h in stalls[:6]:
            print("  stall %-40s %.2f" % (h.split("issue_stalled_")[1].split("_per_issue")[0], v))


if __name__ == "__main__":
    main()
