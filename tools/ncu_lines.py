#!/usr/bin/env python3
"""Per-source-line instruction counts of the first kernel in an ncu report (needs -lineinfo and --import-source on).
   python tools/ncu_lines.py <file.ncu-rep> [top]"""
import collections
import csv
import subprocess
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
    hdr, cur, curfile, tot = None, None, None, 0
    agg = collections.OrderedDict()
    for r in csv.reader(out.splitlines()):
        if len(r) == 2 and r[0] == "File Path":
            curfile = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            ii, ti, si = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
            continue
        if hdr is None or len(r) <= ti:
            continue
        if r[0] != "":
            cur = (curfile, int(r[0]), r[1].strip()[:110])
            agg.setdefault(cur, [0, 0, 0])
        else:
            try:
                agg[cur][0] += int(r[ii]); agg[cur][1] += int(r[ti]); agg[cur][2] += int(r[si]); tot += int(r[ii])
            except (ValueError, TypeError):
                pass
    print("total warp instructions", tot)
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%5.1f%% thr %4.1f samp %5d  %s:%d  %s" % (100 * v[0] / tot, v[1] / max(v[0], 1), v[2], k[0], k[1], k[2]))


if __name__ == "__main__":
    main()
