#!/usr/bin/env python3
"""Compares the SASS of every kernel in two builds of libgpualloc.so (cuobjdump -sass dumps), instruction by instruction
with addresses and encodings stripped.  Used to check, without a GPU, that a refactor left the device code unchanged.
   cuobjdump -sass old.so > a.txt; cuobjdump -sass new.so > b.txt; python tools/sass_diff.py a.txt b.txt"""
import difflib
import re
import sys


def funcs(path):
    d, cur = {}, None
    for line in open(path):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            d[cur] = []
            continue
        if cur is not None:
            m2 = re.search(r"/\*[0-9a-f]{4,}\*/\s+(.*?)\s*;", line)
            if m2:
                d[cur].append(m2.group(1))
    return d


def main():
    a, b = funcs(sys.argv[1]), funcs(sys.argv[2])
    rc = 0
    for k in sorted(set(a) | set(b)):
        if a.get(k) == b.get(k):
            print("same  %5d  %s" % (len(a[k]), k[:100]))
            continue
        rc = 1
        print("DIFF  %5d -> %5d  %s" % (len(a.get(k, [])), len(b.get(k, [])), k[:100]))
        diff = [l for l in difflib.unified_diff(a.get(k, []), b.get(k, []), lineterm="", n=0) if not l.startswith(("---", "+++", "@@"))]
        print("\n".join("      " + l for l in diff[:20]))
    return rc


if __name__ == "__main__":
    sys.exit(main())
