#!/usr/bin/env python3
"""One single-request search, for ncu: python tools/search_one.py N K [enum|dp] [reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import _lib as gpa  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import synthetic as syn  # noqa: E402

n, k = int(sys.argv[1]), int(sys.argv[2])
flag = {"enum": gpa.MODE_FORCE_ENUM, "dp": gpa.MODE_FORCE_DP}.get(sys.argv[3] if len(sys.argv) > 3 else "dp", 0)
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
ctx = gpa.Context(0)
ctx.set_topology_links(syn.topo_random(n, 77))
for _ in range(reps):
    print(ctx.besteffort(list(range(n)), [], k, flag), ctx.search_stats())
ctx.close()
