#!/usr/bin/env python3
"""Single-request search latency over (n, k): enumeration vs subset DP vs the oracle (one core).
   python tools/search_sweep.py [--oracle]   -> one JSON line per case"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import _lib as gpa  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import synthetic as syn  # noqa: E402
from oracle import oracle as orc  # noqa: E402

with_oracle = "--oracle" in sys.argv
ctx = gpa.Context(0)
for n, k in [(8, 2), (8, 3), (8, 4), (10, 2), (10, 3), (12, 2), (12, 3), (12, 4), (13, 3), (14, 2), (14, 4), (16, 2), (16, 3), (16, 4),
             (16, 5), (16, 6), (16, 8)]:
    t = syn.topo_random(n, 77)
    W, nl = syn.links_to_weights(t)
    ctx.set_topology_links(t)
    row = {"n": n, "k": k, "partitions": orc.partition_count(n, k)}
    for name, flag in (("enum", gpa.MODE_FORCE_ENUM), ("dp", gpa.MODE_FORCE_DP)):
        kms, call = [], []
        for i in range(8):
            t0 = time.perf_counter()
            rc, res = ctx.besteffort(list(range(n)), [], k, flag)
            call.append(time.perf_counter() - t0)
            kms.append(ctx.search_stats()["kernel_ms"])
        row[name + "_kernel_us"] = round(float(np.median(kms[2:])) * 1e3, 1)
        row[name + "_call_us"] = round(float(np.median(call[2:])) * 1e6, 1)
        row[name + "_answer"] = res
    assert row["enum_answer"] == row["dp_answer"]
    if with_oracle and row["partitions"] <= 30000000:
        t0 = time.perf_counter()
        want = orc.besteffort(W, nl, list(range(n)), [], k)
        row["oracle_call_us"] = round((time.perf_counter() - t0) * 1e6, 1)
        assert want[1] == row["dp_answer"]
    del row["enum_answer"]
    print(json.dumps(row))
