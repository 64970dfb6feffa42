"""Runs the literal replica-set kernel on the largest case the configs define (4 of 256 replicas: C(256, 4) combinations)
a few times (profiling target).   python tools/subset_literal.py [--reps 5]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import _lib as gpa  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import synthetic as syn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    ctx = gpa.Context(0)
    ctx.set_topology_links(syn.topo_dgx1v_like())
    phys = [g for g in range(8) for _ in range(32)]
    for i in range(a.reps):
        rc, pos = ctx.subset_besteffort(phys, [], 4, gpa.MODE_FORCE_ENUM)
        st = ctx.search_stats()
        print("rep %d: rc %d %r, %d combinations, %.3f ms, %.1f G/s" % (i, rc, pos, st["partitions"], st["kernel_ms"],
                                                                        st["partitions"] / max(st["kernel_ms"], 1e-9) / 1e6))
    ctx.close()


if __name__ == "__main__":
    main()
