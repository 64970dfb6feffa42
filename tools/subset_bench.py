"""Runs the count-form replica-set batch kernel a few times on device-resident records (profiling target).
   python tools/subset_bench.py [--replicas 32] [--count 1048576] [--reps 5]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
import torch  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import _lib as gpa  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import synthetic as syn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--replicas", type=int, default=32)
    ap.add_argument("--count", type=int, default=1 << 20)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    ctx = gpa.Context(0)
    ctx.set_topology_links(syn.topo_two_sockets(8))
    ctx.set_device_order(syn.sorted_positions(syn.node_uuids(8, seed=1)))
    reqs = syn.random_subset_batch(8, a.replicas, a.count, seed=300 + a.replicas)
    d_in = torch.from_numpy(reqs.view(np.uint8).reshape(-1, 48)).cuda()
    d_out = torch.empty((len(reqs), 24), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(a.reps):
        e0.record(stream)
        ctx.subset_batch_device(d_in.data_ptr(), len(reqs), d_out.data_ptr(), stream=stream.cuda_stream)
        e1.record(stream)
        torch.cuda.synchronize()
        print("rep %d: %.3f ms, %.1f M decisions/s" % (i, e0.elapsed_time(e1), len(reqs) / e0.elapsed_time(e1) / 1e3))
    ctx.close()


if __name__ == "__main__":
    main()
