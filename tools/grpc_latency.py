#!/usr/bin/env python3
"""p50 / p99 of the kubelet-side gRPC round trip for Allocate and GetPreferredAllocation over the unix
socket, served by wire.py from the C++ plugin bodies.  replicas = 4 (32 virtual devices): the host branch, no GPU needed.
The transport is grpcio in Python on both ends, so these are upper bounds for a Go daemon; the C ABI call itself is in
profiles/r01_config2_stream_native_cpu.json.      python tools/grpc_latency.py [calls]"""
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import _lib as gpa  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import synthetic as syn  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import wire  # noqa: E402


def main():
    calls = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
    uu = syn.node_uuids(8, seed=1)
    devs = [{"id": uu[i], "index": str(i), "paths": ["/dev/nvidia%d" % i], "total_memory": 183359, "numa": i // 4, "nvml_index": i}
            for i in range(8)]
    plug = gpa.Plugin(devs, replicas=4, has_policy=True)
    out = {"what": "gRPC round trip over the unix socket (grpcio/Python on both ends), 8 GPUs x 4 replicas, microseconds",
           "calls": calls}
    with tempfile.TemporaryDirectory() as d:
        server, servicer = wire.serve(plug, os.path.join(d, "nvidia-gpu.sock"))
        c = wire.DevicePluginClient(os.path.join(d, "nvidia-gpu.sock"))
        ids = [x.ID for x in next(iter(c.ListAndWatch(wire.MSG["Empty"]()))).devices]
        rng = np.random.default_rng(2)
        lat_a, lat_p, lat_abi_a, lat_abi_p = [], [], [], []
        for i in range(calls + 200):
            free = [str(x) for x in rng.permutation(ids)[: int(rng.integers(8, 33))]]
            size = int(rng.choice([1, 2, 4, 8]))
            req = wire.MSG["PreferredAllocationRequest"]()
            cr = req.container_requests.add()
            cr.available_deviceIDs.extend(free)
            cr.allocation_size = size
            t0 = time.perf_counter()
            pref = list(c.GetPreferredAllocation(req).container_responses[0].deviceIDs)
            t1 = time.perf_counter()
            areq = wire.MSG["AllocateRequest"]()
            areq.container_requests.add().devicesIDs.extend(pref)
            t2 = time.perf_counter()
            c.Allocate(areq)
            t3 = time.perf_counter()
            plug.get_preferred_allocation(free, [], size)
            t4 = time.perf_counter()
            plug.allocate(pref)
            t5 = time.perf_counter()
            if i >= 200:
                lat_p.append(t1 - t0); lat_a.append(t3 - t2); lat_abi_p.append(t4 - t3); lat_abi_a.append(t5 - t4)
        for name, v in (("get_preferred_allocation_grpc", lat_p), ("allocate_grpc", lat_a),
                        ("get_preferred_allocation_ctypes", lat_abi_p), ("allocate_ctypes", lat_abi_a)):
            v = np.array(v) * 1e6
            out[name] = {"p50": round(float(np.percentile(v, 50)), 1), "p99": round(float(np.percentile(v, 99)), 1)}
        c.close()
        servicer.stop()
        server.stop(0)
    plug.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
