#!/usr/bin/env python3
"""Topology validation probes (BASELINE config 5): per-GPU HBM stream GB/s and the n x n P2P
bandwidth matrix, next to what NVML reports.  Prints one JSON document.

    python tools/probe.py [--gpus N] [--hbm-bytes B] [--p2p-bytes B] [--reps R]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import _lib as gpa  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--hbm-bytes", type=int, default=4 << 30)
    ap.add_argument("--p2p-bytes", type=int, default=1 << 30)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    peak = 6650.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    out = {"hbm_peak_gbs_measured": peak, "hbm": [], "nvlink_nominal_gbs_per_direction": 900.0,
           "nvlink_measured_reference_gbs": 770.0}
    for d in range(args.gpus):
        rd = gpa.probe_hbm(d, args.hbm_bytes, args.reps, mode=0)
        cp = gpa.probe_hbm(d, args.hbm_bytes, args.reps, mode=1)
        out["hbm"].append({"gpu": d, "bytes": args.hbm_bytes, "read_best_gbs": rd[0], "read_median_gbs": rd[1],
                           "copy_best_gbs": cp[0], "copy_median_gbs": cp[1], "copy_frac_of_measured_peak": cp[0] / peak,
                           "read_frac_of_measured_peak": rd[0] / peak})
    if args.gpus > 1:
        m, acc = gpa.probe_p2p(args.gpus, args.p2p_bytes, max(args.reps // 2, 3), mode=0)
        m2, _ = gpa.probe_p2p(args.gpus, args.p2p_bytes, max(args.reps // 2, 3), mode=1)
        m3, _ = gpa.probe_p2p(args.gpus, args.p2p_bytes, max(args.reps // 2, 3), mode=2)
        m4, _ = gpa.probe_p2p(args.gpus, args.p2p_bytes, max(args.reps // 2, 3), mode=3)
        off = ~np.eye(args.gpus, dtype=bool)
        out["p2p"] = {"bytes": args.p2p_bytes, "can_access_peer": acc.tolist(),
                      "kernel_peer_read_gbs": np.round(m, 1).tolist(), "memcpy_peer_gbs": np.round(m2, 1).tolist(),
                      "kernel_offdiag_min_mean_max": [float(m[off].min()), float(m[off].mean()), float(m[off].max())],
                      "memcpy_offdiag_min_mean_max": [float(m2[off].min()), float(m2[off].mean()), float(m2[off].max())],
                      "all_gpus_at_once": {
                          "what": "round s: every GPU c pulls from GPU (c + s) mod n at the same time (all NVLink ports busy both ways)",
                          "kernel_peer_read_gbs": np.round(m3, 1).tolist(), "memcpy_peer_gbs": np.round(m4, 1).tolist(),
                          "kernel_offdiag_min_mean_max": [float(m3[off].min()), float(m3[off].mean()), float(m3[off].max())],
                          "memcpy_offdiag_min_mean_max": [float(m4[off].min()), float(m4[off].mean()), float(m4[off].max())],
                          "aggregate_kernel_gbs_per_round_mean": float(m3[off].sum() / max(args.gpus - 1, 1))}}
    try:
        nv = gpa.Nvml()
        n = nv.device_count()
        parity, calls = nv.read_topology(gpa.TOPO_PARITY)
        aware, _ = nv.read_topology(gpa.TOPO_NVSWITCH_AWARE)
        out["nvml"] = {"devices": n, "uuids": [nv.device_uuid(i) for i in range(n)], "nvml_calls_per_build": calls,
                       "reference_nvml_calls_per_request": 2 * (1 + 13 * n + n * (n - 1) * (1 + 2 * 12)),
                       "parity_pcie_link": parity[:, :, 0].tolist(), "parity_nvlink": parity[:, :, 1].tolist(),
                       "aware_nvlink": aware[:, :, 1].tolist()}
        nv.close()
    except Exception as e:  # noqa: BLE001
        out["nvml"] = {"error": str(e)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
