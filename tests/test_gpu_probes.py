"""The probe kernels run and report sane bandwidths (small sizes; the full-size numbers come from
tools/probe.py and live under profiles/)."""
import json
import os

import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa

pytestmark = pytest.mark.gpu


def test_hbm_stream_probe_reads_and_copies_at_hbm_rates():
    peak = 6650.0
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    read_best, read_med = gpa.probe_hbm(0, 1 << 30, 10, mode=0)
    copy_best, copy_med = gpa.probe_hbm(0, 1 << 30, 10, mode=1)
    # > L2 (126 MB) working sets, so these are DRAM rates; a B200 streams several TB/s
    assert 0.5 * peak < read_best < 1.3 * peak and read_med <= read_best
    assert 0.5 * peak < copy_best < 1.3 * peak and copy_med <= copy_best


def test_p2p_probe_single_gpu_diagonal():
    m, acc = gpa.probe_p2p(1, 256 << 20, 3, mode=0)
    assert m.shape == (1, 1) and m[0, 0] > 500 and acc[0, 0] == 1
    m2, _ = gpa.probe_p2p(1, 256 << 20, 3, mode=1)
    assert m2[0, 0] > 500
    for mode in (2, 3):  # the all-GPUs-at-once rounds degenerate to the local copy
        m3, _ = gpa.probe_p2p(1, 256 << 20, 3, mode=mode)
        assert m3[0, 0] > 500


def test_probe_argument_checks():
    with pytest.raises(gpa.GpaError):
        gpa.probe_hbm(0, 1024, 1, mode=0)
    with pytest.raises(gpa.GpaError):
        gpa.probe_p2p(99, 1 << 30, 1, mode=0)


def test_lut_probe_reports_a_plausible_ceiling():
    g = gpa.probe_lut(0, 8, 2)
    # 148 SMs x 32 shared-memory lanes x ~1.9 GHz = ~9 000 G lookups/s at best
    assert 500 < g < 12000
    with pytest.raises(gpa.GpaError):
        gpa.probe_lut(0, 20, 1)


def test_p2p_matrix_against_nvml_on_two_or_more_gpus():
    """config 5 on a multi-GPU box (skipped on one GPU): the n x n peer bandwidth matrix, both ways of moving the bytes,
    against what NVML reports.  On an NVSwitch machine every off-diagonal pair runs at NVLink speed (the pool measures
    ~770 GB/s one way of 900 nominal); nvml.GetNVLink as the reference calls it sees NO link there (the far end of every
    link is the switch, nvml.go:617-658 with its TODO at :655), while GPA_TOPO_NVSWITCH_AWARE reports links for exactly
    the pairs the probe finds fast."""
    import numpy as np
    import torch
    n = min(torch.cuda.device_count(), 8)
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    kern, acc = gpa.probe_p2p(n, 1 << 30, 5, mode=0)
    copy, _ = gpa.probe_p2p(n, 1 << 30, 5, mode=1)
    off = ~np.eye(n, dtype=bool)
    assert acc[off].all()
    nvlink_measured = 770.0
    assert kern[off].min() >= 0.8 * nvlink_measured, kern
    assert copy[off].min() >= 0.8 * nvlink_measured, copy
    assert abs(kern[off].mean() - copy[off].mean()) <= 0.03 * copy[off].mean()      # the two modes agree
    assert (np.abs(kern[off] - copy[off]) <= 0.10 * copy[off]).all()
    # all GPUs pulling at once (round s: GPU c reads GPU (c + s) mod n): the switch gives every pair its full rate
    # Peer LOADS send a request packet for every line they pull (268 MB of nvltx per GiB, profiles/r02_p2p_nvlink_counters_*),
    # and with every port busy in both directions those requests share the wire with the other GPU's data: the pull rate
    # settles at 776 x 1.208 / (1.208 + 0.268) = 636 GB/s, while cudaMemcpyPeerAsync keeps its 774.
    busy, _ = gpa.probe_p2p(n, 1 << 30, 5, mode=2)
    assert busy[off].min() >= 0.75 * nvlink_measured, busy
    # (mode 3, cudaMemcpyPeerAsync from every GPU at once, is reported by tools/probe.py but not asserted: on 8 GPUs the
    # driver runs some of those copies one after the other - 110 to 775 GB/s per pair - which says nothing about a link)
    # the same pairs as NVML sees them; CUDA device i -> NVML index by UUID
    nv = gpa.Nvml()
    try:
        nvml_uuids = [nv.device_uuid(j) for j in range(nv.device_count())]
        to_nvml = [nvml_uuids.index("GPU-" + str(torch.cuda.get_device_properties(i).uuid)) for i in range(n)]
        parity, _ = nv.read_topology(gpa.TOPO_PARITY)
        aware, _ = nv.read_topology(gpa.TOPO_NVSWITCH_AWARE)
        for i in range(n):
            for j in range(n):
                if i == j:
                    continue
                fast = kern[i, j] > 300.0          # PCIe peer traffic would stay below ~60 GB/s
                assert (aware[to_nvml[i], to_nvml[j], 1] != 0) == fast, (i, j, kern[i, j])
                assert parity[to_nvml[i], to_nvml[j], 1] == 0
    finally:
        nv.close()
