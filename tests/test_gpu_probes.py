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
