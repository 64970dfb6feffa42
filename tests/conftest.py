import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402

entry.load_package()

if os.environ.get("GPA_TEST_LIB"):  # a differently built libgpualloc.so (tools/sanitize_cpu.sh: ASan + UBSan on the host code)
    from k8s_gpu_sharing_plugin_b200 import _lib as _gpa_lib
    _gpa_lib.LIB_PATH = os.environ["GPA_TEST_LIB"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def fake_nvml():
    """tests/csrc/fake_nvml.c built and loaded (a scriptable stand-in for libnvidia-ml.so.1) -> (ctypes handle, path)"""
    import ctypes as C
    import subprocess
    entry.build_native()
    here = os.path.dirname(os.path.abspath(__file__))
    src = os.path.join(here, "csrc", "fake_nvml.c")
    so = os.path.join(here, "csrc", "libfake-nvidia-ml.so.1")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O1", "-fPIC", "-shared", "-I/usr/local/cuda/include", "-o", so, src, "-lpthread"])
    F = C.CDLL(so, mode=C.RTLD_GLOBAL)
    F.fake_nvml_inject.argtypes = [C.c_int, C.c_ulonglong, C.c_ulonglong]
    F.fake_nvml_set_device.argtypes = [C.c_int, C.c_char_p, C.c_char_p]
    return F, so
