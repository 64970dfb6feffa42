"""ThreadSanitizer over the host C++ of the library (replica.cpp, plugin.cpp, nvml_host.cpp): config 4's scenario - eight
threads calling Allocate while the XID watcher marks devices unhealthy and ListAndWatch data is rebuilt - must report no
data race.  The reference has exactly this race (d.Health written at server.go:260 while Allocate reads);
here the plugin bodies take a mutex.  The CUDA-side entry points are stubbed (tests/csrc/cuda_stubs.cpp): host-only build."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "k8s-gpu-sharing-plugin_b200", "csrc")


def test_allocate_under_health_churn_is_race_free(tmp_path, fake_nvml):
    _, fake_so = fake_nvml
    flags = ["-std=c++17", "-O1", "-g", "-fPIC", "-fsanitize=thread", "-I", os.path.join(ROOT, "include"), "-I/usr/local/cuda/include"]
    objs = []
    for src in (os.path.join(CSRC, "replica.cpp"), os.path.join(CSRC, "plugin.cpp"), os.path.join(CSRC, "nvml_host.cpp"),
                os.path.join(ROOT, "tests", "csrc", "cuda_stubs.cpp")):
        obj = str(tmp_path / (os.path.basename(src) + ".o"))
        r = subprocess.run(["g++"] + flags + ["-c", src, "-o", obj], capture_output=True, text=True)
        if r.returncode != 0 and "sanitize" in r.stderr:
            pytest.skip("no ThreadSanitizer runtime in this toolchain")
        assert r.returncode == 0, r.stderr[-2000:]
        objs.append(obj)
    lib = str(tmp_path / "libgpualloc.so")
    r = subprocess.run(["g++", "-shared", "-fsanitize=thread", "-o", lib] + objs + ["-ldl", "-lpthread"], capture_output=True, text=True)
    if r.returncode != 0 and "tsan" in r.stderr:
        pytest.skip("no ThreadSanitizer runtime in this toolchain")
    assert r.returncode == 0, r.stderr[-2000:]
    exe = str(tmp_path / "churn")
    r = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-pthread", "-fsanitize=thread", os.path.join(ROOT, "tools", "churn.cpp"), "-o", exe,
                        "-I", os.path.join(ROOT, "include"), "-L", str(tmp_path), "-lgpualloc", "-ldl", "-Wl,-rpath," + str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    run = subprocess.run([exe, fake_so, "2"], capture_output=True, text=True, timeout=300)
    assert "ThreadSanitizer" not in run.stderr and "ThreadSanitizer" not in run.stdout, run.stderr[-3000:]
    assert run.returncode == 0, run.stderr[-2000:]
    assert '"allocate_errors": 0' in run.stdout
