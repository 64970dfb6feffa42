"""Lifecycle shell: --resource-config grammar, Serve / Register against a fake kubelet, restart
when kubelet.sock is re-created.  CPU only: replicas > 1 takes the host branch of GetPreferredAllocation
(prioritizeDevices), so no CUDA context is needed.  The reference has no test for any of this (parity unpinned);
the parser is checked against the literal restatement in oracle/ref_server.py, the rest against the cited behaviour."""
import ctypes as C
import os
import tempfile
import threading
import time
from concurrent import futures

import grpc
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import lifecycle, wire
from oracle import ref_server


CASES = ["", "gpu:gpu:1", "gpu:shared-gpu:4", " gpu:shared:16 , mig-1g.5gb:small:-1 ,,", "gpu:a:1,gpu:b:2", "gpu:x", "gpu:x:y",
         "a:b:3000000000", "a:b:2147483647", "a:b:-2147483648", "a:b:-2147483649", "a:b:", "a:b:1:2", "a:b:+7", "a:b:-2",
         "a:b: 3", "a :b:3", "\tgpu:g:2\n", "a:b:1_0", "a:b:0x10", "::0", "gpu:gpu:1,bad", " gpu:g:1 ", "a:b:--1"]


@pytest.mark.parametrize("config", CASES)
def test_resource_config_grammar_matches_the_restatement(config):
    try:
        want = ref_server.parse_resource_config(config)
    except ValueError as e:
        with pytest.raises(ValueError) as got:
            gpa.parse_resource_config(config)
        assert str(got.value) == str(e)
        return
    got = gpa.parse_resource_config(config)
    assert {k: (v["name"], v["replicas"], v["auto_replicas"]) for k, v in got.items()} == want
    for name in list(want) + ["gpu", "absent"]:
        g = gpa.resource_config_get(got, name)
        assert (g["name"], g["replicas"], g["auto_replicas"]) == ref_server.resource_config_get(want, name)


def test_no_resource_config_means_zero_replicas():
    """the daemon started without --resource-config advertises no device"""
    assert gpa.resource_config_get(gpa.parse_resource_config(""), "gpu") == {"name": "gpu", "replicas": 0, "auto_replicas": False}


class FakeKubelet:
    """Registration service on kubelet.sock (api.proto:22-24): records what plugins send"""

    def __init__(self, plugin_dir):
        self.path = os.path.join(plugin_dir, lifecycle.KUBELET_SOCKET_NAME)
        self.requests = []
        self.cv = threading.Condition()
        self.server = None

    def _register(self, request, context):
        with self.cv:
            self.requests.append(request)
            self.cv.notify_all()
        return wire.MSG["Empty"]()

    def start(self):
        h = grpc.unary_unary_rpc_method_handler(self._register, request_deserializer=wire.MSG["RegisterRequest"].FromString,
                                                response_serializer=wire.MSG["Empty"].SerializeToString)
        self.server = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
        self.server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(wire.PACKAGE + ".Registration", {"Register": h}),))
        self.server.add_insecure_port("unix:" + self.path)
        self.server.start()

    def stop(self):
        self.server.stop(0)
        try:
            os.remove(self.path)
        except FileNotFoundError:
            pass

    def wait_for(self, n, timeout=10):
        with self.cv:
            return self.cv.wait_for(lambda: len(self.requests) >= n, timeout)


def make_devices(n):
    return [{"id": "GPU-%08d-aaaa-bbbb-cccc-dddddddddddd" % i, "index": str(i), "paths": ["/dev/nvidia%d" % i],
             "total_memory": 183359, "numa": i // 4, "nvml_index": i} for i in range(n)]


def test_start_serves_and_registers_then_restarts_with_kubelet():
    with tempfile.TemporaryDirectory() as d:
        kubelet = FakeKubelet(d)
        kubelet.start()
        cfg = gpa.parse_resource_config("gpu:shared-gpu:4")
        (resource, sock, kw), = lifecycle.plugins_for("none", cfg, make_devices(2))
        assert (resource, sock) == ("nvidia.com/shared-gpu", "nvidia-gpu.sock")        # mig-strategy.go:91-105
        daemon = lifecycle.PluginDaemon(kw, plugin_dir=d, socket_name=sock)
        stop, logs = threading.Event(), []
        t = threading.Thread(target=lifecycle.run, args=([daemon],), kwargs=dict(stop_event=stop, poll_s=0.05, log=logs.append))
        t.start()
        try:
            assert kubelet.wait_for(1)
            r = kubelet.requests[0]                                                       # server.go:224-232
            assert (r.version, r.endpoint, r.resource_name) == ("v1beta1", "nvidia-gpu.sock", "nvidia.com/shared-gpu")
            assert r.options.get_preferred_allocation_available and not r.options.pre_start_required
            # what kubelet does next: dial the endpoint it was given
            c = wire.DevicePluginClient(os.path.join(d, r.endpoint))
            devs = next(iter(c.ListAndWatch(wire.MSG["Empty"]()))).devices
            ids = [x.ID for x in devs]
            assert len(ids) == 8 and ids[0].endswith("-replica-0") and all(x.health == "Healthy" for x in devs)
            req = wire.MSG["PreferredAllocationRequest"]()
            cr = req.container_requests.add()
            cr.available_deviceIDs.extend(ids)
            cr.allocation_size = 2
            pref = list(c.GetPreferredAllocation(req).container_responses[0].deviceIDs)
            assert pref == gpa.prioritize_devices(ids, [], 2)[0]
            areq = wire.MSG["AllocateRequest"]()
            areq.container_requests.add().devicesIDs.extend(pref)
            envs = dict(c.Allocate(areq).container_responses[0].envs)
            assert set(envs["NVIDIA_VISIBLE_DEVICES"].split(",")) == {x.split("-replica-")[0] for x in pref}
            c.close()
            # kubelet restarts: its socket is re-created, every plugin restarts and registers again (main.go:296-300)
            kubelet.stop()
            time.sleep(0.2)
            kubelet.start()
            assert kubelet.wait_for(2)
            assert kubelet.requests[1].endpoint == "nvidia-gpu.sock"
            assert any("created, restarting" in m for m in logs)
        finally:
            stop.set()
            t.join(timeout=10)
            kubelet.stop()
        assert not t.is_alive() and not os.path.exists(os.path.join(d, sock))           # Stop() removes the socket (:159-163)


def test_start_fails_without_kubelet_and_retries():
    with tempfile.TemporaryDirectory() as d:
        (_, sock, kw), = lifecycle.plugins_for("none", gpa.parse_resource_config("gpu:gpu:2"), make_devices(1))
        daemon = lifecycle.PluginDaemon(kw, plugin_dir=d, socket_name=sock)
        old = lifecycle.DIAL_TIMEOUT_S
        lifecycle.DIAL_TIMEOUT_S = 0.3
        try:
            with pytest.raises(Exception):
                daemon.start()                       # Register cannot dial kubelet.sock (server.go:218-221) -> Stop(), error
            assert daemon.plugin is None and not os.path.exists(daemon.socket)
            stop, logs = threading.Event(), []
            t = threading.Thread(target=lifecycle.run, args=([daemon],), kwargs=dict(stop_event=stop, poll_s=0.05, retry_s=0.1, log=logs.append))
            t.start()
            time.sleep(1.0)
            kubelet = FakeKubelet(d)
            kubelet.start()                          # kubelet appears: the next retry registers
            assert kubelet.wait_for(1, timeout=15)
            stop.set()
            t.join(timeout=10)
            kubelet.stop()
            assert any("Could not contact Kubelet, retrying" in m for m in logs)
        finally:
            lifecycle.DIAL_TIMEOUT_S = old


def test_zero_replicas_still_registers_and_lists_nothing():
    """D2 end to end: physical GPUs exist, no --resource-config: the plugin starts, registers and advertises nothing"""
    with tempfile.TemporaryDirectory() as d:
        kubelet = FakeKubelet(d)
        kubelet.start()
        (resource, sock, kw), = lifecycle.plugins_for("none", gpa.parse_resource_config(""), make_devices(2))
        assert resource == "nvidia.com/gpu" and kw["replicas"] == 0
        daemon = lifecycle.PluginDaemon(kw, plugin_dir=d, socket_name=sock)
        try:
            daemon.start()
            assert kubelet.wait_for(1)
            c = wire.DevicePluginClient(daemon.socket)
            assert len(next(iter(c.ListAndWatch(wire.MSG["Empty"]()))).devices) == 0
            c.close()
        finally:
            daemon.stop()
            kubelet.stop()


XID = 0x8  # nvmlEventTypeXidCriticalError


def test_daemon_on_a_scripted_nvml_node_resends_the_list_on_an_xid(fake_nvml):
    """node_from_nvml + nvml_health_source against tests/csrc/fake_nvml.c: devices come from NVML, a critical XID makes
    ListAndWatch send the list again (server.go:258-262); in parity mode the replicas stay Healthy, with
    GPA_FIX_HEALTH_PROPAGATION every replica of the GPU flips."""
    F, FAKE_SO = fake_nvml
    F.fake_nvml_reset(2)
    uuids = ["GPU-%08x-1111-2222-3333-444444444444" % i for i in range(2)]
    for i, u in enumerate(uuids):
        F.fake_nvml_set_device(i, u.encode(), None)
    nv = gpa.Nvml(FAKE_SO)
    for fixes, want_health in ((0, "Healthy"), (gpa.FIX_HEALTH_PROPAGATION, "Unhealthy")):
        with tempfile.TemporaryDirectory() as d:
            kubelet = FakeKubelet(d)
            kubelet.start()
            devices = lifecycle.node_from_nvml(nv)
            assert [x["id"] for x in devices] == uuids and devices[1]["paths"] == ["/dev/nvidia1"]
            (_, sock, kw), = lifecycle.plugins_for("none", gpa.parse_resource_config("gpu:gpu:2"), devices, fixes=fixes)
            daemon = lifecycle.PluginDaemon(kw, plugin_dir=d, socket_name=sock,
                                            health_source=lifecycle.nvml_health_source(nv, uuids, disable_env="", wait_timeout_ms=50))
            try:
                daemon.start()
                assert kubelet.wait_for(1)
                c = wire.DevicePluginClient(daemon.socket)
                stream = c.ListAndWatch(wire.MSG["Empty"]())
                first = next(stream)
                assert [x.health for x in first.devices] == ["Healthy"] * 4
                time.sleep(0.2)                       # the watcher thread has registered for events by now
                F.fake_nvml_inject(1, XID, 79)        # "GPU has fallen off the bus": critical
                second = next(stream)
                by_gpu = {x.ID: x.health for x in second.devices}
                assert all(h == "Healthy" for i, h in by_gpu.items() if i.startswith(uuids[0]))
                assert [h for i, h in by_gpu.items() if i.startswith(uuids[1])] == [want_health] * 2
                stream.cancel()
                c.close()
            finally:
                daemon.stop()
                kubelet.stop()
    nv.close()


def test_node_from_nvml_takes_memory_minor_and_numa_node_from_nvml(fake_nvml, monkeypatch, tmp_path):
    """buildDevice (cmd/nvidia-device-plugin/nvidia.go:162-178): TotalMemory = free + used MiB, path /dev/nvidia<MINOR>,
    Topology from the bus id's numa_node.  With auto replicas (`gpu:gpu:-1`) the daemon must advertise
    TotalMemory / 1000 replicas per GPU (server.go:100-103), each with its TopologyInfo."""
    F, FAKE_SO = fake_nvml
    F.fake_nvml_set_memory.argtypes = [C.c_int, C.c_ulonglong, C.c_ulonglong]
    F.fake_nvml_set_minor.argtypes = [C.c_int, C.c_int]
    F.fake_nvml_reset(2)
    uuids = ["GPU-%08x-1111-2222-3333-444444444444" % i for i in range(2)]
    buses = ["00000000:1B:00.0", "00000000:9C:00.0"]
    for i, u in enumerate(uuids):
        F.fake_nvml_set_device(i, u.encode(), buses[i].encode())
    MiB = 1024 * 1024
    F.fake_nvml_set_memory(0, 4000 * MiB + 5, 999 * MiB + MiB - 1)   # 4000 + 999 MiB: each part rounds down on its own
    F.fake_nvml_set_memory(1, 2500 * MiB, 0)
    F.fake_nvml_set_minor(0, 3)
    F.fake_nvml_set_minor(1, 0)
    # /sys/bus/pci/devices/<bus id without its first four characters, lower case>/numa_node (nvml.go:314-332)
    for bus, node in ((buses[0], "1\n"), (buses[1], "-1\n")):
        d = tmp_path / bus[4:].lower()
        d.mkdir()
        (d / "numa_node").write_text(node)
    monkeypatch.setenv("GPA_SYSFS_PCI_DEVICES", str(tmp_path))
    nv = gpa.Nvml(FAKE_SO)
    try:
        assert nv.device_info(0) == {"total_memory": 4999, "minor": 3, "numa": 1, "bus_id": buses[0]}
        devices = lifecycle.node_from_nvml(nv)
        assert [(x["paths"], x["total_memory"], x["numa"]) for x in devices] == [(["/dev/nvidia3"], 4999, 1), (["/dev/nvidia0"], 2500, -1)]
        (_, sock, kw), = lifecycle.plugins_for("none", gpa.parse_resource_config("gpu:gpu:-1"), devices)
        plug = gpa.Plugin(**kw)
        try:
            listed = plug.api_devices()
            assert len(listed) == 4 + 2                                # 4999 / 1000 and 2500 / 1000 replicas
            assert [x["ID"] for x in listed][:2] == [uuids[0] + "-replica-0", uuids[0] + "-replica-1"]
            assert all(x.get("topology") == {"nodes": [{"ID": 1}]} for x in listed[:4])
            assert all("topology" not in x for x in listed[4:])        # a negative numa_node means none (nvml.go:325-328)
        finally:
            plug.close()
        F.fake_nvml_set_minor(1, -1)                                   # NVML_ERROR_NOT_SUPPORTED: ErrUnsupportedGPU (nvml.go:467-469)
        with pytest.raises(RuntimeError):
            lifecycle.node_from_nvml(nv)
    finally:
        nv.close()


def test_health_event_refreshes_the_link_matrix_before_the_list_is_resent(fake_nvml):
    """the daemon reads the NVML link matrix once, so an XID is where that cache is invalidated:
    refresh_topology runs for every event, before device_unhealthy; a failing re-read keeps the old matrix and is logged"""
    F, FAKE_SO = fake_nvml
    F.fake_nvml_reset(2)
    uuids = ["GPU-%08x-1111-2222-3333-444444444444" % i for i in range(2)]
    for i, u in enumerate(uuids):
        F.fake_nvml_set_device(i, u.encode(), None)
    nv = gpa.Nvml(FAKE_SO)
    order, logs = [], []

    def refresh():
        order.append("refresh")
        if len([x for x in order if x == "refresh"]) == 2:
            raise RuntimeError("GPU is lost")

    src = lifecycle.nvml_health_source(nv, uuids, disable_env="", wait_timeout_ms=50, refresh_topology=refresh, log=logs.append)
    stop = threading.Event()
    t = threading.Thread(target=src, args=(stop, lambda i: order.append(("unhealthy", i))))
    t.start()
    try:
        time.sleep(0.2)
        F.fake_nvml_inject(1, XID, 79)
        F.fake_nvml_inject(0, XID, 48)
        deadline = time.time() + 5
        while len(order) < 4 and time.time() < deadline:
            time.sleep(0.02)
    finally:
        stop.set()
        t.join()
        nv.close()
    assert order == ["refresh", ("unhealthy", 1), "refresh", ("unhealthy", 0)]
    assert len(logs) == 1 and "keeping the cached one" in logs[0] and "GPU is lost" in logs[0]
