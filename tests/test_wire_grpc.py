"""Wire-level check of the kubelet device-plugin v1beta1 surface (SURVEY.md §8b "outer wire API"): a gRPC client
talks to the shim over a unix socket with messages built from the reference's field numbers; the answers must be
what the reference restatement gives, and the bytes must decode as the reference's messages do."""
import os
import tempfile

import grpc
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import synthetic as syn
from k8s_gpu_sharing_plugin_b200 import wire
from oracle import ref_server


def make_devices(n):
    uuids = syn.node_uuids(n, seed=1)
    return [{"id": uuids[i], "index": str(i), "paths": ["/dev/nvidia%d" % i], "total_memory": 183359, "numa": i // 4,
             "nvml_index": i} for i in range(n)]


@pytest.fixture()
def served():
    import __graft_entry__ as entry
    entry.build_native()
    devices = make_devices(8)
    plugin = gpa.Plugin(devices, replicas=4, has_policy=True, fixes=gpa.FIX_HEALTH_PROPAGATION,
                        resource_name="nvidia.com/sharedgpu")
    ref = ref_server.Plugin(devices, replicas=4, has_policy=True, resource_name="nvidia.com/sharedgpu")
    d = tempfile.mkdtemp()
    sock = os.path.join(d, "nvidia-sharedgpu.sock")
    server, servicer = wire.serve(plugin, sock)
    client = wire.DevicePluginClient(sock)
    yield client, servicer, ref, devices
    client.close()
    servicer.stop()
    server.stop(0)
    plugin.close()


def test_field_numbers_are_the_references():
    """spot-check the encoded bytes: tag = (field number << 3) | wire type, as api.proto:23-211 numbers them"""
    r = wire.MSG["ContainerPreferredAllocationRequest"](available_deviceIDs=["a"], must_include_deviceIDs=["b"], allocation_size=3)
    assert r.SerializeToString() == b"\x0a\x01a" + b"\x12\x01b" + b"\x18\x03"
    o = wire.MSG["DevicePluginOptions"](pre_start_required=True, get_preferred_allocation_available=True)
    assert o.SerializeToString() == b"\x08\x01\x10\x01"
    d = wire.MSG["Device"](ID="x", health="Healthy")
    d.topology.nodes.add().ID = 1
    assert d.SerializeToString() == b"\x0a\x01x" + b"\x12\x07Healthy" + b"\x1a\x04\x0a\x02\x08\x01"
    a = wire.MSG["ContainerAllocateResponse"]()
    a.envs["K"] = "V"
    a.mounts.add(container_path="/c", host_path="/h", read_only=True)
    a.devices.add(container_path="/d", host_path="/e", permissions="rw")
    assert a.SerializeToString() == (b"\x0a\x06\x0a\x01K\x12\x01V" + b"\x12\x0a\x0a\x02/c\x12\x02/h\x18\x01" +
                                     b"\x1a\x0c\x0a\x02/d\x12\x02/e\x1a\x02rw")


def test_options_preferred_allocation_and_allocate_over_the_socket(served):
    client, servicer, ref, devices = served
    o = client.GetDevicePluginOptions(wire.MSG["Empty"]())
    assert (o.pre_start_required, o.get_preferred_allocation_available) == (False, True)
    ids = [d["ID"] for d in ref.api_devices()]
    req = wire.MSG["PreferredAllocationRequest"]()
    cases = [(ids, [], 4), (ids[:9], [ids[3]], 2), (list(reversed(ids)), [], 8), (ids[:2], [], 1)]
    for avail, must, size in cases:
        c = req.container_requests.add()
        c.available_deviceIDs.extend(avail)
        c.must_include_deviceIDs.extend(must)
        c.allocation_size = size
    resp = client.GetPreferredAllocation(req)
    assert len(resp.container_responses) == len(cases)
    for (avail, must, size), got in zip(cases, resp.container_responses):
        assert list(got.deviceIDs) == ref.get_preferred_allocation(avail, must, size)
    areq = wire.MSG["AllocateRequest"]()
    for got in resp.container_responses:
        areq.container_requests.add().devicesIDs.extend(got.deviceIDs)
    aresp = client.Allocate(areq)
    for got, want_ids in zip(aresp.container_responses, resp.container_responses):
        want = ref.allocate(list(want_ids.deviceIDs))
        assert dict(got.envs) == want["envs"] and len(got.mounts) == len(want["mounts"])
    assert client.PreStartContainer(wire.MSG["PreStartContainerRequest"](devicesIDs=ids[:1])) == wire.MSG["PreStartContainerResponse"]()


def test_errors_travel_as_unknown_status_with_the_reference_text(served):
    client, servicer, ref, devices = served
    areq = wire.MSG["AllocateRequest"]()
    areq.container_requests.add().devicesIDs.extend(["nope-replica-0"])
    with pytest.raises(grpc.RpcError) as e:
        client.Allocate(areq)
    assert e.value.code() == grpc.StatusCode.UNKNOWN
    assert e.value.details() == "invalid allocation request for 'nvidia.com/sharedgpu': unknown device: nope-replica-0"
    req = wire.MSG["PreferredAllocationRequest"]()
    c = req.container_requests.add()
    c.available_deviceIDs.extend(["GPU-unknown-replica-0"])
    c.allocation_size = 1
    with pytest.raises(grpc.RpcError) as e:
        client.GetPreferredAllocation(req)
    assert e.value.details() == "unable to retrieve list of available devices: no device with uuid: GPU-unknown"


def test_list_and_watch_streams_an_update_per_health_event(served):
    client, servicer, ref, devices = served
    stream = client.ListAndWatch(wire.MSG["Empty"]())
    first = next(stream)
    assert [d.ID for d in first.devices] == [d["ID"] for d in ref.api_devices()]
    assert all(d.health == "Healthy" for d in first.devices)
    assert first.devices[0].topology.nodes[0].ID == 0 and first.devices[-1].topology.nodes[0].ID == 1
    servicer.device_unhealthy(2)
    second = next(stream)
    bad = [d.ID for d in second.devices if d.health == "Unhealthy"]
    assert bad == [devices[2]["id"] + "-replica-%d" % i for i in range(4)]
    stream.cancel()
