"""The daemon end to end on the GPU box: real NVML for devices and the link matrix, a fake kubelet for registration,
GetPreferredAllocation over the unix socket answered by the CUDA policy (replicas = 1) and by the replica-set search
(replicas = 4 with --replica-topology), each checked against the oracle on the matrix NVML gave."""
import tempfile

import grpc
import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import lifecycle, wire
from oracle import oracle as orc
from test_lifecycle import FakeKubelet

pytestmark = pytest.mark.gpu


def preferred(client, avail, must, size):
    req = wire.MSG["PreferredAllocationRequest"]()
    cr = req.container_requests.add()
    cr.available_deviceIDs.extend(avail)
    cr.must_include_deviceIDs.extend(must)
    cr.allocation_size = size
    return list(client.GetPreferredAllocation(req).container_responses[0].deviceIDs)


@pytest.mark.parametrize("config,fixes", [("gpu:gpu:1", 0), ("gpu:shared-gpu:4", gpa.OPT_REPLICA_TOPOLOGY)])
def test_daemon_with_real_nvml_and_cuda_policy(config, fixes):
    nv = gpa.Nvml()
    ctx = gpa.Context(0)
    try:
        devices = lifecycle.node_from_nvml(nv, ctx)
        n = len(devices)
        assert n >= 1 and all(d["id"].startswith("GPU-") for d in devices)
        links, _ = nv.read_topology()
        W, nl = orc.links_to_weights(links)
        with tempfile.TemporaryDirectory() as d:
            kubelet = FakeKubelet(d)
            kubelet.start()
            (resource, sock, kw), = lifecycle.plugins_for("none", gpa.parse_resource_config(config), devices, ctx=ctx, fixes=fixes)
            daemon = lifecycle.PluginDaemon(kw, plugin_dir=d, socket_name=sock,
                                            health_source=lifecycle.nvml_health_source(nv, [x["id"] for x in devices],
                                                                                       wait_timeout_ms=100))
            try:
                daemon.start()
                assert kubelet.wait_for(1) and kubelet.requests[0].resource_name == resource
                c = wire.DevicePluginClient(daemon.socket)
                ids = [x.ID for x in next(iter(c.ListAndWatch(wire.MSG["Empty"]()))).devices]
                reps = kw["replicas"]
                assert len(ids) == n * reps
                index_of = {x["id"]: x["nvml_index"] for x in devices}
                rng = np.random.default_rng(n)
                for _ in range(50):
                    avail = [str(x) for x in rng.permutation(ids)[: int(rng.integers(1, len(ids) + 1))]]
                    size = int(rng.integers(1, min(len(avail), 4) + 1))
                    if reps == 1:   # server.go:297-301: bare UUIDs of Policy.Allocate over the sorted physical ids
                        phys = sorted(x.split("-replica-")[0] for x in avail)
                        rc, want = orc.besteffort(W, nl, [index_of[u] for u in phys], [], size)
                        by_index = {v: k for k, v in index_of.items()}
                        if orc.NIL in want:   # the Go handler dereferences nil; the RPC fails
                            with pytest.raises(grpc.RpcError):
                                preferred(c, avail, [], size)
                            continue
                        assert rc == 0 and preferred(c, avail, [], size) == [by_index[i] for i in want]
                        continue
                    got = preferred(c, avail, [], size)
                    srt = sorted(avail)     # GPA_OPT_REPLICA_TOPOLOGY: the replica-set search over the sorted ids
                    rc, pos = orc.subset_besteffort(W, nl, [index_of[x.split("-replica-")[0]] for x in srt], [], size)
                    assert rc == 0 and got == [srt[p] for p in pos]
                areq = wire.MSG["AllocateRequest"]()
                areq.container_requests.add().devicesIDs.extend(ids[:1])
                envs = dict(c.Allocate(areq).container_responses[0].envs)
                assert envs["NVIDIA_VISIBLE_DEVICES"] == ids[0].split("-replica-")[0]
                c.close()
            finally:
                daemon.stop()
                kubelet.stop()
    finally:
        ctx.close()
        nv.close()


def test_xid_event_rereads_the_matrix_and_rebuilds_the_answer_table(fake_nvml):
    """SURVEY.md 8f rank 2 on the device: a scripted NVML whose topology changes under the daemon; the XID makes the
    health source re-install the matrix in the CUDA context, which rebuilds the answer table (one more build kernel),
    and decisions after the event follow the new matrix like the reference's per-request re-enumeration would."""
    import threading
    import time
    F, FAKE_SO = fake_nvml
    n = 4
    F.fake_nvml_reset(n)
    uuids = ["GPU-%08x-1111-2222-3333-444444444444" % i for i in range(n)]
    for i, u in enumerate(uuids):
        F.fake_nvml_set_device(i, u.encode(), None)
    for i in range(n):
        for j in range(n):
            F.fake_nvml_set_level(i, j, 50)          # NVML_TOPOLOGY_SYSTEM: CrossCPU, 10 per pair
    nv = gpa.Nvml(FAKE_SO)
    ctx = gpa.Context(0)
    try:
        nv.install_topology(ctx)
        W0, _ = ctx.get_topology()
        table0 = ctx.answer_table().copy()
        assert ctx.besteffort([0, 1, 2, 3], [], 2) == (gpa.OK, [0, 1])
        launches0 = ctx.launch_count
        F.fake_nvml_set_level(2, 3, 10)              # NVML_TOPOLOGY_SINGLE: GPUs 2 and 3 now share a switch (50)
        F.fake_nvml_set_level(3, 2, 10)
        src = lifecycle.nvml_health_source(nv, uuids, disable_env="", wait_timeout_ms=50,
                                           refresh_topology=lambda: nv.install_topology(ctx))
        stop, seen = threading.Event(), []
        t = threading.Thread(target=src, args=(stop, seen.append))
        t.start()
        try:
            time.sleep(0.2)
            F.fake_nvml_inject(1, 8, 79)             # nvmlEventTypeXidCriticalError on GPU 1
            deadline = time.time() + 5
            while not seen and time.time() < deadline:
                time.sleep(0.02)
        finally:
            stop.set()
            t.join()
        assert seen == [1]
        W1, nl1 = ctx.get_topology()
        assert W0[2, 3] == 10 and W1[2, 3] == 50 and (W1 != W0).sum() == 2
        assert ctx.launch_count > launches0 and (ctx.answer_table() != table0).any()  # the build kernel ran again
        want = orc.besteffort(W1, nl1, [0, 1, 2, 3], [], 2)
        assert want == (orc.OK, [2, 3]) and ctx.besteffort([0, 1, 2, 3], [], 2) == (gpa.OK, [2, 3])
    finally:
        ctx.close()
        nv.close()
