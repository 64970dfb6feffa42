"""The best-effort branch of GetPreferredAllocation through the plugin mirror (server.go:297-301 ->
Policy.Allocate -> the CUDA search), and the NVML code against the box's real libnvidia-ml.so.1."""
import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import synthetic as syn
from oracle import ref_server

pytestmark = pytest.mark.gpu


def make_devices(n, seed=1):
    uuids = syn.node_uuids(n, seed=seed)
    return [{"id": uuids[i], "index": str(i), "paths": ["/dev/nvidia%d" % i], "total_memory": 183359, "numa": i // 4,
             "nvml_index": i} for i in range(n)]


def call(f, *a):
    try:
        return ("ok", f(*a))
    except Exception as e:  # noqa: BLE001
        return ("err", str(e))


@pytest.mark.parametrize("topo_i", [2, 3, 6])
def test_policy_branch_matches_reference_restatement(topo_i):
    name, t = syn.standard_topologies()[topo_i]
    W, nl = syn.links_to_weights(t)
    devices = make_devices(8)
    ctx = gpa.Context(0)
    ctx.set_topology_links(t)
    got_p = gpa.Plugin(devices, ctx=ctx, replicas=1, has_policy=True)
    ref_p = ref_server.Plugin(devices, replicas=1, has_policy=True, W=W, nlinks=nl)
    try:
        ids = [d["ID"] for d in ref_p.api_devices()]
        assert ids == [d["id"] + "-replica-0" for d in devices]
        rng = np.random.default_rng(topo_i)
        for _ in range(300):
            avail = [str(x) for x in rng.permutation(ids)[: int(rng.integers(1, 9))]]  # kubelet order is arbitrary
            size = int(rng.integers(1, len(avail) + 1))
            must = [avail[0]] if rng.random() < 0.15 else []
            a = call(got_p.get_preferred_allocation, avail, must, size)
            assert a == call(ref_p.get_preferred_allocation, avail, must, size), (avail, must, size)
            if must:
                assert a == ("ok", [])  # D4: MustIncludeDeviceIDs can never be satisfied on this branch
            elif a[0] == "ok":
                assert len(a[1]) == size and all("-replica-" not in x for x in a[1])  # D3: bare UUIDs
        # an nil-padded winner makes the Go handler dereference nil (D5): reported as a panic, not as an answer
    finally:
        got_p.close()
        ctx.close()


def test_policy_branch_with_fixes():
    t = syn.topo_dgx1v_like()
    W, nl = syn.links_to_weights(t)
    devices = make_devices(8)
    ctx = gpa.Context(0)
    ctx.set_topology_links(t)
    p = gpa.Plugin(devices, ctx=ctx, replicas=1, has_policy=True, fixes=gpa.FIX_REPLICA_SUFFIX | gpa.FIX_REQUIRED_IDENTITY)
    try:
        ids = [d["id"] + "-replica-0" for d in devices]
        got = p.get_preferred_allocation(ids, [ids[5]], 2)
        assert len(got) == 2 and ids[5] in got and all(x in ids for x in got)
        # the same decision through the Policy seam, required matched by identity
        order = sorted(range(8), key=lambda i: devices[i]["id"].encode())
        from oracle import oracle as orc
        rc, want = orc.besteffort(W, nl, order, [5], 2)
        assert sorted(got) == sorted(ids[i] for i in want)
    finally:
        p.close()
        ctx.close()


def test_nil_padded_winner_is_a_panic_like_in_go():
    t = syn.topo_uniform(5)
    for i, j in ((2, 3), (2, 4), (3, 4)):
        t[i, j, 0] = t[j, i, 0] = 5
    t[0, 1] = t[1, 0] = (5, syn.nvlink_type(1))
    # make NVML order == sorted-uuid order so that the D5 vector keeps its shape
    uuids = sorted(syn.node_uuids(5, seed=3), key=lambda u: u.encode())
    devices = [{"id": uuids[i], "index": str(i), "paths": [], "total_memory": 1000, "numa": -1, "nvml_index": i} for i in range(5)]
    ctx = gpa.Context(0)
    ctx.set_topology_links(t)
    p = gpa.Plugin(devices, ctx=ctx, replicas=1, has_policy=True)
    try:
        with pytest.raises(gpa.PluginError) as e:
            p.get_preferred_allocation([u + "-replica-0" for u in uuids], [], 3)
        assert e.value.code == gpa.ERR_PANIC and "nil pointer dereference" in str(e.value)
    finally:
        p.close()
        ctx.close()


def test_real_nvml_topology_and_health_open():
    """the box's own NVML: the matrix builds, installs, and the watcher registers for XID events"""
    nv = gpa.Nvml()  # libnvidia-ml.so.1
    try:
        n = nv.device_count()
        assert n >= 1
        uuids = [nv.device_uuid(i) for i in range(n)]
        assert all(u.startswith("GPU-") for u in uuids)
        links, calls = nv.read_topology(gpa.TOPO_PARITY)
        assert links.shape == (n, n, 2) and calls > 0
        off = ~np.eye(n, dtype=bool)
        assert (links[:, :, 0][off] >= 1).all() and (links[:, :, 0][off] <= 6).all()
        aware, _ = nv.read_topology(gpa.TOPO_NVSWITCH_AWARE)
        assert (aware[:, :, 1] >= links[:, :, 1]).all()
        ctx = gpa.Context(0)
        nv.install_topology(ctx, gpa.TOPO_PARITY)
        W, nl = ctx.get_topology()
        assert W.shape == (n, n)
        assert ctx.besteffort(list(range(n)), [], 1) == (gpa.OK, [0])
        ctx.close()
        h = gpa.HealthWatcher(nv, uuids, disable_env="", wait_timeout_ms=100)
        assert h.poll(150) is None and h.stats()["disabled"] is False
        h.close()
    finally:
        nv.close()


def test_replica_requests_by_topology_option():
    """GPA_OPT_REPLICA_TOPOLOGY (no reference counterpart): with replicas > 1 GetPreferredAllocation returns the
    best-connected replicas - checked against the replica-set oracle on the byte-wise sorted ids - while the default stays
    prioritizeDevices; error kinds and texts stay the reference's (replica.go:121,136,183)."""
    from oracle import oracle as orc
    t = syn.topo_dgx1v_like()
    W, nl = syn.links_to_weights(t)
    devices = make_devices(8)
    ctx = gpa.Context(0)
    ctx.set_topology_links(t)
    plug = gpa.Plugin(devices, ctx=ctx, replicas=4, has_policy=True, fixes=gpa.OPT_REPLICA_TOPOLOGY)
    plain = gpa.Plugin(devices, ctx=ctx, replicas=4, has_policy=True)
    index_of = {d["id"]: d["nvml_index"] for d in devices}
    try:
        ids = [d["ID"] for d in plug.api_devices()]
        assert len(ids) == 32
        rng = np.random.default_rng(5)
        differs = 0
        for _ in range(200):
            avail = [str(x) for x in rng.permutation(ids)[: int(rng.integers(2, 33))]]
            size = int(rng.integers(1, min(len(avail), 6) + 1))
            must = [avail[int(rng.integers(len(avail)))]] if rng.random() < 0.3 else []
            got = plug.get_preferred_allocation(avail, must, size)
            srt = sorted(avail)                                   # sort.Strings order: byte-wise
            phys = [index_of[x.split("-replica-")[0]] for x in srt]
            rc, pos = orc.subset_besteffort(W, nl, phys, [srt.index(m) for m in must], size)
            assert rc == 0 and got == [srt[p] for p in pos], (avail, must, size)
            assert got == sorted(got) and set(must) <= set(got)
            assert got == plug.get_preferred_allocation(avail[::-1], must, size)      # order kubelet lists ids in: irrelevant
            differs += got != plain.get_preferred_allocation(avail, must, size)
        assert differs > 20                                       # it IS a different policy from prioritizeDevices
        a = [ids[0], ids[5]]
        assert call(plug.get_preferred_allocation, a, [ids[9]], 1) == call(plain.get_preferred_allocation, a, [ids[9]], 1)
        assert "missing from availableDeviceIDs" in call(plug.get_preferred_allocation, a, [ids[9]], 1)[1]
        assert call(plug.get_preferred_allocation, a, [], 3) == call(plain.get_preferred_allocation, a, [], 3)
        assert "no devices left" in call(plug.get_preferred_allocation, a, [], 3)[1]
        assert call(plug.get_preferred_allocation, a, a, 1) == call(plain.get_preferred_allocation, a, a, 1)   # :121 panic
        assert plug.get_preferred_allocation(a, [], 0) == []
    finally:
        plug.close()
        plain.close()
        ctx.close()


def test_config2_stream_through_the_best_effort_branch():
    """tools/stream.cpp with replicas = 1 (8 virtual devices): the stateful GetPreferredAllocation -> Allocate / free stream
    through the handler bodies on the best-effort branch, every answer compared with ref_besteffort_fast; the decisions
    come from the host copy of the GPU-built answer table, so the stream itself launches nothing."""
    import json
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["make", "-C", os.path.join(root, "tools"), "-s", "bin/stream"])
    out = subprocess.run([os.path.join(root, "tools", "bin", "stream"), "1", "20000"], capture_output=True, text=True, timeout=300)
    r = json.loads(out.stdout)
    assert out.returncode == 0 and r["mismatches_vs_oracle"] == 0 and r["virtual_devices"] == 8, r
    assert r["decisions"] > 5000 and r["allocations"] > 1000 and r["kernel_launches_during_the_stream"] == 0
