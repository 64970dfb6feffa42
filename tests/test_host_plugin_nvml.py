"""Host-side rows of SURVEY.md §8 that need no GPU: the NVML link-matrix builder (A7/A8), the
XID health watcher (A13), and the RPC bodies around the decision in replica mode (A11, A12,
ListAndWatch), against the Python restatement oracle/ref_server.py and a scripted NVML
(tests/csrc/fake_nvml.c, a stand-in for libnvidia-ml.so.1)."""
import ctypes as C
import os
import subprocess
import threading
import time

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import synthetic as syn
from oracle import ref_server

HERE = os.path.dirname(os.path.abspath(__file__))
FAKE_SRC = os.path.join(HERE, "csrc", "fake_nvml.c")
FAKE_SO = os.path.join(HERE, "csrc", "libfake-nvidia-ml.so.1")
XID = 0x8  # nvmlEventTypeXidCriticalError


@pytest.fixture(scope="module")
def fake():
    import __graft_entry__ as entry
    entry.build_native()
    if not os.path.exists(FAKE_SO) or os.path.getmtime(FAKE_SO) < os.path.getmtime(FAKE_SRC):
        subprocess.check_call(["gcc", "-O1", "-fPIC", "-shared", "-I/usr/local/cuda/include", "-o", FAKE_SO, FAKE_SRC,
                               "-lpthread"])
    F = C.CDLL(FAKE_SO, mode=C.RTLD_GLOBAL)
    F.fake_nvml_bus.restype = C.c_char_p
    F.fake_nvml_calls.restype = C.c_ulonglong
    F.fake_nvml_inject.argtypes = [C.c_int, C.c_ulonglong, C.c_ulonglong]
    F.fake_nvml_set_device.argtypes = [C.c_int, C.c_char_p, C.c_char_p]
    F.fake_nvml_set_link.argtypes = [C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
    return F


def script_topology(F, n, seed, nvswitch=False):
    """scripts the stand-in and returns what the reference's NewDevices would record"""
    rng = np.random.default_rng(seed)
    F.fake_nvml_reset(n)
    uuids = syn.node_uuids(n, seed=seed + 1)
    for i in range(n):
        F.fake_nvml_set_device(i, uuids[i].encode(), None)
    bus = [F.fake_nvml_bus(i).decode() for i in range(n)]
    level = [[None] * n for _ in range(n)]
    for i in range(n):
        for j in range(i + 1, n):
            lv = int(rng.choice([0, 10, 20, 30, 40, 50, -1], p=[.05, .15, .15, .15, .25, .2, .05]))
            level[i][j] = level[j][i] = None if lv < 0 else lv
            F.fake_nvml_set_level(i, j, lv)
            F.fake_nvml_set_level(j, i, lv)
    links = [[(False, "")] * 18 for _ in range(n)]
    for i in range(n):
        for l in range(18):
            if nvswitch:
                links[i][l] = (True, "00000000:%02X:00.0" % (0xA0 + l % 4))  # far end is a switch
                F.fake_nvml_set_link(i, l, 1, links[i][l][1].encode(), 2)
            elif rng.random() < 0.7:
                peer = int(rng.integers(n))
                if peer != i:
                    links[i][l] = (True, bus[peer])
                    F.fake_nvml_set_link(i, l, 1, bus[peer].encode(), 0)
    return uuids, bus, level, links


@pytest.mark.parametrize("n,seed", [(1, 0), (2, 1), (4, 2), (8, 3), (8, 4), (16, 5)])
def test_topology_builder_matches_reference_restatement(fake, n, seed):
    uuids, bus, level, links = script_topology(fake, n, seed)
    nv = gpa.Nvml(FAKE_SO)
    try:
        assert nv.device_count() == n
        assert [nv.device_uuid(i) for i in range(n)] == uuids
        got, calls = nv.read_topology(gpa.TOPO_PARITY)
        want = np.array(ref_server.new_devices_links(n, level, bus, links), dtype=np.int32).reshape(n, n, 2)
        assert (got == want).all()
        # one build here costs O(n * links + n^2) NVML calls; the reference pays n^2 * (1 + 2 * 12) + 13 n per
        # NewDevices() and runs it twice per request (server.go:274,279)
        assert calls <= 1 + 2 * n + 2 * 12 * n + n * (n - 1)
    finally:
        nv.close()


def test_nvswitch_box_has_no_nvlink_in_parity_mode_and_18_links_when_aware(fake):
    """HGX B200: every link ends in a switch, so nvml.GetNVLink finds none (nvml.go:655 TODO);
    the opt-in aware mode counts min(switch links) = 18 -> capped at the 12-link enum"""
    uuids, bus, level, links = script_topology(fake, 8, 9, nvswitch=True)
    nv = gpa.Nvml(FAKE_SO)
    try:
        parity, _ = nv.read_topology(gpa.TOPO_PARITY)
        assert (parity[:, :, 1] == 0).all()
        want = np.array(ref_server.new_devices_links(8, level, bus, links), dtype=np.int32).reshape(8, 8, 2)
        assert (parity == want).all()
        aware, _ = nv.read_topology(gpa.TOPO_NVSWITCH_AWARE)
        off = ~np.eye(8, dtype=bool)
        assert (aware[:, :, 1][off] == 18).all() and (aware[:, :, 0] == parity[:, :, 0]).all()
    finally:
        nv.close()


def test_nvml_missing_library_fails_loudly():
    with pytest.raises(gpa.GpaError) as e:
        gpa.Nvml("/nonexistent/libnvidia-ml.so.1")
    assert e.value.code == gpa.ERR_NVML


def drain(h, timeout_ms=300):
    out = []
    while True:
        r = h.poll(timeout_ms)
        if r is None:
            return out
        out.append(r)


def test_health_watcher_follows_check_health(fake):
    n = 4
    uuids, *_ = script_topology(fake, n, 21)
    nv = gpa.Nvml(FAKE_SO)
    h = gpa.HealthWatcher(nv, uuids, disable_env="", wait_timeout_ms=50)
    try:
        events = [(1, XID, 48), (2, XID, 13), (0, XID, 31), (3, XID, 79), (-1, XID, 63), (2, XID, 45), (2, XID, 94)]
        for dev, et, xid in events:
            fake.fake_nvml_inject(dev, et, xid)
        got = drain(h)
        want = ref_server.check_health_events(uuids, [(uuids[d] if d >= 0 else "", et, x) for d, et, x in events])
        assert [g[0] for g in got] == want == [1, 3, 0, 1, 2, 3, 2]
        assert [g[1] for g in got] == [48, 79, 63, 63, 63, 63, 94]
        st = h.stats()
        assert st == {"events_seen": 7, "events_skipped": 3, "disabled": False}
    finally:
        h.close()
        nv.close()


@pytest.mark.parametrize("env,expect_disabled,extra", [("all", True, []), ("XIDS", True, []), ("foo,xids", True, []),
                                                         ("48, 94", False, [48, 94]), ("", False, [])])
def test_health_disable_env(fake, env, expect_disabled, extra):
    uuids, *_ = script_topology(fake, 2, 22)
    nv = gpa.Nvml(FAKE_SO)
    h = gpa.HealthWatcher(nv, uuids, disable_env=env, wait_timeout_ms=50)
    try:
        events = [(0, XID, 48), (1, XID, 94), (1, XID, 79)]
        if not expect_disabled:
            for dev, et, xid in events:
                fake.fake_nvml_inject(dev, et, xid)
        got = [g[0] for g in drain(h, 200)]
        want = ref_server.check_health_events(uuids, [(uuids[d], et, x) for d, et, x in events], env)
        assert got == want
        assert h.stats()["disabled"] == expect_disabled
    finally:
        h.close()
        nv.close()
        fake.fake_nvml_reset(2)


def test_health_unsupported_device_is_unhealthy_at_once_and_unknown_device_is_fatal(fake):
    uuids, *_ = script_topology(fake, 3, 23)
    fake.fake_nvml_set_events_supported(1, 0)
    nv = gpa.Nvml(FAKE_SO)
    h = gpa.HealthWatcher(nv, uuids, disable_env="", wait_timeout_ms=50)
    try:
        assert drain(h, 100) == [(1, 0)]  # nvidia.go:220-224
    finally:
        h.close()
    with pytest.raises(gpa.GpaError) as e:  # RegisterEventForDevice: "nvml: device not found" -> check() panics
        gpa.HealthWatcher(nv, uuids + ["GPU-not-there"], disable_env="", wait_timeout_ms=50)
    assert e.value.code == gpa.ERR_UNKNOWN_DEVICE
    nv.close()


# ---------------------------------------------------------------------------------------
def make_devices(n, seed=1, mem_mib=183359):
    uuids = syn.node_uuids(n, seed=seed)
    return [{"id": uuids[i], "index": str(i), "paths": ["/dev/nvidia%d" % i], "total_memory": mem_mib,
             "numa": i // (n // 2 or 1), "nvml_index": i} for i in range(n)]


def both(devices, **kw):
    ref_kw = dict(kw)
    ref_kw["list_strategy"] = {gpa.LIST_ENVVAR: "envvar", gpa.LIST_VOLUME_MOUNTS: "volume-mounts"}[kw.get("list_strategy", 0)]
    ref_kw["id_strategy"] = {gpa.ID_UUID: "uuid", gpa.ID_INDEX: "index"}[kw.get("id_strategy", 0)]
    ref_kw.pop("fixes", None)
    return gpa.Plugin(devices, **kw), ref_server.Plugin(devices, **ref_kw)


def call(f, *a):
    try:
        return ("ok", f(*a))
    except Exception as e:  # noqa: BLE001 - compare the message text
        return ("err", str(e))


@pytest.mark.parametrize("replicas,auto", [(4, False), (16, False), (32, False), (1, True)])
def test_replica_mode_stream_matches_reference_restatement(replicas, auto):
    """config 2 of SURVEY.md §8d in miniature: a stateful stream of GetPreferredAllocation -> Allocate / free at
    32, 128, 256 and 1 464 virtual devices; the reference answers with prioritizeDevices (server.go:285-296)"""
    devices = make_devices(8)
    got_p, ref_p = both(devices, replicas=replicas, auto_replicas=auto, has_policy=True)
    try:
        assert got_p.options() == ref_p.options() == {"pre_start_required": False, "get_preferred_allocation_available": True}
        assert got_p.api_devices() == ref_p.api_devices()
        n_virtual = 8 * (183 if auto else replicas)
        assert got_p.device_count() == (8, n_virtual)
        rng = np.random.default_rng(2)
        free = [d["ID"] for d in ref_p.api_devices()]
        held = []
        for step in range(300 if n_virtual <= 256 else 60):
            if rng.random() < 0.6 and free:
                size = int(rng.choice([1, 2, 4, 8]))
                avail = [str(x) for x in rng.permutation(free)]
                must = [avail[0]] if rng.random() < 0.1 else []
                a = call(got_p.get_preferred_allocation, avail, must, size)
                b = call(ref_p.get_preferred_allocation, avail, must, size)
                assert a == b, (step, size, must)
                if a[0] == "ok" and a[1]:
                    assert call(got_p.allocate, a[1]) == call(ref_p.allocate, a[1])
                    for x in a[1]:
                        free.remove(x)
                    held.append(a[1])
            elif held:
                free += held.pop(int(rng.integers(len(held))))
    finally:
        got_p.close()


def test_get_preferred_allocation_error_strings():
    devices = make_devices(4)
    got_p, ref_p = both(devices, replicas=2, has_policy=True)
    ids = [d["ID"] for d in ref_p.api_devices()]
    cases = [
        (ids + ["GPU-unknown-replica-0"], [], 1),                 # unable to retrieve list of available devices
        (ids, ["GPU-unknown-replica-0"], 1),                      # ... of required devices
        (ids[:3], [ids[5]], 1),                                   # missing from availableDeviceIDs
        (ids[:2], [], 5),                                         # no devices left to allocate
        ([], [], 1),
        (ids, [], 0),
    ]
    for avail, must, size in cases:
        assert call(got_p.get_preferred_allocation, avail, must, size) == call(ref_p.get_preferred_allocation, avail, must, size)
    a = call(got_p.get_preferred_allocation, ids + ["GPU-unknown-replica-0"], [], 1)
    assert a == ("err", "unable to retrieve list of available devices: no device with uuid: GPU-unknown")
    got_p.close()
    # replicas <= 1 and no policy: the RPC is "not implemented" (server.go:302-304)
    got_p, ref_p = both(devices, replicas=1, has_policy=False)
    ids = [d["ID"] for d in ref_p.api_devices()]
    assert call(got_p.get_preferred_allocation, ids, [], 1) == call(ref_p.get_preferred_allocation, ids, [], 1) == (
        "err", "GetPreferredAllocation() not implemented in this case")
    assert got_p.options()["get_preferred_allocation_available"] is False
    got_p.close()


@pytest.mark.parametrize("list_strategy", [gpa.LIST_ENVVAR, gpa.LIST_VOLUME_MOUNTS])
@pytest.mark.parametrize("id_strategy", [gpa.ID_UUID, gpa.ID_INDEX])
@pytest.mark.parametrize("specs", [False, True])
def test_allocate_response_matches_reference_restatement(list_strategy, id_strategy, specs):
    devices = make_devices(8)
    got_p, ref_p = both(devices, replicas=4, list_strategy=list_strategy, id_strategy=id_strategy, pass_device_specs=specs,
                        driver_root="/run/nvidia/driver", resource_name="nvidia.com/sharedgpu")
    try:
        ids = [d["ID"] for d in ref_p.api_devices()]
        rng = np.random.default_rng(7)
        for _ in range(50):
            req = [str(x) for x in rng.choice(ids, size=int(rng.integers(1, 9)), replace=False)]
            assert call(got_p.allocate, req) == call(ref_p.allocate, req)
        # two replicas of one GPU collapse to the same physical id: that IS the sharing mechanism
        r = got_p.allocate([ids[0], ids[1]])
        if list_strategy == gpa.LIST_ENVVAR:
            assert r["envs"]["NVIDIA_VISIBLE_DEVICES"] in (devices[0]["id"], "0")
        bad = call(got_p.allocate, [ids[0], "nope-replica-0"])
        assert bad == call(ref_p.allocate, [ids[0], "nope-replica-0"])
        assert bad == ("err", "invalid allocation request for 'nvidia.com/sharedgpu': unknown device: nope-replica-0")
        assert call(got_p.allocate, [devices[0]["id"]])[0] == "err"   # a bare uuid is not an advertised replica id
    finally:
        got_p.close()


def test_list_and_watch_health_propagation_parity_and_fixed():
    devices = make_devices(2)
    got_p, ref_p = both(devices, replicas=3)
    before = got_p.api_devices()
    assert got_p.mark_unhealthy(1) == 0          # D1: the re-sent list is unchanged in the reference
    ref_p.mark_unhealthy(1)
    assert got_p.api_devices() == ref_p.api_devices() == before
    got_p.close()
    fixed = gpa.Plugin(devices, replicas=3, fixes=gpa.FIX_HEALTH_PROPAGATION)
    assert fixed.mark_unhealthy(1) == 3
    lst = fixed.api_devices()
    assert [d["health"] for d in lst] == ["Healthy"] * 3 + ["Unhealthy"] * 3
    assert fixed.mark_unhealthy(1) == 0          # idempotent: no recovery path exists (FIXME at server.go:259)
    fixed.close()


def test_health_churn_under_concurrent_allocate(fake):
    """config 4 in miniature: 4 client threads call Allocate back to back while XIDs arrive at ~200 Hz; every
    Allocate succeeds, and the watcher + ListAndWatch bookkeeping ends where the reference's would"""
    devices = make_devices(8)
    uuids = [d["id"] for d in devices]
    fake.fake_nvml_reset(8)
    for i, u in enumerate(uuids):
        fake.fake_nvml_set_device(i, u.encode(), None)
    nv = gpa.Nvml(FAKE_SO)
    h = gpa.HealthWatcher(nv, uuids, disable_env="", wait_timeout_ms=20)
    plugin = gpa.Plugin(devices, replicas=4, fixes=gpa.FIX_HEALTH_PROPAGATION)
    ids = [d["ID"] for d in plugin.api_devices()]
    stop = threading.Event()
    errors, counts, lat = [], [0] * 4, [[] for _ in range(4)]

    def client(k):
        rng = np.random.default_rng(k)
        while not stop.is_set():
            req = [str(x) for x in rng.choice(ids, size=2, replace=False)]
            t0 = time.perf_counter()
            try:
                plugin.allocate(req)
            except Exception as e:  # noqa: BLE001
                errors.append(e)
            lat[k].append(time.perf_counter() - t0)
            counts[k] += 1

    threads = [threading.Thread(target=client, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    rng = np.random.default_rng(4)
    events = []
    for _ in range(200):
        xid = int(rng.choice([13, 31, 43, 45, 68, 48, 63, 79, 94]))
        dev = int(rng.integers(-1, 8)) if rng.random() < 0.95 else -1
        events.append((dev, XID, xid))
        fake.fake_nvml_inject(dev, XID, xid)
        time.sleep(0.005)
    unhealthy = [g[0] for g in drain(h, 300)]
    stop.set()
    for t in threads:
        t.join()
    want = ref_server.check_health_events(uuids, [(uuids[d] if d >= 0 else "", et, x) for d, et, x in events])
    assert unhealthy == want
    changed = sum(plugin.mark_unhealthy(i) for i in unhealthy)
    assert changed == 4 * len(set(unhealthy))
    assert not errors and sum(counts) > 100
    h.close()
    nv.close()
    plugin.close()


@pytest.mark.parametrize("mode", ["parity", "fixed"])
def test_config4_churn_counts_in_parity_and_fixed_mode(fake, mode):
    """SURVEY.md 8d config 4 (tools/churn.cpp, shortened from 30 s to 2 s): 8 Allocate clients while XIDs arrive at 50 Hz.
    One ListAndWatch re-send per critical event (an event without a UUID counts once per device); the re-sent list
    never changes in parity mode (D1) and loses every replica of a hit GPU in fixed mode; no Allocate error."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["make", "-C", os.path.join(root, "tools"), "-s", "bin/churn"])
    out = subprocess.run([os.path.join(root, "tools", "bin", "churn"), FAKE_SO, "2", mode], capture_output=True, text=True, timeout=120)
    r = json.loads(out.stdout)
    assert out.returncode == 0 and r["as_specified"] and r["allocate_errors"] == 0, r
    assert r["unhealthy_notifications"] == r["list_resends"] == r["expected_notifications"] > 0
    assert r["list_entries_changed"] == (0 if mode == "parity" else r["expected_entries_changed"])
    assert mode == "parity" or r["list_entries_changed"] > 0
