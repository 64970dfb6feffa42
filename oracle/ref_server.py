"""ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/ref_besteffort.c for the rule).

Pure-Python restatement of the reference's request handlers around the allocation decision
(small cases only), used by tests/ to check the C++ plugin mirror and the NVML topology
builder.  PARITY STATUS: unpinned — the reference has no test for these handlers; authority
is the cited source.  Paths: cmd/ = cmd/nvidia-device-plugin/, alloc/ =
vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/, nvml/ =
vendor/github.com/NVIDIA/gpu-monitoring-tools/bindings/go/nvml/.
"""
import os

from . import oracle as orc

# nvml.P2PLinkType, nvml/nvml.go:132-154
UNKNOWN, CROSS_CPU, SAME_CPU, HOST_BRIDGE, MULTI_SWITCH, SINGLE_SWITCH, SAME_BOARD, ONE_NVLINK = 0, 1, 2, 3, 4, 5, 6, 7
# nvmlGpuTopologyLevel_t
LEVEL_TO_LINK = {0: SAME_BOARD, 10: SINGLE_SWITCH, 20: MULTI_SWITCH, 30: HOST_BRIDGE, 40: SAME_CPU, 50: CROSS_CPU}
REF_MAX_LINKS = 12  # NVML_NVLINK_MAX_LINKS in the reference's nvml.h:273


def new_devices_links(n, level, bus, links):
    """gpuallocator.NewDevices (alloc/device.go:33-72) over a scripted NVML.
    level[i][j]: topology level or None (NOT_SUPPORTED); bus[i]: PCI bus id;
    links[i]: list of (enabled, remote_bus) per link index.
    -> links2[i][j] = [pcie type, nvlink type] (0 = not recorded)"""
    out = [[[0, 0] for _ in range(n)] for _ in range(n)]
    for i in range(n):
        for j in range(n):
            if i == j:  # device.go:51
                continue
            # nvml.GetP2PLink, nvml.go:592-615
            lv = level[i][j]
            p2p = UNKNOWN if lv is None else LEVEL_TO_LINK[lv]
            if p2p != UNKNOWN:  # device.go:56-58
                out[i][j][0] = p2p
            # nvml.GetNVLink, nvml.go:617-658 over deviceGetAllNvLinkRemotePciInfo, bindings.go:358-386
            nv = UNKNOWN
            for l, (enabled, remote) in enumerate(links[i]):
                if l >= REF_MAX_LINKS:
                    break
                if enabled and remote == bus[j]:
                    nv = ONE_NVLINK if nv == UNKNOWN else min(nv + 1, ONE_NVLINK + 11)
            if nv != UNKNOWN:  # device.go:64-66
                out[i][j][1] = nv
    return out


class Plugin:
    """NvidiaDevicePlugin, cmd/server.go:56-125, reduced to what the handlers read."""

    def __init__(self, devices, resource_name="nvidia.com/gpu", envvar="NVIDIA_VISIBLE_DEVICES", driver_root="/",
                 replicas=1, auto_replicas=False, list_strategy="envvar", id_strategy="uuid", pass_device_specs=False,
                 has_policy=True, W=None, nlinks=None):
        # devices: dicts {id, index, paths, total_memory, numa, nvml_index}
        self.cached = [dict(d, health="Healthy") for d in devices]
        self.resource_name, self.envvar, self.driver_root = resource_name, envvar, driver_root
        self.replicas, self.auto_replicas = replicas, auto_replicas
        self.list_strategy, self.id_strategy, self.pass_device_specs = list_strategy, id_strategy, pass_device_specs
        self.has_policy, self.W, self.nlinks = has_policy, W, nlinks
        # initialize(), server.go:95-111
        self.device_replicas = []
        for dev in self.cached:
            reps = dev["total_memory"] // 1000 if auto_replicas else replicas
            for i in range(reps):
                r = dict(dev)  # a struct copy (:107)
                r["id"] = "%s-replica-%d" % (dev["id"], i)
                self.device_replicas.append(r)

    # GetDevicePluginOptions, server.go:243-248
    def options(self):
        return {"pre_start_required": False,
                "get_preferred_allocation_available": bool(self.has_policy or self.replicas > 1 or self.auto_replicas)}

    # apiDevices, server.go:415-421
    def api_devices(self):
        out = []
        for d in self.device_replicas:
            e = {"ID": d["id"], "health": d["health"]}
            if d.get("numa", -1) >= 0:
                e["topology"] = {"nodes": [{"ID": d["numa"]}]}
            out.append(e)
        return out

    # case d := <-m.health, server.go:258-262: the pointer is into cachedDevices
    def mark_unhealthy(self, cached_index):
        self.cached[cached_index]["health"] = "Unhealthy"

    def _new_devices_from(self, uuids, what):
        # alloc/device.go:75-95
        res = []
        for u in uuids:
            hit = [d for d in self.cached if d["id"] == u]
            if not hit:
                raise RuntimeError("unable to retrieve list of %s devices: no device with uuid: %s" % (what, u))
            res.append(hit[0])
        return res

    # one iteration of the loop in GetPreferredAllocation, server.go:273-311
    def get_preferred_allocation(self, available_ids, must_ids, size):
        available = self._new_devices_from(orc.strip_replicas(available_ids), "available")  # :274
        required = self._new_devices_from(orc.strip_replicas(must_ids), "required")         # :279
        if self.replicas > 1 or self.auto_replicas:  # :285
            ids, kind, arg = orc.prioritize_devices(available_ids, must_ids, size)
            if kind == "no_devices_left":
                raise RuntimeError("no devices left to allocate")
            if kind == "must_missing":
                raise RuntimeError("device '%s' in mustIncludeDeviceIDs is missing from availableDeviceIDs" % arg)
            if kind == "panic":
                raise RuntimeError("panic")
            return ids  # NonUniqueError is ignored (:288-291)
        if not self.has_policy:  # :302-304
            raise RuntimeError("GetPreferredAllocation() not implemented in this case")
        # :298 — `required` holds pointers from a second NewDevices(): foreign to `available`
        rc, out = orc.besteffort(self.W, self.nlinks, [d["nvml_index"] for d in available],
                                 [orc.FOREIGN for _ in required], size)
        if rc != orc.OK:
            raise RuntimeError("panic")
        if any(x == orc.NIL for x in out):
            raise RuntimeError("panic")  # device.UUID on nil (:299-301)
        by_index = {d["nvml_index"]: d for d in self.cached}
        return [by_index[x]["id"] for x in out]  # bare UUIDs (:300)

    # one iteration of the loop in Allocate, server.go:318-350
    def allocate(self, ids):
        known = {d["id"] for d in self.device_replicas}
        for i in ids:  # :319-323
            if i not in known:
                raise RuntimeError("invalid allocation request for '%s': unknown device: %s" % (self.resource_name, i))
        uuids = orc.strip_replicas(ids)  # :325
        for u in uuids:  # :328-332
            if not any(d["id"] == u for d in self.cached):
                raise RuntimeError("invalid allocation request for '%s': unknown device: %s" % (self.resource_name, u))
        # deviceIDsFromUUIDs, :397-413
        if self.id_strategy == "uuid":
            device_ids = list(uuids)
        else:
            device_ids = [d["index"] for d in self.cached for u in uuids if d["id"] == u]
        resp = {"envs": {}, "mounts": [], "devices": []}
        root = "/var/run/nvidia-container-devices"
        if self.list_strategy == "envvar":  # :338-340
            resp["envs"] = {self.envvar: ",".join(device_ids)}
        if self.list_strategy == "volume-mounts":  # :341-344
            resp["envs"] = {self.envvar: root}
            resp["mounts"] = [{"container_path": os.path.join(root, i), "host_path": "/dev/null"} for i in device_ids]
        if self.pass_device_specs:  # :345-347, apiDeviceSpecs :443-480
            def spec(p):
                return {"container_path": p, "host_path": os.path.normpath(os.path.join(self.driver_root, p.lstrip("/"))),
                        "permissions": "rw"}
            for p in ("/dev/nvidiactl", "/dev/nvidia-uvm", "/dev/nvidia-uvm-tools", "/dev/nvidia-modeset"):
                if os.path.exists(p):
                    resp["devices"].append(spec(p))
            for d in self.cached:
                for u in uuids:
                    if d["id"] == u:
                        resp["devices"] += [spec(p) for p in d["paths"]]
        return resp


def check_health_events(device_ids, events, disable_env=""):
    """checkHealth, cmd/nvidia.go:181-269, as a pure function of the event stream.
    events: [(uuid or "" / None, etype, xid)] as nvml.WaitForEvent delivers them.
    -> list of indices into device_ids, in the order `unhealthy <- d` fires."""
    dis = disable_env.lower()
    if dis == "all":
        dis = "xids"
    if "xids" in dis:
        return []
    skipped = {13, 31, 43, 45, 68} | set(orc.get_additional_xids(dis))
    out = []
    for uuid, etype, xid in events:
        # (a wait that returns without error is processed whatever its type, :236-238)
        if xid in skipped:
            continue
        if not uuid:
            out += list(range(len(device_ids)))
            continue
        out += [i for i, d in enumerate(device_ids) if d == uuid]
    return out


def _go_parse_int32(s):
    """strconv.ParseInt(s, 10, 32): optional sign, decimal digits, int32 range; anything else is an error"""
    body = s[1:] if s[:1] in ("+", "-") else s
    if not body or any(c not in "0123456789" for c in body):
        raise ValueError
    v = int(body) * (-1 if s[:1] == "-" else 1)
    if not -(1 << 31) <= v < (1 << 31):
        raise ValueError
    return v


_GO_SPACE = " \t\n\v\f\r\x85\xa0"


def parse_resource_config(config):
    """parseResourceConfig, cmd/main.go:171-203 -> {origName: (newName, replicas, autoReplicas)} or raises ValueError(text)"""
    out = {}
    for entry in config.split(","):          # :174
        entry = entry.strip(_GO_SPACE)        # :175
        if len(entry) == 0:                   # :176-178
            continue
        parts = entry.split(":")              # :179
        if len(parts) != 3:                   # :180-182
            raise ValueError("an entry must have three parts separated by a semicolon")
        try:
            replicas = _go_parse_int32(parts[2])  # :185
        except ValueError:
            raise ValueError("replica must be an integer") from None
        auto = False
        if replicas == -1:                    # :189-193
            auto, replicas = True, 1
        out[parts[0]] = (parts[1], replicas % (1 << 64), auto)   # uint(replicas), :195-199
    return out


def resource_config_get(cfg, name):
    """resourceConfiguration.Get, cmd/mig-strategy.go:66-76"""
    return cfg.get(name, (name, 0, False))
