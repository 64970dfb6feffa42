"""Lifecycle shell of the device plugin daemon over the C++ plugin bodies: --resource-config, Serve, Register with
kubelet, restart on a new kubelet.sock.

The reference does this in Go (cmd/nvidia-device-plugin/main.go:205-324, server.go:127-241, watchers.go); Go is absent
from this image, so the shell is Python over `wire.py` (the v1beta1 messages) and `_lib.Plugin` (C++ behind the C ABI).
Kept to what a drop-in needs:

  parse_resource_config      main.go:171-203 (C++: gpa_parse_resource_config)
  plugins_for("none")        migStrategyNone.GetPlugins, mig-strategy.go:91-105: one plugin "nvidia.com/<name>" on
                             nvidia-gpu.sock with the variant's replicas; an absent "gpu" entry means ZERO replicas (D2)
  PluginDaemon.start()       NvidiaDevicePlugin.Start, server.go:127-150: initialize, Serve (remove stale socket, listen,
                             blocking self-dial, 5 s), Register (dial kubelet.sock, 5 s; RegisterRequest{version "v1beta1",
                             endpoint = basename(socket), resource_name, options}), then the health loop
  PluginDaemon.stop()        server.go:153-165: stop the server, remove the socket
  run()                      main.go:246-321: start every plugin that has devices; if one fails, retry all; restart all
                             when kubelet.sock is (re)created (inotify there, a 0.2 s stat poll here) or on SIGHUP; stop on
                             SIGINT / SIGTERM / SIGQUIT
MIG strategies "single" / "mixed" need MIG enumeration through NVML and are out of scope (nil policy).
"""
import os
import signal
import threading

import grpc

from . import _lib as gpa
from . import wire

DEVICE_PLUGIN_PATH = "/var/lib/kubelet/device-plugins/"  # dpapi/constants.go:29
KUBELET_SOCKET_NAME = "kubelet.sock"                       # dpapi/constants.go:31
DIAL_TIMEOUT_S = 5.0                                       # server.go:205,220


def plugins_for(strategy, resource_config, devices, ctx=None, fixes=0, **plugin_kwargs):
    """-> [(resource name, socket basename, plugin kwargs)] as MigStrategy.GetPlugins() would build them"""
    if strategy != "none":
        raise NotImplementedError("mig.strategy=%s needs MIG enumeration (out of scope)" % strategy)
    rc = gpa.resource_config_get(resource_config, "gpu")  # mig-strategy.go:92
    kw = dict(plugin_kwargs, resource_name="nvidia.com/" + rc["name"], envvar="NVIDIA_VISIBLE_DEVICES", replicas=rc["replicas"],
              auto_replicas=rc["auto_replicas"], has_policy=True, fixes=fixes)
    return [("nvidia.com/" + rc["name"], "nvidia-gpu.sock", dict(devices=devices, ctx=ctx, **kw))]


class PluginDaemon:
    """One NvidiaDevicePlugin with its socket: Start / Stop / Register (server.go:127-241)."""

    def __init__(self, plugin_kwargs, plugin_dir=DEVICE_PLUGIN_PATH, socket_name="nvidia-gpu.sock", health_source=None):
        self.plugin_kwargs = plugin_kwargs
        self.plugin_dir = plugin_dir
        self.socket = os.path.join(plugin_dir, socket_name)
        self.kubelet_socket = os.path.join(plugin_dir, KUBELET_SOCKET_NAME)
        self.health_source = health_source  # callable(stop_event, on_unhealthy(cached_index)) run on a thread, or None
        self.plugin = self.server = self.servicer = None
        self._stop = None
        self._health_thread = None

    # -- server.go:127-150
    def start(self):
        self.plugin = gpa.Plugin(**self.plugin_kwargs)  # initialize(), :95-111
        try:
            self._serve()
            self._register()
        except Exception:
            self.stop()
            raise
        self._stop = threading.Event()
        if self.health_source is not None:  # go m.CheckHealth(m.stop, m.cachedDevices, m.health), :148
            self._health_thread = threading.Thread(target=self.health_source, args=(self._stop, self.servicer.device_unhealthy),
                                                   daemon=True)
            self._health_thread.start()

    def device_count(self):
        return len(self.plugin.api_devices()) if self.plugin else 0

    # -- server.go:168-214
    def _serve(self):
        try:
            os.remove(self.socket)  # :169
        except FileNotFoundError:
            pass
        self.server, self.servicer = wire.serve(self.plugin, self.socket)
        ch = grpc.insecure_channel("unix:" + self.socket)  # "wait for server to start by launching a blocking connexion"
        try:
            grpc.channel_ready_future(ch).result(timeout=DIAL_TIMEOUT_S)
        finally:
            ch.close()

    # -- server.go:217-241
    def _register(self):
        ch = grpc.insecure_channel("unix:" + self.kubelet_socket)
        try:
            grpc.channel_ready_future(ch).result(timeout=DIAL_TIMEOUT_S)
            call = ch.unary_unary("/%s.Registration/Register" % wire.PACKAGE,
                                  request_serializer=wire.MSG["RegisterRequest"].SerializeToString,
                                  response_deserializer=wire.MSG["Empty"].FromString)
            o = self.plugin.options()
            req = wire.MSG["RegisterRequest"](version=wire.VERSION, endpoint=os.path.basename(self.socket),
                                              resource_name=self.plugin_kwargs["resource_name"])
            req.options.get_preferred_allocation_available = o["get_preferred_allocation_available"]
            call(req, timeout=DIAL_TIMEOUT_S)
        finally:
            ch.close()

    # -- server.go:153-165
    def stop(self):
        if self._stop is not None:
            self._stop.set()
        if self.servicer is not None:
            self.servicer.stop()
        if self.server is not None:
            self.server.stop(0)
            try:
                os.remove(self.socket)
            except FileNotFoundError:
                pass
        if self._health_thread is not None:
            self._health_thread.join(timeout=2)
        if self.plugin is not None:
            self.plugin.close()
        self.plugin = self.server = self.servicer = self._stop = self._health_thread = None


def _socket_identity(path):
    try:
        st = os.stat(path)
        return (st.st_ino, st.st_ctime_ns)
    except FileNotFoundError:
        return None


def run(daemons, stop_event=None, restart_event=None, poll_s=0.2, retry_s=1.0, log=print, install_signals=False):
    """The restart loop of main.go:246-321 over a list of PluginDaemon.  Returns when stop_event is set (or on
    SIGINT / SIGTERM / SIGQUIT with install_signals).  restart_event (or SIGHUP) restarts every plugin."""
    stop_event = stop_event or threading.Event()
    restart_event = restart_event or threading.Event()
    if install_signals:
        signal.signal(signal.SIGHUP, lambda *_: restart_event.set())
        for s in (signal.SIGINT, signal.SIGTERM, signal.SIGQUIT):
            signal.signal(s, lambda *_: stop_event.set())
    kubelet = daemons[0].kubelet_socket if daemons else None
    starts = 0
    while not stop_event.is_set():
        for d in daemons:  # "idempotently stop any running plugins before recreating them"
            d.stop()
        seen = _socket_identity(kubelet) if kubelet else None
        start_error = False
        started = 0
        for d in daemons:
            if len(d.plugin_kwargs.get("devices", [])) == 0:  # main.go:268-270: no (physical) devices to serve for this plugin
                continue
            try:
                d.start()
            except Exception as e:  # noqa: BLE001 - main.go:275-281: any failure => retry all
                log("Could not contact Kubelet, retrying. Did you enable the device plugin feature gate? (%s)" % e)
                start_error = True
                break
            started += 1
        starts += 1
        if started == 0 and not start_error:
            log("No devices found. Waiting indefinitely.")
        if start_error:
            if stop_event.wait(retry_s):
                break
            continue
        while not stop_event.is_set():  # events:
            if restart_event.is_set():
                restart_event.clear()
                log("Received SIGHUP, restarting.")
                break
            now = _socket_identity(kubelet) if kubelet else None
            if now is not None and now != seen:  # fsnotify.Create on pluginapi.KubeletSocket, main.go:296-300
                log("inotify: %s created, restarting." % kubelet)
                break
            seen = now if now is None else seen
            stop_event.wait(poll_s)
    for d in daemons:
        d.stop()
    return starts


# ---------------------------------------------------------------------------------------------------------------------
# A runnable daemon on a real node: NVML for devices, link matrix and health events; CUDA for the decisions.

def node_from_nvml(nvml, ctx=None, topology_flags=gpa.TOPO_PARITY):
    """-> device dicts for Plugin (NewGpuDeviceManager(false).Devices(), cmd/nvidia.go:70-99: one per GPU, id = UUID,
    path /dev/nvidia<minor>) and, when a context is given, the NVML link matrix installed in it ONCE (the reference
    re-enumerates NVML twice per request, server.go:274,279)."""
    n = nvml.device_count()
    devices = []
    for i in range(n):
        info = nvml.device_info(i)
        if info["minor"] < 0:
            raise RuntimeError("unsupported GPU device")  # nvml.ErrUnsupportedGPU, nvml.go:467-469: no minor number
        # buildDevice, nvidia.go:162-178: TotalMemory = free + used MiB (sizes auto replicas), Topology from the NUMA node
        devices.append({"id": nvml.device_uuid(i), "index": str(i), "paths": ["/dev/nvidia%d" % info["minor"]],
                        "total_memory": info["total_memory"], "numa": info["numa"], "nvml_index": i})
    if ctx is not None and n > 0:
        nvml.install_topology(ctx, topology_flags)
    return devices


def nvml_health_source(nvml, device_ids, disable_env=None, wait_timeout_ms=500, refresh_topology=None, log=print):
    """checkHealth (cmd/nvidia.go:181-269) as a PluginDaemon health source: XID events -> device_unhealthy(index).

    refresh_topology: called after every event, before the list is re-sent.  The reference rebuilds its link matrix from
    NVML inside every request (server.go:274,279), so the request after an XID already sees whatever the event did to the
    links (an NVLink that went down no longer counts); this daemon reads the matrix once, so the event is where its cache
    - the matrix and the answer table built from it - is invalidated and re-read.  A GPU that fell off the bus can make
    the re-read fail; the old matrix then stays (kubelet no longer offers that GPU once its replicas are Unhealthy, so
    its rows are not consulted), which is logged."""
    def loop(stop_event, on_unhealthy):
        watcher = gpa.HealthWatcher(nvml, device_ids, disable_env=disable_env, wait_timeout_ms=wait_timeout_ms)
        try:
            while not stop_event.is_set():
                ev = watcher.poll(wait_timeout_ms)
                if ev is None:
                    continue
                idx, _xid = ev
                if refresh_topology is not None:
                    try:
                        refresh_topology()
                    except Exception as e:  # noqa: BLE001
                        log("XID %s: link matrix could not be re-read, keeping the cached one (%s)" % (_xid, e))
                for i in (range(len(device_ids)) if idx < 0 else [idx]):  # an event without a UUID marks every device (:243-250)
                    on_unhealthy(i)
        finally:
            watcher.close()
    return loop


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser(description="kubelet device plugin daemon over libgpualloc (drop-in for nvidia-device-plugin)")
    ap.add_argument("--resource-config", default=os.environ.get("RESOURCE_CONFIG", ""), help="orig:new:replicas,... (main.go:123-128)")
    ap.add_argument("--mig-strategy", default=os.environ.get("MIG_STRATEGY", "none"), choices=["none"])
    ap.add_argument("--fail-on-init-error", action="store_true", default=os.environ.get("FAIL_ON_INIT_ERROR", "true") == "true")
    ap.add_argument("--plugin-dir", default=DEVICE_PLUGIN_PATH)
    ap.add_argument("--cuda-device", type=int, default=0, help="GPU that runs the allocation kernels")
    ap.add_argument("--nvswitch-aware", action="store_true", help="count NVSwitch links (not what the reference sees)")
    ap.add_argument("--replica-topology", action="store_true", help="GPA_OPT_REPLICA_TOPOLOGY: best-connected replicas")
    ap.add_argument("--fixes", type=int, default=0, help="GPA_FIX_* bits (default: bit-exact with the reference, defects included)")
    a = ap.parse_args(argv)
    cfg = gpa.parse_resource_config(a.resource_config)
    print("Running with resource config: %r" % cfg)
    try:
        nvml = gpa.Nvml()
    except gpa.GpaError as e:
        print("Failed to initialize NVML: %s." % e)
        if a.fail_on_init_error:
            return 1
        threading.Event().wait()  # `select {}`, main.go:228
    ctx = gpa.Context(a.cuda_device)  # no CPU fall-back: fails loudly without a CUDA device
    flags = gpa.TOPO_NVSWITCH_AWARE if a.nvswitch_aware else gpa.TOPO_PARITY
    devices = node_from_nvml(nvml, ctx, flags)
    daemons = []
    for _resource, sock, kw in plugins_for(a.mig_strategy, cfg, devices, ctx=ctx,
                                           fixes=a.fixes | (gpa.OPT_REPLICA_TOPOLOGY if a.replica_topology else 0)):
        daemons.append(PluginDaemon(kw, plugin_dir=a.plugin_dir, socket_name=sock,
                                    health_source=nvml_health_source(nvml, [d["id"] for d in devices],
                                                                     disable_env=os.environ.get("DP_DISABLE_HEALTHCHECKS"),
                                                                     refresh_topology=lambda: nvml.install_topology(ctx, flags))))
    try:
        run(daemons, install_signals=True)
    finally:
        ctx.close()
        nvml.close()
    return 0
