"""kubelet device-plugin v1beta1 wire surface over the C++ plugin bodies (test / demo shim).

The reference serves `v1beta1.DevicePlugin` over a unix socket with generated gogo-protobuf
code (vendor/k8s.io/kubelet/pkg/apis/deviceplugin/v1beta1/api.proto:50-76, api.pb.go).  The
drop-in claim of this repository is about what travels on that wire, so this module rebuilds
the message types at run time from a FileDescriptorProto with the SAME package, message
names, field names and field numbers (there is no protoc in the image), and serves the four
RPCs of the path from a `_lib.Plugin` (C++ behind the C ABI).  Registration with kubelet and
the restart loop of cmd/main.go are in lifecycle.py.

Errors: the reference returns plain `fmt.Errorf` values, which grpc-go maps to status
UNKNOWN with the message text (server.go:276,281,303,321,330); this shim does the same.
"""
import threading
from concurrent import futures

import grpc
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

from . import _lib as gpa

PACKAGE = "v1beta1"
VERSION = "v1beta1"  # dpapi/constants.go:27


def _build_messages():
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "gpualloc_v1beta1_api.proto"
    fd.package = PACKAGE
    fd.syntax = "proto3"
    T = descriptor_pb2.FieldDescriptorProto

    def msg(name, fields):
        m = fd.message_type.add()
        m.name = name
        for fname, num, ftype, label, tname in fields:
            f = m.field.add()
            f.name, f.number, f.type = fname, num, ftype
            f.label = T.LABEL_REPEATED if label == "rep" else T.LABEL_OPTIONAL
            if tname:
                f.type_name = "." + PACKAGE + "." + tname
        return m

    S, B, I32, I64, M = T.TYPE_STRING, T.TYPE_BOOL, T.TYPE_INT32, T.TYPE_INT64, T.TYPE_MESSAGE
    msg("DevicePluginOptions", [("pre_start_required", 1, B, "", None), ("get_preferred_allocation_available", 2, B, "", None)])
    msg("RegisterRequest", [("version", 1, S, "", None), ("endpoint", 2, S, "", None), ("resource_name", 3, S, "", None),
                            ("options", 4, M, "", "DevicePluginOptions")])
    msg("Empty", [])
    msg("ListAndWatchResponse", [("devices", 1, M, "rep", "Device")])
    msg("TopologyInfo", [("nodes", 1, M, "rep", "NUMANode")])
    msg("NUMANode", [("ID", 1, I64, "", None)])
    msg("Device", [("ID", 1, S, "", None), ("health", 2, S, "", None), ("topology", 3, M, "", "TopologyInfo")])
    msg("PreStartContainerRequest", [("devicesIDs", 1, S, "rep", None)])
    msg("PreStartContainerResponse", [])
    msg("PreferredAllocationRequest", [("container_requests", 1, M, "rep", "ContainerPreferredAllocationRequest")])
    msg("ContainerPreferredAllocationRequest", [("available_deviceIDs", 1, S, "rep", None), ("must_include_deviceIDs", 2, S, "rep", None),
                                                ("allocation_size", 3, I32, "", None)])
    msg("PreferredAllocationResponse", [("container_responses", 1, M, "rep", "ContainerPreferredAllocationResponse")])
    msg("ContainerPreferredAllocationResponse", [("deviceIDs", 1, S, "rep", None)])
    msg("AllocateRequest", [("container_requests", 1, M, "rep", "ContainerAllocateRequest")])
    msg("ContainerAllocateRequest", [("devicesIDs", 1, S, "rep", None)])
    msg("AllocateResponse", [("container_responses", 1, M, "rep", "ContainerAllocateResponse")])
    car = msg("ContainerAllocateResponse", [("envs", 1, M, "rep", "ContainerAllocateResponse.EnvsEntry"),
                                            ("mounts", 2, M, "rep", "Mount"), ("devices", 3, M, "rep", "DeviceSpec"),
                                            ("annotations", 4, M, "rep", "ContainerAllocateResponse.AnnotationsEntry")])
    for ename in ("EnvsEntry", "AnnotationsEntry"):  # map<string,string> = repeated nested entry with map_entry set
        e = car.nested_type.add()
        e.name = ename
        e.options.map_entry = True
        for fname, num in (("key", 1), ("value", 2)):
            f = e.field.add()
            f.name, f.number, f.type, f.label = fname, num, S, T.LABEL_OPTIONAL
    msg("Mount", [("container_path", 1, S, "", None), ("host_path", 2, S, "", None), ("read_only", 3, B, "", None)])
    msg("DeviceSpec", [("container_path", 1, S, "", None), ("host_path", 2, S, "", None), ("permissions", 3, S, "", None)])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    names = [m.name for m in fd.message_type]
    return {n: message_factory.GetMessageClass(pool.FindMessageTypeByName(PACKAGE + "." + n)) for n in names}


MSG = _build_messages()


class DevicePluginServicer:
    """The five RPCs of v1beta1.DevicePlugin (api.proto:50-76) over a _lib.Plugin."""

    def __init__(self, plugin):
        self.plugin = plugin
        self._cv = threading.Condition()
        self._health_events = 0
        self._stopped = False

    # ref: `unhealthy <- d` feeding `case d := <-m.health` (nvidia.go:222,248,265 -> server.go:258-262)
    def device_unhealthy(self, cached_index):
        self.plugin.mark_unhealthy(cached_index)
        with self._cv:
            self._health_events += 1
            self._cv.notify_all()

    def stop(self):
        with self._cv:
            self._stopped = True
            self._cv.notify_all()

    def _devices(self):
        resp = MSG["ListAndWatchResponse"]()
        for d in self.plugin.api_devices():
            dev = resp.devices.add()
            dev.ID, dev.health = d["ID"], d["health"]
            for node in d.get("topology", {}).get("nodes", []):
                dev.topology.nodes.add().ID = node["ID"]
        return resp

    def GetDevicePluginOptions(self, request, context):
        o = self.plugin.options()
        return MSG["DevicePluginOptions"](pre_start_required=o["pre_start_required"],
                                          get_preferred_allocation_available=o["get_preferred_allocation_available"])

    def ListAndWatch(self, request, context):
        yield self._devices()  # server.go:252
        seen = 0
        while True:
            with self._cv:
                while self._health_events == seen and not self._stopped and context.is_active():
                    self._cv.wait(0.05)
                if self._stopped or not context.is_active():
                    return
                seen += 1
            yield self._devices()  # one re-send per health event (server.go:262)

    def GetPreferredAllocation(self, request, context):
        resp = MSG["PreferredAllocationResponse"]()
        for req in request.container_requests:
            try:
                ids = self.plugin.get_preferred_allocation(list(req.available_deviceIDs), list(req.must_include_deviceIDs),
                                                           req.allocation_size)
            except gpa.PluginError as e:
                context.abort(grpc.StatusCode.UNKNOWN, str(e))  # any error aborts the whole RPC (server.go:276,281,293)
            resp.container_responses.add().deviceIDs.extend(ids)
        return resp

    def Allocate(self, request, context):
        resp = MSG["AllocateResponse"]()
        for req in request.container_requests:
            try:
                r = self.plugin.allocate(list(req.devicesIDs))
            except gpa.PluginError as e:
                context.abort(grpc.StatusCode.UNKNOWN, str(e))
            c = resp.container_responses.add()
            for k, v in r["envs"].items():
                c.envs[k] = v
            for m in r["mounts"]:
                c.mounts.add(container_path=m["container_path"], host_path=m["host_path"])
            for d in r["devices"]:
                c.devices.add(container_path=d["container_path"], host_path=d["host_path"], permissions=d["permissions"])
        return resp

    def PreStartContainer(self, request, context):
        return MSG["PreStartContainerResponse"]()  # server.go:356-358


_METHODS = {
    "GetDevicePluginOptions": ("Empty", "DevicePluginOptions", False),
    "ListAndWatch": ("Empty", "ListAndWatchResponse", True),
    "GetPreferredAllocation": ("PreferredAllocationRequest", "PreferredAllocationResponse", False),
    "Allocate": ("AllocateRequest", "AllocateResponse", False),
    "PreStartContainer": ("PreStartContainerRequest", "PreStartContainerResponse", False),
}


def serve(plugin, socket_path, max_workers=8):
    """Serve v1beta1.DevicePlugin on unix:<socket_path> (server.go:168-175).  -> (grpc server, servicer)"""
    servicer = DevicePluginServicer(plugin)
    handlers = {}
    for name, (req, resp, streaming) in _METHODS.items():
        fn = getattr(servicer, name)
        make = grpc.unary_stream_rpc_method_handler if streaming else grpc.unary_unary_rpc_method_handler
        handlers[name] = make(fn, request_deserializer=MSG[req].FromString, response_serializer=MSG[resp].SerializeToString)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(PACKAGE + ".DevicePlugin", handlers),))
    server.add_insecure_port("unix:" + socket_path)
    server.start()
    return server, servicer


class DevicePluginClient:
    """What kubelet's devicemanager does on the other end of the socket."""

    def __init__(self, socket_path):
        self.channel = grpc.insecure_channel("unix:" + socket_path)
        self._calls = {}
        for name, (req, resp, streaming) in _METHODS.items():
            make = self.channel.unary_stream if streaming else self.channel.unary_unary
            self._calls[name] = make("/%s.DevicePlugin/%s" % (PACKAGE, name), request_serializer=MSG[req].SerializeToString,
                                     response_deserializer=MSG[resp].FromString)

    def __getattr__(self, name):
        try:
            return self._calls[name]
        except KeyError:
            raise AttributeError(name)

    def close(self):
        self.channel.close()
