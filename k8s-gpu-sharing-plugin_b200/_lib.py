"""ctypes binding of libgpualloc.so (the C ABI in include/gpualloc.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C csrc``.  There is no
Python or CPU implementation behind these calls: without the shared object, or without a
CUDA device, they raise.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgpualloc.so")

OK = 0
ERR_INVALID, ERR_CUDA, ERR_NO_TOPOLOGY, ERR_ASYMMETRIC, ERR_TOO_LARGE, ERR_WEIGHT_RANGE, ERR_BUFFER = -1, -2, -3, -4, -5, -6, -7
ERR_NVML, ERR_PANIC, ERR_NO_DEVICES_LEFT, ERR_MUST_MISSING, ERR_UNKNOWN_DEVICE, ERR_NOT_IMPLEMENTED = -8, -9, -10, -11, -12, -13
MODE_GO_FAITHFUL, MODE_INTENDED = 0, 1
MODE_FORCE_ENUM, MODE_FORCE_DP, MODE_FORCE_KERNEL = 2, 4, 8
NIL, FOREIGN = -1, -2
RES_END, RES_ERROR = -128, -127
MAX_AVAILABLE, BATCH_MAX_AVAILABLE = 16, 12


class GpaError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        msg = lib().gpa_strerror(code).decode()
        super().__init__("%s (%d)%s" % (msg, code, (": " + detail) if detail else ""))


class SearchStats(C.Structure):
    _fields_ = [("partitions", C.c_ulonglong), ("items", C.c_ulonglong), ("cut_depth", C.c_int), ("sets", C.c_int),
                ("padding", C.c_int), ("best_score", C.c_int), ("score_unit", C.c_int), ("grid", C.c_int),
                ("block", C.c_int), ("smem_bytes", C.c_int), ("kernel_ms", C.c_float), ("algorithm", C.c_int)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libgpualloc.so is not built (%s); run `python -c 'import __graft_entry__ as e; e.build()'`"
                           % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    i32p, u8p, vp = C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.c_void_p
    L.gpa_strerror.restype = C.c_char_p
    L.gpa_strerror.argtypes = [C.c_int]
    L.gpa_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.gpa_destroy.restype = None
    L.gpa_destroy.argtypes = [vp]
    L.gpa_last_error.restype = C.c_char_p
    L.gpa_last_error.argtypes = [vp]
    L.gpa_launch_count.restype = C.c_ulonglong
    L.gpa_launch_count.argtypes = [vp]
    L.gpa_set_topology.argtypes = [vp, C.c_int, i32p, u8p]
    L.gpa_set_topology_links.argtypes = [vp, C.c_int, i32p]
    L.gpa_link_type_score.argtypes = [C.c_int]
    L.gpa_set_device_order.restype = C.c_int
    L.gpa_set_device_order.argtypes = [vp, i32p, C.c_int]
    L.gpa_get_topology.argtypes = [vp, C.POINTER(C.c_int), i32p, u8p, C.c_int]
    L.gpa_get_answer_table.restype = C.c_int
    L.gpa_get_answer_table.argtypes = [vp, C.POINTER(C.c_uint16), C.c_int, C.POINTER(C.c_int)]
    L.gpa_besteffort.argtypes = [vp, i32p, C.c_int, i32p, C.c_int, C.c_int, C.c_int, i32p, C.POINTER(C.c_int)]
    L.gpa_last_search_stats.argtypes = [vp, C.POINTER(SearchStats)]
    L.gpa_subset_batch.argtypes = [vp, vp, C.c_int, vp]
    L.gpa_subset_batch_device.argtypes = [vp, vp, C.c_int, vp, vp]
    L.gpa_subset_besteffort.argtypes = [vp, i32p, C.c_int, i32p, C.c_int, C.c_int, C.c_int, i32p, C.POINTER(C.c_int)]
    L.gpa_besteffort_batch.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp]
    L.gpa_besteffort_batch_device.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp]
    L.gpa_besteffort_batch_masks.restype = C.c_int
    L.gpa_besteffort_batch_masks.argtypes = [vp, vp, C.c_int, C.c_int, vp]
    L.gpa_besteffort_batch_masks_device.restype = C.c_int
    L.gpa_besteffort_batch_masks_device.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp]
    L.gpa_besteffort_batch_masks8.restype = C.c_int
    L.gpa_besteffort_batch_masks8.argtypes = [vp, vp, C.c_int, vp]
    L.gpa_besteffort_batch_masks8_device.restype = C.c_int
    L.gpa_besteffort_batch_masks8_device.argtypes = [vp, vp, C.c_int, vp, vp]
    L.gpa_pack_mask_requests.restype = C.c_int
    L.gpa_pack_mask_requests.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.POINTER(C.c_int)]
    L.gpa_unpack_mask_results.restype = C.c_int
    L.gpa_unpack_mask_results.argtypes = [vp, vp, C.c_int, vp, C.c_int]
    L.gpa_alloc_host.restype = vp
    L.gpa_alloc_host.argtypes = [C.c_size_t]
    L.gpa_free_host.restype = None
    L.gpa_free_host.argtypes = [vp]
    L.gpa_strip_replicas.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int)]
    L.gpa_prioritize_devices.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_char_p), C.c_int, C.c_int,
                                         C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_char_p, C.c_int]
    L.gpa_get_additional_xids.argtypes = [C.c_char_p, C.POINTER(C.c_uint64), C.c_int]
    for name in ("gpa_create", "gpa_set_topology", "gpa_set_topology_links", "gpa_link_type_score", "gpa_get_topology",
                 "gpa_besteffort", "gpa_last_search_stats", "gpa_besteffort_batch", "gpa_besteffort_batch_device",
                 "gpa_strip_replicas", "gpa_prioritize_devices", "gpa_get_additional_xids"):
        getattr(L, name).restype = C.c_int
    _lib = L
    return L


def _strs(xs):
    arr = (C.c_char_p * max(len(xs), 1))()
    for i, x in enumerate(xs):
        arr[i] = x.encode() if isinstance(x, str) else x
    return arr


def _unpack(buf, n):
    out, off, raw = [], 0, buf.raw
    for _ in range(n):
        end = raw.index(b"\0", off)
        out.append(raw[off:end].decode())
        off = end + 1
    return out


def strip_replicas(ids):
    """ref: stripReplicas, cmd/nvidia-device-plugin/replica.go:32-45"""
    buf = C.create_string_buffer(sum(len(x) + 1 for x in ids) + 16)
    n = C.c_int(0)
    rc = lib().gpa_strip_replicas(_strs(ids), len(ids), buf, len(buf), C.byref(n))
    if rc != OK:
        raise GpaError(rc)
    return _unpack(buf, n.value)


def prioritize_devices(avail, must, size):
    """ref: prioritizeDevices, cmd/nvidia-device-plugin/replica.go:96-198.
    -> (ids or None, error kind in {none, non_unique, no_devices_left, must_missing, panic}, offending id)"""
    buf = C.create_string_buffer(sum(len(x) + 1 for x in list(avail) + list(must)) + 16)
    err = C.create_string_buffer(512)
    n, nu = C.c_int(0), C.c_int(0)
    rc = lib().gpa_prioritize_devices(_strs(avail), len(avail), _strs(must), len(must), int(size), buf, len(buf),
                                      C.byref(n), C.byref(nu), err, len(err))
    if rc == OK:
        return _unpack(buf, n.value), ("non_unique" if nu.value else "none"), ""
    kind = {ERR_NO_DEVICES_LEFT: "no_devices_left", ERR_MUST_MISSING: "must_missing", ERR_PANIC: "panic"}.get(rc)
    if kind is None:
        raise GpaError(rc)
    return None, kind, err.value.decode()


def get_additional_xids(s):
    """ref: getAdditionalXids, cmd/nvidia-device-plugin/nvidia.go:274-294"""
    out = (C.c_uint64 * 256)()
    n = lib().gpa_get_additional_xids(s.encode() if isinstance(s, str) else s, out, 256)
    if n < 0:
        raise GpaError(n)
    return [int(out[i]) for i in range(n)]


class ResourceVariant(C.Structure):
    _fields_ = [("orig", C.c_char * 64), ("name", C.c_char * 64), ("replicas", C.c_ulonglong), ("auto_replicas", C.c_int)]


def parse_resource_config(config):
    """ref: parseResourceConfig, cmd/nvidia-device-plugin/main.go:171-203.
    -> {orig: {"name", "replicas", "auto_replicas"}} in first-seen order; raises ValueError with the reference's text"""
    L = lib()
    L.gpa_parse_resource_config.argtypes = [C.c_char_p, C.POINTER(ResourceVariant), C.c_int, C.POINTER(C.c_int), C.c_char_p, C.c_int]
    out = (ResourceVariant * 64)()
    n = C.c_int(0)
    err = C.create_string_buffer(256)
    rc = L.gpa_parse_resource_config(config.encode() if isinstance(config, str) else config, out, 64, C.byref(n), err, len(err))
    if rc == ERR_INVALID and err.value:
        raise ValueError(err.value.decode())
    if rc != OK:
        raise GpaError(rc)
    return {out[i].orig.decode(): {"name": out[i].name.decode(), "replicas": int(out[i].replicas),
                                   "auto_replicas": bool(out[i].auto_replicas)} for i in range(n.value)}


def resource_config_get(variants, orig):
    """ref: resourceConfiguration.Get, mig-strategy.go:66-76 (absent => same name, ZERO replicas)"""
    L = lib()
    L.gpa_resource_config_get.restype = None
    L.gpa_resource_config_get.argtypes = [C.POINTER(ResourceVariant), C.c_int, C.c_char_p, C.POINTER(ResourceVariant)]
    arr = (ResourceVariant * max(len(variants), 1))()
    for i, (k, v) in enumerate(variants.items()):
        arr[i].orig, arr[i].name = k.encode(), v["name"].encode()
        arr[i].replicas, arr[i].auto_replicas = v["replicas"], int(v["auto_replicas"])
    out = ResourceVariant()
    L.gpa_resource_config_get(arr, len(variants), orig.encode(), C.byref(out))
    return {"name": out.name.decode(), "replicas": int(out.replicas), "auto_replicas": bool(out.auto_replicas)}


def link_type_score(t):
    return int(lib().gpa_link_type_score(int(t)))


class Context:
    """One gpa_ctx bound to a CUDA device."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().gpa_create(int(device), C.byref(self._h))
        if rc != OK:
            raise GpaError(rc, "gpa_create(device=%d) — a CUDA device is required, there is no CPU fall-back" % device)

    def close(self):
        if self._h:
            lib().gpa_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != OK:
            raise GpaError(rc, lib().gpa_last_error(self._h).decode())

    @property
    def launch_count(self):
        return int(lib().gpa_launch_count(self._h))

    def set_topology(self, W, nlinks):
        W = np.ascontiguousarray(W, dtype=np.int32)
        nl = np.ascontiguousarray(nlinks, dtype=np.uint8)
        self._check(lib().gpa_set_topology(self._h, W.shape[0], W.ctypes.data_as(C.POINTER(C.c_int32)),
                                           nl.ctypes.data_as(C.POINTER(C.c_uint8))))

    def set_topology_links(self, links2):
        l2 = np.ascontiguousarray(links2, dtype=np.int32)
        self._check(lib().gpa_set_topology_links(self._h, l2.shape[0], l2.ctypes.data_as(C.POINTER(C.c_int32))))

    def set_device_order(self, order):
        """order[label] = NVML index: the node-wide order requests list their devices in (sorted-UUID order)"""
        o = np.ascontiguousarray(order, dtype=np.int32)
        self._check(lib().gpa_set_device_order(self._h, o.ctypes.data_as(C.POINTER(C.c_int32)), len(o)))

    def get_topology(self):
        n = C.c_int(0)
        self._check(lib().gpa_get_topology(self._h, C.byref(n), None, None, 0))
        W = np.zeros((n.value, n.value), dtype=np.int32)
        nl = np.zeros((n.value, n.value), dtype=np.uint8)
        self._check(lib().gpa_get_topology(self._h, C.byref(n), W.ctypes.data_as(C.POINTER(C.c_int32)),
                                           nl.ctypes.data_as(C.POINTER(C.c_uint8)), n.value * n.value))
        return W, nl

    def answer_table(self):
        """the GPU-built answer table of the installed topology (nodes of up to 8 GPUs) as a uint16 array, or None"""
        n = C.c_int(0)
        rc = lib().gpa_get_answer_table(self._h, None, 0, C.byref(n))
        if rc == ERR_TOO_LARGE:
            return None
        self._check(rc)
        tab = np.zeros(n.value, dtype=np.uint16)
        self._check(lib().gpa_get_answer_table(self._h, tab.ctypes.data_as(C.POINTER(C.c_uint16)), n.value, C.byref(n)))
        return tab

    def besteffort(self, avail, req, size, mode=MODE_GO_FAITHFUL):
        """Policy.Allocate(available, required, size) -> (rc, [NVML indices, NIL for a pad])"""
        av = np.ascontiguousarray(avail, dtype=np.int32)
        rq = np.ascontiguousarray(req, dtype=np.int32)
        out = np.full(max(int(size), 1) + 32, -9, dtype=np.int32)
        n_out = C.c_int(0)
        rc = lib().gpa_besteffort(self._h, av.ctypes.data_as(C.POINTER(C.c_int32)), len(av),
                                  rq.ctypes.data_as(C.POINTER(C.c_int32)), len(rq), int(size), int(mode),
                                  out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out))
        return rc, [int(x) for x in out[:n_out.value]]

    def subset_besteffort(self, phys, req_pos, size, mode=MODE_GO_FAITHFUL):
        """Replica-set search (no reference counterpart, see gpualloc.h): phys[p] = NVML index of the GPU behind
        available virtual device p, req_pos = positions that must be chosen -> (rc, [positions, ascending])"""
        ph = np.ascontiguousarray(phys, dtype=np.int32)
        rq = np.ascontiguousarray(req_pos, dtype=np.int32)
        out = np.full(max(int(size), 1) + 8, -9, dtype=np.int32)
        n_out = C.c_int(0)
        rc = lib().gpa_subset_besteffort(self._h, ph.ctypes.data_as(C.POINTER(C.c_int32)), len(ph),
                                         rq.ctypes.data_as(C.POINTER(C.c_int32)), len(rq), int(size), int(mode),
                                         out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out))
        return rc, [int(x) for x in out[:n_out.value]]

    def subset_batch(self, reqs, out=None):
        """reqs: numpy array of synthetic.SUBSET_REQ_DTYPE (host memory) -> array of SUBSET_RES_DTYPE"""
        from . import synthetic as syn
        reqs = np.ascontiguousarray(reqs)
        assert reqs.dtype == syn.SUBSET_REQ_DTYPE
        if out is None:
            out = np.zeros(len(reqs), dtype=syn.SUBSET_RES_DTYPE)
        self._check(lib().gpa_subset_batch(self._h, reqs.ctypes.data, len(reqs), out.ctypes.data))
        return out

    def subset_batch_device(self, d_reqs_ptr, count, d_out_ptr, stream=None):
        self._check(lib().gpa_subset_batch_device(self._h, C.c_void_p(d_reqs_ptr), int(count), C.c_void_p(d_out_ptr),
                                                  C.c_void_p(stream or 0)))

    def search_stats(self):
        st = SearchStats()
        self._check(lib().gpa_last_search_stats(self._h, C.byref(st)))
        return {f: getattr(st, f) for f, _ in SearchStats._fields_}

    def besteffort_batch(self, reqs, mode=MODE_GO_FAITHFUL, out=None, max_avail=0):
        """reqs: numpy array of synthetic.REQ_DTYPE (host memory) -> array of RES_DTYPE"""
        from . import synthetic as syn
        reqs = np.ascontiguousarray(reqs)
        assert reqs.dtype == syn.REQ_DTYPE
        if out is None:
            out = np.zeros(len(reqs), dtype=syn.RES_DTYPE)
        self._check(lib().gpa_besteffort_batch(self._h, reqs.ctypes.data, len(reqs), int(max_avail), int(mode),
                                               out.ctypes.data))
        return out

    def besteffort_batch_masks(self, reqs, mode=MODE_GO_FAITHFUL, out=None):
        """reqs: numpy array of synthetic.MASK_REQ_DTYPE (host memory) -> array of MASK_RES_DTYPE"""
        from . import synthetic as syn
        reqs = np.ascontiguousarray(reqs)
        assert reqs.dtype == syn.MASK_REQ_DTYPE
        if out is None:
            out = np.zeros(len(reqs), dtype=syn.MASK_RES_DTYPE)
        self._check(lib().gpa_besteffort_batch_masks(self._h, reqs.ctypes.data, len(reqs), int(mode), out.ctypes.data))
        return out

    def besteffort_batch_masks8(self, reqs, out=None):
        """reqs: numpy array of synthetic.MASK8_REQ_DTYPE (host memory) -> array of MASK8_RES_DTYPE; nodes of up to 8 GPUs"""
        from . import synthetic as syn
        reqs = np.ascontiguousarray(reqs)
        assert reqs.dtype == syn.MASK8_REQ_DTYPE
        if out is None:
            out = np.zeros(len(reqs), dtype=syn.MASK8_RES_DTYPE)
        self._check(lib().gpa_besteffort_batch_masks8(self._h, reqs.ctypes.data, len(reqs), out.ctypes.data))
        return out

    def besteffort_batch_masks8_device(self, d_reqs_ptr, count, d_out_ptr, stream=None):
        self._check(lib().gpa_besteffort_batch_masks8_device(self._h, C.c_void_p(d_reqs_ptr), int(count), C.c_void_p(d_out_ptr),
                                                             C.c_void_p(stream or 0)))

    def pack_mask_requests(self, reqs, out=None, threads=0):
        """REQ_DTYPE id-list records -> (MASK_REQ_DTYPE records, number that cannot be masks); host code"""
        from . import synthetic as syn
        reqs = np.ascontiguousarray(reqs)
        assert reqs.dtype == syn.REQ_DTYPE
        if out is None:
            out = np.zeros(len(reqs), dtype=syn.MASK_REQ_DTYPE)
        bad = C.c_int(0)
        self._check(lib().gpa_pack_mask_requests(self._h, reqs.ctypes.data, len(reqs), out.ctypes.data, int(threads), C.byref(bad)))
        return out, bad.value

    def unpack_mask_results(self, res, out=None, threads=0):
        """MASK_RES_DTYPE results -> RES_DTYPE device lists; host code"""
        from . import synthetic as syn
        res = np.ascontiguousarray(res)
        assert res.dtype == syn.MASK_RES_DTYPE
        if out is None:
            out = np.zeros(len(res), dtype=syn.RES_DTYPE)
        self._check(lib().gpa_unpack_mask_results(self._h, res.ctypes.data, len(res), out.ctypes.data, int(threads)))
        return out

    def besteffort_batch_masks_device(self, d_reqs_ptr, count, d_out_ptr, mode=MODE_GO_FAITHFUL, stream=None):
        self._check(lib().gpa_besteffort_batch_masks_device(self._h, C.c_void_p(d_reqs_ptr), int(count), int(mode),
                                                            C.c_void_p(d_out_ptr), C.c_void_p(stream or 0)))

    def besteffort_batch_device(self, d_reqs_ptr, count, max_avail, d_out_ptr, mode=MODE_GO_FAITHFUL, stream=None):
        self._check(lib().gpa_besteffort_batch_device(self._h, C.c_void_p(d_reqs_ptr), int(count), int(max_avail),
                                                      int(mode), C.c_void_p(d_out_ptr), C.c_void_p(stream or 0)))


def decode_result(rec):
    """RES_DTYPE record -> (rc, [devices])"""
    dev = rec["dev"]
    if dev[0] == RES_ERROR:
        return int(dev[1]), []
    out = []
    for d in dev:
        if d == RES_END:
            break
        out.append(int(d))
    return OK, out


class PinnedBuffer:
    """Page-locked host memory from gpa_alloc_host, viewed as a numpy array (``.array``)."""

    def __init__(self, count, dtype):
        dtype = np.dtype(dtype)
        self.nbytes = max(int(count) * dtype.itemsize, 1)
        self.ptr = lib().gpa_alloc_host(self.nbytes)
        if not self.ptr:
            raise GpaError(ERR_CUDA, "gpa_alloc_host(%d)" % self.nbytes)
        self._buf = (C.c_uint8 * self.nbytes).from_address(self.ptr)
        self.array = np.frombuffer(self._buf, dtype=dtype, count=int(count))

    def free(self):
        if self.ptr:
            self.array = None
            self._buf = None
            lib().gpa_free_host(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------
# NVML topology + health, plugin server bodies (host code of the path)

TOPO_PARITY, TOPO_NVSWITCH_AWARE = 0, 1
LIST_ENVVAR, LIST_VOLUME_MOUNTS = 0, 1
ID_UUID, ID_INDEX = 0, 1
FIX_HEALTH_PROPAGATION, FIX_REPLICA_SUFFIX, FIX_REQUIRED_IDENTITY = 1, 2, 4
OPT_REPLICA_TOPOLOGY = 8  # not a fix: replica requests answered by the replica-set search (gpualloc.h)


class PluginConfig(C.Structure):
    _fields_ = [("resource_name", C.c_char_p), ("device_list_envvar", C.c_char_p), ("driver_root", C.c_char_p),
                ("replicas", C.c_uint32), ("auto_replicas", C.c_int), ("device_list_strategy", C.c_int),
                ("device_id_strategy", C.c_int), ("pass_device_specs", C.c_int), ("has_policy", C.c_int),
                ("fixes", C.c_int)]


class PluginDevice(C.Structure):
    _fields_ = [("id", C.c_char_p), ("index", C.c_char_p), ("paths", C.POINTER(C.c_char_p)), ("n_paths", C.c_int),
                ("total_memory_mib", C.c_ulonglong), ("numa_node", C.c_int), ("nvml_index", C.c_int)]


def _bind_host(L):
    vp = C.c_void_p
    if getattr(L, "_gpa_host_bound", False):
        return L
    L.gpa_nvml_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.gpa_nvml_close.restype = None
    L.gpa_nvml_close.argtypes = [vp]
    L.gpa_nvml_last_error.restype = C.c_char_p
    L.gpa_nvml_last_error.argtypes = [vp]
    L.gpa_nvml_device_count.argtypes = [vp, C.POINTER(C.c_int)]
    L.gpa_nvml_device_uuid.argtypes = [vp, C.c_int, C.c_char_p, C.c_int]
    L.gpa_nvml_read_topology.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int32), C.c_int,
                                         C.POINTER(C.c_ulonglong)]
    L.gpa_set_topology_from_nvml.argtypes = [vp, vp, C.c_int]
    L.gpa_health_open.argtypes = [vp, C.POINTER(C.c_char_p), C.c_int, C.c_char_p, C.c_int, C.POINTER(vp)]
    L.gpa_health_poll.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_ulonglong)]
    L.gpa_health_stats.argtypes = [vp, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong), C.POINTER(C.c_int)]
    L.gpa_health_close.restype = None
    L.gpa_health_close.argtypes = [vp]
    L.gpa_plugin_create.argtypes = [vp, C.POINTER(PluginConfig), C.POINTER(PluginDevice), C.c_int, C.POINTER(vp)]
    L.gpa_plugin_destroy.restype = None
    L.gpa_plugin_destroy.argtypes = [vp]
    L.gpa_plugin_get_options.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.gpa_plugin_device_count.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.gpa_plugin_list_devices.argtypes = [vp, C.c_char_p, C.c_int]
    L.gpa_plugin_mark_unhealthy.argtypes = [vp, C.c_int]
    L.gpa_plugin_get_preferred_allocation.argtypes = [vp, C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_char_p), C.c_int,
                                                      C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.c_char_p, C.c_int]
    L.gpa_plugin_allocate.argtypes = [vp, C.POINTER(C.c_char_p), C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    for name in ("gpa_nvml_open", "gpa_nvml_device_count", "gpa_nvml_device_uuid", "gpa_nvml_read_topology",
                 "gpa_set_topology_from_nvml", "gpa_health_open", "gpa_health_poll", "gpa_health_stats", "gpa_plugin_create",
                 "gpa_plugin_get_options", "gpa_plugin_device_count", "gpa_plugin_list_devices", "gpa_plugin_mark_unhealthy",
                 "gpa_plugin_get_preferred_allocation", "gpa_plugin_allocate"):
        getattr(L, name).restype = C.c_int
    L._gpa_host_bound = True
    return L


class Nvml:
    """gpa_nvml: NVML reached through dlopen (libpath=None -> libnvidia-ml.so.1)."""

    def __init__(self, libpath=None):
        L = _bind_host(lib())
        self._h = C.c_void_p()
        rc = L.gpa_nvml_open(libpath.encode() if libpath else None, C.byref(self._h))
        if rc != OK:
            raise GpaError(rc, "gpa_nvml_open(%s)" % (libpath or "libnvidia-ml.so.1"))

    def close(self):
        if self._h:
            lib().gpa_nvml_close(self._h)
            self._h = C.c_void_p()

    def _check(self, rc):
        if rc != OK:
            raise GpaError(rc, lib().gpa_nvml_last_error(self._h).decode())

    def device_count(self):
        n = C.c_int(0)
        self._check(lib().gpa_nvml_device_count(self._h, C.byref(n)))
        return n.value

    def device_uuid(self, i):
        buf = C.create_string_buffer(96)
        self._check(lib().gpa_nvml_device_uuid(self._h, i, buf, len(buf)))
        return buf.value.decode()

    def device_info(self, i):
        """-> {"total_memory": MiB, "minor": n or -1, "numa": node or -1, "bus_id": str} (gpa_nvml_device_info)"""
        L = lib()
        L.gpa_nvml_device_info.restype = C.c_int
        L.gpa_nvml_device_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_ulonglong), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                           C.c_char_p, C.c_int]
        mem, minor, numa = C.c_ulonglong(0), C.c_int(-1), C.c_int(-1)
        bus = C.create_string_buffer(64)
        self._check(L.gpa_nvml_device_info(self._h, int(i), C.byref(mem), C.byref(minor), C.byref(numa), bus, len(bus)))
        return {"total_memory": int(mem.value), "minor": minor.value, "numa": numa.value, "bus_id": bus.value.decode()}

    def read_topology(self, flags=TOPO_PARITY):
        """-> (links2[n, n, 2] int32, nvml_calls)"""
        n = C.c_int(0)
        self._check(lib().gpa_nvml_read_topology(self._h, flags, C.byref(n), None, 0, None))
        links = np.zeros((n.value, n.value, 2), dtype=np.int32)
        calls = C.c_ulonglong(0)
        self._check(lib().gpa_nvml_read_topology(self._h, flags, C.byref(n), links.ctypes.data_as(C.POINTER(C.c_int32)),
                                                 links.size, C.byref(calls)))
        return links, int(calls.value)

    def install_topology(self, ctx, flags=TOPO_PARITY):
        self._check(lib().gpa_set_topology_from_nvml(ctx._h, self._h, flags))


class HealthWatcher:
    """gpa_health: the XID watcher of checkHealth (cmd/nvidia-device-plugin/nvidia.go:181-269)."""

    def __init__(self, nvml, device_ids, disable_env="", wait_timeout_ms=0):
        L = _bind_host(lib())
        self._nv = nvml
        self._h = C.c_void_p()
        rc = L.gpa_health_open(nvml._h, _strs(device_ids), len(device_ids),
                               None if disable_env is None else disable_env.encode(), wait_timeout_ms, C.byref(self._h))
        if rc != OK:
            raise GpaError(rc, lib().gpa_nvml_last_error(nvml._h).decode())

    def poll(self, timeout_ms=0):
        """-> (device index, xid) or None"""
        idx, xid = C.c_int(-1), C.c_ulonglong(0)
        rc = lib().gpa_health_poll(self._h, timeout_ms, C.byref(idx), C.byref(xid))
        if rc < 0:
            raise GpaError(rc)
        return (idx.value, int(xid.value)) if rc == 1 else None

    def stats(self):
        a, b, d = C.c_ulonglong(0), C.c_ulonglong(0), C.c_int(0)
        lib().gpa_health_stats(self._h, C.byref(a), C.byref(b), C.byref(d))
        return {"events_seen": int(a.value), "events_skipped": int(b.value), "disabled": bool(d.value)}

    def close(self):
        if self._h:
            lib().gpa_health_close(self._h)
            self._h = C.c_void_p()


class PluginError(RuntimeError):
    def __init__(self, code, message):
        self.code = code
        super().__init__(message)


class Plugin:
    """gpa_plugin: the RPC bodies of NvidiaDevicePlugin (cmd/nvidia-device-plugin/server.go) without the transport."""

    def __init__(self, devices, ctx=None, resource_name="nvidia.com/gpu", envvar="NVIDIA_VISIBLE_DEVICES", driver_root="/",
                 replicas=1, auto_replicas=False, list_strategy=LIST_ENVVAR, id_strategy=ID_UUID, pass_device_specs=False,
                 has_policy=True, fixes=0):
        import json
        self._json = json
        L = _bind_host(lib())
        self._keep = []
        cfg = PluginConfig(resource_name.encode(), envvar.encode(), driver_root.encode(), replicas, int(auto_replicas),
                           list_strategy, id_strategy, int(pass_device_specs), int(has_policy), fixes)
        arr = (PluginDevice * max(len(devices), 1))()
        for i, d in enumerate(devices):
            paths = _strs(d.get("paths", []))
            self._keep.append(paths)
            arr[i] = PluginDevice(d["id"].encode(), str(d.get("index", i)).encode(), paths, len(d.get("paths", [])),
                                  int(d.get("total_memory", 0)), int(d.get("numa", -1)), int(d.get("nvml_index", i)))
        self._h = C.c_void_p()
        rc = L.gpa_plugin_create(ctx._h if ctx is not None else None, C.byref(cfg), arr, len(devices), C.byref(self._h))
        if rc != OK:
            raise GpaError(rc)

    def close(self):
        if self._h:
            lib().gpa_plugin_destroy(self._h)
            self._h = C.c_void_p()

    def options(self):
        a, b = C.c_int(0), C.c_int(0)
        lib().gpa_plugin_get_options(self._h, C.byref(a), C.byref(b))
        return {"pre_start_required": bool(a.value), "get_preferred_allocation_available": bool(b.value)}

    def device_count(self):
        a, b = C.c_int(0), C.c_int(0)
        lib().gpa_plugin_device_count(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    def api_devices(self):
        _, nrep = self.device_count()
        buf = C.create_string_buffer(256 * nrep + 64)
        rc = lib().gpa_plugin_list_devices(self._h, buf, len(buf))
        if rc != OK:
            raise GpaError(rc)
        return self._json.loads(buf.value.decode())

    def mark_unhealthy(self, cached_index):
        rc = lib().gpa_plugin_mark_unhealthy(self._h, cached_index)
        if rc < 0:
            raise GpaError(rc)
        return rc

    def get_preferred_allocation(self, available_ids, must_ids, size):
        buf = C.create_string_buffer(sum(len(x) + 16 for x in list(available_ids) + list(must_ids)) + 64)
        err = C.create_string_buffer(1024)
        n = C.c_int(0)
        rc = lib().gpa_plugin_get_preferred_allocation(self._h, _strs(available_ids), len(available_ids), _strs(must_ids),
                                                       len(must_ids), int(size), buf, len(buf), C.byref(n), err, len(err))
        if rc != OK:
            raise PluginError(rc, err.value.decode() or lib().gpa_strerror(rc).decode())
        return _unpack(buf, n.value)

    def allocate(self, ids):
        buf = C.create_string_buffer(4096 + 512 * len(ids))
        err = C.create_string_buffer(1024)
        rc = lib().gpa_plugin_allocate(self._h, _strs(ids), len(ids), buf, len(buf), err, len(err))
        if rc != OK:
            raise PluginError(rc, err.value.decode() or lib().gpa_strerror(rc).decode())
        return self._json.loads(buf.value.decode())


def probe_hbm(device=0, nbytes=4 << 30, reps=20, mode=1):
    """-> (best GB/s, median GB/s); mode 0 = read-only, 1 = copy (read + write bytes)"""
    L = lib()
    L.gpa_probe_hbm.restype = C.c_int
    L.gpa_probe_hbm.argtypes = [C.c_int, C.c_ulonglong, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                C.c_char_p, C.c_int]
    best, med = C.c_double(0), C.c_double(0)
    err = C.create_string_buffer(256)
    rc = L.gpa_probe_hbm(device, nbytes, reps, mode, C.byref(best), C.byref(med), err, len(err))
    if rc != OK:
        raise GpaError(rc, err.value.decode())
    return best.value, med.value


def probe_p2p(n_devices, nbytes=1 << 30, reps=10, mode=0):
    """-> (matrix[n, n] GB/s one way: row pulls from column, can_access[n, n])"""
    L = lib()
    L.gpa_probe_p2p.restype = C.c_int
    L.gpa_probe_p2p.argtypes = [C.c_int, C.c_ulonglong, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int),
                                C.c_char_p, C.c_int]
    m = np.zeros((n_devices, n_devices), dtype=np.float64)
    acc = np.zeros((n_devices, n_devices), dtype=np.int32)
    err = C.create_string_buffer(256)
    rc = L.gpa_probe_p2p(n_devices, nbytes, reps, mode, m.ctypes.data_as(C.POINTER(C.c_double)),
                         acc.ctypes.data_as(C.POINTER(C.c_int)), err, len(err))
    if rc != OK:
        raise GpaError(rc, err.value.decode())
    return m, acc


def probe_lut(device=0, log2_entries=8, reps=5):
    """shared-memory table gather + integer add ceiling, 10^9 lookups/s over the whole GPU"""
    L = lib()
    L.gpa_probe_lut.restype = C.c_int
    L.gpa_probe_lut.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_char_p, C.c_int]
    best = C.c_double(0)
    err = C.create_string_buffer(256)
    rc = L.gpa_probe_lut(device, log2_entries, reps, C.byref(best), err, len(err))
    if rc != OK:
        raise GpaError(rc, err.value.decode())
    return best.value
