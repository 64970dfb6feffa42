"""ctypes front end of the test oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")

NIL = -1      # a nil *Device in a returned set (defect D5)
FOREIGN = -2  # a `required` pointer that equals no `available` pointer (defect D4)

OK, PANIC_ASYMMETRIC, PANIC_INTERNAL, ERR_NOMEM, ERR_ARG = 0, 1, 2, 3, 4
GO_FAITHFUL, INTENDED = 0, 1

ERR_NONE, ERR_NON_UNIQUE, ERR_NO_DEVICES_LEFT, ERR_MUST_MISSING, ERR_PANIC = 0, 1, 2, 3, 4
ERR_KIND_NAMES = {0: "none", 1: "non_unique", 2: "no_devices_left", 3: "must_missing", 4: "panic"}


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in ("ref_besteffort.c", "ref_replica.c", "ref_batch.c", "ref_subset.c")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.check_call(["make", "-C", HERE, "-s", "-B", "liboracle.so"])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = C.CDLL(LIB)
        i32p = C.POINTER(C.c_int32)
        u8p = C.POINTER(C.c_uint8)
        for name in ("ref_besteffort_literal", "ref_besteffort_fast"):
            f = getattr(L, name)
            f.restype = C.c_int
            f.argtypes = [C.c_int, i32p, u8p, i32p, C.c_int, i32p, C.c_int, C.c_int, C.c_int, i32p,
                          C.POINTER(C.c_int), C.POINTER(C.c_ulonglong)]
        L.ref_partition_count.restype = C.c_ulonglong
        L.ref_partition_count.argtypes = [C.c_int, C.c_int]
        L.ref_go_growcap.restype = C.c_int
        L.ref_go_growcap.argtypes = [C.c_int, C.c_int, C.c_int]
        L.ref_link_type_score.restype = C.c_int
        L.ref_link_type_score.argtypes = [C.c_int]
        L.ref_links_to_weights.restype = None
        L.ref_links_to_weights.argtypes = [C.c_int, i32p, i32p, u8p]
        L.ref_strip_replicas.restype = C.c_int
        L.ref_strip_replicas.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int)]
        L.ref_prioritize_devices.restype = C.c_int
        L.ref_prioritize_devices.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_char_p), C.c_int, C.c_int,
                                             C.c_char_p, C.c_int, C.POINTER(C.c_int), C.c_char_p, C.c_int]
        L.ref_besteffort_batch.restype = C.c_ulonglong
        L.ref_besteffort_batch.argtypes = [C.c_int, i32p, u8p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_int]
        L.ref_get_additional_xids.restype = C.c_int
        L.ref_get_additional_xids.argtypes = [C.c_char_p, C.POINTER(C.c_uint64), C.c_int]
        L.ref_subset_besteffort.restype = C.c_int
        L.ref_subset_besteffort.argtypes = [C.c_int, i32p, u8p, i32p, C.c_int, i32p, C.c_int, C.c_int, i32p,
                                            C.POINTER(C.c_int), C.POINTER(C.c_longlong), C.POINTER(C.c_ulonglong)]
        _lib = L
    return _lib


def _i32(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(C.POINTER(C.c_int32))


def besteffort(W, nlinks, avail, req, size, mode=GO_FAITHFUL, variant="fast", want_visited=False):
    """-> (rc, [device indices, NIL for a nil pad]) [, visited]"""
    W = np.ascontiguousarray(W, dtype=np.int32)
    nt = W.shape[0]
    nl = np.ascontiguousarray(nlinks, dtype=np.uint8)
    av, avp = _i32(avail)
    rq, rqp = _i32(req)
    out = np.full(max(int(size), 1) + 64, -9, dtype=np.int32)
    n_out = C.c_int(0)
    visited = C.c_ulonglong(0)
    f = lib().ref_besteffort_literal if variant == "literal" else lib().ref_besteffort_fast
    rc = f(nt, W.ctypes.data_as(C.POINTER(C.c_int32)), nl.ctypes.data_as(C.POINTER(C.c_uint8)), avp, len(av), rqp,
           len(rq), int(size), int(mode), out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out), C.byref(visited))
    res = (rc, [int(x) for x in out[:n_out.value]])
    return res + (visited.value,) if want_visited else res


def go_growcap(oldcap, needed, elemsize):
    """capacity after an append() that had to grow (Go 1.17 runtime.growslice + roundupsize, restated)"""
    return int(lib().ref_go_growcap(oldcap, needed, elemsize))


def partition_count(n, k):
    return int(lib().ref_partition_count(n, k))


def link_type_score(t):
    return int(lib().ref_link_type_score(t))


def links_to_weights(links2):
    links2 = np.ascontiguousarray(links2, dtype=np.int32)
    nt = links2.shape[0]
    W = np.zeros((nt, nt), dtype=np.int32)
    nl = np.zeros((nt, nt), dtype=np.uint8)
    lib().ref_links_to_weights(nt, links2.ctypes.data_as(C.POINTER(C.c_int32)),
                               W.ctypes.data_as(C.POINTER(C.c_int32)), nl.ctypes.data_as(C.POINTER(C.c_uint8)))
    return W, nl


def _strs(xs):
    arr = (C.c_char_p * max(len(xs), 1))()
    for i, x in enumerate(xs):
        arr[i] = x.encode() if isinstance(x, str) else x
    return arr


def _unpack(buf, n):
    out, off = [], 0
    raw = buf.raw
    for _ in range(n):
        end = raw.index(b"\0", off)
        out.append(raw[off:end].decode())
        off = end + 1
    return out


def strip_replicas(ids):
    buf = C.create_string_buffer(sum(len(x) + 1 for x in ids) + 16)
    n = C.c_int(0)
    rc = lib().ref_strip_replicas(_strs(ids), len(ids), buf, len(buf), C.byref(n))
    assert rc == 0
    return _unpack(buf, n.value)


def prioritize_devices(avail, must, size):
    """-> (result list or None, error kind name, error arg)"""
    buf = C.create_string_buffer(sum(len(x) + 1 for x in list(avail) + list(must)) + 16)
    err = C.create_string_buffer(512)
    n = C.c_int(0)
    rc = lib().ref_prioritize_devices(_strs(avail), len(avail), _strs(must), len(must), int(size), buf, len(buf),
                                      C.byref(n), err, len(err))
    assert rc >= 0
    kind = ERR_KIND_NAMES[rc]
    res = _unpack(buf, n.value) if rc in (ERR_NONE, ERR_NON_UNIQUE) else None
    return res, kind, err.value.decode()


def get_additional_xids(s):
    out = (C.c_uint64 * 256)()
    n = lib().ref_get_additional_xids(s.encode() if isinstance(s, str) else s, out, 256)
    return [int(out[i]) for i in range(n)]


def besteffort_batch(W, nlinks, reqs, mode=GO_FAITHFUL, threads=1, out=None):
    """packed gpa_request records (numpy, 32 B each) -> packed results (16 B each), visited"""
    W = np.ascontiguousarray(W, dtype=np.int32)
    nl = np.ascontiguousarray(nlinks, dtype=np.uint8)
    reqs = np.ascontiguousarray(reqs)
    assert reqs.dtype.itemsize == 32
    if out is None:
        out = np.zeros(len(reqs), dtype=np.dtype([("dev", np.int8, 16)]))
    visited = lib().ref_besteffort_batch(W.shape[0], W.ctypes.data_as(C.POINTER(C.c_int32)),
                                         nl.ctypes.data_as(C.POINTER(C.c_uint8)), reqs.ctypes.data, len(reqs), int(mode),
                                         out.ctypes.data, int(threads))
    return out, int(visited)


def subset_besteffort(W, nlinks, phys, req_pos, size, want_stats=False):
    """Replica-set search, literal form (oracle/ref_subset.c) -> (rc, [positions]) [, best score, candidates seen]"""
    W = np.ascontiguousarray(W, dtype=np.int32)
    nl = np.ascontiguousarray(nlinks, dtype=np.uint8)
    ph, php = _i32(phys)
    rq, rqp = _i32(req_pos)
    out = np.full(max(int(size), 1), -9, dtype=np.int32)
    n_out = C.c_int(0)
    score = C.c_longlong(0)
    seen = C.c_ulonglong(0)
    rc = lib().ref_subset_besteffort(W.shape[0], W.ctypes.data_as(C.POINTER(C.c_int32)), nl.ctypes.data_as(C.POINTER(C.c_uint8)),
                                     php, len(ph), rqp, len(rq), int(size), out.ctypes.data_as(C.POINTER(C.c_int32)),
                                     C.byref(n_out), C.byref(score), C.byref(seen))
    res = (rc, [int(x) for x in out[:n_out.value]])
    return res + (score.value, seen.value) if want_stats else res
