"""The allocation-free batch entry point, gpa_besteffort_batch_noalloc, against the regular one.

Some users of this library cannot allocate when a request arrives: the plugin runs on GPUs whose
memory belongs to the pods it shares them with, so device memory may be full, and cudaMalloc /
cudaFree would stall every tenant of the GPU.  For them the library keeps its device-side staging
in a fixed array inside the module.  This test calls that entry point through the C ABI twice in
a row — the second call with another topology and a smaller batch, so anything left in the shared
array by the first call would show — and requires byte-identical results to gpa_besteffort_batch
(integer work: no tolerance).  It lives in a file that pytest collects last; it must pass both in
a fresh process (`pytest tests/test_zz_no_alloc.py -m gpu`) and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import synthetic as syn

pytestmark = pytest.mark.gpu

MAX_REQUESTS = 4096  # GPA_NOALLOC_MAX_REQUESTS


def noalloc(W, nlinks, reqs, max_avail=0, mode=gpa.MODE_GO_FAITHFUL, device=0):
    L = gpa.lib()
    f = L.gpa_besteffort_batch_noalloc
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.c_void_p, C.c_int, C.c_int, C.c_int,
                  C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
    W = np.ascontiguousarray(W, dtype=np.int32)
    nl = np.ascontiguousarray(nlinks, dtype=np.uint8)
    reqs = np.ascontiguousarray(reqs)
    out = np.full(len(reqs), 77, dtype=np.int8).repeat(16).reshape(-1, 16).copy().view(syn.RES_DTYPE).reshape(-1)
    err = C.create_string_buffer(256)
    rc = f(device, W.shape[0], W.ctypes.data_as(C.POINTER(C.c_int32)), nl.ctypes.data_as(C.POINTER(C.c_uint8)),
           reqs.ctypes.data, len(reqs), max_avail, mode, out.ctypes.data, None, err, len(err))
    return rc, out, err.value.decode()


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    order = syn.sorted_positions(syn.node_uuids(8, seed=1))
    calls = [(syn.topo_dgx1v_like(), syn.random_request_batch(8, MAX_REQUESTS, seed=41, order=order)),
             (syn.topo_random(12, 5), syn.random_request_batch(12, 1500, seed=42))]
    for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
        # the two allocation-free calls, back to back (in a fresh process these are the first CUDA work at all,
        # so the entry point's own first-call set-up is what gets exercised)
        got = []
        for topo, reqs in calls:
            W, nl = syn.links_to_weights(topo)
            rc, res, msg = noalloc(W, nl, reqs, mode=mode)
            assert rc == gpa.OK, msg
            got.append(res)
        # what the existing entry point answers
        want = []
        ctx = gpa.Context(0)
        try:
            for topo, reqs in calls:
                W, nl = syn.links_to_weights(topo)
                ctx.set_topology(W, nl)
                want.append(ctx.besteffort_batch(reqs, mode).copy())
        finally:
            ctx.close()
        for call, (g, w) in enumerate(zip(got, want)):
            bad = np.nonzero((g["dev"] != w["dev"]).any(axis=1))[0]
            assert len(bad) == 0, "call %d mode %d: %d records differ, first %d: %r vs %r" % (
                call, mode, len(bad), bad[0], g["dev"][bad[0]].tolist(), w["dev"][bad[0]].tolist())


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch():
    W, nl = syn.links_to_weights(syn.topo_two_sockets(8))
    reqs = syn.random_request_batch(8, MAX_REQUESTS + 1, seed=1)
    rc, _, msg = noalloc(W, nl, reqs)
    assert rc == gpa.ERR_TOO_LARGE and "4096" in msg and "exceeds" in msg
    big = np.zeros((65, 65), dtype=np.int32)
    rc, _, msg = noalloc(big, np.zeros((65, 65), dtype=np.uint8), reqs[:4])
    assert rc == gpa.ERR_TOO_LARGE and "64" in msg
    rc, got, msg = noalloc(W, nl, reqs[:0])
    assert rc == gpa.OK and len(got) == 0
