"""The allocation-free batch entry point, gpa_besteffort_batch_noalloc, against the regular one.

Some users of this library cannot allocate when a request arrives: the plugin runs on GPUs whose
memory belongs to the pods it shares them with, so device memory may be full, and cudaMalloc /
cudaFree would stall every tenant of the GPU.  For them the library keeps its device-side staging
in a fixed array inside the module.  This test calls that entry point through the C ABI twice in
a row — the second call with another topology and a smaller batch, so anything left in the shared
array by the first call would show — and requires byte-identical results to gpa_besteffort_batch
AND to the oracle (integer work: no tolerance); a second test makes the very first call of a fresh
process on a non-blocking stream, where the entry point's own set-up has to order itself with the
caller's stream.  It lives in a file that pytest collects last; it must pass both in
a fresh process (`pytest tests/test_zz_no_alloc.py -m gpu`) and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import synthetic as syn

pytestmark = pytest.mark.gpu

MAX_REQUESTS = 4096  # GPA_NOALLOC_MAX_REQUESTS


def noalloc(W, nlinks, reqs, max_avail=0, mode=gpa.MODE_GO_FAITHFUL, device=0, stream=None):
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
           reqs.ctypes.data, len(reqs), max_avail, mode, out.ctypes.data, C.c_void_p(stream or 0), err, len(err))
    return rc, out, err.value.decode()


def test_noalloc_matches_regular_entry_point_twice_in_a_row(oracle):
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
        # and what the oracle answers: the two CUDA paths are not only equal to each other
        for call, ((topo, reqs), g) in enumerate(zip(calls, got)):
            W, nl = syn.links_to_weights(topo)
            ref, _ = oracle.besteffort_batch(W, nl, reqs, mode=mode, threads=8)
            assert (g["dev"] == ref["dev"]).all(), "call %d mode %d differs from the oracle" % (call, mode)


FRESH_PROCESS = """
import sys
sys.path.insert(0, %r)
import __graft_entry__ as entry
entry.load_package()
import importlib.util, os
import numpy as np, torch
spec = importlib.util.spec_from_file_location("noalloc_helpers", os.path.join(sys.path[0], "tests", "test_zz_no_alloc.py"))
t = importlib.util.module_from_spec(spec)
spec.loader.exec_module(t)
from k8s_gpu_sharing_plugin_b200 import synthetic as syn
from oracle import oracle as orc
orc.build()
torch.cuda.init()
st = torch.cuda.Stream()            # torch streams are cudaStreamNonBlocking
topo = syn.topo_random(10, 9)
W, nl = syn.links_to_weights(topo)
reqs = syn.random_request_batch(10, 3000, seed=77)
rc, got, msg = t.noalloc(W, nl, reqs, stream=st.cuda_stream)   # the first CUDA work of this process besides the stream itself
assert rc == 0, msg
want, _ = orc.besteffort_batch(W, nl, reqs, threads=4)
assert (got["dev"] == want["dev"]).all()
print("fresh-process first call on a non-blocking stream: ok")
"""


def test_noalloc_first_call_of_a_process_on_a_non_blocking_stream():
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", FRESH_PROCESS % root], capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


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
