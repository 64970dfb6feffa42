#!/usr/bin/env python3
"""Small run of every kernel for compute-sanitizer (one tool per gpurun call):
   compute-sanitizer --tool memcheck python tools/sanitize_smoke.py"""
import os
import sys


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import _lib as gpa  # noqa: E402
from k8s_gpu_sharing_plugin_b200 import synthetic as syn  # noqa: E402
from oracle import oracle as orc  # noqa: E402

ctx = gpa.Context(0)
order = syn.sorted_positions(syn.node_uuids(8, seed=1))
t = syn.topo_dgx1v_like()
W, nl = syn.links_to_weights(t)
ctx.set_topology_links(t)
ctx.set_device_order(order)
reqs = syn.random_request_batch(8, 20000, seed=3, order=order)
want, _ = orc.besteffort_batch(W, nl, reqs)
assert (ctx.besteffort_batch(reqs)["dev"] == want["dev"]).all()                       # fast kernel
ctx.set_device_order(list(range(8)))
assert (ctx.besteffort_batch(reqs)["dev"] == want["dev"]).all()                       # general kernel (deferred pass)
ctx.set_device_order(order)
m = syn.records_to_masks(reqs, order)
assert (syn.mask_results_to_records(ctx.besteffort_batch_masks(m), order)["dev"] == want["dev"]).all()  # mask format
m8 = syn.masks_to_masks8(m)
assert (syn.mask_results_to_records(syn.masks8_results_to_masks(ctx.besteffort_batch_masks8(m8)), order)["dev"] == want["dev"]).all()
assert (ctx.besteffort_batch(reqs, gpa.MODE_FORCE_ENUM)["dev"] == want["dev"]).all()  # tile-binned kernel + deferred pass
mixed = reqs.copy()
mixed[::3] = syn.random_request_batch(8, 20000, seed=4)[::3]                           # identity order: foreign to this context
want_mixed, _ = orc.besteffort_batch(W, nl, mixed)
assert (ctx.besteffort_batch(mixed)["dev"] == want_mixed["dev"]).all()                 # table kernel answering its deferred records
for n, k in ((8, 3), (12, 3), (13, 2), (14, 5), (16, 8)):
    tt = syn.topo_random(n, 5)
    Wn, nln = syn.links_to_weights(tt)
    ctx.set_topology_links(tt)
    assert ctx.besteffort(list(range(n)), [], k) == orc.besteffort(Wn, nln, list(range(n)), [], k)  # search kernel
print("hbm", gpa.probe_hbm(0, 64 << 20, 2, 1), "p2p", gpa.probe_p2p(1, 64 << 20, 2, 0)[0])
ctx.close()
print("sanitize smoke ok")
