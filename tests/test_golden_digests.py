"""Oracle and product against the committed digests of tests/golden/oracle_digests.json (written by
tests/golden/make_digests.py).  The digests pin the ORACLE to itself across rounds - they are not reference outputs
(no Go toolchain) - and the GPU test holds the CUDA path to the same bytes."""
import hashlib
import importlib.util
import json
import os

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import synthetic as syn

HERE = os.path.dirname(os.path.abspath(__file__))
DIGESTS = json.load(open(os.path.join(HERE, "golden", "oracle_digests.json")))
spec = importlib.util.spec_from_file_location("make_digests", os.path.join(HERE, "golden", "make_digests.py"))
make_digests = importlib.util.module_from_spec(spec)
spec.loader.exec_module(make_digests)


def test_oracle_still_gives_the_committed_answers(oracle):
    assert make_digests.besteffort_digests() == DIGESTS["besteffort"]
    assert make_digests.subset_digests() == DIGESTS["subset"]


@pytest.mark.gpu
def test_cuda_path_gives_the_committed_answers():
    from k8s_gpu_sharing_plugin_b200 import _lib as gpa
    ctx = gpa.Context(0)
    try:
        for ti, (name, t) in enumerate(syn.standard_topologies()):
            ctx.set_topology_links(t)
            ctx.set_device_order(make_digests.ORDER8)
            reqs = syn.pack_requests(list(syn.exhaustive_requests(make_digests.ORDER8, seed=ti, n_pairs=8)))
            want = DIGESTS["besteffort"][name]
            assert hashlib.sha256(reqs.tobytes()).hexdigest() == want["requests_sha256"]
            for mode, key in ((gpa.MODE_GO_FAITHFUL, "go_faithful"), (gpa.MODE_INTENDED, "intended")):
                got = ctx.besteffort_batch(reqs, mode)
                assert hashlib.sha256(np.ascontiguousarray(got["dev"]).tobytes()).hexdigest() == want[key], (name, key)
            W, nl = syn.links_to_weights(t)
            h = hashlib.sha256()
            for phys, req, size in make_digests.subset_requests(ti):
                rc, pos = ctx.subset_besteffort(phys, req, size)
                h.update(json.dumps([phys, req, size, rc, pos]).encode())
            assert h.hexdigest() == DIGESTS["subset"][name], name
    finally:
        ctx.close()
