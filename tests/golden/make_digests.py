#!/usr/bin/env python3
"""Regression digests of the ORACLE on the config-1 request sets: for each of the 14 eight-GPU topologies,
the SHA-256 of the packed oracle answers to the exhaustive request list (every mask x size x required), in both modes,
plus the answers of the replica-set oracle on a seeded list.

These are NOT reference outputs - there is no Go toolchain to produce any, the best-effort oracle is "parity unpinned"
 - they pin the oracle to ITSELF across rounds: a change that moves the oracle and the kernels together in
the same wrong direction shows up here.  Needs no reference tree:

    python tests/golden/make_digests.py          # rewrites tests/golden/oracle_digests.json
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import synthetic as syn  # noqa: E402
from oracle import oracle as orc  # noqa: E402

ORDER8 = syn.sorted_positions(syn.node_uuids(8, seed=1))


def besteffort_digests():
    out = {}
    for ti, (name, t) in enumerate(syn.standard_topologies()):
        W, nl = syn.links_to_weights(t)
        reqs = syn.pack_requests(list(syn.exhaustive_requests(ORDER8, seed=ti, n_pairs=8)))
        row = {"requests": int(len(reqs)), "requests_sha256": hashlib.sha256(reqs.tobytes()).hexdigest()}
        for mode, key in ((orc.GO_FAITHFUL, "go_faithful"), (orc.INTENDED, "intended")):
            res, _ = orc.besteffort_batch(W, nl, reqs, mode=mode, threads=os.cpu_count() or 1)
            row[key] = hashlib.sha256(np.ascontiguousarray(res["dev"]).tobytes()).hexdigest()
        out[name] = row
    return out


def subset_requests(ti):
    rng = np.random.default_rng(9000 + ti)
    reqs = []
    for _ in range(60):
        reqs.append(syn.random_subset_request(rng, 8, int(rng.choice([1, 2, 4])), p_free=float(rng.choice([0.3, 0.6, 0.9])),
                                              grouped=bool(rng.random() < 0.7), max_size=5))
    return reqs


def subset_digests():
    out = {}
    for ti, (name, t) in enumerate(syn.standard_topologies()):
        W, nl = syn.links_to_weights(t)
        h = hashlib.sha256()
        for phys, req, size in subset_requests(ti):
            rc, pos = orc.subset_besteffort(W, nl, phys, req, size)
            h.update(json.dumps([phys, req, size, rc, pos]).encode())
        out[name] = h.hexdigest()
    return out


def main():
    orc.build()
    doc = {"what": "SHA-256 digests of oracle answers (self-pinning, not reference outputs); see make_digests.py",
           "besteffort": besteffort_digests(), "subset": subset_digests()}
    with open(os.path.join(HERE, "oracle_digests.json"), "w") as f:
        json.dump(doc, f, indent=1, sort_keys=True)
    print("wrote oracle_digests.json:", len(doc["besteffort"]), "topologies")


if __name__ == "__main__":
    main()
