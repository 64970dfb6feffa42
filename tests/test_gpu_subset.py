"""Parity of the CUDA replica-set search (gpa_subset_besteffort) with its literal oracle (oracle/ref_subset.c), through
the C ABI.  Bit-exact positions on every request.  There is no reference function for this search as a whole - it is
BASELINE.json north_star's "k-subsets of the available replica set", built from the reference's own combination order,
pair score and first-best rule (see include/gpualloc.h) - so the oracle is the definition and parity is UNPINNED."""
import os

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import synthetic as syn

pytestmark = pytest.mark.gpu

FORCE_ENUM = gpa.MODE_FORCE_ENUM


@pytest.fixture(scope="module")
def ctx():
    c = gpa.Context(0)
    yield c
    c.close()


def install(ctx, links2):
    ctx.set_topology_links(links2)
    return syn.links_to_weights(links2)


def test_known_answers_and_stats(ctx, oracle):
    t = syn.topo_uniform(8)
    t[0, 1, 1] = t[1, 0, 1] = syn.nvlink_type(1)   # one NVLink between GPUs 0 and 1
    W, nl = install(ctx, t)
    phys = [i // 4 for i in range(32)]
    want = oracle.subset_besteffort(W, nl, phys, [], 4, True)
    assert want[1] == [0, 1, 4, 5]                 # two replicas on each end of the one strong link
    for mode, algo, total in ((0, 4, 330), (FORCE_ENUM, 5, 35960)):
        before = ctx.launch_count
        assert ctx.subset_besteffort(phys, [], 4, mode) == (0, want[1])
        st = ctx.search_stats()
        assert ctx.launch_count == before + 1
        assert (st["algorithm"], st["partitions"]) == (algo, total)
        assert st["best_score"] * st["score_unit"] == want[2]
    assert ctx.subset_besteffort(phys, [30], 3) == oracle.subset_besteffort(W, nl, phys, [30], 3)
    # guards (besteffort_policy.go:35-45) and the no-search case answer without a launch
    before = ctx.launch_count
    for size, req in ((0, []), (-3, []), (33, []), (1, [0, 1])):
        assert ctx.subset_besteffort(phys, req, size) == (0, [])
    assert ctx.subset_besteffort(phys, [7, 2], 2) == (0, [2, 7])
    assert ctx.launch_count == before
    assert ctx.subset_besteffort(phys, [40], 2)[0] == gpa.ERR_INVALID
    assert ctx.subset_besteffort([0, 9], [], 1)[0] == gpa.ERR_INVALID
    assert ctx.subset_besteffort(list(range(8)) * 40, [], 9, FORCE_ENUM)[0] == gpa.ERR_TOO_LARGE


def test_asymmetric_links_are_the_reference_panic(ctx, oracle):
    W, nl = syn.links_to_weights(syn.topo_uniform(8))
    anl = nl.copy()
    anl[2, 5] = 2
    ctx.set_topology(W, anl)
    phys = [0, 0, 2, 2, 5]
    assert ctx.subset_besteffort(phys, [], 2)[0] == gpa.ERR_ASYMMETRIC
    assert oracle.subset_besteffort(W, anl, phys, [], 2)[0] == oracle.PANIC_ASYMMETRIC
    assert ctx.subset_besteffort(phys, [], 1) == (0, [0]) == oracle.subset_besteffort(W, anl, phys, [], 1)
    assert ctx.subset_besteffort([0, 0, 2, 2], [], 2) == oracle.subset_besteffort(W, anl, [0, 0, 2, 2], [], 2)


@pytest.mark.parametrize("topo_i", range(14))
def test_both_kernels_match_the_oracle_on_sorted_ids(ctx, oracle, topo_i):
    """config 1 topologies x replicas 1 / 2 / 4, positions in GPU order: count-vector kernel by default, literal kernel
    when forced; every answer against the oracle"""
    name, t = syn.standard_topologies()[topo_i]
    W, nl = install(ctx, t)
    rng = np.random.default_rng(500 + topo_i)
    algos = set()
    for replicas in (1, 2, 4):
        for _ in range(25):
            phys, req, size = syn.random_subset_request(rng, 8, replicas, p_free=float(rng.choice([0.3, 0.6, 0.9])), max_size=6)
            want = oracle.subset_besteffort(W, nl, phys, req, size)
            for mode in (0, FORCE_ENUM):
                got = ctx.subset_besteffort(phys, req, size, mode)
                if got[0] == gpa.ERR_TOO_LARGE:
                    assert mode == FORCE_ENUM
                    continue
                assert got == want, (name, phys, req, size, mode)
                algos.add(ctx.search_stats()["algorithm"])
    assert {4, 5} <= algos


def test_literal_kernel_on_any_order(ctx, oracle):
    W, nl = install(ctx, syn.topo_dgx1v_like())
    rng = np.random.default_rng(8)
    for _ in range(150):
        phys, req, size = syn.random_subset_request(rng, 8, 3, p_free=0.6, grouped=False, max_size=5)
        assert ctx.subset_besteffort(phys, req, size) == oracle.subset_besteffort(W, nl, phys, req, size), (phys, req, size)


@pytest.mark.parametrize("replicas,size,with_oracle", [(4, 4, True), (16, 3, True), (32, 3, True), (16, 4, False), (32, 4, False)])
def test_config_sizes_32_128_256_virtual_devices(ctx, oracle, replicas, size, with_oracle):
    """SURVEY.md 8d configs 2-3 (fragmented availability, p = 0.1 / 0.5 / 0.9): the count-vector kernel against the oracle
    where the oracle finishes in seconds, and against the literal kernel everywhere (C(256, 4) = 1.7e8 candidates)"""
    W, nl = install(ctx, syn.topo_random(8, 1000 + replicas))
    rng = np.random.default_rng(replicas * 10 + size)
    for p_free in (0.1, 0.5, 0.9):
        free = [g for g in range(8) for _ in range(replicas) if rng.random() < p_free]
        if len(free) < size + 1:
            continue
        for req in ([], [int(rng.integers(len(free)))]):
            rc, pos = ctx.subset_besteffort(free, req, size)
            st = ctx.search_stats()
            assert rc == 0 and st["algorithm"] == 4
            rc2, pos2 = ctx.subset_besteffort(free, req, size, FORCE_ENUM)
            st2 = ctx.search_stats()
            assert (rc2, pos2, st2["algorithm"]) == (0, pos, 5)
            assert st2["best_score"] == st["best_score"]
            if with_oracle:
                assert oracle.subset_besteffort(W, nl, free, req, size) == (0, pos)


def test_large_requests_only_the_count_vectors_can_answer(ctx):
    """16 / 32 of ~180 free replicas: properties (required kept, exact score replay, no improving swap)"""
    t = syn.topo_random(8, 1003)
    W, nl = install(ctx, t)
    rng = np.random.default_rng(13)
    free = [g for g in range(8) for _ in range(32) if rng.random() < 0.7]
    for size in (8, 16, 32):
        req = sorted(int(x) for x in rng.choice(len(free), size=3, replace=False))
        rc, pos = ctx.subset_besteffort(free, req, size)
        st = ctx.search_stats()
        assert (rc, len(pos), st["algorithm"]) == (0, size, 4) and set(req) <= set(pos) and pos == sorted(pos)
        score = sum(int(W[free[a], free[b]]) for i, a in enumerate(pos) for b in pos[i + 1:] if free[a] != free[b])
        assert score == st["best_score"] * st["score_unit"]
        chosen = set(pos)
        for a in [p for p in pos if p not in req][:4]:
            for b in [p for p in range(len(free)) if p not in chosen][::17]:
                alt = sorted((chosen - {a}) | {b})
                assert sum(int(W[free[x], free[y]]) for i, x in enumerate(alt) for y in alt[i + 1:] if free[x] != free[y]) <= score
    assert ctx.subset_besteffort(free, [], 70)[0] == gpa.ERR_TOO_LARGE
    # the full candidate space of a 16-GPU node with deep replicas: still one launch
    t16 = syn.topo_random(16, 77)
    W16, _ = install(ctx, t16)
    free16 = [g for g in range(16) for _ in range(8)]
    rc, pos = ctx.subset_besteffort(free16, [], 8)
    st = ctx.search_stats()
    assert rc == 0 and st["algorithm"] == 4 and st["partitions"] == 490314     # C(23, 15) count vectors
    score = sum(int(W16[free16[a], free16[b]]) for i, a in enumerate(pos) for b in pos[i + 1:] if free16[a] != free16[b])
    assert score == st["best_score"] * st["score_unit"]


def test_batch_in_count_form_matches_the_oracle(ctx, oracle):
    """gpa_subset_batch: fragmented availability at 8 GPUs x 4 replicas (config 2), every record against the literal oracle
    on its expansion to positions, and against the single-request call"""
    order = syn.sorted_positions(syn.node_uuids(8, seed=1))
    for ti in (1, 3, 6, 11):
        name, t = syn.standard_topologies()[ti]
        ctx.set_topology_links(t)
        ctx.set_device_order(order)
        W, nl = syn.links_to_weights(t)
        reqs = syn.random_subset_batch(8, 4, 400, seed=20 + ti, sizes=(1, 2, 3, 4, 5), p_must=0.3)
        reqs["size"][:4] = (0, -1, 40, 33)                       # guards
        got = ctx.subset_batch(reqs)
        unit = None
        rng = np.random.default_rng(ti)
        for i, rec in enumerate(reqs):
            phys_lab, req = syn.subset_record_to_positions(rec, rng)
            phys = [order[g] for g in phys_lab]                   # label -> NVML index
            rc, pos, score, _ = oracle.subset_besteffort(W, nl, phys, req, int(rec["size"]), True)
            assert rc == 0 and got["status"][i] == 0
            assert int(got["n_out"][i]) == len(pos), (name, i)
            if pos:
                assert got["take"][i].tolist() == syn.positions_to_take(phys_lab, set(req), pos), (name, i, rec)
                if unit is None and score > 0:
                    assert score % int(got["score"][i]) == 0
                    unit = score // int(got["score"][i])
                if unit is not None:
                    assert int(got["score"][i]) * unit == score
            else:
                assert not got["take"][i].any() and int(got["score"][i]) == -1
            if i < 60:
                assert ctx.subset_besteffort(phys, req, int(rec["size"])) == (0, pos)


def test_batch_statuses_and_larger_nodes(ctx, oracle):
    W, nl = syn.links_to_weights(syn.topo_uniform(8))
    anl = nl.copy()
    anl[2, 5] = 2
    ctx.set_topology(W, anl)
    reqs = np.zeros(6, dtype=syn.SUBSET_REQ_DTYPE)
    reqs["free"][0, [2, 5]] = 2; reqs["size"][0] = 2             # scores the asymmetric pair: the reference panics
    reqs["free"][1, [2, 5]] = 2; reqs["size"][1] = 1             # size 1 scores nothing
    reqs["free"][2, [0, 2]] = 2; reqs["size"][2] = 3
    reqs["free"][3, 9] = 1; reqs["size"][3] = 1                  # the node has 8 GPUs
    reqs["free"][4, :8] = 30; reqs["size"][4] = 17               # beyond the batch form
    reqs["must"][5, :8] = 20; reqs["size"][5] = 160              # beyond GPA_SUBSET_MAX_SIZE
    got = ctx.subset_batch(reqs)
    assert got["status"].tolist() == [gpa.ERR_ASYMMETRIC, 0, 0, gpa.ERR_INVALID, gpa.ERR_TOO_LARGE, gpa.ERR_TOO_LARGE]
    assert got["n_out"].tolist() == [0, 1, 3, 0, 0, 0]
    assert got["take"][1].tolist()[:8] == [0, 0, 1, 0, 0, 0, 0, 0] and got["take"][2].tolist()[:3] == [2, 0, 1]
    # a 16-GPU node, deep replicas, sizes up to 16: against the single-request kernel (same classes, one request a launch)
    t16 = syn.topo_random(16, 78)
    ctx.set_topology_links(t16)
    big = syn.random_subset_batch(16, 8, 64, seed=5, sizes=(2, 4, 8, 12), p_must=0.5)
    got = ctx.subset_batch(big)
    from math import comb
    n_large = 0
    for i, rec in enumerate(big):
        phys, req = syn.subset_record_to_positions(rec)
        K, g = int(rec["size"]) - len(req), int((rec["free"] > 0).sum())
        if len(phys) >= int(rec["size"]) and comb(K + g - 1, g - 1) > (1 << 20):     # the batch form's work limit
            assert got["status"][i] == gpa.ERR_TOO_LARGE
            n_large += 1
            continue
        rc, pos = ctx.subset_besteffort(phys, req, int(rec["size"]))
        assert rc == 0 and got["status"][i] == 0 and int(got["n_out"][i]) == len(pos)
        if pos:
            assert got["take"][i].tolist() == syn.positions_to_take(phys, set(req), pos), i
            assert int(got["score"][i]) == ctx.search_stats()["best_score"] or len(req) == int(rec["size"])
    assert 0 < n_large < len(big) // 2


def test_batch_two_halves_walk_equals_depth_first_walk_and_single_call(ctx):
    """nodes of up to 8 GPUs answer the count-form batch with the two-halves walk (K <= 8; the depth-first walk beyond):
    the same take[], score, n_out and status as the depth-first kernel (GPA_SUBSET_WALK=dfs) on every record, at 4 / 16 / 32
    replicas per GPU, sizes up to 15 and up to three must-include replicas; a sample against the single-request kernel"""
    os.environ["GPA_SUBSET_WALK"] = "dfs"
    try:
        dfs = gpa.Context(0)
    finally:
        del os.environ["GPA_SUBSET_WALK"]
    try:
        for n_gpus, topo in ((8, syn.topo_random(8, 31)), (8, syn.topo_uniform(8)), (5, syn.topo_random(5, 32)),
                             (8, syn.standard_topologies()[3][1])):
            order = syn.sorted_positions(syn.node_uuids(n_gpus, seed=n_gpus))
            for c in (ctx, dfs):
                c.set_topology_links(topo)
                c.set_device_order(order)
            for reps in (4, 16, 32):
                reqs = syn.random_subset_batch(n_gpus, reps, 6000, seed=900 + reps + n_gpus, sizes=(1, 2, 3, 4, 6, 8, 9, 12, 15), p_must=0.3)
                rng = np.random.default_rng(reps)
                extra = rng.integers(0, len(reqs), 600)          # second and third must-include replicas
                for i in extra:
                    g = int(rng.integers(0, n_gpus))
                    reqs["must"][i, g] += 1
                reqs["size"][:3] = (0, -2, 200)
                a, b = ctx.subset_batch(reqs), dfs.subset_batch(reqs)
                for field in ("status", "n_out", "score", "take"):
                    assert (a[field] == b[field]).all(), (n_gpus, reps, field, np.nonzero((a[field] != b[field]).reshape(len(reqs), -1).any(axis=1))[0][:5])
                assert (a["status"] == 0).sum() > 5000 and (reps < 16 or (a["n_out"] > 8).sum() > 300)
                for i in range(40):
                    phys, req = syn.subset_record_to_positions(reqs[i])
                    if a["status"][i] != 0:
                        continue
                    rc, pos = ctx.subset_besteffort([order[g] for g in phys], req, int(reqs["size"][i]))
                    assert rc == 0 and len(pos) == int(a["n_out"][i])
                    if pos:
                        assert a["take"][i].tolist() == syn.positions_to_take(phys, set(req), pos), (n_gpus, reps, i)
        # requests no free count binds (every GPU has at least K free, no must-include): the kernel hands out the answer it
        # worked out for the empty node when the topology was installed - the same as searching
        for n_gpus, topo in ((8, syn.topo_random(8, 33)), (6, syn.topo_random(6, 34)), (12, syn.topo_random(12, 35))):
            for c in (ctx, dfs):
                c.set_topology_links(topo)
                c.set_device_order(list(range(n_gpus))[::-1])
            reqs = np.zeros(45, dtype=syn.SUBSET_REQ_DTYPE)
            for i in range(45):
                K = i % 15 + 1
                reqs["size"][i] = K
                reqs["free"][i, :n_gpus] = K + i // 15 * 7       # exactly K, K + 7, K + 14 free on every GPU
            reqs["free"][44, 0] = 3                              # ... but not here: GPU 0 has 3 of the 15 wanted
            a, b = ctx.subset_batch(reqs), dfs.subset_batch(reqs)
            for field in ("status", "n_out", "score", "take"):
                assert (a[field] == b[field]).all(), (n_gpus, field)
            ok = a["status"] == 0                                # 12 GPUs: the larger requests are beyond the batch form
            assert ok.sum() >= 20 and (a["n_out"][ok] == reqs["size"][ok]).all() and (ok.all() or n_gpus > 8)
        # weights at the score table's limit (546 units a pair) and 124 must-include replicas on one GPU: one more replica
        # elsewhere adds more than 65 535 units against them, which takes the kernel's second pair of dot products
        rng = np.random.default_rng(5)
        W = np.zeros((8, 8), dtype=np.int32)
        iu = np.triu_indices(8, 1)
        W[iu] = rng.integers(530, 547, size=len(iu[0]))
        W[0, 1], W[0, 2] = 541, 546                              # no common unit to divide away
        W = W + W.T
        nl = np.ones((8, 8), dtype=np.uint8) - np.eye(8, dtype=np.uint8)
        for c in (ctx, dfs):
            c.set_topology(W, nl)
            c.set_device_order(list(range(8)))
        reqs = syn.random_subset_batch(8, 16, 3000, seed=77, sizes=(1, 2, 3, 4), p_must=0.0)
        for i in range(0, len(reqs), 2):
            reqs["must"][i, int(rng.integers(0, 8))] = 124
        reqs["size"] = reqs["size"] + reqs["must"].sum(axis=1)          # K = 1 .. 4 further replicas
        a, b = ctx.subset_batch(reqs), dfs.subset_batch(reqs)
        for field in ("status", "n_out", "score", "take"):
            assert (a[field] == b[field]).all(), field
        assert (a["status"] == 0).all() and int(a["score"].max()) > 124 * 541 * 2
    finally:
        dfs.close()


def test_batch_launches_in_flight_on_several_streams_and_in_a_graph(ctx):
    """every launch of the two-halves kernel draws its requests from a counter of its own; with more launches in flight than
    there are counters, or while a stream is being captured, the warps stride instead.  Same answers every time."""
    import torch
    ctx.set_topology_links(syn.topo_two_sockets(8))
    ctx.set_device_order(syn.sorted_positions(syn.node_uuids(8, seed=1)))
    reqs = syn.random_subset_batch(8, 16, 1 << 17, seed=41)
    want = ctx.subset_batch(reqs)
    d_in = torch.from_numpy(reqs.view(np.uint8).reshape(-1, 48)).cuda()
    streams = [torch.cuda.Stream() for _ in range(3)]
    outs = [torch.zeros((len(reqs), 24), dtype=torch.uint8, device="cuda") for _ in range(24)]
    torch.cuda.synchronize()
    for i, o in enumerate(outs):                                  # 24 launches, no synchronisation in between
        ctx.subset_batch_device(d_in.data_ptr(), len(reqs), o.data_ptr(), stream=streams[i % 3].cuda_stream)
    torch.cuda.synchronize()
    for o in outs:
        got = o.cpu().numpy().view(syn.SUBSET_RES_DTYPE).reshape(-1)
        for field in ("status", "n_out", "score", "take"):
            assert (got[field] == want[field]).all(), field
    # captured into a CUDA graph and replayed
    out_g = torch.zeros((len(reqs), 24), dtype=torch.uint8, device="cuda")
    s = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        ctx.subset_batch_device(d_in.data_ptr(), len(reqs), out_g.data_ptr(), stream=s.cuda_stream)
    for _ in range(2):
        out_g.zero_()
        g.replay()
        torch.cuda.synchronize()
        got = out_g.cpu().numpy().view(syn.SUBSET_RES_DTYPE).reshape(-1)
        for field in ("status", "n_out", "score", "take"):
            assert (got[field] == want[field]).all(), field
