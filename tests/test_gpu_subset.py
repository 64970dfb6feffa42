"""Parity of the CUDA replica-set search (gpa_subset_besteffort) with its literal oracle (oracle/ref_subset.c), through
the C ABI.  Bit-exact positions on every request.  There is no reference function for this search as a whole - it is
BASELINE.json north_star's "k-subsets of the available replica set", built from the reference's own combination order,
pair score and first-best rule (see include/gpualloc.h) - so the oracle is the definition and parity is UNPINNED."""
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
