"""Self-consistency of the best-effort oracle (oracle/ref_besteffort.c).

The reference holds no test for this policy, so these tests cannot pin the
oracle to reference outputs; they check the literal slice-emulating restatement against
closed forms, against hand-derived vectors, and against the allocation-free variant that
is used for larger sizes.
"""
import math

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import synthetic as syn


def closed_form(n, k):
    if k == 1:
        return n
    N = -(-n // k) * k
    P, m = N - n, N // k
    f = math.factorial
    if P == 0:
        return f(n) // (f(k) ** m * f(m))
    return math.comb(n, k - P) * f((m - 1) * k) // (f(k) ** (m - 1) * f(m - 1))


@pytest.mark.parametrize("n", range(1, 11))
def test_partition_counts_match_closed_form(oracle, n):
    W, nl = syn.links_to_weights(syn.topo_uniform(n))
    for k in range(1, n + 1):
        assert oracle.partition_count(n, k) == closed_form(n, k)
        for variant in ("literal", "fast"):
            rc, out, visited = oracle.besteffort(W, nl, list(range(n)), [], k, variant=variant, want_visited=True)
            assert rc == oracle.OK and visited == closed_form(n, k), (n, k, variant)
            assert out == list(range(k))  # uniform weights: the first partition wins, its first set too


def test_known_partition_counts_n8_n16(oracle):
    assert [oracle.partition_count(8, k) for k in range(1, 9)] == [8, 105, 280, 35, 56, 28, 8, 1]
    assert [oracle.partition_count(16, k) for k in range(2, 9)] == [2027025, 22422400, 2627625, 2018016, 840840, 205920, 6435]


def test_nil_pad_vector_D5(oracle):
    # 5 GPUs, k=3, W01=150, W23=W24=W34=50, rest 10 -> [0, 1, nil]
    t = syn.topo_uniform(5)
    for i, j in ((2, 3), (2, 4), (3, 4)):
        t[i, j, 0] = t[j, i, 0] = 5
    t[0, 1] = t[1, 0] = (5, syn.nvlink_type(1))
    W, nl = syn.links_to_weights(t)
    assert W[0, 1] == 150 and W[2, 3] == 50 and W[0, 2] == 10
    for variant in ("literal", "fast"):
        for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
            assert oracle.besteffort(W, nl, [0, 1, 2, 3, 4], [], 3, mode, variant) == (oracle.OK, [0, 1, oracle.NIL])


def test_guards_and_trivial_answers(oracle):
    W, nl = syn.links_to_weights(syn.topo_zero(8))
    av = [3, 1, 4, 0, 5]
    assert oracle.besteffort(W, nl, av, [], 0)[1] == []            # size <= 0
    assert oracle.besteffort(W, nl, av, [], 6)[1] == []            # len(available) < size
    assert oracle.besteffort(W, nl, av, [3, 1, 4], 2)[1] == []     # len(required) > size
    for k in range(1, 6):
        assert oracle.besteffort(W, nl, av, [], k, variant="literal")[1] == av[:k]  # all-zero W: first k
    assert oracle.besteffort(W, nl, av, [], 1)[1] == [3]           # size 1 -> available[0]
    assert oracle.besteffort(W, nl, av, [0], 1)[1] == [0]          # ... or the required device
    assert oracle.besteffort(W, nl, [], [], 1)[1] == []


def test_foreign_required_pointer_yields_empty_D4(oracle):
    W, nl = syn.links_to_weights(syn.topo_dgx1v_like())
    for variant in ("literal", "fast"):
        assert oracle.besteffort(W, nl, list(range(8)), [oracle.FOREIGN], 2, variant=variant) == (oracle.OK, [])


def test_asymmetric_link_count_panics_only_when_scored(oracle):
    W, nl = syn.links_to_weights(syn.topo_uniform(4))
    nl = nl.copy()
    nl[0, 1] = 2
    for variant in ("literal", "fast"):
        assert oracle.besteffort(W, nl, [0, 1, 2, 3], [], 2, variant=variant)[0] == oracle.PANIC_ASYMMETRIC
        assert oracle.besteffort(W, nl, [0, 1, 2, 3], [], 1, variant=variant)[0] == oracle.OK  # no pair is scored
        assert oracle.besteffort(W, nl, [2, 3], [], 2, variant=variant)[0] == oracle.OK


def test_asymmetric_weights_use_the_earlier_devices_view(oracle):
    # calculateGPUPairScore sums gpu0.Links[gpu1.Index] with gpu0 = earlier in `available`
    W = np.zeros((3, 3), dtype=np.int32)
    nl = np.ones((3, 3), dtype=np.uint8)
    W[0, 1], W[1, 0] = 10, 500
    W[0, 2], W[2, 0] = 100, 100
    assert oracle.besteffort(W, nl, [0, 1, 2], [], 2, variant="literal")[1] == [0, 2]
    assert oracle.besteffort(W, nl, [1, 0, 2], [], 2, variant="literal")[1] == [1, 0]


@pytest.mark.parametrize("topo_i", [2, 3, 4, 7])
def test_literal_equals_fast_exhaustive_n8(oracle, topo_i):
    name, t = syn.standard_topologies()[topo_i]
    W, nl = syn.links_to_weights(t)
    order = syn.sorted_positions(syn.node_uuids(8, seed=1))
    n = 0
    for avail, req, k in syn.exhaustive_requests(order, seed=topo_i, n_pairs=2):
        if len(avail) < 5 and topo_i != 3:
            continue
        a = oracle.besteffort(W, nl, avail, req, k, oracle.GO_FAITHFUL, "literal")
        b = oracle.besteffort(W, nl, avail, req, k, oracle.GO_FAITHFUL, "fast")
        assert a == b, (name, avail, req, k)
        n += 1
    assert n > 1000


@pytest.mark.parametrize("n,k", [(9, 2), (9, 4), (10, 3), (11, 2), (11, 5), (12, 3), (12, 5), (13, 2), (14, 2), (14, 4)])
def test_literal_equals_fast_seeded_large(oracle, n, k):
    for seed in range(3):
        W, nl = syn.links_to_weights(syn.topo_random(n, 50 + seed))
        rng = np.random.default_rng(seed)
        avail = [int(x) for x in rng.permutation(n)]
        for req in ([], [avail[int(rng.integers(n))]]):
            for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
                a = oracle.besteffort(W, nl, avail, req, k, mode, "literal", want_visited=True)
                b = oracle.besteffort(W, nl, avail, req, k, mode, "fast", want_visited=True)
                assert a == b and a[0] == oracle.OK and len(a[1]) == k, (n, k, seed, req, mode)


@pytest.mark.parametrize("n", [13, 14, 15, 16])
def test_slice_aliasing_D6_changes_the_answer(oracle, n):
    """m = 7 or 8 sets: Go appends the last sets of a candidate in place, so the saved
    bestPartition is overwritten by later siblings.  `intended` always
    finds the spiked pair; `go_faithful` usually does not.  The literal slice emulation and
    the direct emulation in the fast variant must agree on what Go returns."""
    differs = 0
    seeds = range(4) if n <= 14 else range(1)
    for seed in seeds:
        W, nl = syn.links_to_weights(syn.topo_spiked(n, seed))
        avail = list(range(n))
        lit_f = oracle.besteffort(W, nl, avail, [], 2, oracle.GO_FAITHFUL, "literal")
        fast_f = oracle.besteffort(W, nl, avail, [], 2, oracle.GO_FAITHFUL, "fast")
        lit_i = oracle.besteffort(W, nl, avail, [], 2, oracle.INTENDED, "literal")
        fast_i = oracle.besteffort(W, nl, avail, [], 2, oracle.INTENDED, "fast")
        assert lit_f == fast_f and lit_i == fast_i
        assert lit_i == (oracle.OK, [n - 2, n - 1])
        differs += lit_f != lit_i
    assert differs >= 1


def test_alias_free_set_counts_agree_between_modes(oracle):
    # m in {1..6, 9, 10}: no candidate slot is both shared and rewritten
    for n, k in [(8, 2), (8, 3), (10, 2), (12, 2), (12, 3), (12, 4), (10, 5)]:
        for seed in range(3):
            W, nl = syn.links_to_weights(syn.topo_random(n, 900 + seed))
            a = oracle.besteffort(W, nl, list(range(n)), [], k, oracle.GO_FAITHFUL, "fast")
            b = oracle.besteffort(W, nl, list(range(n)), [], k, oracle.INTENDED, "fast")
            assert a == b


def test_go_growslice_capacity_sequence(oracle):
    """The D6 analysis rests on how append() grows a [][]*Device (24-byte slice headers) that starts empty: cap 1, 2, 4,
    8, 16, 32 (Go 1.17 runtime.growslice doubles below 1024 elements and rounds the byte size up to a malloc size class;
    24, 48, 96, 192, 384 and 768 bytes are all classes, so nothing is added).  Two published examples of the rounding
    pin the size-class table itself: append to an empty []int64 of 5 elements gives cap 6 (40 -> 48 bytes), and 3
    24-byte elements give cap 3 (72 -> 80 bytes)."""
    cap, seq = 0, []
    for length in range(1, 33):
        if length > cap:
            cap = oracle.go_growcap(cap, length, 24)
            seq.append(cap)
    assert seq == [1, 2, 4, 8, 16, 32]
    cap, seq = 0, []
    for length in range(1, 17):
        if length > cap:
            cap = oracle.go_growcap(cap, length, 8)
            seq.append(cap)
    assert seq == [1, 2, 4, 8, 16]
    assert oracle.go_growcap(0, 5, 8) == 6
    assert oracle.go_growcap(0, 3, 24) == 3


HAND_DERIVED_D6 = syn.HAND_DERIVED_D6
spiked_uniform = syn.spiked_uniform


@pytest.mark.parametrize("n,pair,faithful,intended", HAND_DERIVED_D6)
def test_hand_derived_slice_aliasing_vectors(oracle, n, pair, faithful, intended):
    W, nl = spiked_uniform(n, pair)
    for variant in ("literal", "fast"):
        assert oracle.besteffort(W, nl, list(range(n)), [], 2, oracle.GO_FAITHFUL, variant) == (oracle.OK, faithful)
        assert oracle.besteffort(W, nl, list(range(n)), [], 2, oracle.INTENDED, variant) == (oracle.OK, intended)
