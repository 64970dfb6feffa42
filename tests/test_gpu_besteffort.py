"""Parity of the CUDA best-effort path with the oracle, through the C ABI (ctypes).

Bit-exact on every request: integer/index work, no tolerance.  Nothing here reads
/root/reference; inputs are seeded synthetic topologies."""
import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa
from k8s_gpu_sharing_plugin_b200 import synthetic as syn

pytestmark = pytest.mark.gpu

ORDER8 = syn.sorted_positions(syn.node_uuids(8, seed=1))


@pytest.fixture(scope="module")
def ctx():
    c = gpa.Context(0)
    yield c
    c.close()


def install(ctx, links2, order=None):
    """installs a topology; order = the node-wide device order (None = NVML index order)"""
    W, nl = syn.links_to_weights(links2)
    ctx.set_topology_links(links2)
    ctx.set_device_order(list(range(links2.shape[0])) if order is None else order)
    W2, nl2 = ctx.get_topology()
    assert (W2 == W).all() and (nl2 == nl).all()  # the ABI's score switch == the numpy statement of it
    return W, nl


def oracle_batch(oracle, W, nl, reqs, mode):
    out, _ = oracle.besteffort_batch(W, nl, reqs, mode=mode, threads=8)
    return out


def assert_batch_equal(got, want, reqs, what):
    bad = np.nonzero((got["dev"] != want["dev"]).any(axis=1))[0]
    if len(bad):
        i = int(bad[0])
        raise AssertionError("%s: %d mismatches; first at %d: request %r got %r want %r" % (
            what, len(bad), i, syn.unpack_request(reqs[i]), got["dev"][i].tolist(), want["dev"][i].tolist()))


@pytest.mark.parametrize("canonical", [True, False], ids=["canonical-order", "foreign-order"])
@pytest.mark.parametrize("topo_i", range(14))
def test_batch_exhaustive_n8_all_masks_all_sizes(ctx, oracle, topo_i, canonical):
    """config 1: every non-empty mask x k = 1..|mask| x required in {none, each single, seeded pairs}.
    canonical: the context knows the node's sorted-UUID order, so every request takes the fast kernel;
    otherwise the requests' order is foreign to the context and the general kernel answers them."""
    name, t = syn.standard_topologies()[topo_i]
    W, nl = install(ctx, t, ORDER8 if canonical else None)
    reqs = syn.pack_requests(list(syn.exhaustive_requests(ORDER8, seed=topo_i, n_pairs=8)))
    assert len(reqs) > 9000
    for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
        got = ctx.besteffort_batch(reqs, mode)
        assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, mode), reqs, name)


def test_single_request_api_matches_batch_and_oracle_n8(ctx, oracle):
    name, t = syn.standard_topologies()[3]
    W, nl = install(ctx, t)
    rng = np.random.default_rng(5)
    all_reqs = list(syn.exhaustive_requests(ORDER8, seed=3, n_pairs=2))
    for i in rng.choice(len(all_reqs), 400, replace=False):
        avail, req, k = all_reqs[int(i)]
        want = oracle.besteffort(W, nl, avail, req, k)
        assert ctx.besteffort(avail, req, k) == (gpa.OK, want[1]), (avail, req, k)
        assert ctx.besteffort(avail, req, k, gpa.MODE_FORCE_DP) == (gpa.OK, want[1]), ("dp", avail, req, k)


@pytest.mark.parametrize("topo_i", [0, 3, 8])
def test_single_request_on_a_small_node_takes_the_one_warp_kernel(ctx, oracle, topo_i):
    """8-GPU node, devices listed in node order, GPA_MODE_FORCE_KERNEL (without it the host copy of the answer table
    answers, tests/test_gpu_table.py): the one-warp partition-table kernel answers (algorithm 3), both modes, with and
    without required devices, every mask and size"""
    name, t = syn.standard_topologies()[topo_i]
    W, nl = install(ctx, t, ORDER8)
    n = 0
    for avail, req, k in syn.exhaustive_requests(ORDER8, seed=topo_i, n_pairs=2):
        if k < 2 or len(req) > k or (n % 5 and len(avail) < 7):
            n += 1
            continue
        for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
            want = oracle.besteffort(W, nl, avail, req, k, mode)
            assert ctx.besteffort(avail, req, k, mode | gpa.MODE_FORCE_KERNEL) == (gpa.OK, want[1]), (name, avail, req, k, mode)
            assert ctx.search_stats()["algorithm"] == 3
        n += 1
    # the same node, devices listed in another order: falls back to the general search, same answer
    avail = [int(x) for x in reversed(ORDER8)]
    assert ctx.besteffort(avail, [], 3) == (gpa.OK, oracle.besteffort(W, nl, avail, [], 3)[1])
    assert ctx.search_stats()["algorithm"] == 1
    # D5 on this path: a nil-padded winner
    t5 = syn.topo_uniform(5)
    for i, j in ((2, 3), (2, 4), (3, 4)):
        t5[i, j, 0] = t5[j, i, 0] = 5
    t5[0, 1] = t5[1, 0] = (5, syn.nvlink_type(1))
    install(ctx, t5)
    assert ctx.besteffort([0, 1, 2, 3, 4], [], 3, gpa.MODE_FORCE_KERNEL) == (gpa.OK, [0, 1, gpa.NIL]) and ctx.search_stats()["algorithm"] == 3
    assert ctx.besteffort([0, 1, 2, 3, 4], [], 3) == (gpa.OK, [0, 1, gpa.NIL]) and ctx.search_stats()["algorithm"] == 6


def test_guards_and_trivial_answers(ctx, oracle):
    W, nl = install(ctx, syn.topo_zero(8))
    av = [3, 1, 4, 0, 5]
    assert ctx.besteffort(av, [], 0) == (gpa.OK, [])           # size <= 0
    assert ctx.besteffort(av, [], -3) == (gpa.OK, [])
    assert ctx.besteffort(av, [], 6) == (gpa.OK, [])           # len(available) < size
    assert ctx.besteffort(av, [3, 1, 4], 2) == (gpa.OK, [])    # len(required) > size
    for k in range(1, 6):
        assert ctx.besteffort(av, [], k) == (gpa.OK, av[:k])   # all-zero W: first k devices
    assert ctx.besteffort(av, [0], 1) == (gpa.OK, [0])         # size 1: the required device
    assert ctx.besteffort([], [], 1) == (gpa.OK, [])
    assert ctx.besteffort(av, [gpa.FOREIGN], 2) == (gpa.OK, [])  # D4: a pointer that is in no set
    assert ctx.besteffort(av, [7], 2) == (gpa.OK, [])          # required but not available
    assert ctx.besteffort([0, 0, 1], [], 2)[0] == gpa.ERR_INVALID
    assert ctx.besteffort([0, 9], [], 2)[0] == gpa.ERR_INVALID
    # the same through the packed batch format
    reqs = syn.pack_requests([(av, [], 0), (av, [], 6), (av, [3, 1, 4], 2), (av, [], 3), (av, [0], 1), (av, [-2], 2),
                              (av, [7], 2), ([2], [], 1), (av, [5, 4], 2)])
    got = ctx.besteffort_batch(reqs)
    want = oracle_batch(oracle, W, nl, reqs, 0)
    assert_batch_equal(got, want, reqs, "guards")
    assert gpa.decode_result(got[3]) == (gpa.OK, [3, 1, 4]) and gpa.decode_result(got[8]) == (gpa.OK, [4, 5])


def test_nil_pad_vector_D5(ctx, oracle):
    t = syn.topo_uniform(5)
    for i, j in ((2, 3), (2, 4), (3, 4)):
        t[i, j, 0] = t[j, i, 0] = 5
    t[0, 1] = t[1, 0] = (5, syn.nvlink_type(1))
    W, nl = install(ctx, t)
    assert ctx.besteffort([0, 1, 2, 3, 4], [], 3) == (gpa.OK, [0, 1, gpa.NIL])
    got = ctx.besteffort_batch(syn.pack_requests([([0, 1, 2, 3, 4], [], 3)]))
    assert gpa.decode_result(got[0]) == (gpa.OK, [0, 1, gpa.NIL])


def test_asymmetric_link_counts_are_reported_where_go_panics(ctx, oracle):
    W, nl = syn.links_to_weights(syn.topo_uniform(4))
    nl = nl.copy()
    nl[0, 1] = 2
    ctx.set_topology(W, nl)
    assert ctx.besteffort([0, 1, 2, 3], [], 2)[0] == gpa.ERR_ASYMMETRIC
    assert ctx.besteffort([0, 1, 2, 3], [], 1) == (gpa.OK, [0])   # size 1 scores no pair
    assert ctx.besteffort([2, 3], [], 2) == (gpa.OK, [2, 3])
    reqs = syn.pack_requests([([0, 1, 2, 3], [], 2), ([0, 1, 2, 3], [], 1), ([2, 3], [], 2)])
    got = ctx.besteffort_batch(reqs)
    assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, 0), reqs, "asym")
    assert gpa.decode_result(got[0])[0] == gpa.ERR_ASYMMETRIC


def test_asymmetric_pair_panics_only_where_go_scores_it(ctx, oracle):
    """calculateGPUPairScore panics lazily (besteffort_policy.go:307-310): only for a pair inside a candidate that passed
    the `required` filter (:53-56), or inside a set the final loop scores (:77-85).  Hand-derived: 4 devices, size 2,
    link counts of (0, 2) differ.  required = {0, 1}: the only accepted partition is (0,1)(2,3), the pair (0, 2) is never
    scored, Go answers [0, 1].  required = {0}: (0,2)(1,3) is accepted and scored -> panic.  required = {} -> panic.
    Then seeded topologies with several such pairs, every mask / size / required, single and batched, against the
    oracle (which restates the lazy rule, ref_besteffort.c fast_leaf / final loop)."""
    W, nl = syn.links_to_weights(syn.topo_uniform(4))
    nl = nl.copy()
    nl[0, 2] = 2
    ctx.set_topology(W, nl)
    assert ctx.besteffort([0, 1, 2, 3], [0, 1], 2) == (gpa.OK, [0, 1])
    assert ctx.besteffort([0, 1, 2, 3], [2, 3], 2) == (gpa.OK, [2, 3])
    assert ctx.besteffort([0, 1, 2, 3], [0], 2)[0] == gpa.ERR_ASYMMETRIC
    assert ctx.besteffort([0, 1, 2, 3], [], 2)[0] == gpa.ERR_ASYMMETRIC
    assert ctx.besteffort([0, 1, 2, 3], [], 4)[0] == gpa.ERR_ASYMMETRIC      # one set: every pair is scored
    assert ctx.besteffort([0, 1, 3], [], 2) == (gpa.OK, [0, 1])             # the pair is not even available
    got = ctx.besteffort_batch(syn.pack_requests([([0, 1, 2, 3], [0, 1], 2), ([0, 1, 2, 3], [0], 2)]))
    assert gpa.decode_result(got[0]) == (gpa.OK, [0, 1]) and gpa.decode_result(got[1])[0] == gpa.ERR_ASYMMETRIC
    rng = np.random.default_rng(17)
    for n, seed in ((6, 1), (7, 2), (8, 3)):
        W, nl = syn.links_to_weights(syn.topo_random(n, 60 + seed))
        nl = nl.copy()
        for _ in range(2):
            a, b = (int(x) for x in rng.choice(n, 2, replace=False))
            nl[a, b] += 1
        ctx.set_topology(W, nl)
        assert ctx.answer_table() is None
        order = [int(x) for x in rng.permutation(n)]
        all_reqs = list(syn.exhaustive_requests(order, seed=seed, n_pairs=4))
        reqs = syn.pack_requests(all_reqs)
        n_ok = n_panic = 0
        for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
            want = oracle_batch(oracle, W, nl, reqs, mode)
            assert_batch_equal(ctx.besteffort_batch(reqs, mode), want, reqs, "lazy-asym n=%d" % n)
            n_panic += int((want["dev"][:, 0] == gpa.RES_ERROR).sum())
            n_ok += int((want["dev"][:, 0] != gpa.RES_ERROR).sum())
            for i in rng.choice(len(all_reqs), 150, replace=False):
                avail, req, k = all_reqs[int(i)]
                rc_o, out_o = oracle.besteffort(W, nl, avail, req, k, mode)
                rc, out = ctx.besteffort(avail, req, k, mode)
                assert (rc == gpa.ERR_ASYMMETRIC) == (rc_o == oracle.PANIC_ASYMMETRIC), (n, avail, req, k, mode)
                assert rc != gpa.OK or out == out_o, (n, avail, req, k, mode)
        assert n_ok > 100 and n_panic > 100   # both outcomes are well represented


def test_asymmetric_weights_use_the_earlier_devices_view(ctx, oracle):
    W = np.zeros((3, 3), dtype=np.int32)
    nl = np.ones((3, 3), dtype=np.uint8)
    W[0, 1], W[1, 0] = 10, 500
    W[0, 2], W[2, 0] = 100, 100
    ctx.set_topology(W, nl)
    assert ctx.besteffort([0, 1, 2], [], 2) == (gpa.OK, [0, 2])
    assert ctx.besteffort([1, 0, 2], [], 2) == (gpa.OK, [1, 0])


def test_weight_range_is_checked(ctx):
    W = np.zeros((4, 4), dtype=np.int32)
    nl = np.ones((4, 4), dtype=np.uint8)
    W[0, 1] = W[1, 0] = 1 << 20
    W[2, 3] = W[3, 2] = 7
    with pytest.raises(gpa.GpaError) as e:
        ctx.set_topology(W, nl)
    assert e.value.code == gpa.ERR_WEIGHT_RANGE
    W[0, 1] = W[1, 0] = -5
    with pytest.raises(gpa.GpaError):
        ctx.set_topology(W, nl)


@pytest.mark.parametrize("n,k", [(9, 2), (9, 4), (10, 3), (10, 4), (11, 2), (11, 3), (11, 5), (12, 3), (12, 5), (12, 7),
                                 (12, 11), (13, 2), (13, 3), (13, 6), (14, 2), (14, 4), (14, 5), (15, 2), (15, 4),
                                 (16, 2), (16, 4), (16, 5), (16, 6), (16, 7), (16, 8), (16, 9), (16, 13), (16, 16)])
def test_single_search_seeded_large(ctx, oracle, n, k):
    """config 3b: synthetic n = 9..16 physical devices, both alias modes, with and without required devices"""
    for seed in range(2):
        W, nl = install(ctx, syn.topo_random(n, 300 + seed))
        rng = np.random.default_rng(seed)
        avail = [int(x) for x in rng.permutation(n)]
        reqs = [[], [avail[int(rng.integers(n))]]]
        if k >= 3:
            reqs.append([avail[int(x)] for x in rng.choice(n, 2, replace=False)])
        for req in reqs:
            for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
                want = oracle.besteffort(W, nl, avail, req, k, mode, "fast")
                # both search algorithms: every partition visited / subset dynamic programme
                for algo, code in ((gpa.MODE_FORCE_ENUM, 1), (gpa.MODE_FORCE_DP, 2)):
                    got = ctx.besteffort(avail, req, k, mode | algo)
                    st = ctx.search_stats()
                    assert got == (gpa.OK, want[1]), (n, k, seed, req, mode, st)
                    assert st["partitions"] == oracle.partition_count(n, k) and st["algorithm"] == code


def test_full_size_n16_k3_matches_oracle(ctx, oracle):
    """the largest search of BASELINE config 3b: 22 422 400 candidate partitions"""
    W, nl = install(ctx, syn.topo_random(16, 77))
    avail = list(range(16))
    want = oracle.besteffort(W, nl, avail, [], 3, oracle.GO_FAITHFUL, "fast")
    assert ctx.besteffort(avail, [], 3, gpa.MODE_FORCE_ENUM) == (gpa.OK, want[1])
    st = ctx.search_stats()
    assert st["partitions"] == 22422400 and st["sets"] == 6 and st["padding"] == 2 and st["algorithm"] == 1
    assert ctx.besteffort(avail, [], 3) == (gpa.OK, want[1])   # default: the subset DP at this size
    assert ctx.search_stats()["algorithm"] == 2


@pytest.mark.parametrize("n", [13, 14, 15, 16])
def test_slice_aliasing_D6(ctx, oracle, n):
    """m = 7 / 8 sets: go_faithful must return what the aliased Go slices return, intended
    must find the spiked pair"""
    for seed in range(3):
        W, nl = install(ctx, syn.topo_spiked(n, seed))
        avail = list(range(n))
        for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
            want = oracle.besteffort(W, nl, avail, [], 2, mode, "fast")
            for algo in (gpa.MODE_FORCE_ENUM, gpa.MODE_FORCE_DP):
                assert ctx.besteffort(avail, [], 2, mode | algo) == (gpa.OK, want[1]), (n, seed, mode, algo)
        assert ctx.besteffort(avail, [], 2, gpa.MODE_INTENDED)[1] == [n - 2, n - 1]


def test_hand_derived_slice_aliasing_vectors(ctx):
    """The known answers of synthetic.HAND_DERIVED_D6 - derived on paper from
    besteffort_policy.go:57,247,285 and Go's append growth, NOT from the oracle - against the CUDA path itself."""
    for n, pair, faithful, intended in syn.HAND_DERIVED_D6:
        W, nl = syn.spiked_uniform(n, pair)
        ctx.set_topology(W, nl)
        ctx.set_device_order(list(range(n)))
        for algo in (gpa.MODE_FORCE_ENUM, gpa.MODE_FORCE_DP, 0):
            assert ctx.besteffort(list(range(n)), [], 2, gpa.MODE_GO_FAITHFUL | algo) == (gpa.OK, faithful), (n, pair, algo)
            assert ctx.besteffort(list(range(n)), [], 2, gpa.MODE_INTENDED | algo) == (gpa.OK, intended), (n, pair, algo)
    # SURVEY.md 8c (1), also derived by hand: 5 GPUs, k = 3, W01 = 150, W23 = W24 = W34 = 50, rest 10 -> [0, 1, nil]
    W = np.full((5, 5), 10, dtype=np.int32)
    np.fill_diagonal(W, 0)
    W[0, 1] = W[1, 0] = 150
    for i, j in ((2, 3), (2, 4), (3, 4)):
        W[i, j] = W[j, i] = 50
    ctx.set_topology(W, np.ones((5, 5), dtype=np.uint8))
    ctx.set_device_order(list(range(5)))
    for algo in (gpa.MODE_FORCE_ENUM, gpa.MODE_FORCE_DP, gpa.MODE_FORCE_KERNEL, 0):
        assert ctx.besteffort([0, 1, 2, 3, 4], [], 3, algo) == (gpa.OK, [0, 1, gpa.NIL])


@pytest.mark.parametrize("canonical", [True, False])
def test_batch_n9_to_n12(ctx, oracle, canonical):
    """the batch kernels up to GPA_BATCH_MAX_AVAILABLE devices per request"""
    order = [int(x) for x in np.random.default_rng(3).permutation(12)]
    W, nl = install(ctx, syn.topo_random(12, 11), order if canonical else None)
    reqs = syn.random_request_batch(12, 3000, seed=12, order=order)
    for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
        got = ctx.besteffort_batch(reqs, mode)
        assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, mode), reqs, "n12")


def test_batch_mixed_canonical_and_foreign_orders_in_one_launch(ctx, oracle):
    """half the records list devices in the node order (fast kernel), half in a shuffled order (deferred to the
    general kernel inside the same call); plus records with an unknown device and with a duplicate"""
    W, nl = install(ctx, syn.topo_dgx1v_like(), ORDER8)
    a = syn.random_request_batch(8, 20000, seed=31, order=ORDER8)
    rng = np.random.default_rng(32)
    b_list = []
    for _ in range(5000):
        n = int(rng.integers(1, 9))
        avail = [int(x) for x in rng.permutation(8)[:n]]
        req = [avail[int(rng.integers(n))]] if rng.random() < 0.3 else []
        b_list.append((avail, req, int(rng.integers(1, n + 1))))
    b = syn.pack_requests(b_list)
    reqs = np.concatenate([a, b])
    reqs = reqs[rng.permutation(len(reqs))]
    got = ctx.besteffort_batch(reqs)
    assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, 0), reqs, "mixed")
    assert (got["dev"][:, 0] != -126).all()  # the internal "deferred" mark never leaves the library
    odd = syn.pack_requests([([0, 1, 200], [], 2), ([0, 1, 1], [], 2), ([5, 3], [], 2)])
    res = ctx.besteffort_batch(odd)
    assert gpa.decode_result(res[0])[0] == gpa.ERR_INVALID and gpa.decode_result(res[1])[0] == gpa.ERR_INVALID
    assert gpa.decode_result(res[2]) == (gpa.OK, [5, 3])


def test_fast_path_is_bypassed_for_big_or_asymmetric_nodes(ctx, oracle):
    """more than 12 devices on the node, or asymmetric link counts: the general kernel takes the whole batch"""
    W, nl = install(ctx, syn.topo_random(16, 2))
    reqs = syn.random_request_batch(16, 2000, seed=5)
    reqs = reqs[reqs["n_avail"] <= 12]
    assert_batch_equal(ctx.besteffort_batch(reqs), oracle_batch(oracle, W, nl, reqs, 0), reqs, "nt16")
    W, nl = syn.links_to_weights(syn.topo_random(8, 3))
    nl = nl.copy()
    nl[2, 5] += 1
    ctx.set_topology(W, nl)
    reqs = syn.random_request_batch(8, 4000, seed=6)
    got = ctx.besteffort_batch(reqs)
    assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, 0), reqs, "asym")
    assert (got["dev"][:, 0] == gpa.RES_ERROR).sum() > 50


def test_batch_stream_scale_independent_properties(ctx, oracle):
    """full bench-size batch (2^20 requests): (1) equals the oracle on a seeded sample,
    (2) every answer is empty or `size` distinct available devices containing the required
    one, (3) relabelling invariance: scaling all weights by a constant changes nothing,
    (4) a second identical call gives identical bytes (no state leaks between launches)"""
    W, nl = install(ctx, syn.topo_dgx1v_like(), ORDER8)
    reqs = syn.random_request_batch(8, 1 << 20, seed=2, order=ORDER8)
    got = ctx.besteffort_batch(reqs)
    idx = np.random.default_rng(0).choice(len(reqs), 50000, replace=False)
    want = oracle_batch(oracle, W, nl, reqs[idx], 0)
    assert_batch_equal(got[idx], want, reqs[idx], "sample")
    dev = got["dev"].astype(np.int16)
    n_out = (dev != gpa.RES_END).sum(axis=1)
    assert ((n_out == 0) | (n_out == reqs["size"])).all()
    assert (dev[:, 0] != gpa.RES_ERROR).all()
    # each answer is a subset of the available devices and holds the required one
    avail_bits = np.zeros(len(reqs), dtype=np.int64)
    for p in range(8):
        avail_bits |= np.where(p < reqs["n_avail"], 1 << reqs["avail"][:, p].astype(np.int64), 0)
    ans_bits = np.zeros(len(reqs), dtype=np.int64)
    for s in range(16):
        d = dev[:, s]
        ans_bits |= np.where(d >= 0, 1 << np.clip(d, 0, 62).astype(np.int64), 0)
    assert ((ans_bits & ~avail_bits) == 0).all()
    ctx.set_topology(W * 3, nl)
    assert (ctx.besteffort_batch(reqs)["dev"] == got["dev"]).all()
    ctx.set_topology(W, nl)
    assert (ctx.besteffort_batch(reqs)["dev"] == got["dev"]).all()


def test_device_resident_batch_on_a_torch_stream(ctx, oracle):
    """gpa_besteffort_batch_device: inputs and outputs already in HBM, caller's stream"""
    import torch
    W, nl = install(ctx, syn.topo_two_sockets(8), ORDER8)
    reqs = syn.random_request_batch(8, 100000, seed=7, order=ORDER8)
    d_in = torch.from_numpy(reqs.view(np.uint8).reshape(-1, 32)).cuda()
    d_out = torch.zeros((len(reqs), 16), dtype=torch.int8, device="cuda")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ctx.besteffort_batch_device(d_in.data_ptr(), len(reqs), 8, d_out.data_ptr(), stream=s.cuda_stream)
    s.synchronize()
    got = np.zeros(len(reqs), dtype=syn.RES_DTYPE)
    got["dev"] = d_out.cpu().numpy()
    assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, 0), reqs, "device-resident")


def test_pinned_host_buffers(ctx, oracle):
    W, nl = install(ctx, syn.topo_two_sockets(8))
    reqs = syn.random_request_batch(8, 50000, seed=8, order=ORDER8)
    pin_in = gpa.PinnedBuffer(len(reqs), syn.REQ_DTYPE)
    pin_out = gpa.PinnedBuffer(len(reqs), syn.RES_DTYPE)
    pin_in.array[:] = reqs
    got = ctx.besteffort_batch(pin_in.array, out=pin_out.array)
    assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, 0), reqs, "pinned")
    pin_in.free()
    pin_out.free()


def test_batch_max_avail_hint(ctx, oracle):
    W, nl = install(ctx, syn.topo_two_sockets(8))
    reqs = syn.pack_requests([([0, 1, 2, 3, 4, 5], [], 2), ([0, 1, 2], [], 2)])
    want0 = oracle.besteffort(W, nl, [0, 1, 2, 3, 4, 5], [], 2)[1]
    got = ctx.besteffort_batch(reqs, max_avail=4)      # the caller lied: 6 devices in the first record
    # the fast kernel does not need the hint and answers; the general kernel sizes its tables by it and refuses
    assert gpa.decode_result(got[0]) in ((gpa.ERR_TOO_LARGE, []), (gpa.OK, want0))
    ctx.set_device_order([5, 4, 3, 2, 1, 0, 6, 7])     # now the records are not in node order: general kernel
    got = ctx.besteffort_batch(reqs, max_avail=4)
    assert gpa.decode_result(got[0])[0] == gpa.ERR_TOO_LARGE
    assert gpa.decode_result(got[1]) == (gpa.OK, oracle.besteffort(W, nl, [0, 1, 2], [], 2)[1])
    got = ctx.besteffort_batch(reqs, max_avail=8)
    assert_batch_equal(got, oracle_batch(oracle, W, nl, reqs, 0), reqs, "hint")


def test_mask_format_matches_record_format_and_oracle(ctx, oracle):
    """gpa_besteffort_batch_masks: 8-byte label-mask records; same decisions as the 32-byte records"""
    for nt, topo, order in ((8, syn.topo_dgx1v_like(), ORDER8),
                            (12, syn.topo_random(12, 11), [int(x) for x in np.random.default_rng(3).permutation(12)])):
        W, nl = install(ctx, topo, order)
        reqs = syn.random_request_batch(nt, 60000, seed=50 + nt, order=order)
        reqs = reqs[reqs["n_avail"] <= 12]
        masks = syn.records_to_masks(reqs, order)
        for mode in (gpa.MODE_GO_FAITHFUL, gpa.MODE_INTENDED):
            want = oracle_batch(oracle, W, nl, reqs, mode)
            got = syn.mask_results_to_records(ctx.besteffort_batch_masks(masks, mode), order)
            # the record path lists the chosen devices in `available` order = node order, like the mask path
            assert_batch_equal(got, want, reqs, "masks nt=%d" % nt)
            assert_batch_equal(ctx.besteffort_batch(reqs, mode), want, reqs, "records nt=%d" % nt)
    # guards, foreign / unavailable required devices, size 1, bad masks
    W, nl = install(ctx, syn.topo_dgx1v_like(), ORDER8)
    m = np.zeros(7, dtype=syn.MASK_REQ_DTYPE)
    m["avail_mask"] = [0b1111, 0b1111, 0b1111, 0b0110, 0b0110, 0b1 << 9, 0b1111]
    m["req_mask"] = [0, 0, 0b10000, 0b0100, 0, 0, 0]
    m["size"] = [0, 5, 2, 1, 1, 1, 2]
    m["flags"] = [0, 0, 0, 0, 0, 0, 1]
    r = ctx.besteffort_batch_masks(m)
    assert r["chosen_mask"].tolist()[:5] == [0, 0, 0, 0b0100, 0b0010] and r["status"].tolist()[:5] == [0] * 5
    assert r["status"][5] == gpa.ERR_INVALID and r["chosen_mask"][6] == 0 and r["status"][6] == 0


def test_maximum_sizes(ctx, oracle):
    """64-device topology (GPA_MAX_TOPOLOGY), 16 available devices (GPA_MAX_AVAILABLE), one too many of each"""
    t = syn.topo_random(64, 9)
    W, nl = install(ctx, t)
    rng = np.random.default_rng(1)
    for n, k in ((16, 8), (16, 16), (12, 4), (5, 2)):
        avail = [int(x) for x in rng.choice(64, n, replace=False)]
        want = oracle.besteffort(W, nl, avail, [], k)
        assert ctx.besteffort(avail, [], k) == (gpa.OK, want[1])
    avail17 = [int(x) for x in rng.choice(64, 17, replace=False)]
    assert ctx.besteffort(avail17, [], 2)[0] == gpa.ERR_TOO_LARGE
    assert ctx.besteffort(avail17, [], 18) == (gpa.OK, [])       # the guard len(available) < size answers first
    reqs = syn.pack_requests([([int(x) for x in rng.choice(64, int(rng.integers(1, 13)), replace=False)], [], 1 + int(rng.integers(0, 3)))
                              for _ in range(2000)])
    assert_batch_equal(ctx.besteffort_batch(reqs), oracle_batch(oracle, W, nl, reqs, 0), reqs, "nt64")
    with pytest.raises(gpa.GpaError):
        ctx.set_topology(np.zeros((65, 65), dtype=np.int32), np.zeros((65, 65), dtype=np.uint8))
    # sizes up to 255 travel in the packed record
    big = syn.pack_requests([([1, 2, 3], [], 255), ([1, 2, 3], [], 3)])
    got = ctx.besteffort_batch(big)
    assert gpa.decode_result(got[0]) == (gpa.OK, []) and gpa.decode_result(got[1]) == (gpa.OK, [1, 2, 3])


def test_concurrent_callers_share_one_context(ctx, oracle):
    """grpc-go runs every RPC on its own goroutine (SURVEY.md §8b threading): calls on one context are serialised
    inside the library and every caller gets its own answer"""
    import threading
    t = syn.topo_random(12, 21)
    W, nl = install(ctx, t)
    rng = np.random.default_rng(2)
    cases = []
    for _ in range(64):
        n = int(rng.integers(4, 13))
        avail = [int(x) for x in rng.permutation(12)[:n]]
        k = int(rng.integers(2, n + 1))
        cases.append((avail, k, oracle.besteffort(W, nl, avail, [], k)[1]))
    batch = syn.random_request_batch(12, 5000, seed=77)
    want_batch = oracle_batch(oracle, W, nl, batch, 0)
    errors = []

    def single(tid):
        for i in range(tid, len(cases), 4):
            avail, k, want = cases[i]
            for _ in range(5):
                if ctx.besteffort(avail, [], k) != (gpa.OK, want):
                    errors.append(("single", avail, k))

    def batched():
        for _ in range(5):
            if (ctx.besteffort_batch(batch)["dev"] != want_batch["dev"]).any():
                errors.append(("batch",))

    threads = [threading.Thread(target=single, args=(i,)) for i in range(4)] + [threading.Thread(target=batched) for _ in range(2)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:3]


def test_two_contexts_are_independent(oracle):
    a, b = gpa.Context(0), gpa.Context(0)
    try:
        ta, tb = syn.topo_dgx1v_like(), syn.topo_two_sockets(8)
        a.set_topology_links(ta)
        b.set_topology_links(tb)
        Wa, nla = syn.links_to_weights(ta)
        Wb, nlb = syn.links_to_weights(tb)
        for k in range(2, 8):
            assert a.besteffort(list(range(8)), [], k)[1] == oracle.besteffort(Wa, nla, list(range(8)), [], k)[1]
            assert b.besteffort(list(range(8)), [], k)[1] == oracle.besteffort(Wb, nlb, list(range(8)), [], k)[1]
    finally:
        a.close()
        b.close()


def test_launch_counter_counts_kernels(ctx):
    install(ctx, syn.topo_two_sockets(8))
    before = ctx.launch_count
    ctx.besteffort(list(range(8)), [], 4, gpa.MODE_FORCE_KERNEL)
    ctx.besteffort(list(range(8)), [], 4)  # answered from the host copy of the answer table: no kernel
    ctx.besteffort_batch(syn.random_request_batch(8, 100, seed=1))  # one launch: the table kernel also answers what it cannot look up
    ctx.besteffort(list(range(8)), [], 1)  # answered by the guards: no kernel
    assert ctx.launch_count == before + 2
    ctx.besteffort_batch(syn.random_request_batch(8, 100, seed=1), gpa.MODE_FORCE_ENUM)  # tile-binned kernel + deferred pass
    assert ctx.launch_count == before + 4
