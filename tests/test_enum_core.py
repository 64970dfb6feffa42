"""CPU check of the enumeration core the CUDA kernels are built from
(k8s-gpu-sharing-plugin_b200/csrc/gpa_enum.cuh), run sequentially by the test harness
tests/csrc/enum_check.cpp and compared with the oracle.  This is a unit test of shared
header logic; the product path itself only exists as CUDA kernels."""
import ctypes as C
import itertools
import os
import subprocess

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import synthetic as syn

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "enum_check.cpp")
HDR = os.path.join(HERE, "..", "k8s-gpu-sharing-plugin_b200", "csrc", "gpa_enum.cuh")
HDR2 = os.path.join(HERE, "..", "k8s-gpu-sharing-plugin_b200", "csrc", "part_table.cuh")
HDR3 = os.path.join(HERE, "..", "k8s-gpu-sharing-plugin_b200", "csrc", "ans_table.cuh")
# GPA_HARNESS_DIR / GPA_HARNESS_CXXFLAGS: build the harness elsewhere with other flags (tools/sanitize_cpu.sh: ASan + UBSan)
SO = os.path.join(os.environ.get("GPA_HARNESS_DIR", os.path.join(HERE, "csrc")), "libenum_check.so")
XFLAGS = os.environ.get("GPA_HARNESS_CXXFLAGS", "-O2").split()


@pytest.fixture(scope="module")
def harness():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(HDR), os.path.getmtime(HDR2), os.path.getmtime(HDR3)):
        subprocess.check_call(["g++"] + XFLAGS + ["-std=c++17", "-fPIC", "-shared", "-o", SO, SRC])
    L = C.CDLL(SO)
    L.enum_check_search.restype = C.c_int
    L.enum_check_search.argtypes = [C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, C.c_uint32, C.c_int,
                                    C.c_int, C.c_ulonglong, C.POINTER(C.c_int32), C.POINTER(C.c_int),
                                    C.POINTER(C.c_ulonglong), C.POINTER(C.c_int)]
    L.enum_check_search_sparse.restype = C.c_int
    L.enum_check_search_sparse.argtypes = [C.c_int, C.POINTER(C.c_int32), C.c_uint32, C.c_uint32, C.c_int, C.c_int,
                                           C.c_ulonglong, C.POINTER(C.c_int32), C.POINTER(C.c_int)]
    L.enum_check_tracked_sparse.restype = C.c_int
    L.enum_check_tracked_sparse.argtypes = [C.c_int, C.POINTER(C.c_int32), C.c_uint32, C.c_uint32, C.c_int, C.c_int,
                                            C.POINTER(C.c_int32), C.POINTER(C.c_int)]
    L.enum_check_dp.restype = C.c_int
    L.enum_check_dp.argtypes = [C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, C.c_uint32, C.c_int, C.c_int,
                                C.POINTER(C.c_int32), C.POINTER(C.c_int)]
    L.enum_check_level_sets.restype = C.c_int
    L.enum_check_level_sets.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint32, C.POINTER(C.c_uint32), C.c_int]
    return L


def run(L, W, avail, req, k, intended, want_items):
    W = np.ascontiguousarray(W, dtype=np.int32)
    av = np.ascontiguousarray(avail, dtype=np.int32)
    rmask = 0
    for d in req:
        rmask |= 1 << list(avail).index(d)
    out = np.zeros(64, dtype=np.int32)
    n_out, n_items, depth = C.c_int(0), C.c_ulonglong(0), C.c_int(0)
    rc = L.enum_check_search(W.shape[0], W.ctypes.data_as(C.POINTER(C.c_int32)), av.ctypes.data_as(C.POINTER(C.c_int32)),
                             len(av), rmask, k, int(intended), want_items, out.ctypes.data_as(C.POINTER(C.c_int32)),
                             C.byref(n_out), C.byref(n_items), C.byref(depth))
    assert rc == 0
    return [int(x) for x in out[:n_out.value]], n_items.value, depth.value


def test_stepper_emits_the_reference_level_order(harness):
    """one iterate() level: anchor + every (k-1)-combination of the literal remainder in
    lexicographic order, keeping those with 0 or all P nil pads (besteffort_policy.go:266-276)"""
    for n, k in [(5, 3), (8, 3), (7, 4), (9, 4), (10, 3), (6, 5), (11, 6), (7, 2), (13, 5), (5, 5), (4, 3)]:
        N = -(-n // k) * k
        P = N - n
        for tok_present in ([0, 1] if P else [0]):
            # any remainder that a real walk can reach: r reals with r = j*k (no token) or j*k - P (token)
            for j in range(1, N // k + 1):
                r = j * k - (P if tok_present else 0)
                if r < 1 or r > n:
                    continue
                rng = np.random.default_rng(n * 100 + k * 10 + j)
                reals = sorted(int(x) for x in rng.choice(n, r, replace=False))
                lit = reals + ([n + i for i in range(P)] if tok_present else [])
                want = []
                for combo in itertools.combinations(lit[1:], k - 1):
                    pads = sum(1 for p in combo if p >= n)
                    if pads not in (0, P):
                        continue
                    mask = 1 << lit[0]
                    for p in combo:
                        mask |= (1 << p) if p < n else (1 << n)
                    want.append(mask)
                buf = (C.c_uint32 * 20000)()
                cnt = harness.enum_check_level_sets(n, k, tok_present, sum(1 << p for p in reals), buf, 20000)
                assert [buf[i] for i in range(cnt)] == want, (n, k, tok_present, reals)


@pytest.mark.parametrize("want_items", [1, 40, 10**12])
@pytest.mark.parametrize("topo_i", [0, 2, 3, 5])
def test_core_matches_oracle_exhaustive_n8(harness, oracle, topo_i, want_items):
    name, t = syn.standard_topologies()[topo_i]
    W, nl = syn.links_to_weights(t)
    order = syn.sorted_positions(syn.node_uuids(8, seed=1))
    cnt = 0
    for avail, req, k in syn.exhaustive_requests(order, seed=topo_i, n_pairs=3):
        if k < 2 or len(req) > k:
            continue
        if len(avail) < 6 and (cnt % 3):
            cnt += 1
            continue
        want = oracle.besteffort(W, nl, avail, req, k, oracle.GO_FAITHFUL, "fast")
        got, _, _ = run(harness, W, avail, req, k, False, want_items)
        assert (oracle.OK, got) == want, (name, avail, req, k)
        cnt += 1
    assert cnt > 2000


@pytest.mark.parametrize("n,k", [(9, 2), (9, 4), (10, 3), (10, 4), (11, 2), (11, 3), (11, 5), (12, 3), (12, 5), (12, 7),
                                 (13, 2), (13, 3), (13, 6), (14, 2), (14, 4), (14, 5), (15, 2), (15, 4), (16, 2), (16, 5),
                                 (16, 6), (16, 7), (16, 8), (16, 9), (16, 13), (12, 11), (16, 16)])
def test_core_matches_oracle_seeded_large(harness, oracle, n, k):
    big = oracle.partition_count(n, k) > 100000  # keep the CPU suite to minutes
    for seed in range(1 if big else 2):
        W, nl = syn.links_to_weights(syn.topo_random(n, 300 + seed))
        rng = np.random.default_rng(seed)
        avail = [int(x) for x in rng.permutation(n)]
        reqs = [[], [avail[int(rng.integers(n))]]]
        if k >= 3 and not big:
            reqs.append([avail[int(x)] for x in rng.choice(n, 2, replace=False)])
        for req in reqs:
            for mode in ((oracle.GO_FAITHFUL,) if big and req else (oracle.GO_FAITHFUL, oracle.INTENDED)):
                want = oracle.besteffort(W, nl, avail, req, k, mode, "fast")
                for want_items in ((5000,) if big else (1, 5000, 10**12)):
                    got, n_items, depth = run(harness, W, avail, req, k, mode == oracle.INTENDED, want_items)
                    assert (oracle.OK, got) == want, (n, k, seed, req, mode, want_items, depth)


@pytest.mark.parametrize("n", [13, 14, 15, 16])
def test_core_reproduces_slice_aliasing_D6(harness, oracle, n):
    for seed in range(3 if n <= 14 else 1):
        W, nl = syn.links_to_weights(syn.topo_spiked(n, seed))
        avail = list(range(n))
        for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
            want = oracle.besteffort(W, nl, avail, [], 2, mode, "fast")
            got, _, _ = run(harness, W, avail, [], 2, mode == oracle.INTENDED, 20000)
            assert (oracle.OK, got) == want, (n, seed, mode)


def test_core_nil_pad_vector_D5(harness, oracle):
    t = syn.topo_uniform(5)
    for i, j in ((2, 3), (2, 4), (3, 4)):
        t[i, j, 0] = t[j, i, 0] = 5
    t[0, 1] = t[1, 0] = (5, syn.nvlink_type(1))
    W, nl = syn.links_to_weights(t)
    got, _, _ = run(harness, W, [0, 1, 2, 3, 4], [], 3, False, 1)
    assert got == [0, 1, -1]


@pytest.mark.parametrize("nt", [8, 12])
def test_core_sparse_label_masks_match_oracle(harness, oracle, nt):
    """the canonical-order layout: devices = set bits of a label mask, one score table per topology"""
    rng = np.random.default_rng(nt)
    W, nl = syn.links_to_weights(syn.topo_random(nt, 40 + nt))
    # asymmetric weights too: the earlier label's view must be used
    W = W.copy()
    W[np.triu_indices(nt, 1)] += 10
    Wc = np.ascontiguousarray(W, dtype=np.int32)
    done = 0
    for _ in range(600):
        mask = int(rng.integers(1, 1 << nt))
        labels = [b for b in range(nt) if mask >> b & 1]
        if len(labels) < 2 or oracle.partition_count(len(labels), 2) > 20000:
            continue
        k = int(rng.integers(2, len(labels) + 1))
        req = [int(x) for x in rng.choice(labels, size=int(rng.integers(0, min(k, 2) + 1)), replace=False)]
        rmask = sum(1 << r for r in req)
        for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
            want = oracle.besteffort(W, nl, labels, req, k, mode, "fast")
            out = np.zeros(64, dtype=np.int32)
            n_out = C.c_int(0)
            rc = harness.enum_check_search_sparse(nt, Wc.ctypes.data_as(C.POINTER(C.c_int32)), mask, rmask, k,
                                                  int(mode == oracle.INTENDED), int(rng.choice([1, 50, 10**9])),
                                                  out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out))
            assert rc == 0 and (oracle.OK, [int(x) for x in out[:n_out.value]]) == want, (mask, k, req, mode)
            if -(-len(labels) // k) <= 4:  # the tracked whole-tree walk covers trees of up to 4 sets
                out2 = np.zeros(64, dtype=np.int32)
                n2 = C.c_int(0)
                rc = harness.enum_check_tracked_sparse(nt, Wc.ctypes.data_as(C.POINTER(C.c_int32)), mask, rmask, k,
                                                       int(mode == oracle.INTENDED), out2.ctypes.data_as(C.POINTER(C.c_int32)),
                                                       C.byref(n2))
                assert rc == 0 and (oracle.OK, [int(x) for x in out2[:n2.value]]) == want, ("tracked", mask, k, req, mode)
        done += 1
    assert done > 300


def run_dp(L, W, avail, req, k, intended):
    W = np.ascontiguousarray(W, dtype=np.int32)
    av = np.ascontiguousarray(avail, dtype=np.int32)
    rmask = 0
    for d in req:
        rmask |= 1 << list(avail).index(d)
    out = np.zeros(64, dtype=np.int32)
    n_out = C.c_int(0)
    rc = L.enum_check_dp(W.shape[0], W.ctypes.data_as(C.POINTER(C.c_int32)), av.ctypes.data_as(C.POINTER(C.c_int32)), len(av),
                         rmask, k, int(intended), out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out))
    assert rc == 0
    return [int(x) for x in out[:n_out.value]]


@pytest.mark.parametrize("topo_i", [0, 2, 3, 5, 9])
def test_subset_dp_matches_oracle_exhaustive_n8(harness, oracle, topo_i):
    """the subset dynamic programme must return the FIRST best partition's set, like the enumeration"""
    name, t = syn.standard_topologies()[topo_i]
    W, nl = syn.links_to_weights(t)
    order = syn.sorted_positions(syn.node_uuids(8, seed=1))
    cnt = 0
    for avail, req, k in syn.exhaustive_requests(order, seed=topo_i, n_pairs=4):
        if k < 2 or len(req) > k:
            continue
        want = oracle.besteffort(W, nl, avail, req, k, oracle.GO_FAITHFUL, "fast")
        assert (oracle.OK, run_dp(harness, W, avail, req, k, False)) == want, (name, avail, req, k)
        cnt += 1
    assert cnt > 4000


@pytest.mark.parametrize("n,k", [(9, 2), (9, 4), (10, 3), (10, 4), (11, 2), (11, 3), (11, 5), (12, 3), (12, 5), (12, 7), (12, 11),
                                 (13, 2), (13, 3), (13, 6), (14, 2), (14, 4), (14, 5), (15, 2), (15, 4), (16, 2), (16, 3), (16, 4),
                                 (16, 5), (16, 6), (16, 7), (16, 8), (16, 9), (16, 13), (16, 16)])
def test_subset_dp_matches_oracle_seeded_large(harness, oracle, n, k):
    big = oracle.partition_count(n, k) > 3000000
    for seed in range(1 if big else 2):
        W, nl = syn.links_to_weights(syn.topo_random(n, 300 + seed))
        rng = np.random.default_rng(seed)
        avail = [int(x) for x in rng.permutation(n)]
        reqs = [[], [avail[int(rng.integers(n))]]]
        if k >= 3 and not big:
            reqs.append([avail[int(x)] for x in rng.choice(n, 2, replace=False)])
        for req in reqs:
            for mode in ((oracle.GO_FAITHFUL,) if big else (oracle.GO_FAITHFUL, oracle.INTENDED)):
                want = oracle.besteffort(W, nl, avail, req, k, mode, "fast")
                assert (oracle.OK, run_dp(harness, W, avail, req, k, mode == oracle.INTENDED)) == want, (n, k, seed, req, mode)


def test_subset_dp_handles_ties_spikes_and_pads(harness, oracle):
    for n in (13, 14):  # D6 vectors: go_faithful differs from intended
        for seed in range(3):
            W, nl = syn.links_to_weights(syn.topo_spiked(n, seed))
            for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
                want = oracle.besteffort(W, nl, list(range(n)), [], 2, mode, "fast")
                assert (oracle.OK, run_dp(harness, W, list(range(n)), [], 2, mode == oracle.INTENDED)) == want
    for n in range(2, 11):  # all-equal weights: every partition ties, the very first must win
        W, nl = syn.links_to_weights(syn.topo_uniform(n))
        for k in range(2, n + 1):
            for req in ([], [n - 1]):
                want = oracle.besteffort(W, nl, list(range(n)), req, k, oracle.GO_FAITHFUL, "fast")
                assert (oracle.OK, run_dp(harness, W, list(range(n)), req, k, False)) == want, (n, k, req)


@pytest.mark.parametrize("topo_i", [0, 1, 3, 6, 12])
def test_partition_tables_match_oracle_exhaustive_n8(harness, oracle, topo_i):
    """part_table.cuh (the fast batch kernel's path on nodes of up to 8 GPUs): words, index and keys built as the library
    builds them, every (mask, size, required) answered by part_table_search, at every 16-byte phase of the key table"""
    harness.enum_check_tables_new.restype = C.c_void_p
    harness.enum_check_tables_new.argtypes = [C.c_int, C.POINTER(C.c_int32)]
    harness.enum_check_tables_free.argtypes = [C.c_void_p]
    harness.enum_check_tables_words.argtypes = [C.c_void_p]
    harness.enum_check_table_search.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int,
                                                C.POINTER(C.c_int32), C.POINTER(C.c_int)]
    name, t = syn.standard_topologies()[topo_i]
    W, nl = syn.links_to_weights(t)
    Wc = np.ascontiguousarray(W, dtype=np.int32)
    tab = harness.enum_check_tables_new(8, Wc.ctypes.data_as(C.POINTER(C.c_int32)))
    assert tab and harness.enum_check_tables_words(tab) == 6289
    rng = np.random.default_rng(topo_i)
    out = np.zeros(64, dtype=np.int32)
    n_out = C.c_int(0)
    checked = 0
    try:
        for mask in range(1, 256):
            labels = [b for b in range(8) if mask >> b & 1]
            for k in range(2, len(labels) + 1):
                reqs = [[]] + [[d] for d in labels]
                if k >= 2 and len(labels) >= 2:
                    reqs.append([int(x) for x in rng.choice(labels, size=2, replace=False)])
                for req in reqs:
                    rmask = sum(1 << r for r in req)
                    for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
                        want = oracle.besteffort(W, nl, labels, req, k, mode, "fast")
                        rc = harness.enum_check_table_search(tab, mask, rmask, k, int(mode == oracle.INTENDED), checked & 3,
                                                             out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out))
                        assert rc == 0 and (oracle.OK, [int(x) for x in out[:n_out.value]]) == want, (name, mask, k, req, mode)
                        checked += 1
    finally:
        harness.enum_check_tables_free(tab)
    assert checked > 10000


@pytest.mark.parametrize("nt", [2, 3, 5, 6, 7])
def test_partition_tables_on_smaller_nodes(harness, oracle, nt):
    """the tables are built per node size: every (mask, size, required) of a node of 2..7 GPUs, asymmetric weights too"""
    harness.enum_check_tables_new.restype = C.c_void_p
    harness.enum_check_tables_new.argtypes = [C.c_int, C.POINTER(C.c_int32)]
    harness.enum_check_tables_free.argtypes = [C.c_void_p]
    harness.enum_check_table_search.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int,
                                                C.POINTER(C.c_int32), C.POINTER(C.c_int)]
    W, nl = syn.links_to_weights(syn.topo_random(nt, 70 + nt))
    W = W.copy()
    W[np.triu_indices(nt, 1)] += 10          # the earlier label's view must be the one used (:314)
    Wc = np.ascontiguousarray(W, dtype=np.int32)
    tab = harness.enum_check_tables_new(nt, Wc.ctypes.data_as(C.POINTER(C.c_int32)))
    assert tab
    out = np.zeros(64, dtype=np.int32)
    n_out = C.c_int(0)
    checked = 0
    try:
        for mask in range(1, 1 << nt):
            labels = [b for b in range(nt) if mask >> b & 1]
            for k in range(2, len(labels) + 1):
                for req in [[]] + [[d] for d in labels] + ([labels[:2]] if len(labels) >= 2 else []):
                    rmask = sum(1 << r for r in req)
                    for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
                        want = oracle.besteffort(W, nl, labels, req, k, mode, "fast")
                        rc = harness.enum_check_table_search(tab, mask, rmask, k, int(mode == oracle.INTENDED), checked & 3,
                                                             out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out))
                        assert rc == 0 and (oracle.OK, [int(x) for x in out[:n_out.value]]) == want, (nt, mask, k, req, mode)
                        checked += 1
    finally:
        harness.enum_check_tables_free(tab)
    assert checked >= (2 if nt == 2 else 20)


@pytest.mark.parametrize("nt,topo_i", [(8, 3), (8, 9), (5, 1), (3, 2)])
def test_answer_table_every_entry_matches_oracle(harness, oracle, nt, topo_i):
    """ans_table.cuh: the table the GPU builds at topology install, computed here entry by entry with the same function.
    EVERY (mask, required subset, size) entry is compared with the oracle in both alias modes (the table is shared by
    both: no partition of a node of up to 8 GPUs has the 7 sets the aliasing needs), and the index arithmetic of the
    lookup is checked to be a bijection onto the table."""
    harness.enum_check_tables_new.restype = C.c_void_p
    harness.enum_check_tables_new.argtypes = [C.c_int, C.POINTER(C.c_int32)]
    harness.enum_check_tables_free.argtypes = [C.c_void_p]
    harness.enum_check_answer_table.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint16), C.c_int]
    harness.enum_check_answer_index.restype = C.c_uint32
    harness.enum_check_answer_index.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
    if nt == 8:
        name, t = syn.standard_topologies()[topo_i]
    else:
        name, t = "random-%d" % nt, syn.topo_random(nt, 40 + topo_i)
    W, nl = syn.links_to_weights(t)
    Wc = np.ascontiguousarray(W, dtype=np.int32)
    tab = harness.enum_check_tables_new(nt, Wc.ctypes.data_as(C.POINTER(C.c_int32)))
    assert tab
    n_ent = 3 ** nt * 8
    try:
        tables = []
        for intended in (0, 1):
            buf = np.zeros(n_ent, dtype=np.uint16)
            assert harness.enum_check_answer_table(tab, intended, buf.ctypes.data_as(C.POINTER(C.c_uint16)), n_ent) == n_ent
            tables.append(buf)
        assert (tables[0] == tables[1]).all()
    finally:
        harness.enum_check_tables_free(tab)
    table = tables[0]
    seen = np.zeros(n_ent, dtype=bool)
    # the oracle on every entry, batched: requests over labels 0..nt-1 in label order
    reqs, where = [], []
    for mask in range(1 << nt):
        labels = [b for b in range(nt) if mask >> b & 1]
        sub = mask
        while True:  # every subset of mask, 0 included
            req = [b for b in range(nt) if sub >> b & 1]
            for size in range(1, 9):
                idx = harness.enum_check_answer_index(mask, sub, size)
                assert idx < n_ent and not seen[idx]
                seen[idx] = True
                if mask:
                    reqs.append((labels, req, size))
                    where.append(idx)
                else:
                    assert table[idx] == 0
            if sub == 0:
                break
            sub = (sub - 1) & mask
    assert seen.all()
    packed = syn.pack_requests(reqs)
    for mode in (oracle.GO_FAITHFUL, oracle.INTENDED):
        want, _ = oracle.besteffort_batch(W, nl, packed, mode=mode, threads=8)
        for i, idx in enumerate(where):
            e = int(table[idx])
            got = [b for b in range(8) if e >> b & 1] + [-1] * (e >> 8)
            row = want["dev"][i].tolist()
            exp = row[:row.index(-128)] if -128 in row else row
            assert got == exp, (name, reqs[i], mode, got, exp)
