"""CPU check of the replica-set search core (csrc/subset_core.cuh, csrc/subset_plan.h): the product's plan, ranking,
scoring and decode code driven sequentially by tests/csrc/subset_check.cpp, against the literal oracle
(oracle/ref_subset.c).  Both forms - per-GPU count vectors and literal combinations - must give the oracle's positions
bit for bit.  The product path itself only exists as CUDA kernels (tests/test_gpu_subset.py)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import synthetic as syn
from oracle import oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "subset_check.cpp")
CSRC = os.path.join(HERE, "..", "k8s-gpu-sharing-plugin_b200", "csrc")
DEPS = [SRC] + [os.path.join(CSRC, f) for f in ("subset_core.cuh", "subset_plan.h", "gpa_enum.cuh")]
# GPA_HARNESS_DIR / GPA_HARNESS_CXXFLAGS: build the harness elsewhere with other flags (tools/sanitize_cpu.sh: ASan + UBSan)
SO = os.path.join(os.environ.get("GPA_HARNESS_DIR", os.path.join(HERE, "csrc")), "libsubset_check.so")
XFLAGS = os.environ.get("GPA_HARNESS_CXXFLAGS", "-O2").split()

FORCE_ENUM = 2
ERR_INVALID, ERR_ASYMMETRIC, ERR_TOO_LARGE = -1, -4, -5


@pytest.fixture(scope="module")
def harness():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(d) for d in DEPS):
        subprocess.check_call(["g++"] + XFLAGS + ["-std=c++17", "-fPIC", "-shared", "-o", SO, SRC])
    L = C.CDLL(SO)
    L.subset_check.restype = C.c_int
    L.subset_check.argtypes = [C.c_int, C.POINTER(C.c_uint16), C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.c_int,
                               C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.c_uint, C.POINTER(C.c_int32),
                               C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_ulonglong), C.POINTER(C.c_int)]
    return L


def run(L, W, nl, phys, req, size, mode=0, chunk=7):
    wq = np.ascontiguousarray(W, dtype=np.uint16)
    asym = np.ascontiguousarray(nl != nl.T, dtype=np.uint8)
    ph = np.ascontiguousarray(phys, dtype=np.int32)
    rq = np.ascontiguousarray(req, dtype=np.int32)
    out = np.full(max(size, 1) + 8, -9, dtype=np.int32)
    n_out, form, total, score = C.c_int(0), C.c_int(0), C.c_ulonglong(0), C.c_int(0)
    rc = L.subset_check(W.shape[0], wq.ctypes.data_as(C.POINTER(C.c_uint16)), asym.ctypes.data_as(C.POINTER(C.c_uint8)),
                        ph.ctypes.data_as(C.POINTER(C.c_int32)), len(ph), rq.ctypes.data_as(C.POINTER(C.c_int32)), len(rq),
                        size, mode, chunk, out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_out), C.byref(form),
                        C.byref(total), C.byref(score))
    return rc, [int(x) for x in out[:n_out.value]], form.value, total.value, score.value


def topologies():
    return [(name, *syn.links_to_weights(t)) for name, t in syn.standard_topologies()]


def test_known_answers(harness):
    """hand-checked: two replicas on each end of the one strong link beat four distinct GPUs"""
    W = np.full((8, 8), 20, np.int32)
    np.fill_diagonal(W, 0)
    W[0, 1] = W[1, 0] = 150
    nl = (W > 0).astype(np.uint8)
    phys = [i // 4 for i in range(32)]
    assert orc.subset_besteffort(W, nl, phys, [], 4, True) == (0, [0, 1, 4, 5], 600, 35960)
    for mode in (0, FORCE_ENUM):
        rc, pos, form, total, score = run(harness, W, nl, phys, [], 4, mode)
        assert (rc, pos, score) == (0, [0, 1, 4, 5], 600)
        assert (form, total) == ((5, 35960) if mode else (4, 330))
    # a required replica of a far GPU drags the answer to it
    assert orc.subset_besteffort(W, nl, phys, [30], 3) == (0, [0, 4, 30])
    for mode in (0, FORCE_ENUM):
        assert run(harness, W, nl, phys, [30], 3, mode)[:2] == (0, [0, 4, 30])
    # all-zero weights: the first candidate (SURVEY.md 8c vector 4)
    Z = np.zeros((8, 8), np.int32)
    for mode in (0, FORCE_ENUM):
        assert run(harness, Z, nl, phys, [], 5, mode)[:2] == (0, [0, 1, 2, 3, 4])
        assert run(harness, Z, nl, phys, [9], 3, mode)[:2] == (0, [0, 1, 9])


def test_guards_and_argument_errors(harness):
    W, nl = syn.links_to_weights(syn.topo_uniform(8))
    phys = [0, 0, 1, 1]
    for size, req in ((0, []), (-1, []), (5, []), (1, [0, 1])):   # besteffort_policy.go:35-45
        assert run(harness, W, nl, phys, req, size)[:2] == (0, [])
        assert orc.subset_besteffort(W, nl, phys, req, size) == (0, [])
    assert run(harness, W, nl, [0, 8], [], 1)[0] == ERR_INVALID        # no such GPU
    assert run(harness, W, nl, phys, [4], 2)[0] == ERR_INVALID         # no such position
    assert run(harness, W, nl, phys, [1, 1], 2)[0] == ERR_INVALID      # twice the same position
    assert run(harness, W, nl, phys, [3, 0], 2)[:2] == (0, [0, 3])     # required fills the request: no search
    assert run(harness, W, nl, list(range(8)) * 40, [], 9, FORCE_ENUM)[0] == ERR_TOO_LARGE   # 320 free positions
    big, bnl = syn.links_to_weights(syn.topo_random(17, 5))
    assert run(harness, big, bnl, list(range(17)), [], 2)[0] == ERR_TOO_LARGE                 # 17 GPUs in one request
    # where calculateGPUPairScore panics (:307-310)
    anl = nl.copy()
    anl[0, 1] = 2
    assert run(harness, W, anl, phys, [], 2)[0] == ERR_ASYMMETRIC
    assert orc.subset_besteffort(W, anl, phys, [], 2)[0] == orc.PANIC_ASYMMETRIC
    assert run(harness, W, anl, phys, [], 1)[:2] == (0, [0])           # no pair is scored for size 1
    assert orc.subset_besteffort(W, anl, phys, [], 1) == (0, [0])


@pytest.mark.parametrize("replicas", [1, 2, 4])
def test_both_forms_match_the_oracle_on_sorted_ids(harness, replicas):
    """positions in GPU order (the plugin's sorted replica ids): count vectors by default, literal when forced"""
    rng = np.random.default_rng(100 + replicas)
    seen_forms = set()
    for name, W, nl in topologies():
        for _ in range(40):
            phys, req, size = syn.random_subset_request(rng, 8, replicas, p_free=float(rng.choice([0.3, 0.6, 0.9])), max_size=6)
            want = orc.subset_besteffort(W, nl, phys, req, size, True)
            for mode in (0, FORCE_ENUM):
                rc, pos, form, total, score = run(harness, W, nl, phys, req, size, mode, chunk=int(rng.integers(1, 40)))
                if rc == ERR_TOO_LARGE:
                    assert mode == FORCE_ENUM
                    continue
                assert (rc, pos) == want[:2], (name, phys, req, size, mode)
                if form:
                    assert score == want[2]
                    seen_forms.add(form)
                    if form == 5:
                        assert total == want[3] or req   # with no required positions the literal walk IS the oracle's
    assert seen_forms == {4, 5}


def test_literal_form_on_any_order(harness):
    """positions in arbitrary order: only the literal form applies, and it still matches"""
    rng = np.random.default_rng(7)
    for name, W, nl in topologies()[1::3]:
        for _ in range(60):
            phys, req, size = syn.random_subset_request(rng, 8, 3, p_free=0.6, grouped=False, max_size=5)
            want = orc.subset_besteffort(W, nl, phys, req, size)
            rc, pos, form, _, _ = run(harness, W, nl, phys, req, size, 0, chunk=int(rng.integers(1, 100)))
            assert (rc, pos) == want, (name, phys, req, size)
            contiguous = all(phys[i] == phys[i - 1] or phys[i] not in phys[:i] for i in range(1, len(phys)))
            assert form == 0 or form == (4 if contiguous else 5)


def test_counts_form_at_the_configs_sizes(harness):
    """32 / 128 / 256 virtual devices: the count-vector form against the literal oracle where
    the oracle finishes in seconds, and against the forced literal walk of the same code elsewhere"""
    rng = np.random.default_rng(11)
    W, nl = syn.links_to_weights(syn.topo_dgx1v_like())
    for replicas, size, oracle_ok in ((4, 4, True), (16, 3, True), (32, 3, True), (16, 4, False), (32, 4, False)):
        free = [g for g in range(8) for _ in range(replicas) if rng.random() < 0.5]
        req = [int(x) for x in rng.choice(len(free), size=1)]
        rc, pos, form, total, score = run(harness, W, nl, free, req, size)
        assert (rc, form) == (0, 4) and total <= 330
        if oracle_ok:
            assert orc.subset_besteffort(W, nl, free, req, size) == (0, pos)
        else:
            rc2, pos2, form2, total2, score2 = run(harness, W, nl, free, req, size, FORCE_ENUM, chunk=512)
            assert (rc2, pos2, form2, score2) == (0, pos, 5, score) and total2 > 100 * total


def test_large_requests_beyond_the_literal_form(harness):
    """sizes only the count vectors can answer; properties instead of the oracle: required kept, size right, score no
    worse than a greedy pick, and input-order invariance of the score under a relabelling of the GPUs"""
    rng = np.random.default_rng(13)
    W, nl = syn.links_to_weights(syn.topo_random(8, 1003))
    free = [g for g in range(8) for _ in range(32) if rng.random() < 0.7]
    for size in (8, 16, 32):
        req = sorted(int(x) for x in rng.choice(len(free), size=3, replace=False))
        rc, pos, form, total, score = run(harness, W, nl, free, req, size)
        assert (rc, form, len(pos)) == (0, 4, size) and set(req) <= set(pos) and pos == sorted(pos)
        got = sum(int(W[free[a], free[b]]) for i, a in enumerate(pos) for b in pos[i + 1:] if free[a] != free[b])
        assert got == score
        # every single swap of a chosen non-required position for an unchosen one is no better (local optimality)
        chosen = set(pos)
        for a in [p for p in pos if p not in req][:4]:
            for b in [p for p in range(len(free)) if p not in chosen][::17]:
                alt = sorted((chosen - {a}) | {b})
                s = sum(int(W[free[x], free[y]]) for i, x in enumerate(alt) for y in alt[i + 1:] if free[x] != free[y])
                assert s <= score
    assert len(free) >= 70 and run(harness, W, nl, free, [], 70)[0] == ERR_TOO_LARGE   # more than 64 beyond the required


def test_fuzz_small_requests_any_order_any_weights(harness):
    """hypothesis: tiny requests with arbitrary (also asymmetric-valued) weights, any position order, any required set;
    both forms where they apply must equal the literal oracle"""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=300, deadline=None)
    @given(st.integers(2, 5), st.integers(0, 2**32 - 1), st.lists(st.integers(0, 4), min_size=1, max_size=9),
           st.integers(1, 5), st.integers(0, 3), st.booleans())
    def check(n_gpus, seed, phys, size, n_req, sort_positions):
        rng = np.random.default_rng(seed)
        W = rng.integers(0, 40, size=(n_gpus, n_gpus)).astype(np.int32) * 10     # W[i][j] != W[j][i] is allowed (:314)
        np.fill_diagonal(W, 0)
        nl = np.ones((n_gpus, n_gpus), np.uint8) - np.eye(n_gpus, dtype=np.uint8)
        phys = [p % n_gpus for p in phys]
        if sort_positions:
            phys.sort()
        req = [int(x) for x in rng.choice(len(phys), size=min(n_req, len(phys)), replace=False)]
        want = orc.subset_besteffort(W, nl, phys, req, size, True)
        for mode in (0, FORCE_ENUM):
            rc, pos, form, total, score = run(harness, W, nl, phys, req, size, mode, chunk=int(rng.integers(1, 9)))
            assert (rc, pos) == want[:2], (W.tolist(), phys, req, size, mode)
            if form:
                assert score == want[2]

    check()


def batch_run(L, W, order, rec):
    """tests/csrc/subset_check.cpp subset_batch_check: the batch kernel's search with its lanes run one after the other"""
    nt = len(order)
    w = np.zeros(256, dtype=np.uint16)
    for a in range(nt):
        for b in range(a + 1, nt):
            w[a * 16 + b] = W[order[a], order[b]]
    free = np.ascontiguousarray(rec["free"], dtype=np.uint8)
    must = np.ascontiguousarray(rec["must"], dtype=np.uint8)
    take = np.zeros(16, dtype=np.uint8)
    score = C.c_int(0)
    L.subset_batch_check.restype = None
    L.subset_batch_check(nt, w.ctypes.data_as(C.POINTER(C.c_uint16)), free.ctypes.data_as(C.POINTER(C.c_uint8)),
                         must.ctypes.data_as(C.POINTER(C.c_uint8)), int(rec["size"]), take.ctypes.data_as(C.POINTER(C.c_uint8)),
                         C.byref(score))
    return take.tolist(), score.value


@pytest.mark.parametrize("n_gpus,replicas", [(8, 4), (8, 2), (5, 3), (12, 2)])
def test_batch_walk_matches_the_oracle(harness, n_gpus, replicas):
    """the depth-first walk of the batch kernel (count prefixes as work items, class-ordered, carried partial scores,
    merge on (score, larger prefix)) against the literal oracle on the record's expansion to positions"""
    rng = np.random.default_rng(n_gpus * 10 + replicas)
    order = [int(x) for x in rng.permutation(n_gpus)]          # label -> NVML index
    W, nl = syn.links_to_weights(syn.topo_random(n_gpus, 1000 + n_gpus))
    reqs = syn.random_subset_batch(n_gpus, replicas, 250, seed=n_gpus + replicas, sizes=(1, 2, 3, 4, 5), p_must=0.4)
    checked = 0
    for rec in reqs:
        n_avail = int(rec["free"].sum()) + int(rec["must"].sum())
        size, n_req = int(rec["size"]), int(rec["must"].sum())
        if not (0 < size <= n_avail and n_req <= size):
            continue                                             # guards: the kernel's own code, GPU-tested
        phys_lab, req = syn.subset_record_to_positions(rec, rng)
        rc, pos, score, _ = orc.subset_besteffort(W, nl, [order[g] for g in phys_lab], req, size, True)
        assert rc == 0 and len(pos) == size
        take, got_score = batch_run(harness, W, order, rec)
        assert take == syn.positions_to_take(phys_lab, set(req), pos), (rec, order)
        assert got_score == score
        checked += 1
    assert checked > 120


def batch_run_form(L, W, order, rec, form):
    """subset_batch_check_form: 1 = depth-first walk, 2 = two-halves walk (nodes of up to 8 GPUs, K <= 8)"""
    nt = len(order)
    w = np.zeros(256, dtype=np.uint16)
    for a in range(nt):
        for b in range(a + 1, nt):
            w[a * 16 + b] = W[order[a], order[b]]
    free = np.ascontiguousarray(rec["free"], dtype=np.uint8)
    must = np.ascontiguousarray(rec["must"], dtype=np.uint8)
    take = np.zeros(16, dtype=np.uint8)
    score = C.c_int(0)
    L.subset_batch_check_form.restype = None
    L.subset_batch_check_form(nt, w.ctypes.data_as(C.POINTER(C.c_uint16)), free.ctypes.data_as(C.POINTER(C.c_uint8)),
                              must.ctypes.data_as(C.POINTER(C.c_uint8)), int(rec["size"]), int(form),
                              take.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(score))
    return take.tolist(), score.value


@pytest.mark.parametrize("n_gpus,replicas", [(8, 4), (8, 2), (5, 3), (3, 6)])
def test_two_halves_walk_matches_the_oracle(harness, n_gpus, replicas):
    """the batch kernel's walk on nodes of up to 8 GPUs (per-node table of half-vectors, per-request lists, pairs of a
    prefix and a suffix class) against the literal oracle on the record's expansion to positions"""
    rng = np.random.default_rng(n_gpus * 100 + replicas)
    order = [int(x) for x in rng.permutation(n_gpus)]
    W, nl = syn.links_to_weights(syn.topo_random(n_gpus, 2000 + n_gpus))
    reqs = syn.random_subset_batch(n_gpus, replicas, 250, seed=7 * n_gpus + replicas, sizes=(1, 2, 3, 4, 5), p_must=0.4)
    checked = 0
    for rec in reqs:
        n_avail = int(rec["free"].sum()) + int(rec["must"].sum())
        size, n_req = int(rec["size"]), int(rec["must"].sum())
        if not (0 < size <= n_avail and n_req <= size):
            continue
        phys_lab, req = syn.subset_record_to_positions(rec, rng)
        rc, pos, score, _ = orc.subset_besteffort(W, nl, [order[g] for g in phys_lab], req, size, True)
        assert rc == 0 and len(pos) == size
        take, got_score = batch_run_form(harness, W, order, rec, 2)
        assert take == syn.positions_to_take(phys_lab, set(req), pos), (rec, order)
        assert got_score == score
        checked += 1
    assert checked > 120


@pytest.mark.parametrize("topo", ["random", "uniform", "two_islands", "heavy"])
def test_two_halves_walk_equals_the_depth_first_walk(harness, topo):
    """larger requests than the literal oracle can take: both walks of the batch kernel give the same take[] and score,
    K up to 8, several must-include replicas, ties everywhere on the uniform node"""
    rng = np.random.default_rng({"random": 1, "uniform": 2, "two_islands": 3, "heavy": 4}[topo])
    n_gpus = 8
    if topo == "random":
        W, _ = syn.links_to_weights(syn.topo_random(n_gpus, 77))
    elif topo == "uniform":
        W = np.full((n_gpus, n_gpus), 40, dtype=np.int64)
        np.fill_diagonal(W, 0)
    elif topo == "heavy":                                        # products beyond 16 bits: the high halves are needed
        W = rng.integers(3000, 30000, size=(n_gpus, n_gpus)).astype(np.int64)
        W = np.triu(W, 1) + np.triu(W, 1).T
    else:
        W = np.full((n_gpus, n_gpus), 10, dtype=np.int64)
        W[:4, :4] = 180
        W[4:, 4:] = 180
        np.fill_diagonal(W, 0)
    order = [int(x) for x in rng.permutation(n_gpus)]
    checked = 0
    for _ in range(600):
        rec = np.zeros((), dtype=syn.SUBSET_REQ_DTYPE)
        reps = int(rng.choice([1, 2, 4, 16, 32]))
        rec["free"][:n_gpus] = rng.binomial(reps, float(rng.choice([0.2, 0.5, 0.9])), size=n_gpus)
        n_must = int(rng.choice([0, 0, 1, 2, 5]))
        for _ in range(n_must):
            rec["must"][int(rng.integers(0, n_gpus))] += 1
        K = int(rng.integers(0, 9))
        size = K + int(rec["must"].sum())
        if not (0 < size and int(rec["free"].sum()) >= K):
            continue
        rec["size"] = size
        a = batch_run_form(harness, W, order, rec, 1)
        b = batch_run_form(harness, W, order, rec, 2)
        assert a == b, (rec, order)
        assert sum(b[0]) == K
        checked += 1
    assert checked > 300
