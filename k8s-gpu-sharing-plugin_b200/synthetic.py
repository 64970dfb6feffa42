"""Synthetic nodes, topologies and request streams.

Pure numpy; no GPU, no oracle.  Shared by tests/ and bench.py so that the CUDA path, the
oracle and the reference arm all consume byte-identical seeded inputs.

Link types follow the enum of the reference's NVML binding
(vendor/github.com/NVIDIA/gpu-monitoring-tools/bindings/go/nvml/nvml.go:132-154):
0 Unknown/absent, 1 CrossCPU, 2 SameCPU, 3 HostBridge, 4 MultiSwitch, 5 SingleSwitch,
6 SameBoard, 7..18 = 1..12 NVLinks.  A topology is an int32 array links2[n, n, 2] holding,
per ordered pair, {PCIe-ancestry link, NVLink-count link} as gpuallocator.NewDevices
records them (vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/device.go:49-69).
"""
import itertools
import uuid as _uuid

import numpy as np

PCIE_SCORE = {1: 10, 2: 20, 3: 30, 4: 40, 5: 50, 6: 60}
SCORE_TO_PCIE = {v: k for k, v in PCIE_SCORE.items()}


def nvlink_type(nlinks):
    """number of NVLinks (1..12) -> link-type enum value; 0 -> absent"""
    return 0 if nlinks == 0 else 6 + int(nlinks)


def _sym(n):
    return np.zeros((n, n, 2), dtype=np.int32)


def _set(links2, i, j, pcie, nvl=0):
    links2[i, j, 0] = links2[j, i, 0] = pcie
    links2[i, j, 1] = links2[j, i, 1] = nvl


def topo_zero(n=8):
    """T0: no links at all (every pair scores 0)."""
    return _sym(n)


def topo_uniform(n=8, pcie=1):
    """T1: every pair CrossCPU (score 10, one link)."""
    t = _sym(n)
    for i, j in itertools.combinations(range(n), 2):
        _set(t, i, j, pcie)
    return t


def topo_two_sockets(n=8):
    """T2: an HGX B200 as the reference sees it — NVSwitch links are invisible to
    nvml.GetNVLink (nvml.go:617-658, TODO at :655), so only PCIe ancestry is left:
    SameCPU (20) inside each half, CrossCPU (10) across."""
    t = _sym(n)
    h = n // 2
    for i, j in itertools.combinations(range(n), 2):
        _set(t, i, j, 2 if (i < h) == (j < h) else 1)
    return t


def topo_dgx1v_like():
    """T3: synthetic hybrid cube mesh in the style of a DGX-1V (not a claim about real
    hardware).  PCIe: SingleSwitch for (0,1),(2,3),(4,5),(6,7); SameCPU inside each
    quad; CrossCPU across.  NVLink: 2 links on the eight 'strong' pairs (which include the
    four pairs hard-coded in gpuallocator/staticdgx_policies.go:76), 1 link on eight more."""
    t = _sym(8)
    for i, j in itertools.combinations(range(8), 2):
        if (i, j) in ((0, 1), (2, 3), (4, 5), (6, 7)):
            p = 5
        elif (i < 4) == (j < 4):
            p = 2
        else:
            p = 1
        _set(t, i, j, p)
    for i, j in ((0, 3), (0, 4), (1, 2), (1, 5), (2, 3), (4, 7), (5, 6), (6, 7)):
        t[i, j, 1] = t[j, i, 1] = nvlink_type(2)
    for i, j in ((0, 1), (0, 2), (1, 3), (2, 6), (3, 7), (4, 5), (4, 6), (5, 7)):
        t[i, j, 1] = t[j, i, 1] = nvlink_type(1)
    return t


def topo_random(n, seed, max_nvlinks=6, p_nvlink=0.5):
    """T4..: seeded random symmetric topology over {10..60} + {0,100,..,100*max_nvlinks}."""
    rng = np.random.default_rng(seed)
    t = _sym(n)
    for i, j in itertools.combinations(range(n), 2):
        pcie = int(rng.integers(1, 7))
        nvl = int(rng.integers(1, max_nvlinks + 1)) if rng.random() < p_nvlink else 0
        _set(t, i, j, pcie, nvlink_type(nvl))
    return t


def topo_spiked(n, seed, spike_pair=None, spike=250):
    """D6 hazard vector: weights in {10,20,30} except one pair that is
    worth `spike` (expressed as SingleSwitch 50 + 2 NVLinks = 250)."""
    rng = np.random.default_rng(seed)
    t = _sym(n)
    for i, j in itertools.combinations(range(n), 2):
        _set(t, i, j, int(rng.integers(1, 4)))
    a, b = spike_pair if spike_pair else (n - 2, n - 1)
    assert spike == 250
    _set(t, a, b, 5, nvlink_type(2))
    return t


def spiked_uniform(n, pair, w=1, heavy=5):
    """uniform PCIe weight 10*w everywhere, 10*heavy on one pair"""
    t = topo_uniform(n, w)
    a, b = pair
    t[a, b, 0] = t[b, a, 0] = heavy
    return links_to_weights(t)


# Known answers derived BY HAND from besteffort_policy.go (not from the oracle), n = 14, k = 2 (m = 7 sets, no padding),
# all pairs weigh 10 except one that weighs 50.  Every partition holding the heavy pair ties for the best score, so the
# first one visited is saved (:57).  accum is grown at lengths 0, 1, 2, 4 (cap 1, 2, 4, 8): the array that holds a
# candidate's sets 5..7 is made when its 5th set is appended, so the candidates sharing the first 5 sets share those two
# slots, written in place by each of them in turn (:285, :247).  Under the first prefix (0,1)(2,3)(4,5)(6,7)(8,9) the
# candidates are, in order, (10,11)(12,13) / (10,12)(11,13) / (10,13)(11,12).
#   heavy pair   first best candidate      what bestPartition holds at :65      go_faithful   intended
#   (12,13)      1st under the prefix      ...(10,13)(11,12): no heavy set      [0, 1]        [12, 13]
#   (10,12)      2nd under the prefix      ...(10,13)(11,12): no heavy set      [0, 1]        [10, 12]
#   (10,13)      3rd (last): not rewritten ...(10,13)(11,12)                    [10, 13]      [10, 13]
#   (0,1)        1st                       tail rewritten, set (0,1) intact     [0, 1]        [0, 1]
#   (2,5)        first with (2,5) is the first candidate under prefix (0,1)(2,5)(3,4)(6,7)(8,9); its tail ends as
#                (10,13)(11,12), the heavy set sits in the private part         [2, 5]        [2, 5]
# With 13 devices (one nil pad, m = 7) and the heavy pair (11,12): the first candidate holding it is
# (0,1)(2,3)(4,5)(6,7)(8,9)(10,nil)(11,12) - (10,nil) passes the padding filter (:273-276) - and it is the LAST
# candidate under its prefix, so nothing rewrites it: [11, 12] in both modes.
HAND_DERIVED_D6 = [
    (14, (12, 13), [0, 1], [12, 13]),
    (14, (10, 12), [0, 1], [10, 12]),
    (14, (10, 13), [10, 13], [10, 13]),
    (14, (0, 1), [0, 1], [0, 1]),
    (14, (2, 5), [2, 5], [2, 5]),
    (13, (11, 12), [11, 12], [11, 12]),
]



def standard_topologies():
    """The 14 eight-GPU topologies of config 1: T0..T3 + ten seeded random ones."""
    topos = [("T0_zero", topo_zero(8)), ("T1_uniform", topo_uniform(8)), ("T2_b200_as_ref_sees_it", topo_two_sockets(8)),
             ("T3_dgx1v_like", topo_dgx1v_like())]
    for s in range(10):
        topos.append(("T%d_random" % (4 + s), topo_random(8, 1000 + s)))
    return topos


def links_to_weights(links2):
    """W[i,j] = sum of link scores, nlinks[i,j] = number of links — numpy statement of
    calculateGPUPairScore's switch (besteffort_policy.go:314-353); used to build inputs."""
    score = np.zeros(19, dtype=np.int32)
    for t, s in PCIE_SCORE.items():
        score[t] = s
    for k in range(1, 13):
        score[6 + k] = 100 * k
    W = score[links2[:, :, 0]] + score[links2[:, :, 1]]
    nl = (links2[:, :, 0] != 0).astype(np.uint8) + (links2[:, :, 1] != 0).astype(np.uint8)
    n = links2.shape[0]
    W[np.arange(n), np.arange(n)] = 0
    nl[np.arange(n), np.arange(n)] = 0
    return W.astype(np.int32), nl.astype(np.uint8)


def node_uuids(n, seed=1):
    """UUIDs in NVML-index order, shaped like 'GPU-fef8089b-4820-abfc-e83e-94318197576e'."""
    rng = np.random.default_rng(seed)
    return ["GPU-" + str(_uuid.UUID(bytes=bytes(rng.integers(0, 256, 16, dtype=np.uint8).tolist()))) for _ in range(n)]


def sorted_positions(uuids):
    """NVML indices in byte-wise sorted-UUID order: the order stripReplicas hands to the
    allocator (cmd/nvidia-device-plugin/replica.go:43, server.go:274)."""
    return sorted(range(len(uuids)), key=lambda i: uuids[i].encode())


def exhaustive_requests(order, seed=0, n_pairs=8):
    """Config 1 request set: every non-empty availability mask x k=1..|mask| x required in
    {none, each single available device, n_pairs seeded pairs}.  `order` = NVML indices in
    allocator order.  Yields (avail list, required list, size)."""
    rng = np.random.default_rng(seed)
    n = len(order)
    for mask in range(1, 1 << n):
        avail = [order[p] for p in range(n) if mask >> p & 1]
        pairs = []
        if len(avail) >= 2:
            for _ in range(n_pairs):
                a, b = rng.choice(len(avail), 2, replace=False)
                pairs.append([avail[int(a)], avail[int(b)]])
        for k in range(1, len(avail) + 1):
            yield avail, [], k
            for d in avail:
                yield avail, [d], k
            for pr in pairs:
                yield avail, pr, k


# --------------------------------------------------------------------------------------
# Packed batch format of gpa_besteffort_batch (include/gpualloc.h): one 32-byte record per
# container request.
REQ_DTYPE = np.dtype([("avail", np.uint8, 16), ("req_mask", np.uint16), ("n_avail", np.uint8), ("size", np.uint8),
                      ("flags", np.uint8), ("n_req_extra", np.uint8), ("reserved", np.uint8, 10)])
assert REQ_DTYPE.itemsize == 32
RES_DTYPE = np.dtype([("dev", np.int8, 16)])
FLAG_REQ_FOREIGN = 1  # some required device is not one of the available pointers (D4)


def pack_requests(reqs):
    """[(avail, required, size)] -> REQ_DTYPE array.  required devices are given as NVML
    indices and stored as a bit mask over positions in `avail`."""
    out = np.zeros(len(reqs), dtype=REQ_DTYPE)
    for i, (avail, req, size) in enumerate(reqs):
        out["avail"][i, :len(avail)] = avail
        out["n_avail"][i] = len(avail)
        out["size"][i] = size
        m, fl, extra = 0, 0, 0
        for d in req:
            if d in avail:
                m |= 1 << list(avail).index(d)
            else:
                fl |= FLAG_REQ_FOREIGN
                extra += 1
        if len(set(x for x in req if x >= 0)) != len([x for x in req if x >= 0]):
            raise ValueError("duplicate required device")
        out["req_mask"][i] = m
        out["flags"][i] = fl
        out["n_req_extra"][i] = extra
    return out


def random_request_batch(n_dev, count, seed, order=None, p_required=0.25):
    """Seeded i.i.d. stream of best-effort container requests on an n_dev-GPU node, built
    with vectorised numpy so that millions of requests take seconds.  Number of free GPUs
    uniform in 1..n_dev, which ones uniform among the subsets of that size (kept in
    allocator order), allocation size uniform in 1..|free|, and with probability p_required
    one required device drawn from the free ones."""
    rng = np.random.default_rng(seed)
    order = np.arange(n_dev, dtype=np.uint8) if order is None else np.asarray(order, dtype=np.uint8)
    n_avail = rng.integers(1, n_dev + 1, size=count)
    # random subset of size n_avail: the n_avail smallest of n_dev random keys, then sorted by position
    keys = rng.random((count, n_dev))
    rank = np.argsort(np.argsort(keys, axis=1), axis=1)              # rank of each position's key
    free = rank < n_avail[:, None]                                    # [count, n_dev] by position
    sel = np.where(free, np.arange(n_dev)[None, :], n_dev + np.arange(n_dev)[None, :])
    pos = np.argsort(sel, axis=1, kind="stable")                      # free positions first, ascending
    out = np.zeros(count, dtype=REQ_DTYPE)
    av = order[pos]
    av[np.arange(n_dev)[None, :] >= n_avail[:, None]] = 0
    out["avail"][:, :n_dev] = av
    out["n_avail"] = n_avail
    out["size"] = 1 + (rng.integers(0, 1 << 30, size=count) % n_avail)
    want_req = rng.random(count) < p_required
    which = rng.integers(0, 1 << 30, size=count) % n_avail
    out["req_mask"] = np.where(want_req, 1 << which, 0).astype(np.uint16)
    return out


def unpack_request(rec):
    """REQ_DTYPE record -> (avail list, required list, size); foreign flag -> [-2]"""
    n = int(rec["n_avail"])
    avail = [int(x) for x in rec["avail"][:n]]
    req = [avail[p] for p in range(n) if int(rec["req_mask"]) >> p & 1]
    if int(rec["flags"]) & FLAG_REQ_FOREIGN:
        req = req + [-2] * max(int(rec["n_req_extra"]), 1)
    return avail, req, int(rec["size"])


# --------------------------------------------------------------------------------------
# Mask format of gpa_besteffort_batch_masks: bit b = the device with rank b in the node order.
MASK_REQ_DTYPE = np.dtype([("avail_mask", np.uint16), ("req_mask", np.uint16), ("size", np.uint8), ("flags", np.uint8),
                           ("n_req_extra", np.uint8), ("reserved", np.uint8)])
MASK_RES_DTYPE = np.dtype([("chosen_mask", np.uint16), ("n_nil", np.uint8), ("status", np.int8)])
assert MASK_REQ_DTYPE.itemsize == 8 and MASK_RES_DTYPE.itemsize == 4


MASK8_REQ_DTYPE = np.dtype([("avail_mask", np.uint8), ("req_mask", np.uint8), ("size", np.uint8), ("flags", np.uint8)])
MASK8_RES_DTYPE = np.dtype([("chosen_mask", np.uint8), ("n_nil", np.uint8)])
assert MASK8_REQ_DTYPE.itemsize == 4 and MASK8_RES_DTYPE.itemsize == 2


def masks_to_masks8(masks):
    """MASK_REQ_DTYPE records of a node of up to 8 GPUs -> the compact MASK8_REQ_DTYPE; required devices that the mask
    format only counts (n_req_extra) become the FOREIGN flag: the answer is empty either way"""
    assert (masks["avail_mask"] < 256).all() and (masks["req_mask"] < 256).all()
    out = np.zeros(len(masks), dtype=MASK8_REQ_DTYPE)
    out["avail_mask"], out["req_mask"], out["size"] = masks["avail_mask"], masks["req_mask"], masks["size"]
    out["flags"] = (masks["flags"] & FLAG_REQ_FOREIGN) | (masks["n_req_extra"] > 0)
    return out


def masks8_results_to_masks(res):
    """MASK8_RES_DTYPE -> MASK_RES_DTYPE (n_nil 0x80 = GPA_ERR_INVALID)"""
    out = np.zeros(len(res), dtype=MASK_RES_DTYPE)
    bad = res["n_nil"] == 0x80
    out["chosen_mask"] = np.where(bad, 0, res["chosen_mask"])
    out["n_nil"] = np.where(bad, 0, res["n_nil"])
    out["status"] = np.where(bad, -1, 0)
    return out


def records_to_masks(reqs, order):
    """REQ_DTYPE records whose device lists follow `order` (label -> NVML index) -> MASK_REQ_DTYPE records"""
    rank = np.full(256, 255, dtype=np.int64)
    rank[np.asarray(order, dtype=np.int64)] = np.arange(len(order))
    out = np.zeros(len(reqs), dtype=MASK_REQ_DTYPE)
    avail_mask = np.zeros(len(reqs), dtype=np.int64)
    req_mask = np.zeros(len(reqs), dtype=np.int64)
    n = reqs["n_avail"].astype(np.int64)
    for p in range(16):
        lab = rank[reqs["avail"][:, p]]
        live = p < n
        assert (lab[live] < 255).all(), "device outside the node order"
        bit = np.where(live, 1 << np.where(live, lab, 0), 0)
        avail_mask |= bit
        req_mask |= np.where((reqs["req_mask"].astype(np.int64) >> p) & 1 == 1, bit, 0)
    out["avail_mask"], out["req_mask"] = avail_mask, req_mask
    out["size"], out["flags"], out["n_req_extra"] = reqs["size"], reqs["flags"], reqs["n_req_extra"]
    return out


def mask_results_to_records(res, order):
    """MASK_RES_DTYPE -> RES_DTYPE (devices in node order, then nil pads), for comparison with the record path"""
    order = np.asarray(order, dtype=np.int64)
    out = np.full((len(res), 16), -128, dtype=np.int8)
    fill = np.zeros(len(res), dtype=np.int64)
    chosen = res["chosen_mask"].astype(np.int64)
    for lab in range(len(order)):
        hit = (chosen >> lab) & 1 == 1
        out[np.nonzero(hit)[0], fill[hit]] = order[lab]
        fill += hit
    for _ in range(int(res["n_nil"].max()) if len(res) else 0):
        more = (fill - np.array([bin(int(c)).count("1") for c in chosen])) < res["n_nil"]
        out[np.nonzero(more)[0], fill[more]] = -1
        fill += more
    err = res["status"] != 0
    out[err, 0] = -127
    out[err, 1] = res["status"][err]
    r = np.zeros(len(res), dtype=RES_DTYPE)
    r["dev"] = out
    return r


def random_subset_request(rng, n_gpus, replicas, p_free=0.5, grouped=True, max_size=8, p_required=0.3, max_required=2):
    """One replica-set request (gpa_subset_besteffort): (phys, req_pos, size).  Each of n_gpus * replicas virtual
    devices is free with probability p_free; positions follow the GPU order (the plugin's sorted ids, SURVEY.md 8d
    config 3) or, with grouped=False, a shuffle of it."""
    phys = [g for g in range(n_gpus) for _ in range(replicas) if rng.random() < p_free]
    if not grouped:
        rng.shuffle(phys)
    size = int(rng.integers(1, max_size + 1))
    req = []
    if phys and rng.random() < p_required:
        n_req = int(rng.integers(1, max_required + 1))
        req = [int(x) for x in rng.choice(len(phys), size=min(n_req, len(phys)), replace=False)]
    return [int(x) for x in phys], req, size


# --------------------------------------------------------------------------------------
# Count form of gpa_subset_batch: index g = the GPU with rank g in the node order.
SUBSET_REQ_DTYPE = np.dtype([("free", np.uint8, 16), ("must", np.uint8, 16), ("size", np.int32), ("reserved", np.uint8, 12)])
SUBSET_RES_DTYPE = np.dtype([("take", np.uint8, 16), ("score", np.int32), ("n_out", np.int16), ("status", np.int16)])
assert SUBSET_REQ_DTYPE.itemsize == 48 and SUBSET_RES_DTYPE.itemsize == 24


def random_subset_batch(n_gpus, replicas, count, seed, p_free=(0.1, 0.5, 0.9), sizes=(1, 2, 4, 8), p_must=0.1):
    """Fragmented availability: every replica free with probability p (one p per request, drawn
    from p_free), size drawn from `sizes`, and with probability p_must one free replica is must-include."""
    rng = np.random.default_rng(seed)
    reqs = np.zeros(count, dtype=SUBSET_REQ_DTYPE)
    p = rng.choice(np.asarray(p_free, dtype=np.float64), size=count)
    free = rng.binomial(replicas, p[:, None], size=(count, n_gpus)).astype(np.uint8)
    reqs["size"] = rng.choice(np.asarray(sizes, dtype=np.int32), size=count)
    must_row = np.nonzero(rng.random(count) < p_must)[0]
    for i in must_row:
        g = np.nonzero(free[i])[0]
        if len(g):
            j = int(rng.choice(g))
            free[i, j] -= 1
            reqs["must"][i, j] = 1
    reqs["free"][:, :n_gpus] = free
    return reqs


def subset_record_to_positions(rec, rng=None):
    """count record -> (phys, req_pos) for gpa_subset_besteffort / the oracle: positions grouped by label; WHICH replicas
    of a GPU are the must-include ones does not matter for the counts (any choice gives the same take[])"""
    phys, req = [], []
    for g in range(16):
        n = int(rec["free"][g]) + int(rec["must"][g])
        pick = list(range(n))
        if rng is not None:
            rng.shuffle(pick)
        must_local = set(pick[:int(rec["must"][g])])
        for j in range(n):
            if j in must_local:
                req.append(len(phys))
            phys.append(g)
    return phys, req


def positions_to_take(phys, req_pos, positions):
    take = [0] * 16
    for p in positions:
        if p not in req_pos:
            take[phys[p]] += 1
    return take
