/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under k8s-gpu-sharing-plugin_b200/ may
 * include, link or call this file; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, and only as the checker.
 *
 * CPU restatement of the go-gpuallocator v0.2.1 best-effort policy that the
 * reference vendors at
 *   vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go
 * (called "alloc/besteffort_policy.go" below).  The reference is Go and there is no Go
 * toolchain in this image, so this is a statement-by-statement restatement in C with
 * Go's slice semantics (pointer, len, cap, append growth) emulated explicitly, because
 * the result of the reference depends on them (see "D6" below).
 *
 * PARITY STATUS: **parity unpinned** for this file.  The reference tree holds no test,
 * golden vector or fixture for the best-effort policy (vendor/ has no _test.go), and the
 * Go code cannot be run here.  Authority = the source text cited line by line below.
 * The prioritizeDevices / stripReplicas / getAdditionalXids restatements in
 * ref_replica.c ARE pinned by the reference's own test tables.
 *
 * Two entry points:
 *   ref_besteffort_literal : literal emulation (heap slices, callbacks, aliasing).
 *   ref_besteffort_fast    : same visiting order, no heap; used for larger sizes and as
 *                            the timed CPU baseline.  Cross-checked against _literal in
 *                            tests/test_oracle_besteffort.py.
 *
 * Device model.  A *gpuallocator.Device pointer is modelled as a small integer, the
 * device's NVML index (alloc/device.go:15-19 `Index`).  Pointer equality == integer
 * equality.  The nil pointer is REF_NIL (-1).  A `required` pointer that is not the same
 * pointer as any `available` one (what the reference's own call site produces, because
 * it builds the two lists from two NewDevices() calls, cmd/nvidia-device-plugin/
 * server.go:274,279) is modelled as REF_FOREIGN (-2): equal to nothing.
 * Links: W[i*nt+j] is what calculateGPUPairScore(gpu_i, gpu_j) adds up over
 * gpu_i.Links[gpu_j.Index] (alloc/besteffort_policy.go:314-353) and nlinks[i*nt+j] is
 * len(gpu_i.Links[gpu_j.Index]) (used by the panic at :307-310).
 */
#include <setjmp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define REF_NIL (-1)
#define REF_FOREIGN (-2)

#define REF_OK 0
#define REF_PANIC_ASYMMETRIC 1 /* alloc/besteffort_policy.go:307-310 */
#define REF_PANIC_INTERNAL 2   /* alloc/besteffort_policy.go:240-242 */
#define REF_ERR_NOMEM 3        /* oracle arena exhausted (not a reference behaviour) */
#define REF_ERR_ARG 4

#define REF_MODE_GO_FAITHFUL 0 /* what the compiled Go binary does (slice aliasing, D6) */
#define REF_MODE_INTENDED 1    /* what the author meant: bestPartition is a private copy */

/* ------------------------------------------------------------------ link scores */

/* alloc/besteffort_policy.go:314-353 with the enum of
 * vendor/github.com/NVIDIA/gpu-monitoring-tools/bindings/go/nvml/nvml.go:132-154
 * (P2PLinkUnknown=0, CrossCPU=1 ... SameBoard=6, SingleNVLINKLink=7 ... Twelve=18). */
int ref_link_type_score(int link_type)
{
    switch (link_type) {
    case 1: return 10;   /* P2PLinkCrossCPU */
    case 2: return 20;   /* P2PLinkSameCPU */
    case 3: return 30;   /* P2PLinkHostBridge */
    case 4: return 40;   /* P2PLinkMultiSwitch */
    case 5: return 50;   /* P2PLinkSingleSwitch */
    case 6: return 60;   /* P2PLinkSameBoard */
    case 7: return 100;  /* SingleNVLINKLink */
    case 8: return 200;
    case 9: return 300;
    case 10: return 400;
    case 11: return 500;
    case 12: return 600;
    case 13: return 700;
    case 14: return 800;
    case 15: return 900;
    case 16: return 1000;
    case 17: return 1100;
    case 18: return 1200; /* TwelveNVLINKLinks */
    default: return 0;    /* the Go switch has no default: unknown types add nothing */
    }
}

/* alloc/device.go:49-69: per ordered pair, Links[j] gets the PCIe-ancestry link if it is
 * not P2PLinkUnknown (:56-58) and then the NVLink-count link if it is not Unknown
 * (:64-66).  links2[(i*nt+j)*2 + {0,1}] = {pcie type, nvlink type}, 0 = absent. */
void ref_links_to_weights(int nt, const int32_t *links2, int32_t *W, uint8_t *nlinks)
{
    for (int i = 0; i < nt; i++)
        for (int j = 0; j < nt; j++) {
            int w = 0, c = 0;
            if (i != j)
                for (int s = 0; s < 2; s++) {
                    int t = links2[(i * nt + j) * 2 + s];
                    if (t != 0) {
                        w += ref_link_type_score(t);
                        c++;
                    }
                }
            W[i * nt + j] = w;
            nlinks[i * nt + j] = (uint8_t)c;
        }
}

/* ------------------------------------------------------------------ Go runtime bits */

typedef struct {
    int32_t *ptr;
    int len, cap;
} devslice; /* []*Device */

typedef struct {
    devslice *ptr;
    int len, cap;
} partslice; /* [][]*Device */

typedef struct {
    const int32_t *W;
    const uint8_t *nlinks;
    int nt;
    int mode;
    /* arena */
    char *arena;
    size_t arena_cap, arena_top;
    jmp_buf jb;
    /* Allocate() locals captured by the closure at alloc/besteffort_policy.go:52-61 */
    devslice required;
    partslice bestPartition;
    int bestPartitionNil;
    long long bestScore;
    unsigned long best_epoch;
    /* iterateGPUPartitions locals */
    int padding;
    unsigned long long visited;
} ctx_t;

static void *aalloc(ctx_t *c, size_t n)
{
    n = (n + 7u) & ~(size_t)7u;
    if (c->arena_top + n > c->arena_cap)
        longjmp(c->jb, REF_ERR_NOMEM);
    void *p = c->arena + c->arena_top;
    c->arena_top += n;
    return p;
}

/* Go 1.17 runtime size classes, enough for the sizes met here.  This table is the ONE part of the oracle that cannot be
 * derived from /root/reference (the Go runtime is not vendored): it restates, from the published Go 1.17 source,
 *   src/runtime/sizeclasses.go  `var class_to_size = [_NumSizeClasses]uint16{0, 8, 16, 24, 32, 48, 64, 80, 96, 112,
 *                               128, 144, ... 28672, 32768}` (68 classes, generated by mksizeclasses.go), and
 *   src/runtime/msize.go        roundupsize(): sizes up to _MaxSmallSize (32768) round up to their class through
 *                               size_to_class8 / size_to_class128, larger ones to a multiple of the 8 KiB page.
 * growslice (src/runtime/slice.go, func growslice, "newcap := old.cap; doublecap := newcap + newcap; if cap >
 * doublecap ... else if old.cap < 1024 { newcap = doublecap }", then `capmem = roundupsize(uintptr(newcap) *
 * et.size); newcap = int(capmem / et.size)`) is restated below in go_growcap.  What the D6 analysis needs from all
 * this is only the capacity sequence of a [][]*Device (24-byte headers) that starts empty: append grows it
 * 0 -> 1 (24 B, class 24) -> 2 (48) -> 4 (96) -> 8 (192) -> 16 (384) -> 32 (768): every doubled capacity x 24 is
 * itself a size class, so roundupsize never adds slack and cap is exactly 1, 2, 4, 8, 16, 32.
 * tests/test_oracle_besteffort.py::test_go_growslice_capacity_sequence pins that sequence. */
static const int go_size_classes[] = {8,    16,   24,   32,   48,   64,   80,   96,   112,  128,  144,  160,
                                      176,  192,  208,  224,  240,  256,  288,  320,  352,  384,  416,  448,
                                      480,  512,  576,  640,  704,  768,  896,  1024, 1152, 1280, 1408, 1536,
                                      1792, 2048, 2304, 2688, 3072, 3200, 3456, 4096, 4864, 5376, 6144, 6528,
                                      6784, 6912, 8192, 9472, 9728, 10240, 10880, 12288, 13568, 14336, 16384, 18432,
                                      19072, 20480, 21760, 24576, 27264, 28672, 32768};

static size_t go_roundupsize(size_t size)
{
    for (size_t i = 0; i < sizeof go_size_classes / sizeof go_size_classes[0]; i++)
        if ((size_t)go_size_classes[i] >= size)
            return (size_t)go_size_classes[i];
    return (size + 8191u) & ~(size_t)8191u; /* large objects: page multiple */
}

/* runtime.growslice (Go 1.17, runtime/slice.go:162-240): capacity after append() had to
 * grow.  The reference builds with Go 1.17.9 (versions.mk:21). */
static int go_growcap(int oldcap, int needed, int elemsize)
{
    int newcap = oldcap;
    int doublecap = newcap + newcap;
    if (needed > doublecap) {
        newcap = needed;
    } else if (oldcap < 1024) {
        newcap = doublecap;
    } else {
        while (0 < newcap && newcap < needed)
            newcap += newcap / 4;
        if (newcap <= 0)
            newcap = needed;
    }
    size_t mem = go_roundupsize((size_t)newcap * (size_t)elemsize);
    return (int)(mem / (size_t)elemsize);
}

/* the capacity rule alone, for tests/test_oracle_besteffort.py */
int ref_go_growcap(int oldcap, int needed, int elemsize) { return go_growcap(oldcap, needed, elemsize); }

/* append(s, v) for []*Device (8-byte elements). */
static devslice append_dev(ctx_t *c, devslice s, int32_t v)
{
    if (s.len < s.cap) {
        s.ptr[s.len] = v; /* in place: shares the backing array */
        s.len++;
        return s;
    }
    int nc = go_growcap(s.cap, s.len + 1, 8);
    int32_t *p = (int32_t *)aalloc(c, sizeof(int32_t) * (size_t)nc);
    if (s.len)
        memcpy(p, s.ptr, sizeof(int32_t) * (size_t)s.len);
    p[s.len] = v;
    devslice r = {p, s.len + 1, nc};
    return r;
}

/* append(s, v) for [][]*Device (24-byte slice headers). */
static partslice append_part(ctx_t *c, partslice s, devslice v)
{
    if (s.len < s.cap) {
        s.ptr[s.len] = v; /* in place: this is the write that D6 is about */
        s.len++;
        /* if that array is the one bestPartition points into, `v` is now reachable from
         * it: keep the arena below alive, as Go's GC would */
        if (!c->bestPartitionNil && s.ptr == c->bestPartition.ptr)
            c->best_epoch++;
        return s;
    }
    int nc = go_growcap(s.cap, s.len + 1, 24);
    devslice *p = (devslice *)aalloc(c, sizeof(devslice) * (size_t)nc);
    if (s.len)
        memcpy(p, s.ptr, sizeof(devslice) * (size_t)s.len);
    p[s.len] = v;
    partslice r = {p, s.len + 1, nc};
    return r;
}

/* ------------------------------------------------------------------ helpers :92-151 */

/* alloc/besteffort_policy.go:92-99 */
static int gpuSetContains(devslice gpuSet, int32_t gpu)
{
    for (int i = 0; i < gpuSet.len; i++)
        if (gpuSet.ptr[i] == gpu && gpu != REF_FOREIGN)
            return 1;
    return 0;
}

/* alloc/besteffort_policy.go:102-109 */
static int gpuSetContainsAll(devslice gpuSet, devslice gpuSubset)
{
    for (int i = 0; i < gpuSubset.len; i++)
        if (!gpuSetContains(gpuSet, gpuSubset.ptr[i]))
            return 0;
    return 1;
}

/* alloc/besteffort_policy.go:141-151 */
static int gpuSetCountPadding(devslice gpuSet)
{
    int count = 0;
    for (int i = 0; i < gpuSet.len; i++)
        if (gpuSet.ptr[i] == REF_NIL)
            count++;
    return count;
}

/* alloc/besteffort_policy.go:112-119 */
static int gpuPartitionContainsSetWithAll(partslice gpuPartition, devslice gpuSubset)
{
    for (int i = 0; i < gpuPartition.len; i++)
        if (gpuSetContainsAll(gpuPartition.ptr[i], gpuSubset) && gpuSetCountPadding(gpuPartition.ptr[i]) == 0)
            return 1;
    return 0;
}

/* alloc/besteffort_policy.go:127-138 */
static devslice gpuSetCopyAndAddPadding(ctx_t *c, devslice gpuSet, int size)
{
    devslice gpus = {0, 0, 0};
    if (size <= 0)
        return gpus;
    /* append([]*Device{}, gpuSet...) : one growth to the rounded-up capacity */
    if (gpuSet.len > 0) {
        int nc = go_growcap(0, gpuSet.len, 8);
        gpus.ptr = (int32_t *)aalloc(c, sizeof(int32_t) * (size_t)nc);
        memcpy(gpus.ptr, gpuSet.ptr, sizeof(int32_t) * (size_t)gpuSet.len);
        gpus.len = gpuSet.len;
        gpus.cap = nc;
    }
    while (gpus.len % size != 0)
        gpus = append_dev(c, gpus, REF_NIL);
    return gpus;
}

/* ------------------------------------------------------------------ scores :298-380 */

/* alloc/besteffort_policy.go:298-356 */
static long long calculateGPUPairScore(ctx_t *c, int32_t gpu0, int32_t gpu1)
{
    if (gpu0 == REF_NIL || gpu1 == REF_NIL)
        return 0;
    if (gpu0 == gpu1)
        return 0;
    if (c->nlinks[gpu0 * c->nt + gpu1] != c->nlinks[gpu1 * c->nt + gpu0])
        longjmp(c->jb, REF_PANIC_ASYMMETRIC);
    /* the loop at :314-353 runs over gpu0.Links[gpu1.Index]: gpu0's view of the pair */
    return c->W[gpu0 * c->nt + gpu1];
}

typedef void (*set_cb)(ctx_t *c, void *env, devslice set);

/* alloc/besteffort_policy.go:155-204.  NOTE :180 — one `set` buffer is reused for every
 * callback, so a callee that keeps it must copy it (the callee at :267 does). */
static void iterateGPUSets(ctx_t *c, devslice devices, int size, set_cb callback, void *env)
{
    if (size <= 0)
        return;
    if (size > devices.len)
        return;
    int level = 0;
    int *indices = (int *)aalloc(c, sizeof(int) * (size_t)size);
    memset(indices, 0, sizeof(int) * (size_t)size);
    devslice set;
    set.ptr = (int32_t *)aalloc(c, sizeof(int32_t) * (size_t)size);
    set.len = size;
    set.cap = size;
    for (;;) {
        if (indices[level] == devices.len) {
            if (level == 0)
                break;
            level--;
            indices[level]++;
            continue;
        }
        set.ptr[level] = devices.ptr[indices[level]];
        if (level < size - 1) {
            level++;
            indices[level] = indices[level - 1] + 1;
            continue;
        }
        callback(c, env, set);
        indices[level]++;
    }
}

static void pair_cb(ctx_t *c, void *env, devslice gpus)
{
    *(long long *)env += calculateGPUPairScore(c, gpus.ptr[0], gpus.ptr[1]);
}

/* alloc/besteffort_policy.go:360-368 */
static long long calculateGPUSetScore(ctx_t *c, devslice gpuSet)
{
    long long score = 0;
    size_t mark = c->arena_top;
    iterateGPUSets(c, gpuSet, 2, pair_cb, &score);
    c->arena_top = mark; /* scratch of the pair iterator is garbage now */
    return score;
}

/* alloc/besteffort_policy.go:372-380 */
static long long calculateGPUPartitionScore(ctx_t *c, partslice gpuPartition)
{
    long long score = 0;
    for (int i = 0; i < gpuPartition.len; i++)
        score += calculateGPUSetScore(c, gpuPartition.ptr[i]);
    return score;
}

/* ------------------------------------------------------------------ Allocate closure :52-61 */

static void allocate_callback(ctx_t *c, partslice candidate)
{
    c->visited++;
    if (!gpuPartitionContainsSetWithAll(candidate, c->required))
        return;
    long long score = calculateGPUPartitionScore(c, candidate);
    if (score > c->bestScore || c->bestPartitionNil) {
        if (c->mode == REF_MODE_INTENDED) {
            /* private copy of the outer slice: immune to later in-place appends */
            partslice cp;
            cp.ptr = (devslice *)aalloc(c, sizeof(devslice) * (size_t)candidate.len);
            memcpy(cp.ptr, candidate.ptr, sizeof(devslice) * (size_t)candidate.len);
            cp.len = cp.cap = candidate.len;
            c->bestPartition = cp;
        } else {
            c->bestPartition = candidate; /* :58 — slice header only; array stays shared */
        }
        c->bestPartitionNil = 0;
        c->bestScore = score;
        c->best_epoch++;
    }
}

/* ------------------------------------------------------------------ partitions :209-290 */

typedef struct {
    devslice devices; /* the `devices` of the enclosing iterate() call */
    int size;
    partslice accum;
} iter_env;

static void iterate(ctx_t *c, devslice devices, int size, partslice accum);

/* body of the closure at alloc/besteffort_policy.go:266-286 */
static void iterate_set_cb(ctx_t *c, void *envp, devslice set_in)
{
    iter_env *e = (iter_env *)envp;
    size_t mark = c->arena_top;
    unsigned long epoch = c->best_epoch;

    /* :267 set = append([]*Device{devices[0]}, set...) — a fresh array every time */
    devslice set;
    set.cap = go_growcap(1, 1 + set_in.len, 8);
    set.ptr = (int32_t *)aalloc(c, sizeof(int32_t) * (size_t)set.cap);
    set.ptr[0] = e->devices.ptr[0];
    memcpy(set.ptr + 1, set_in.ptr, sizeof(int32_t) * (size_t)set_in.len);
    set.len = 1 + set_in.len;

    /* :273-276 */
    int p = gpuSetCountPadding(set);
    if (!(p == 0 || p == c->padding)) {
        c->arena_top = mark;
        return;
    }

    /* :278-283 — membership is pointer equality, so one nil in `set` drops every nil */
    devslice remaining = {0, 0, 0};
    for (int i = 0; i < e->devices.len; i++)
        if (!gpuSetContains(set, e->devices.ptr[i]))
            remaining = append_dev(c, remaining, e->devices.ptr[i]);

    /* :285 */
    iterate(c, remaining, e->size, append_part(c, e->accum, set));

    /* Everything allocated below this node is unreachable unless bestPartition was
     * (re)assigned underneath: Go's GC would free it; we release the arena. */
    if (c->best_epoch == epoch)
        c->arena_top = mark;
}

/* the closure `iterate` at alloc/besteffort_policy.go:238-287 */
static void iterate(ctx_t *c, devslice devices, int size, partslice accum)
{
    /* :240-242 */
    if (size > devices.len)
        longjmp(c->jb, REF_PANIC_INTERNAL);
    /* :245-248 */
    if (size == devices.len) {
        allocate_callback(c, append_part(c, accum, devices));
        return;
    }
    /* :266 iterateGPUSets(devices[1:], size-1, ...) */
    iter_env env = {devices, size, accum};
    devslice tail = {devices.ptr + 1, devices.len - 1, devices.cap - 1};
    iterateGPUSets(c, tail, size - 1, iterate_set_cb, &env);
}

/* alloc/besteffort_policy.go:209-290 */
static void iterateGPUPartitions(ctx_t *c, devslice devices, int size)
{
    if (size <= 0)
        return;
    if (size > devices.len)
        return;
    /* :219-224 */
    if (size == 1) {
        for (int i = 0; i < devices.len; i++) {
            devslice one;
            one.ptr = (int32_t *)aalloc(c, sizeof(int32_t));
            one.ptr[0] = devices.ptr[i];
            one.len = one.cap = 1;
            partslice cand;
            cand.ptr = (devslice *)aalloc(c, sizeof(devslice));
            cand.ptr[0] = one;
            cand.len = cand.cap = 1;
            allocate_callback(c, cand);
        }
        return;
    }
    /* :232-233 */
    devices = gpuSetCopyAndAddPadding(c, devices, size);
    c->padding = gpuSetCountPadding(devices);
    /* :289 */
    partslice empty = {0, 0, 0};
    iterate(c, devices, size, empty);
}

/* ------------------------------------------------------------------ Allocate :34-89 */

/*
 * out[0..*n_out) receives the returned set as NVML indices, REF_NIL (-1) where the Go
 * slice holds a nil pointer (defect D5: cmd/nvidia-device-plugin/server.go:299-301 would
 * then dereference nil).  *visited (optional) = number of candidate partitions the
 * callback saw (self-check against the closed-form counts of SURVEY.md §8a A3).
 */
int ref_besteffort_literal(int nt, const int32_t *W, const uint8_t *nlinks, const int32_t *avail, int n_avail,
                           const int32_t *req, int n_req, int size, int mode, int32_t *out, int *n_out,
                           unsigned long long *visited)
{
    *n_out = 0;
    if (visited)
        *visited = 0;
    if (nt <= 0 || n_avail < 0 || n_req < 0)
        return REF_ERR_ARG;
    for (int i = 0; i < n_avail; i++)
        if (avail[i] < 0 || avail[i] >= nt)
            return REF_ERR_ARG;
    for (int i = 0; i < n_req; i++)
        if (req[i] != REF_FOREIGN && (req[i] < 0 || req[i] >= nt))
            return REF_ERR_ARG;

    /* :35-45 */
    if (size <= 0)
        return REF_OK;
    if (n_avail < size)
        return REF_OK;
    if (n_req > size)
        return REF_OK;

    ctx_t *c = (ctx_t *)calloc(1, sizeof *c);
    if (!c)
        return REF_ERR_NOMEM;
    c->W = W;
    c->nlinks = nlinks;
    c->nt = nt;
    c->mode = mode;
    c->arena_cap = (size_t)3 << 30; /* virtual; touched lazily */
    c->arena = (char *)malloc(c->arena_cap);
    if (!c->arena) {
        free(c);
        return REF_ERR_NOMEM;
    }
    int rc = setjmp(c->jb);
    if (rc != 0) {
        free(c->arena);
        free(c);
        return rc;
    }

    devslice available;
    available.ptr = (int32_t *)aalloc(c, sizeof(int32_t) * (size_t)(n_avail ? n_avail : 1));
    memcpy(available.ptr, avail, sizeof(int32_t) * (size_t)n_avail);
    available.len = available.cap = n_avail;
    c->required.ptr = (int32_t *)aalloc(c, sizeof(int32_t) * (size_t)(n_req ? n_req : 1));
    memcpy(c->required.ptr, req, sizeof(int32_t) * (size_t)n_req);
    c->required.len = c->required.cap = n_req;

    /* :50-61 */
    c->bestPartitionNil = 1;
    c->bestPartition.ptr = 0;
    c->bestPartition.len = c->bestPartition.cap = 0;
    c->bestScore = 0;
    iterateGPUPartitions(c, available, size);
    if (visited)
        *visited = c->visited;

    /* :65-74 — read bestPartition as it is NOW (after any in-place overwrites) */
    int nf = 0;
    devslice *filtered = (devslice *)aalloc(c, sizeof(devslice) * (size_t)(c->bestPartition.len + 1));
    for (int i = 0; i < c->bestPartition.len; i++)
        if (gpuSetContainsAll(c->bestPartition.ptr[i], c->required))
            filtered[nf++] = c->bestPartition.ptr[i];
    if (nf == 0) {
        free(c->arena);
        free(c);
        return REF_OK;
    }

    /* :77-85 */
    devslice bestSet = filtered[0];
    long long bestScore = calculateGPUSetScore(c, bestSet);
    for (int i = 1; i < nf; i++) {
        long long score = calculateGPUSetScore(c, filtered[i]);
        if (score > bestScore) {
            bestSet = filtered[i];
            bestScore = score;
        }
    }

    /* :88 */
    for (int i = 0; i < bestSet.len; i++)
        out[i] = bestSet.ptr[i];
    *n_out = bestSet.len;
    free(c->arena);
    free(c);
    return REF_OK;
}

/* =====================================================================================
 * ref_besteffort_fast — the same algorithm with the same visiting order and the same
 * strict-`>` / first-wins rules, but on fixed-size stack arrays (no heap, no callbacks).
 * D6 is emulated directly: the slots of the candidate that Go appends IN PLACE are
 * shared by every candidate below the same "fresh" prefix, so a saved bestPartition sees
 * them overwritten by each later leaf under that prefix.
 * ===================================================================================== */

#define FAST_MAXN 40

typedef struct {
    const int32_t *W;
    const uint8_t *nlinks;
    int nt;
    int size;
    int padding;
    int m; /* sets per partition */
    int fresh; /* number of leading slots that live in a per-prefix private array */
    int mode;
    const int32_t *req;
    int n_req;
    /* current path */
    int32_t cur[FAST_MAXN][FAST_MAXN];
    long long cur_score[FAST_MAXN];
    int cur_asym[FAST_MAXN]; /* set holds a pair whose scoring would panic (:307-310) */
    unsigned long long prefix_id; /* identity of the current length-`fresh` prefix */
    unsigned long long prefix_ctr;
    /* best */
    int have_best;
    long long bestScore;
    int32_t best[FAST_MAXN][FAST_MAXN];
    unsigned long long best_prefix_id;
    unsigned long long visited;
    int panic_code;
} fast_t;

static inline int fast_contains(const int32_t *set, int len, int32_t g)
{
    if (g == REF_FOREIGN)
        return 0;
    for (int i = 0; i < len; i++)
        if (set[i] == g)
            return 1;
    return 0;
}

static inline int fast_contains_all(const fast_t *f, const int32_t *set, int len)
{
    for (int i = 0; i < f->n_req; i++)
        if (!fast_contains(set, len, f->req[i]))
            return 0;
    return 1;
}

static inline int fast_count_pad(const int32_t *set, int len)
{
    int c = 0;
    for (int i = 0; i < len; i++)
        c += (set[i] == REF_NIL);
    return c;
}

/* Score of one set.  Go only scores the sets of candidates that passed the filter at
 * :53, so a pair that would panic is merely recorded in *asym here and acted on by the
 * caller once it knows the candidate is scored at all. */
static long long fast_set_score(const fast_t *f, const int32_t *set, int len, int *asym)
{
    long long s = 0;
    *asym = 0;
    for (int a = 0; a < len; a++)
        for (int b = a + 1; b < len; b++) {
            int32_t g0 = set[a], g1 = set[b];
            if (g0 == REF_NIL || g1 == REF_NIL || g0 == g1)
                continue;
            if (f->nlinks[g0 * f->nt + g1] != f->nlinks[g1 * f->nt + g0]) {
                *asym = 1;
                continue;
            }
            s += f->W[g0 * f->nt + g1];
        }
    return s;
}

/* number of leading candidate slots that sit in an array private to their prefix:
 * accum grows 0 -> 1 -> 2 -> 4 -> 8 -> 16 ... (go_growcap with 24-byte elements), a new
 * array is made when len == cap, i.e. when the (L+1)-th set is appended for
 * L in {0,1,2,4,8,16,...}.  The candidate of length m lives in the array made at the
 * largest such L < m; its first L+1 slots were copied/written at creation. */
static int fast_fresh_slots(int m)
{
    int L = 0;
    if (m > 1) {
        L = 1;
        while (L * 2 < m)
            L *= 2;
    }
    return L + 1;
}

static void fast_leaf(fast_t *f)
{
    f->visited++;
    int size = f->size, m = f->m;
    /* an in-place append below the best's private prefix rewrote the shared tail */
    if (f->mode == REF_MODE_GO_FAITHFUL && f->have_best && f->prefix_id == f->best_prefix_id)
        for (int s = f->fresh; s < m; s++)
            memcpy(f->best[s], f->cur[s], sizeof(int32_t) * (size_t)size);
    /* :53 */
    int ok = 0;
    for (int s = 0; s < m && !ok; s++)
        ok = fast_contains_all(f, f->cur[s], size) && fast_count_pad(f->cur[s], size) == 0;
    if (!ok)
        return;
    /* :56 */
    long long score = 0;
    for (int s = 0; s < m; s++) {
        if (f->cur_asym[s]) {
            f->panic_code = REF_PANIC_ASYMMETRIC;
            return;
        }
        score += f->cur_score[s];
    }
    /* :57-60 */
    if (score > f->bestScore || !f->have_best) {
        for (int s = 0; s < m; s++)
            memcpy(f->best[s], f->cur[s], sizeof(int32_t) * (size_t)size);
        f->have_best = 1;
        f->bestScore = score;
        f->best_prefix_id = f->prefix_id;
    }
}

static void fast_iterate(fast_t *f, const int32_t *devices, int ndev, int depth)
{
    int size = f->size;
    if (size > ndev) {
        f->panic_code = REF_PANIC_INTERNAL;
        return;
    }
    if (depth + 1 == f->fresh)
        f->prefix_id = ++f->prefix_ctr; /* slot depth is the last one of a private prefix */
    if (size == ndev) {
        memcpy(f->cur[depth], devices, sizeof(int32_t) * (size_t)size);
        f->cur_score[depth] = fast_set_score(f, f->cur[depth], size, &f->cur_asym[depth]);
        fast_leaf(f);
        return;
    }
    int idx[FAST_MAXN];
    int32_t remaining[FAST_MAXN];
    int k1 = size - 1;
    /* odometer over (size-1)-combinations of devices[1:], lexicographic (:178-203) */
    for (int i = 0; i < k1; i++)
        idx[i] = 1 + i;
    for (;;) {
        int32_t *set = f->cur[depth];
        set[0] = devices[0];
        for (int i = 0; i < k1; i++)
            set[1 + i] = devices[idx[i]];
        int p = fast_count_pad(set, size);
        if (p == 0 || p == f->padding) {
            int nr = 0;
            for (int i = 0; i < ndev; i++)
                if (!fast_contains(set, size, devices[i]) && !(devices[i] == REF_NIL && p > 0))
                    remaining[nr++] = devices[i];
            f->cur_score[depth] = fast_set_score(f, set, size, &f->cur_asym[depth]);
            if (depth + 1 == f->fresh)
                f->prefix_id = ++f->prefix_ctr;
            fast_iterate(f, remaining, nr, depth + 1);
            if (f->panic_code)
                return;
        }
        /* next combination */
        int i = k1 - 1;
        while (i >= 0 && idx[i] == ndev - k1 + i)
            i--;
        if (i < 0)
            break;
        idx[i]++;
        for (int j = i + 1; j < k1; j++)
            idx[j] = idx[j - 1] + 1;
    }
}

int ref_besteffort_fast(int nt, const int32_t *W, const uint8_t *nlinks, const int32_t *avail, int n_avail,
                        const int32_t *req, int n_req, int size, int mode, int32_t *out, int *n_out,
                        unsigned long long *visited)
{
    *n_out = 0;
    if (visited)
        *visited = 0;
    if (nt <= 0 || n_avail < 0 || n_req < 0 || n_avail + size > FAST_MAXN)
        return REF_ERR_ARG;
    for (int i = 0; i < n_avail; i++)
        if (avail[i] < 0 || avail[i] >= nt)
            return REF_ERR_ARG;
    for (int i = 0; i < n_req; i++)
        if (req[i] != REF_FOREIGN && (req[i] < 0 || req[i] >= nt))
            return REF_ERR_ARG;
    if (size <= 0 || n_avail < size || n_req > size)
        return REF_OK;

    /* on the stack, and only the scalars are initialised: cur[] / best[] / cur_score[] / cur_asym[] are always
     * written before they are read (cur[d] when level d picks its set, best[] when have_best is set), so the
     * restatement neither allocates nor clears 13 KB per request */
    fast_t fast_storage;
    fast_t *f = &fast_storage;
    f->padding = 0;
    f->prefix_id = f->prefix_ctr = 0;
    f->have_best = 0;
    f->bestScore = 0;
    f->best_prefix_id = 0;
    f->visited = 0;
    f->panic_code = 0;
    f->W = W;
    f->nlinks = nlinks;
    f->nt = nt;
    f->size = size;
    f->mode = mode;
    f->req = req;
    f->n_req = n_req;

    if (size == 1) {
        /* :219-224 — every candidate is a fresh one-set partition */
        f->m = 1;
        f->fresh = 1;
        for (int i = 0; i < n_avail; i++) {
            f->cur[0][0] = avail[i];
            f->cur_score[0] = 0;
            f->cur_asym[0] = 0;
            f->prefix_id = ++f->prefix_ctr;
            fast_leaf(f);
        }
    } else {
        int32_t devices[FAST_MAXN];
        int N = n_avail;
        memcpy(devices, avail, sizeof(int32_t) * (size_t)n_avail);
        while (N % size != 0)
            devices[N++] = REF_NIL;
        f->padding = N - n_avail;
        f->m = N / size;
        f->fresh = fast_fresh_slots(f->m);
        fast_iterate(f, devices, N, 0);
    }
    if (visited)
        *visited = f->visited;
    int rc = f->panic_code;
    if (rc == REF_OK && f->have_best) {
        /* :65-85 */
        int have = 0;
        long long bs = 0;
        const int32_t *bset = 0;
        for (int s = 0; s < f->m; s++) {
            if (!fast_contains_all(f, f->best[s], size))
                continue;
            int asym;
            long long sc = fast_set_score(f, f->best[s], size, &asym);
            if (asym) {
                f->panic_code = REF_PANIC_ASYMMETRIC;
                break;
            }
            if (!have || sc > bs) {
                have = 1;
                bs = sc;
                bset = f->best[s];
            }
        }
        rc = f->panic_code;
        if (rc == REF_OK && have) {
            for (int i = 0; i < size; i++)
                out[i] = bset[i];
            *n_out = size;
        }
    }
    return rc;
}

/* closed-form candidate count of SURVEY.md §8a row A3, as a self-check on enumeration:
 * P=0: n!/((k!)^m m!) ; P>0: C(n,k-P) * ((m-1)k)!/((k!)^(m-1)(m-1)!) ; k=1: n. */
unsigned long long ref_partition_count(int n, int k)
{
    if (k <= 0 || n < k)
        return 0;
    if (k == 1)
        return (unsigned long long)n;
    int N = ((n + k - 1) / k) * k, P = N - n, m = N / k;
    /* count = prod over levels of C(r-1, k-1), computed level by level on the unpadded part */
    unsigned long long binom[64][64];
    for (int i = 0; i < 64; i++) {
        binom[i][0] = 1;
        for (int j = 1; j <= i; j++)
            binom[i][j] = binom[i - 1][j - 1] + (j <= i - 1 ? binom[i - 1][j] : 0);
        for (int j = i + 1; j < 64; j++)
            binom[i][j] = 0;
    }
    unsigned long long full = 1; /* partitions of (m')*k reals into m' full sets */
    int mm = (P == 0) ? m : m - 1;
    for (int r = mm * k; r > 0; r -= k)
        full *= binom[r - 1][k - 1];
    if (P == 0)
        return full;
    return binom[n][k - P] * full;
}
