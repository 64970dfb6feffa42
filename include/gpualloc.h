/*
 * gpualloc.h — C ABI of libgpualloc.so, the B200-native GetPreferredAllocation path of
 * the k8s GPU-sharing device plugin.
 *
 * Every entry point names the reference interface it stands in for ("ref:" lines; paths
 * are relative to the reference repository, iktos/k8s-gpu-sharing-plugin).  Conventions:
 *   - plain C types only; the caller owns every buffer it passes in or receives into;
 *   - return value: GPA_OK (0) or a negative GPA_ERR_* code; no exception crosses the
 *     boundary; gpa_strerror() gives the text;
 *   - devices are identified by their NVML index (ref: gpuallocator.Device.Index,
 *     vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/device.go:15-19);
 *   - all entry points are thread-safe; calls on one context are serialised internally
 *     (ref: grpc-go runs every RPC on its own goroutine and the reference holds no locks,
 *     vendor/google.golang.org/grpc/server.go:742-747).
 *   - the library needs a CUDA device for every allocation decision: there is no CPU
 *     fall-back, and gpa_create() fails with GPA_ERR_CUDA when no device is usable.
 */
#ifndef GPUALLOC_H
#define GPUALLOC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPA_OK 0
#define GPA_ERR_INVALID (-1)       /* bad argument (null pointer, index out of range, duplicate device ...) */
#define GPA_ERR_CUDA (-2)          /* CUDA runtime error; see gpa_last_error() */
#define GPA_ERR_NO_TOPOLOGY (-3)   /* gpa_set_topology*() has not been called on this context */
#define GPA_ERR_ASYMMETRIC (-4)    /* where Go panics: "all P2PLinks between 2 GPUs should be bidirectional"
                                      (ref: gpuallocator/besteffort_policy.go:307-310) */
#define GPA_ERR_TOO_LARGE (-5)     /* more than GPA_MAX_AVAILABLE devices, or a search beyond GPA_MAX_PARTITIONS */
#define GPA_ERR_WEIGHT_RANGE (-6)  /* pair weights negative or too large for the 16-bit score table */
#define GPA_ERR_BUFFER (-7)        /* caller's output buffer is too small */
#define GPA_ERR_NVML (-8)          /* NVML could not be loaded / returned an error */
#define GPA_ERR_PANIC (-9)         /* where Go panics in prioritizeDevices: make([]string, len(must), size) with
                                      len > cap (ref: cmd/nvidia-device-plugin/replica.go:121) */
#define GPA_ERR_NO_DEVICES_LEFT (-10) /* "no devices left to allocate" (ref: replica.go:183) */
#define GPA_ERR_MUST_MISSING (-11) /* "device '%s' in mustIncludeDeviceIDs is missing from availableDeviceIDs"
                                      (ref: replica.go:129,136) */
#define GPA_ERR_UNKNOWN_DEVICE (-12) /* "no device with uuid: %v" (ref: gpuallocator/device.go:89-91) /
                                        "unknown device" (ref: cmd/nvidia-device-plugin/server.go:321,330) */
#define GPA_ERR_NOT_IMPLEMENTED (-13) /* "GetPreferredAllocation() not implemented in this case"
                                         (ref: server.go:302-304) */

#define GPA_MAX_TOPOLOGY 64     /* devices in a topology (NVML indices 0..63) */
#define GPA_MAX_AVAILABLE 16    /* devices in one request's `available` list */
#define GPA_MAX_PARTITIONS (1ull << 36) /* candidate partitions one search may visit */

/* mode_flags */
#define GPA_MODE_GO_FAITHFUL 0 /* bit-exact with the compiled Go reference, including the slice aliasing of
                                  bestPartition — the default */
#define GPA_MODE_INTENDED 1    /* what the Go source means: bestPartition keeps the best candidate intact */
#define GPA_MODE_FORCE_ENUM 2  /* gpa_besteffort(): visit every partition (one launch), whatever the size */
#define GPA_MODE_FORCE_DP 4    /* gpa_besteffort(): subset dynamic programme (one launch per layer), whatever the size;
                                  default: the DP for 15-16 devices or more than 1.5 x 10^5 candidate partitions, else enumeration.
                                  Same answers. */

#define GPA_MODE_FORCE_KERNEL 8 /* gpa_besteffort(): answer with a kernel launch even where the installed answer table
                                  (nodes of up to 8 GPUs, devices listed in node order) would answer without one */

#define GPA_NIL (-1)     /* a nil *Device in a returned set (ref: SURVEY.md D5) */
#define GPA_FOREIGN (-2) /* in `required`: a device pointer that is none of the `available` pointers — what the
                            reference's own call site passes (ref: server.go:274,279) */

typedef struct gpa_ctx gpa_ctx;

/* ---- context ------------------------------------------------------------------------ */

/* Binds a context to one CUDA device: stream, pinned mailbox, device scratch, tables.
 * ref: gpuallocator.NewBestEffortPolicy(), gpuallocator/besteffort_policy.go:14-16 (the
 * Go policy is stateless; the state here is what lets a request cost one launch). */
int gpa_create(int cuda_device, gpa_ctx **out);
void gpa_destroy(gpa_ctx *ctx);
const char *gpa_strerror(int code);
/* text of the last CUDA / NVML failure on this context ("" if none) */
const char *gpa_last_error(gpa_ctx *ctx);
/* number of kernels this context has launched so far (bench.py reports it as gpu_launches) */
unsigned long long gpa_launch_count(gpa_ctx *ctx);

/* ---- topology ----------------------------------------------------------------------- */

/* Installs the pair-weight matrix: W[i*n+j] = what calculateGPUPairScore(gpu_i, gpu_j)
 * adds up over gpu_i.Links[gpu_j.Index]; nlinks[i*n+j] = len(gpu_i.Links[gpu_j.Index]).
 * Copy-on-write: requests already running keep the matrix they started with.
 * ref: gpuallocator.NewDevices(), gpuallocator/device.go:33-72 (rebuilt from NVML on every
 * request there; built once and cached here) and calculateGPUPairScore,
 * gpuallocator/besteffort_policy.go:298-356. */
int gpa_set_topology(gpa_ctx *ctx, int n, const int32_t *W, const uint8_t *nlinks);

/* Same, from link types: links2[(i*n+j)*2 + {0,1}] = {PCIe-ancestry link, NVLink-count link}
 * as nvml.P2PLinkType values (0 = absent).  Applies the score switch of
 * gpuallocator/besteffort_policy.go:314-353 (10..60 for PCIe ancestry, 100 per NVLink).
 * ref: nvml.P2PLinkType, vendor/github.com/NVIDIA/gpu-monitoring-tools/bindings/go/nvml/nvml.go:132-154. */
int gpa_set_topology_links(gpa_ctx *ctx, int n, const int32_t *links2);

/* The score switch alone, for one link type.  ref: besteffort_policy.go:314-353. */
int gpa_link_type_score(int link_type);

/* Declares the node-wide order in which requests list their devices: order[label] = NVML index, a permutation of
 * 0..n-1 (default: identity).  Batch requests whose `avail` list is ascending in this order take a faster kernel
 * path (one score table per node instead of one per request); others are still answered, just slower.
 * ref: the reference always lists devices in byte-wise sorted-UUID order, stripReplicas,
 * cmd/nvidia-device-plugin/replica.go:43 -> server.go:274 — pass that order here. */
int gpa_set_device_order(gpa_ctx *ctx, const int32_t *order, int n);

/* Reads the installed matrix back (n*n entries each); either pointer may be NULL. */
int gpa_get_topology(gpa_ctx *ctx, int *n, int32_t *W, uint8_t *nlinks, int cap);

/* The answer table of the installed topology, for nodes of up to 8 GPUs with symmetric link counts: EVERY answer
 * bestEffortPolicy.Allocate can give on that node, computed by a kernel when gpa_set_topology*() /
 * gpa_set_device_order() ran, and kept both on the device (the batch kernels answer with one lookup) and on the host
 * (gpa_besteffort() answers node-ordered requests without a launch).  Entries are indexed
 *     ((tern[avail_mask] + tern[required_mask]) * 8 + size - 1,  tern[x] = sum of 3^b over the bits b of x
 * (masks over the order of gpa_set_device_order(), required a subset of available, size 1..8) and hold
 *     chosen mask | nil pads << 8, 0 = the empty answer.
 * *n_entries = 8 * 3^n; returns GPA_ERR_TOO_LARGE when the topology has no table, GPA_ERR_BUFFER if cap is short
 * (table may be NULL to query the size).
 * ref: bestEffortPolicy.Allocate, gpuallocator/besteffort_policy.go:34-89 - the whole function, tabulated. */
int gpa_get_answer_table(gpa_ctx *ctx, uint16_t *table, int cap, int *n_entries);

/* ---- best-effort policy ------------------------------------------------------------- */

/* One allocation decision.  avail[0..n_avail): NVML indices in the caller's order (the
 * reference passes them sorted by UUID); req[0..n_req): NVML indices or GPA_FOREIGN.
 * out (>= size entries) receives the chosen devices in `available` order, GPA_NIL for a nil
 * pad; *n_out is `size` or 0 ("cannot allocate": Go returns an empty slice, never an error).
 * ref: gpuallocator.Policy.Allocate(available, required []*Device, size int) []*Device,
 * gpuallocator/allocator.go:24-32, as implemented by bestEffortPolicy.Allocate,
 * gpuallocator/besteffort_policy.go:34-89; called at cmd/nvidia-device-plugin/server.go:298. */
int gpa_besteffort(gpa_ctx *ctx, const int32_t *avail, int n_avail, const int32_t *req, int n_req, int size,
                   int mode_flags, int32_t *out, int *n_out);

/* Search statistics of the last gpa_besteffort() / gpa_subset_besteffort() on this context (all optional). */
typedef struct gpa_search_stats {
  unsigned long long partitions; /* candidate partitions of the request (closed form) */
  unsigned long long items;      /* independent work items the search was cut into */
  int cut_depth;                 /* levels resolved from the item index */
  int sets;                      /* m: sets per partition */
  int padding;                   /* P: nil pads */
  int best_score;                /* score of the winning partition, in units of `score_unit`; -1 if none */
  int score_unit;                /* gcd of the pair weights (scores are kept divided by it on the device) */
  int grid, block, smem_bytes;   /* launch shape of the search kernel, 0 if no kernel was needed */
  float kernel_ms;               /* device time of the search kernel(s) (CUDA events), 0 if none */
  int algorithm;                 /* 0 answered by the guards, 1 enumeration, 2 subset dynamic programme,
                                    3 partition table, one-warp kernel (GPA_MODE_FORCE_KERNEL on a node of up to 8 GPUs),
                                    4 / 5 gpa_subset_besteffort(): count vectors / literal combinations,
                                    6 answer table: built by a kernel when the topology was installed, read on the host
                                      (nodes of up to 8 GPUs, devices listed in node order; best_score is not kept: -1) */
} gpa_search_stats;
int gpa_last_search_stats(gpa_ctx *ctx, gpa_search_stats *out);

/* A batch of container requests against the installed topology, one 32-byte record each.
 * ref: the loop over r.ContainerRequests in GetPreferredAllocation,
 * cmd/nvidia-device-plugin/server.go:273-311, each iteration calling Policy.Allocate (:298). */
typedef struct gpa_request {
  uint8_t avail[GPA_MAX_AVAILABLE]; /* NVML indices in the caller's order */
  uint16_t req_mask;                /* bit p set: avail[p] is required */
  uint8_t n_avail;
  uint8_t size;                     /* allocation size, 0..255 */
  uint8_t flags;                    /* GPA_REQ_* */
  uint8_t n_req_extra;              /* required devices beyond those in req_mask (only their count matters:
                                       guard len(required) > size, besteffort_policy.go:43-45) */
  uint8_t reserved[10];
} gpa_request;
#define GPA_REQ_FOREIGN 1 /* some required device is GPA_FOREIGN */

typedef struct gpa_result {
  int8_t dev[GPA_MAX_AVAILABLE]; /* chosen NVML indices in `available` order, GPA_NIL for a pad, then GPA_RES_END */
} gpa_result;
#define GPA_RES_END (-128)  /* unused tail of gpa_result.dev; dev[0] == GPA_RES_END: empty answer */
#define GPA_RES_ERROR (-127) /* dev[0] == GPA_RES_ERROR: dev[1] holds the GPA_ERR_* code of this request */

/* host buffers: copies in, launches, copies out (in overlapping chunks), synchronises.
 * max_avail = largest n_avail in the batch, e.g. the node's GPU count (<= GPA_BATCH_MAX_AVAILABLE); 0 = unknown,
 * the library scans the records for it.  A record with more than max_avail devices may get GPA_ERR_TOO_LARGE. */
int gpa_besteffort_batch(gpa_ctx *ctx, const gpa_request *reqs, int count, int max_avail, int mode_flags,
                         gpa_result *out);
/* device buffers, asynchronous on `stream` (a cudaStream_t; NULL = the legacy default stream): inputs already
 * resident in HBM, nothing copied, nothing synchronised.  The call only enqueues; gpa_set_topology*() /
 * gpa_set_device_order() wait for the device before they replace the tables such a batch may still be reading.
 * Both buffers must be 16-byte aligned (anything from cudaMalloc is), else GPA_ERR_INVALID. */
int gpa_besteffort_batch_device(gpa_ctx *ctx, const gpa_request *d_reqs, int count, int max_avail, int mode_flags,
                                gpa_result *d_out, void *stream);
#define GPA_BATCH_MAX_AVAILABLE 12 /* larger requests go through gpa_besteffort() one by one */

/* The same batch for callers that already hold device MASKS: bit b of a mask = the device with rank b in the order
 * given to gpa_set_device_order() (sorted-UUID order in the reference).  8 bytes in, 4 bytes out per decision instead
 * of 32 + 16, which matters when the records cross PCIe.  Available for nodes of up to 12 GPUs with symmetric link
 * counts (else GPA_ERR_TOO_LARGE).  `available` is by construction in node order, the reference's own order.
 * ref: the same loop over r.ContainerRequests, server.go:273-311 -> Policy.Allocate (:298). */
typedef struct gpa_mask_request {
  uint16_t avail_mask;   /* available devices */
  uint16_t req_mask;     /* required devices; one that is not available makes the answer empty */
  uint8_t size;          /* allocation size */
  uint8_t flags;         /* GPA_REQ_* */
  uint8_t n_req_extra;   /* required devices not representable in req_mask (count only, for the len(required) > size guard) */
  uint8_t reserved;
} gpa_mask_request;
typedef struct gpa_mask_result {
  uint16_t chosen_mask;  /* chosen devices; 0 with n_nil == 0: the empty answer */
  uint8_t n_nil;         /* nil pads in the returned set */
  int8_t status;         /* GPA_OK or a GPA_ERR_* code for this record */
} gpa_mask_result;
int gpa_besteffort_batch_masks(gpa_ctx *ctx, const gpa_mask_request *reqs, int count, int mode_flags, gpa_mask_result *out);
int gpa_besteffort_batch_masks_device(gpa_ctx *ctx, const gpa_mask_request *d_reqs, int count, int mode_flags,
                                      gpa_mask_result *d_out, void *stream);

/* The compact form of the same batch for nodes of up to 8 GPUs - every node the reference runs on: 4 bytes in, 2 bytes
 * out per decision.  The result record is the answer-table entry itself.  Needs the answer table (gpa_get_answer_table():
 * at most 8 GPUs, symmetric link counts), else GPA_ERR_TOO_LARGE.  Both alias modes give the same answers there, so there
 * is no mode argument.
 * ref: the same loop over r.ContainerRequests, server.go:273-311 -> Policy.Allocate (:298). */
typedef struct gpa_mask8_request {
  uint8_t avail_mask;  /* available devices, bit b = rank b in the order of gpa_set_device_order() */
  uint8_t req_mask;    /* required devices; one that is not available makes the answer empty */
  uint8_t size;        /* allocation size */
  uint8_t flags;       /* GPA_REQ_FOREIGN: some required device cannot be expressed in req_mask (the answer is empty:
                          either len(required) > size, besteffort_policy.go:43-45, or no set holds it, :53) */
} gpa_mask8_request;
typedef struct gpa_mask8_result {
  uint8_t chosen_mask; /* chosen devices; 0 with n_nil == 0: the empty answer */
  uint8_t n_nil;       /* nil pads in the returned set; GPA_MASK8_INVALID: a mask names a device the node lacks */
} gpa_mask8_result;
#define GPA_MASK8_INVALID 0x80
int gpa_besteffort_batch_masks8(gpa_ctx *ctx, const gpa_mask8_request *reqs, int count, gpa_mask8_result *out);
int gpa_besteffort_batch_masks8_device(gpa_ctx *ctx, const gpa_mask8_request *d_reqs, int count, gpa_mask8_result *d_out,
                                       void *stream);

/* Host-side helpers for callers that hold id-list records but want the smaller transfer: gpa_pack_mask_requests()
 * turns 32-byte gpa_request records into mask records (labels = the order of gpa_set_device_order()),
 * gpa_unpack_mask_results() turns mask results back into device lists.  Plain host code on `threads` threads
 * (0 = all hardware threads); nothing is decided here - the decisions are still made by the kernels.  A record whose
 * device list is not ascending in the node order, or names a device the node does not have, cannot be a mask: it is
 * packed as avail_mask 0xFFFF, which gpa_besteffort_batch_masks() answers with status GPA_ERR_INVALID, and counted in
 * *n_unpackable (optional) - send those through gpa_besteffort_batch().
 * ref: what the cgo glue does with available []*Device before the call, cmd/nvidia-device-plugin/
 * server.go:274-279 (NewDevicesFrom keeps the order of the sorted uuids). */
int gpa_pack_mask_requests(gpa_ctx *ctx, const gpa_request *reqs, int count, gpa_mask_request *out, int threads,
                           int *n_unpackable);
int gpa_unpack_mask_results(gpa_ctx *ctx, const gpa_mask_result *res, int count, gpa_result *out, int threads);

/* Page-locked host memory for batch buffers: gpa_besteffort_batch() DMAs such buffers directly and stages pageable
 * ones through an internal pinned mailbox.  (No reference counterpart: Go has no device copies to feed.) */
void *gpa_alloc_host(size_t bytes);
void gpa_free_host(void *p);

/* Allocation-free variant of gpa_besteffort_batch().  No counterpart in the reference (Go never touches device
 * memory); it exists for callers that must not allocate when a request arrives: the plugin runs on GPUs it has
 * handed to pods, so (a) their HBM may be full and a cudaMalloc at decision time would fail when the node is
 * busiest, and (b) cudaMalloc / cudaFree synchronise the whole device and would stall every tenant.
 *
 * It needs no gpa_ctx.  The caller passes the topology (as gpa_set_topology takes it), the request records and the
 * result buffer — all in HOST memory, pageable or pinned — plus the stream to run on (NULL = legacy default
 * stream).  Device-side staging (topology tables, request records, result records) lives in ONE file-scope
 * `static __device__` array of GPA_NOALLOC_SCRATCH_BYTES that every GPU loading this library holds until the
 * process exits.  The call makes no cudaMalloc / cudaMallocHost / cudaFree / stream or event creation; it copies
 * in, launches besteffort_batch_kernel, copies out and synchronises `stream` before returning.
 *
 *   - limits: count <= GPA_NOALLOC_MAX_REQUESTS and n_topology <= GPA_MAX_TOPOLOGY, else GPA_ERR_TOO_LARGE with
 *     the reason in errbuf (optional, may be NULL);
 *   - the first call on a device resolves the array's device address (cudaGetSymbolAddress, with that device
 *     current) and loads a constant table; make that first call before any stream capture begins;
 *   - the call leaves cuda_device as the calling thread's current device;
 *   - all calls share the array and it keeps the last call's contents.  Calls are serialised inside one process
 *     and each waits for its stream, so they do not overlap; work that a caller enqueues by other means and that
 *     touches the array concurrently on the GPU would race.  Result records are cleared before each launch. */
#define GPA_NOALLOC_MAX_REQUESTS 4096
#define GPA_NOALLOC_SCRATCH_BYTES (2 * GPA_MAX_TOPOLOGY * GPA_MAX_TOPOLOGY + GPA_MAX_TOPOLOGY * GPA_MAX_TOPOLOGY + \
                                   (32 + 16) * GPA_NOALLOC_MAX_REQUESTS) /* 208 896 bytes per GPU */
int gpa_besteffort_batch_noalloc(int cuda_device, int n_topology, const int32_t *W, const uint8_t *nlinks,
                                 const gpa_request *reqs, int count, int max_avail, int mode_flags, gpa_result *out,
                                 void *stream, char *errbuf, int errbuf_len);

/* ---- replica-set search (no reference counterpart as a whole) ------------------------- */

/* Which `size` of the available SHARED replicas are best connected.  The reference never asks this: with
 * replicas > 1 it answers GetPreferredAllocation with prioritizeDevices (cmd/nvidia-device-plugin/replica.go:96-198,
 * here gpa_prioritize_devices) and ignores links.  This is BASELINE.json north_star's "all k-subsets of the available
 * replica set scored against the link-weight matrix" (SURVEY.md 8d, config 3c), defined from reference pieces:
 *   candidates  every `size`-combination of positions 0..n_avail-1 in lexicographic index order (iterateGPUSets,
 *               gpuallocator/besteffort_policy.go:155-204) that holds all required positions (:102-110);
 *   score       sum over pairs a < b of calculateGPUPairScore(gpu(a), gpu(b)) (:298-356, :360-368): the earlier
 *               position's links, and 0 for two replicas of one GPU (same *Device, :303-305);
 *   choice      the FIRST candidate with the highest score (strict `>`, :55 / :81);
 *   guards      size <= 0, n_avail < size, n_req > size => *n_out = 0 (:35-45).
 * phys[p] = NVML index of the GPU behind available virtual device p (caller order); req_pos[] = positions into
 * phys[] (distinct); out_pos (>= size entries) receives the chosen positions ascending.
 * Two kernels give the same answer: per-GPU COUNT vectors (replicas of one GPU are interchangeable for the score:
 * C(K+G-1, G-1) classes instead of C(n_avail, size) subsets) when the positions of every GPU are contiguous - the
 * plugin's sorted id order, replica.go:41 - and the literal walk otherwise or with GPA_MODE_FORCE_ENUM.
 * Limits (else GPA_ERR_TOO_LARGE): GPA_SUBSET_MAX_AVAILABLE positions, GPA_SUBSET_MAX_GPUS distinct GPUs,
 * size <= GPA_SUBSET_MAX_SIZE; counts: size - n_req <= 64 and <= GPA_SUBSET_MAX_CLASSES count vectors; literal:
 * n_avail - n_req <= 256, size - n_req <= 8, <= GPA_SUBSET_MAX_CANDIDATES combinations.
 * GPA_ERR_ASYMMETRIC where a candidate would make calculateGPUPairScore panic (checked over every pair of distinct GPUs
 * behind the available replicas when size >= 2).  gpa_last_search_stats(): algorithm 4 = counts, 5 = literal;
 * `partitions` = candidates of the form used. */
#define GPA_SUBSET_MAX_AVAILABLE 4096
#define GPA_SUBSET_MAX_GPUS 16
#define GPA_SUBSET_MAX_SIZE 128
#define GPA_SUBSET_MAX_CLASSES (1ull << 27)
#define GPA_SUBSET_MAX_CANDIDATES (1ull << 33)
int gpa_subset_besteffort(gpa_ctx *ctx, const int32_t *phys, int n_avail, const int32_t *req_pos, int n_req, int size,
                          int mode_flags, int32_t *out_pos, int *n_out);

/* Many replica-set decisions per call, in COUNT form: what matters about a request on a node is how many replicas
 * of each GPU are free and how many must be included, because replicas of one GPU are interchangeable for the score.
 * Index g of the arrays is the LABEL of a GPU in the node order given to gpa_set_device_order() (the plugin's sorted
 * id order), i.e. the request is gpa_subset_besteffort() over positions grouped by label.  The answer is how many
 * FURTHER replicas to take from each GPU; the first candidate of the winning class takes the first `take[g]` free
 * replicas of GPU g in id order next to the must-include ones.  One warp per request, one launch per call (nodes of
 * up to 8 GPUs: pairs of precomputed half-vectors; GPA_SUBSET_WALK=dfs in the environment at gpa_create()
 * selects the depth-first walk that larger nodes use, for A/B measurements).  The _device form may be captured into a
 * CUDA graph and may be in flight on several streams at once.
 * Per-record status: GPA_OK, GPA_ERR_INVALID (replicas of a GPU the node does not have), GPA_ERR_ASYMMETRIC,
 * GPA_ERR_TOO_LARGE (size > GPA_SUBSET_MAX_SIZE, size - sum(must) > GPA_SUBSET_BATCH_MAX_EXTRA, or more than
 * GPA_SUBSET_BATCH_MAX_CLASSES count vectors by the estimate C(K + g - 1, g - 1), K = size - sum(must), g = GPUs with free
 * replicas: use the single call for those).
 * Nodes of up to 16 GPUs (else the call returns GPA_ERR_TOO_LARGE).  No reference counterpart (see above). */
#define GPA_SUBSET_BATCH_MAX_EXTRA 15
#define GPA_SUBSET_BATCH_MAX_CLASSES (1u << 20)
typedef struct gpa_subset_request {
  uint8_t free_replicas[16]; /* free replicas of GPU label g that are NOT in must-include */
  uint8_t must_replicas[16]; /* must-include replicas of GPU label g */
  int32_t size;              /* allocation size */
  uint8_t reserved[12];
} gpa_subset_request;
typedef struct gpa_subset_result {
  uint8_t take[16]; /* further replicas to take from GPU label g */
  int32_t score;    /* score of the answer in units of gpa_search_stats.score_unit; -1 for the empty answer */
  int16_t n_out;    /* `size`, or 0: the empty answer (guards at besteffort_policy.go:35-45) or an error status */
  int16_t status;   /* GPA_OK or a GPA_ERR_* code for this record */
} gpa_subset_result;
int gpa_subset_batch(gpa_ctx *ctx, const gpa_subset_request *reqs, int count, gpa_subset_result *out);
int gpa_subset_batch_device(gpa_ctx *ctx, const gpa_subset_request *d_reqs, int count, gpa_subset_result *d_out, void *stream);

/* ---- replica ids (host code: string work, no parallelism) ---------------------------- */

/* Results are written NUL-separated into outbuf; *n_out = number of strings.
 * ref: stripReplicas(), cmd/nvidia-device-plugin/replica.go:32-45. */
int gpa_strip_replicas(const char *const *ids, int n, char *outbuf, int outbuf_len, int *n_out);

/* ref: prioritizeDevices(), cmd/nvidia-device-plugin/replica.go:96-198.
 * Returns GPA_OK, or GPA_ERR_NO_DEVICES_LEFT / GPA_ERR_MUST_MISSING (offending id in errbuf) /
 * GPA_ERR_PANIC.  *non_unique = 1 stands for &NonUniqueError{} (the result is still valid). */
int gpa_prioritize_devices(const char *const *avail, int n_avail, const char *const *must, int n_must, int size,
                           char *outbuf, int outbuf_len, int *n_out, int *non_unique, char *errbuf, int errbuf_len);

/* ref: getAdditionalXids(), cmd/nvidia-device-plugin/nvidia.go:274-294. */
int gpa_get_additional_xids(const char *input, uint64_t *out, int out_cap);

/* --resource-config "orig:new:replicas,..." -> variants, in first-seen order (a later entry for the same `orig`
 * replaces the earlier one, as the Go map does).  replicas == -1 selects auto replicas (replicas = 1, auto = 1); any
 * other value is converted the way Go's uint(int64) does.  Errors carry the reference's texts in errbuf:
 * GPA_ERR_INVALID "an entry must have three parts separated by a semicolon" / "replica must be an integer"
 * (strconv.ParseInt(s, 10, 32): optional sign, decimal digits, 32-bit range).
 * ref: parseResourceConfig(), cmd/nvidia-device-plugin/main.go:171-203; type variant, mig-strategy.go:57-62. */
typedef struct gpa_resource_variant {
  char orig[64];               /* resource the entry applies to ("gpu", or a MIG profile name) */
  char name[64];               /* advertised as "nvidia.com/<name>" */
  unsigned long long replicas; /* variant.Replicas */
  int auto_replicas;           /* variant.AutoReplicas */
} gpa_resource_variant;
int gpa_parse_resource_config(const char *config, gpa_resource_variant *out, int cap, int *n_out, char *errbuf, int errbuf_len);
/* ref: resourceConfiguration.Get(), mig-strategy.go:66-76: an absent resource keeps its name and gets ZERO replicas
 * (so a daemon started without --resource-config advertises no device). */
void gpa_resource_config_get(const gpa_resource_variant *variants, int n, const char *orig, gpa_resource_variant *out);

/* ---- NVML: link matrix and health events ------------------------------------------------ */

typedef struct gpa_nvml gpa_nvml;

/* dlopen()s NVML and calls nvmlInit_v2.  libpath NULL = "libnvidia-ml.so.1".
 * ref: nvml.Init -> dlhandles.nvmlInit, vendor/github.com/NVIDIA/gpu-monitoring-tools/bindings/go/nvml/
 * nvml_dl.go:30-37, bindings.go:94-106. */
int gpa_nvml_open(const char *libpath, gpa_nvml **out);
void gpa_nvml_close(gpa_nvml *nv);
const char *gpa_nvml_last_error(gpa_nvml *nv);
/* ref: nvml.GetDeviceCount, nvml.go:348 ; Device.UUID, nvml.go:380-447 */
int gpa_nvml_device_count(gpa_nvml *nv, int *n);
int gpa_nvml_device_uuid(gpa_nvml *nv, int index, char *buf, int cap);

/* What the daemon needs to know about GPU `index` besides its UUID: TotalMemory in MiB as the plugin computes it (free +
 * used, each already divided down to MiB; it sizes auto replicas, server.go:100-103), the minor number behind
 * /dev/nvidia<minor>, the NUMA node for TopologyInfo (-1 = none) and the PCI bus id.  Any output may be NULL.
 * ref: GpuDeviceManager.Devices + buildDevice, cmd/nvidia-device-plugin/nvidia.go:85-99,162-178; nvml.NewDeviceLite,
 * nvml/nvml.go:449-486; numaNode, nvml.go:314-332; deviceGetMemoryInfo, nvml/bindings.go:491-520. */
int gpa_nvml_device_info(gpa_nvml *nv, int index, unsigned long long *total_memory_mib, int *minor, int *numa_node, char *bus_id,
                         int bus_id_cap);

#define GPA_TOPO_PARITY 0          /* the matrix the reference builds: NVLinks counted only when the far end of a
                                      link IS the peer GPU (so none through an NVSwitch), links 0..11 only */
#define GPA_TOPO_NVSWITCH_AWARE 1  /* extension, no reference counterpart: all 18 links, and two GPUs that both
                                      hang off the switch fabric count min(their switch links) links */
/* Reads the link matrix: links2[(i*n+j)*2 + {0,1}] = {PCIe-ancestry link, NVLink-count link} as nvml.P2PLinkType
 * values, 0 = absent; feed it to gpa_set_topology_links().  links2 may be NULL to query n.  *nvml_calls (optional)
 * = NVML entry points invoked by this build.
 * ref: gpuallocator.NewDevices, vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/device.go:33-72;
 * nvml.GetP2PLink nvml.go:592-615; nvml.GetNVLink nvml.go:617-658. */
int gpa_nvml_read_topology(gpa_nvml *nv, int flags, int *n, int32_t *links2, int cap, unsigned long long *nvml_calls);
/* read + install: what the daemon does once at start-up instead of once per request */
int gpa_set_topology_from_nvml(gpa_ctx *ctx, gpa_nvml *nv, int flags);

typedef struct gpa_health gpa_health;
/* Starts the XID watcher for the given device ids (full-GPU UUIDs).  disable_env = the value of
 * DP_DISABLE_HEALTHCHECKS (NULL = read the environment): "all" or anything containing "xids" disables the
 * watcher, otherwise it is a comma-separated list of additional XIDs to ignore on top of 13, 31, 43, 45, 68.
 * wait_timeout_ms <= 0 = the reference's 5000 ms.
 * ref: checkHealth, cmd/nvidia-device-plugin/nvidia.go:181-269. */
int gpa_health_open(gpa_nvml *nv, const char *const *device_ids, int n, const char *disable_env, int wait_timeout_ms,
                    gpa_health **out);
/* Next device that went unhealthy (the reference's `unhealthy <- d`, nvidia.go:222,248,265): returns 1 and the
 * index into device_ids (+ the XID, 0 for "health checking not supported"), or 0 after timeout_ms. */
int gpa_health_poll(gpa_health *h, int timeout_ms, int *device_index, unsigned long long *xid);
int gpa_health_stats(gpa_health *h, unsigned long long *events_seen, unsigned long long *events_skipped, int *disabled);
void gpa_health_close(gpa_health *h);

/* ---- device-plugin server bodies (host code) ------------------------------------------------ */

#define GPA_LIST_STRATEGY_ENVVAR 0        /* DeviceListStrategyEnvvar */
#define GPA_LIST_STRATEGY_VOLUME_MOUNTS 1 /* DeviceListStrategyVolumeMounts */
#define GPA_ID_STRATEGY_UUID 0            /* DeviceIDStrategyUUID */
#define GPA_ID_STRATEGY_INDEX 1           /* DeviceIDStrategyIndex */
/* fixes: 0 keeps the reference's behaviour bit for bit (SURVEY.md defects D1, D3, D4) */
#define GPA_FIX_HEALTH_PROPAGATION 1 /* D1: an unhealthy GPU marks all of its replicas Unhealthy in ListAndWatch */
#define GPA_FIX_REPLICA_SUFFIX 2     /* D3: best-effort answers carry the "-replica-0" suffix kubelet knows */
#define GPA_FIX_REQUIRED_IDENTITY 4  /* D4: MustIncludeDeviceIDs are matched by uuid, not by pointer */
#define GPA_OPT_REPLICA_TOPOLOGY 8   /* NOT a fix - an option with no reference counterpart: with replicas > 1,
                                        GetPreferredAllocation picks the best-connected replicas (gpa_subset_besteffort
                                        over the byte-wise sorted ids) instead of prioritizeDevices; needs a policy
                                        context.  Error texts stay those of replica.go:129,136,183. */

typedef struct gpa_plugin gpa_plugin;
typedef struct gpa_plugin_config { /* ref: NvidiaDevicePlugin + config.Flags, server.go:56-71 */
  const char *resource_name;       /* resourceName */
  const char *device_list_envvar;  /* deviceListEnvvar, e.g. "NVIDIA_VISIBLE_DEVICES" */
  const char *driver_root;         /* config.Flags.NvidiaDriverRoot */
  uint32_t replicas;               /* replicas */
  int auto_replicas;               /* autoReplicas: TotalMemory / 1000 replicas per device (server.go:100-103) */
  int device_list_strategy;        /* GPA_LIST_STRATEGY_* */
  int device_id_strategy;          /* GPA_ID_STRATEGY_* */
  int pass_device_specs;           /* config.Flags.PassDeviceSpecs */
  int has_policy;                  /* allocatePolicy != nil */
  int fixes;                       /* GPA_FIX_* */
} gpa_plugin_config;
typedef struct gpa_plugin_device { /* ref: Device, cmd/nvidia-device-plugin/nvidia.go:41-46 */
  const char *id;                  /* pluginapi.Device.ID = GPU UUID */
  const char *index;               /* Index */
  const char *const *paths;        /* Paths */
  int n_paths;
  unsigned long long total_memory_mib; /* TotalMemory */
  int numa_node;                   /* Topology.Nodes[0].ID, -1 = none */
  int nvml_index;                  /* gpuallocator.Device.Index */
} gpa_plugin_device;

/* ref: NewNvidiaDevicePlugin + initialize(), server.go:74-116.  ctx may be NULL when has_policy == 0. */
int gpa_plugin_create(gpa_ctx *ctx, const gpa_plugin_config *cfg, const gpa_plugin_device *devs, int n_devs,
                      gpa_plugin **out);
void gpa_plugin_destroy(gpa_plugin *p);
/* ref: GetDevicePluginOptions, server.go:243-248 */
int gpa_plugin_get_options(gpa_plugin *p, int *pre_start_required, int *get_preferred_allocation_available);
int gpa_plugin_device_count(gpa_plugin *p, int *n_cached, int *n_replicas);
/* ref: apiDevices(), server.go:415-421 — the payload of every ListAndWatchResponse, as JSON
 * [{"ID":..,"health":"Healthy"|"Unhealthy","topology":{"nodes":[{"ID":n}]}}] */
int gpa_plugin_list_devices(gpa_plugin *p, char *json, int cap);
/* ref: `case d := <-m.health`, server.go:258-262.  Returns the number of list entries that changed. */
int gpa_plugin_mark_unhealthy(gpa_plugin *p, int cached_index);
/* ref: one iteration of the ContainerRequests loop of GetPreferredAllocation, server.go:273-311.
 * Device ids are written NUL-separated into outbuf; errbuf receives the reference's error string. */
int gpa_plugin_get_preferred_allocation(gpa_plugin *p, const char *const *avail, int n_avail, const char *const *must,
                                        int n_must, int size, char *outbuf, int outbuf_len, int *n_out, char *errbuf,
                                        int errbuf_len);
/* ref: one iteration of the ContainerRequests loop of Allocate, server.go:318-350.  The ContainerAllocateResponse
 * as JSON {"envs":{..},"mounts":[{"container_path","host_path"}],"devices":[{"container_path","host_path","permissions"}]} */
int gpa_plugin_allocate(gpa_plugin *p, const char *const *ids, int n, char *json, int cap, char *errbuf, int errbuf_len);

/* ---- topology validation probes (no reference counterpart: the reference never touches device memory;
 *      `north_star` asks for them to check the NVML-reported topology against measured bandwidth) ---------- */

/* Per-GPU HBM streaming kernel over `bytes` of device memory, `reps` timed launches after warm-up.
 * mode 0: read-only (bytes counted = bytes read); mode 1: copy (bytes counted = read + written). GB/s. */
int gpa_probe_hbm(int cuda_device, unsigned long long bytes, int reps, int mode, double *best_gbs, double *median_gbs,
                  char *errbuf, int errbuf_len);
/* Ceiling of the allocation kernels' inner operation: dependent gathers from a 2^log2_entries-entry uint16 table in
 * shared memory plus an integer add, in 10^9 lookups/s over the whole GPU ("measure this ceiling"). */
int gpa_probe_lut(int cuda_device, int log2_entries, int reps, double *best_glookups_per_s, char *errbuf, int errbuf_len);
/* n x n bandwidth matrix, GB/s one way: row = GPU that pulls, column = GPU whose HBM is read; the diagonal is the
 * local copy.  mode 0: peer-load kernel (LDG.128 on a peer-mapped pointer, over NVLink/NVSwitch when NVML says
 * so); mode 1: cudaMemcpyPeerAsync - one pair at a time.  Modes 2 / 3: the same two ways of moving the bytes with ALL
 * GPUs pulling at once (in round s GPU c reads GPU (c + s) mod n), i.e. the matrix under full load of the switch.
 * can_access (optional, n x n) = cudaDeviceCanAccessPeer. */
int gpa_probe_p2p(int n_devices, unsigned long long bytes, int reps, int mode, double *matrix_gbs, int *can_access,
                  char *errbuf, int errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* GPUALLOC_H */
