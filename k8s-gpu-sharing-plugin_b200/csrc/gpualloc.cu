// libgpualloc.so, CUDA side: context, topology tables, launch logic and the C ABI (include/gpualloc.h) around the
// allocation kernels.
//
// Replaces the reference's
//   bestEffortPolicy.Allocate   vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:34-89
//   iterateGPUPartitions        ... :209-290
//   calculateGPU{Pair,Set,Partition}Score ... :298-380
// and the per-request weight source gpuallocator.NewDevices (gpuallocator/device.go:33-72), which becomes
// device-resident tables installed once by gpa_set_topology*().
//
// One translation unit; the kernels live in the headers included below and are all built from the enumeration core
// gpa_enum.cuh:
//   search_kernel.cuh  besteffort_search_kernel    ONE request, every partition visited, whole machine
//   dp_kernels.cuh     dp_scores/layer/final       ONE large request by subset dynamic programme (a launch per layer)
//   table_kernels.cuh  answer_table_build_kernel     every answer of a node of up to 8 GPUs, built when a topology is installed
//                      besteffort_batch_table_kernel MANY requests in node order on such a node: one lookup per request
//   batch_kernels.cuh  besteffort_batch_fast_kernel  MANY requests in node order, 9-12 GPUs: tile-binned, one lane per request
//                      besteffort_batch_kernel       MANY requests, any order: one warp per request (also the
//                                                    "deferred" pass after the fast kernel)
//   subset_kernels.cuh subset_counts/enum_kernel   ONE request over shared replicas (no reference counterpart)
// The allocation-free entry point and its static device scratch are at the end of this file.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/gpualloc.h"
#include "gpa_enum.cuh"

#include "kernels_common.cuh"
#include "search_kernel.cuh"
#include "dp_kernels.cuh"
#include "batch_kernels.cuh"
#include "table_kernels.cuh"
#include "subset_kernels.cuh"
#include "subset_plan.h"

namespace {

// ---------------------------------------------------------------------------------------
// Host side

#define CUDA_TRY(ctx, expr)                                                                   \
  do {                                                                                        \
    cudaError_t e_ = (expr);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      (ctx)->last_error = std::string(#expr) + ": " + cudaGetErrorString(e_);                 \
      return GPA_ERR_CUDA;                                                                    \
    }                                                                                         \
  } while (0)

struct Topology {
  int n = 0;
  int unit = 1;  // gcd of the weights; device tables hold W / unit
  std::vector<int32_t> W;
  std::vector<uint8_t> nlinks;
  std::vector<uint16_t> wq;
  std::vector<uint8_t> asym;
  bool any_asym = false;
  std::vector<int> order;  // label -> NVML index: the node-wide order requests list their devices in
};

}  // namespace

struct gpa_ctx {
  int device = 0;
  int sm_count = 0;
  int max_smem_optin = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t pipe[4] = {nullptr, nullptr, nullptr, nullptr};  // copy/compute pipeline of the host-buffer batch call
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::mutex mu;
  std::string last_error;
  unsigned long long launches = 0;
  // scratch + mailbox
  unsigned long long *d_key = nullptr;
  unsigned int *d_ticket = nullptr;
  SearchResult *h_res = nullptr;  // pinned + device-mapped mailbox the search kernels answer into
  SearchResult *d_res = nullptr;  // the same memory as the device sees it
  unsigned int seq = 0;           // launch number of the last single-request search
  bool timing_pending = false;    // ev0/ev1 of the last search not read back yet
  // topology (host copy + device tables)
  Topology topo;
  bool has_topo = false;
  uint16_t *d_wq = nullptr;
  uint8_t *d_asym = nullptr;
  uint16_t *d_dp_any = nullptr;  // subset DP tables of gpa_besteffort(): 2^16 and 2^15 uint16
  uint16_t *d_dp_ok = nullptr;
  uint16_t *d_sfull = nullptr;  // fast batch path: set scores over label masks, nullptr = path unavailable
  uint32_t *d_part_words = nullptr;  // partition table of the fast batch path (nodes of up to 8 GPUs)
  uint32_t *d_part_index = nullptr;
  uint32_t *d_part_key = nullptr;  // per partition word: score under the installed weights << 16 | (0xFFFF - index in group)
  std::vector<uint32_t> h_part_words, h_part_index;
  uint64_t *d_sub_binom = nullptr;   // C(n, j) table of the literal replica-set kernel, built on first use
  std::vector<uint64_t> sub_binom;
  SubsetPlan sub_plan;               // plan of the last gpa_subset_besteffort() (kept for its buffers)
  uint32_t *d_sub_pair = nullptr;  // half-vector table of the replica-set batch kernel on nodes of up to 8 GPUs (sub_pair_fill)
  // request counters the warps of that kernel draw from: one per launch in flight (kSubSlots of them, each guarded by an
  // event; a launch that finds its slot still busy, or is being captured into a graph, strides instead)
  static constexpr int kSubSlots = 8;
  unsigned int *d_sub_next = nullptr;
  cudaEvent_t sub_ev[kSubSlots] = {};
  bool sub_ev_used[kSubSlots] = {};
  int sub_slot = 0;
  SubsetBatchTopo sub_bt;          // label-space weights of the replica-set batch kernels, per topology / order
  bool sub_dfs_only = false;       // GPA_SUBSET_WALK=dfs: the depth-first batch kernel on every node (A/B measurements)
  int part_table_nt = 0;             // node size the table was built for, 0 = none
  bool table_attr_set[3] = {false, false, false};
  bool launch_on_host_memory = false;  // set around the launch of a zero-copy host call
  int zero_copy = -1;  // GPA_ZERO_COPY: 1 / 0 = kernels always / never work on the caller's pinned buffers; -1 = per format
  uint16_t *d_ans = nullptr;         // answer table of the installed topology (ans_table.cuh), built by a kernel
  uint8_t *d_ans8 = nullptr;         // its chosen masks alone (what the batch kernels stage into shared memory)
  std::vector<uint16_t> h_ans;       // its host copy: gpa_besteffort() answers node-ordered requests from it
  uint16_t h_tern[256];              // ans_tern() of every 8-bit mask
  FastTopo fast{};
  // batch staging
  gpa_request *d_reqs = nullptr;
  gpa_result *d_out = nullptr;
  size_t batch_cap = 0;
  gpa_request *h_reqs = nullptr;  // pinned staging
  gpa_result *h_out = nullptr;
  size_t h_batch_cap = 0;
  BinomTable binom;
  gpa_search_stats stats{};
};

namespace {

int gcd_int(int a, int b) {
  while (b) {
    int t = a % b;
    a = b;
    b = t;
  }
  return a;
}

int ensure_batch_capacity(gpa_ctx *ctx, size_t count, bool host_staging) {
  if (count > ctx->batch_cap) {
    if (ctx->d_reqs) cudaFree(ctx->d_reqs);
    if (ctx->d_out) cudaFree(ctx->d_out);
    ctx->d_reqs = nullptr;
    ctx->d_out = nullptr;
    ctx->batch_cap = 0;
    size_t cap = count + count / 4 + 1024;
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_reqs, cap * sizeof(gpa_request)));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_out, cap * sizeof(gpa_result)));
    ctx->batch_cap = cap;
  }
  if (host_staging && count > ctx->h_batch_cap) {
    if (ctx->h_reqs) cudaFreeHost(ctx->h_reqs);
    if (ctx->h_out) cudaFreeHost(ctx->h_out);
    ctx->h_reqs = nullptr;
    ctx->h_out = nullptr;
    ctx->h_batch_cap = 0;
    size_t cap = count + count / 4 + 1024;
    CUDA_TRY(ctx, cudaMallocHost(&ctx->h_reqs, cap * sizeof(gpa_request)));
    CUDA_TRY(ctx, cudaMallocHost(&ctx->h_out, cap * sizeof(gpa_result)));
    ctx->h_batch_cap = cap;
  }
  return GPA_OK;
}

int launch_general(gpa_ctx *ctx, const gpa_request *d_reqs, int count, int max_avail, int mode_flags, gpa_result *d_out,
                   cudaStream_t stream, int deferred_only) {
  int n_cap = max_avail < 2 ? 2 : max_avail;
  if (n_cap > GPA_BATCH_MAX_AVAILABLE) n_cap = GPA_BATCH_MAX_AVAILABLE;
  size_t smem = (size_t)kBatchWarps * (((size_t)1 << n_cap) + kMaxDevices * kMaxDevices) * sizeof(uint16_t);
  if (smem > 48 * 1024)
    CUDA_TRY(ctx, cudaFuncSetAttribute(besteffort_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 1;
  CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, besteffort_batch_kernel, kBatchWarps * 32, smem));
  if (occ < 1) occ = 1;
  long long per_block = deferred_only ? (long long)kBatchWarps * 32 : kBatchWarps;
  long long want = ((long long)count + per_block - 1) / per_block;
  long long cap = (long long)ctx->sm_count * occ;  // persistent: one wave, warps stride over the requests
  int grid = (int)(want < cap ? want : cap);
  BatchTopo t{ctx->d_wq, ctx->d_asym, ctx->topo.n};
  besteffort_batch_kernel<<<grid, kBatchWarps * 32, smem, stream>>>(d_reqs, count, n_cap, mode_flags & GPA_MODE_INTENDED, t,
                                                                    d_out, deferred_only);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return GPA_OK;
}

int launch_fast(gpa_ctx *ctx, const void *d_reqs, int count, int mode_flags, void *d_out, cudaStream_t stream, bool masks) {
  auto kern = masks ? besteffort_batch_fast_kernel<true> : besteffort_batch_fast_kernel<false>;
  size_t smem = fast_smem_bytes(ctx->fast.nt);
  if (smem > 48 * 1024) CUDA_TRY(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 1;
  CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kFastThreads, smem));
  if (occ < 1) occ = 1;
  long long want = ((long long)count + kTile - 1) / kTile;
  long long cap = (long long)ctx->sm_count * occ;
  int grid = (int)(want < cap ? want : cap);
  kern<<<grid, kFastThreads, smem, stream>>>(d_reqs, count, mode_flags & GPA_MODE_INTENDED, ctx->fast, d_out);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return GPA_OK;
}

// Nodes of up to 8 GPUs: one lookup in the answer table per request (table_kernels.cuh).  The table holds the same
// answers for both alias modes (no partition of such a node has the 7 sets the aliasing needs); the mode only reaches the
// records the kernel cannot look up and answers through general_request().
template <int kThreads, int kStages, int kRec>
int launch_table_records(gpa_ctx *ctx, const void *d_reqs, int count, int n_cap, int intended, void *d_out, cudaStream_t stream, int slot) {
  auto kern = besteffort_batch_table_kernel<kThreads, kStages, kRec>;
  if (!ctx->table_attr_set[slot]) {
    CUDA_TRY(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)table_smem_bytes(ans_entries(kAnsMaxTopo), kThreads * kRec, kStages)));
    ctx->table_attr_set[slot] = true;
  }
  // persistent: every CTA stages the answer table once and streams its share of the request tiles
  const long long want = ((long long)count + kThreads * kRec - 1) / (kThreads * kRec);
  const long long cap = (long long)ctx->sm_count * (kThreads >= 1024 ? 1 : 2);
  const int grid = (int)(want < cap ? want : cap);
  const BatchTopo t{ctx->d_wq, ctx->d_asym, ctx->topo.n};
  kern<<<grid, kThreads, table_smem_bytes(ctx->fast.ans_entries, kThreads * kRec, kStages), stream>>>(
      static_cast<const gpa_request *>(d_reqs), count, ctx->fast, static_cast<gpa_result *>(d_out), t, n_cap, intended);
  return GPA_OK;
}

int launch_table(gpa_ctx *ctx, const void *d_reqs, int count, int max_avail, int mode_flags, void *d_out, cudaStream_t stream, bool masks) {
  if (masks) {
    if (!ctx->table_attr_set[1]) {
      CUDA_TRY(ctx, cudaFuncSetAttribute(besteffort_masks_table_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)masks_smem_bytes(ans_entries(kAnsMaxTopo))));
      ctx->table_attr_set[1] = true;
    }
    // persistent: every CTA stages the table once
    long long want = ((long long)count + 2 * kMaskThreads - 1) / (2 * kMaskThreads);
    long long cap = (long long)ctx->sm_count * (ctx->launch_on_host_memory ? 1 : 2);  // as in launch_masks8()
    int grid = (int)(want < cap ? want : cap);
    besteffort_masks_table_kernel<<<grid, kMaskThreads, masks_smem_bytes(ctx->fast.ans_entries), stream>>>(
        static_cast<const uint2 *>(d_reqs), count, ctx->fast, static_cast<uint32_t *>(d_out));
  } else {
    // the bulk copies that feed the kernel need 16-byte aligned records (anything from cudaMalloc is 256-byte aligned)
    if ((reinterpret_cast<uintptr_t>(d_reqs) & 15u) || (reinterpret_cast<uintptr_t>(d_out) & 15u)) {
      ctx->last_error = "device request / result buffers must be 16-byte aligned";
      return GPA_ERR_INVALID;
    }
    int n_cap = max_avail < 2 ? 2 : max_avail;  // as the general kernel sees it: a longer list is GPA_ERR_TOO_LARGE
    if (n_cap > GPA_BATCH_MAX_AVAILABLE) n_cap = GPA_BATCH_MAX_AVAILABLE;
    const int intended = mode_flags & GPA_MODE_INTENDED;
    // One 1024-thread CTA per SM, TWO records per thread and tile (2 048 records = 64 KB), a 2-stage ring.  Measured on
    // B200 (4 Mi requests, event-timed median of the kernel).  With a block-wide barrier per tile: one record per thread
    // with 3 / 4 / 5 stages 41.4 / 45.4 / 45.5 us, two 512-thread CTAs per SM with 3 stages each 41.4 us, this shape
    // 39.4 us, 512 threads x 4 records 39.2 us, 256 x 8 records 45.4 us.  With the warps releasing the stages one by one
    // (table_kernels.cuh) the shape stops mattering: 1024 x 2, 1024 x 1 with 3 or 4 stages and 512 x 2 with 4 stages all
    // land on 37.2-37.4 us (512 x 4: 39.1) - 201 MB of requests and answers at 5.4 TB/s, against ~33 us for a plain copy
    // of the same bytes at the measured 6.55 TB/s plus the launch's ramp.
    int rc = launch_table_records<1024, 2, 2>(ctx, d_reqs, count, n_cap, intended, d_out, stream, 0);
    if (rc != GPA_OK) return rc;
  }
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return GPA_OK;
}

int launch_batch(gpa_ctx *ctx, const gpa_request *d_reqs, int count, int max_avail, int mode_flags, gpa_result *d_out,
                 cudaStream_t stream) {
  if (count == 0) return GPA_OK;
  if (!ctx->d_sfull) return launch_general(ctx, d_reqs, count, max_avail, mode_flags, d_out, stream, 0);
  // nodes of up to 8 GPUs: one launch - the table kernel also answers the records it cannot look up
  if (ctx->fast.ans != nullptr && !(mode_flags & GPA_MODE_FORCE_ENUM))
    return launch_table(ctx, d_reqs, count, max_avail, mode_flags, d_out, stream, false);
  // 9-12 GPUs: tile-binned kernel for the records in node order, then the general kernel mops up whatever was deferred
  int rc = launch_fast(ctx, d_reqs, count, mode_flags, d_out, stream, false);
  if (rc != GPA_OK) return rc;
  return launch_general(ctx, d_reqs, count, max_avail, mode_flags, d_out, stream, 1);
}

int launch_masks8(gpa_ctx *ctx, const void *d_reqs, int count, void *d_out, cudaStream_t stream) {
  if (!ctx->table_attr_set[2]) {
    CUDA_TRY(ctx, cudaFuncSetAttribute(besteffort_masks8_table_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)masks_smem_bytes(ans_entries(kAnsMaxTopo))));
    ctx->table_attr_set[2] = true;
  }
  long long want = ((long long)count + 4 * kMaskThreads - 1) / (4 * kMaskThreads);
  // records in the caller's pinned memory: one CTA per SM keeps PCIe as busy as two (tools/pcie_sm_read: 48.0 vs 47.2 GB/s)
  long long cap = (long long)ctx->sm_count * (ctx->launch_on_host_memory ? 1 : 2);
  int grid = (int)(want < cap ? want : cap);
  besteffort_masks8_table_kernel<<<grid, kMaskThreads, masks_smem_bytes(ctx->fast.ans_entries), stream>>>(
      static_cast<const uint32_t *>(d_reqs), count, ctx->fast, static_cast<uint16_t *>(d_out));
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return GPA_OK;
}

int launch_masks(gpa_ctx *ctx, const void *d_reqs, int count, int mode_flags, void *d_out, cudaStream_t stream) {
  if (ctx->fast.ans != nullptr && !(mode_flags & GPA_MODE_FORCE_ENUM)) return launch_table(ctx, d_reqs, count, 0, mode_flags, d_out, stream, true);
  return launch_fast(ctx, d_reqs, count, mode_flags, d_out, stream, true);
}

// Partition words and index of a node of nt <= 8 GPUs on the device (part_table.cuh); rebuilt only when nt changes.
int ensure_partition_table(gpa_ctx *ctx, int nt) {
  if (ctx->part_table_nt == nt) return GPA_OK;
  if (ctx->d_part_words) cudaFree(ctx->d_part_words);
  if (ctx->d_part_index) cudaFree(ctx->d_part_index);
  ctx->d_part_words = ctx->d_part_index = nullptr;
  ctx->part_table_nt = 0;
  std::vector<uint32_t> words, index;
  if (!build_partition_words(nt, words, index)) return GPA_ERR_TOO_LARGE;  // cannot happen for nt <= 8 (6 289 words)
  CUDA_TRY(ctx, cudaMalloc(&ctx->d_part_words, words.size() * sizeof(uint32_t)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->d_part_index, index.size() * sizeof(uint32_t)));
  CUDA_TRY(ctx, cudaMemcpy(ctx->d_part_words, words.data(), words.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(ctx, cudaMemcpy(ctx->d_part_index, index.data(), index.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  ctx->h_part_words = words;
  ctx->h_part_index = index;
  ctx->part_table_nt = nt;
  return GPA_OK;
}

// (Re)builds what the fast batch path needs from the installed topology and device order: the
// label <-> NVML index maps and the score table over all label masks.  The path is available
// for nodes of up to kFastMaxTopo devices whose link counts are symmetric.
// Label-space tables of the replica-set batch kernels (subset_kernels.cuh): the weights every launch passes by value and,
// on nodes of up to 8 GPUs, the half-vector table of the two-halves walk.  Called with the device drained.
// device blob of the two-halves kernel: half-vector table | table of all small count vectors | the empty node's answers for
// K = 0..15 further replicas (gpa_subset_result each) | the 16 requests they are computed from
constexpr int kSubFreeNodeOff = kSubPairBlobBytes + kSubFlatBytes;
constexpr int kSubFreeNodeReqOff = kSubFreeNodeOff + 16 * (int)sizeof(gpa_subset_result);
constexpr int kSubBlobTotal = kSubFreeNodeReqOff + 16 * (int)sizeof(gpa_subset_request);
static_assert(kSubFreeNodeOff % 16 == 0 && kSubFreeNodeReqOff % 16 == 0, "records are read with 8- and 16-byte accesses");

int rebuild_subset_tables(gpa_ctx *ctx) {
  const Topology &T = ctx->topo;
  SubsetBatchTopo &bt = ctx->sub_bt;
  memset(&bt, 0, sizeof bt);
  bt.nt = T.n;
  if (T.n > kSubGroups) return GPA_OK;
  for (int a = 0; a < T.n; a++) {
    const int da = (int)T.order.size() == T.n ? T.order[a] : a;
    for (int b = 0; b < T.n; b++) {
      const int db = (int)T.order.size() == T.n ? T.order[b] : b;
      if (a < b) bt.w[a * kSubGroups + b] = T.wq[da * T.n + db];
      if (a != b && T.asym[da * T.n + db]) bt.asym_row[a] |= (uint16_t)(1u << b);
    }
  }
  // nodes of up to 8 GPUs: the half-vector table of the two-halves walk, then the table of all count vectors of up to 4
  // further replicas (zeros on larger nodes, whose kernel does not read them)
  std::vector<uint32_t> blob((kSubPairBlobBytes + kSubFlatBytes) / 4, 0u);
  if (T.n <= 8) {
    sub_pair_fill(bt.w, blob.data());
    sub_flat_fill(bt.w, blob.data() + kSubPairBlobBytes / 4);
  }
  if (!ctx->d_sub_pair) CUDA_TRY(ctx, cudaMalloc(&ctx->d_sub_pair, kSubBlobTotal));
  if (!ctx->d_sub_next) {
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_sub_next, gpa_ctx::kSubSlots * sizeof(unsigned int)));
    for (auto &ev : ctx->sub_ev) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  }
  CUDA_TRY(ctx, cudaMemcpy(ctx->d_sub_pair, blob.data(), kSubPairBlobBytes + kSubFlatBytes, cudaMemcpyHostToDevice));
  // the empty node's answers: "K replicas, every GPU has K free", K = 0..15, by the kernel itself (no short cut: nullptr)
  gpa_subset_request free_node[16];
  memset(free_node, 0, sizeof free_node);
  for (int K = 0; K < 16; K++) {
    for (int g = 0; g < T.n; g++) free_node[K].free_replicas[g] = (uint8_t)K;
    free_node[K].size = K;
  }
  unsigned char *base = reinterpret_cast<unsigned char *>(ctx->d_sub_pair);
  CUDA_TRY(ctx, cudaMemcpy(base + kSubFreeNodeReqOff, free_node, sizeof free_node, cudaMemcpyHostToDevice));
  const gpa_subset_request *d_free_reqs = reinterpret_cast<const gpa_subset_request *>(base + kSubFreeNodeReqOff);
  gpa_subset_result *d_free_node = reinterpret_cast<gpa_subset_result *>(base + kSubFreeNodeOff);
  if (T.n <= 8)
    subset_batch_pair_kernel<<<2, kSubThreads, 0, ctx->stream>>>(d_free_reqs, 16, bt, reinterpret_cast<const uint4 *>(ctx->d_sub_pair),
                                                                ctx->d_sub_pair + kSubPairBlobBytes / 4, nullptr, nullptr, d_free_node);
  else
    subset_batch_kernel<16><<<2, kSubThreads, 0, ctx->stream>>>(d_free_reqs, 16, bt, nullptr, nullptr, d_free_node);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return GPA_OK;
}

int rebuild_fast(gpa_ctx *ctx) {
  Topology &T = ctx->topo;
  {
    const int rc = rebuild_subset_tables(ctx);
    if (rc != GPA_OK) return rc;
  }
  if (ctx->d_sfull) {
    cudaFree(ctx->d_sfull);
    ctx->d_sfull = nullptr;
  }
  ctx->h_ans.clear();  // no table until the new one is built: gpa_besteffort() then asks a kernel
  ctx->fast.ans = nullptr;
  if (T.n > kFastMaxTopo || T.any_asym) return GPA_OK;
  if ((int)T.order.size() != T.n) {
    T.order.resize(T.n);
    for (int i = 0; i < T.n; i++) T.order[i] = i;
  }
  FastTopo f;
  memset(&f, 0, sizeof f);
  f.nt = T.n;
  memset(f.rank, 0xFF, sizeof f.rank);
  for (int lab = 0; lab < T.n; lab++) {
    f.order[lab] = (uint8_t)T.order[lab];
    f.rank[T.order[lab]] = (uint8_t)lab;
  }
  std::vector<uint16_t> sfull((size_t)1 << T.n, 0);
  for (uint32_t mask = 1; mask < (1u << T.n); mask++) {
    // score of the set without its lowest label, plus that label's pairs (earlier label's view, :314)
    const int a = __builtin_ctz(mask);
    const uint32_t rest = mask & (mask - 1);
    uint32_t sum = sfull[rest];
    for (uint32_t mb = rest; mb; mb &= mb - 1) sum += T.wq[T.order[a] * T.n + T.order[__builtin_ctz(mb)]];
    sfull[mask] = (uint16_t)(sum > 65535u ? 65535u : sum);
  }
  uint16_t *d = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&d, sfull.size() * sizeof(uint16_t)));
  CUDA_TRY(ctx, cudaMemcpy(d, sfull.data(), sfull.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
  f.s_full = d;
  ctx->d_sfull = d;
  if (T.n <= kPartTableMaxTopo) {
    int rc = ensure_partition_table(ctx, T.n);
    if (rc != GPA_OK) return rc;
    f.part_words = ctx->d_part_words;
    f.part_index = ctx->d_part_index;
    // partition keys under these weights: score << 16 | (0xFFFF - index in the (mask, size) group).  The score is the sum
    // of the set scores of the word's bytes and never exceeds 65 535: the sets of a partition are disjoint, so its pairs
    // are among the node's 120 at most, which gpa_set_topology bounds.
    std::vector<uint32_t> pkey;
    build_partition_keys(T.n, ctx->h_part_words, ctx->h_part_index, sfull.data(), pkey);
    if (ctx->d_part_key) cudaFree(ctx->d_part_key);
    ctx->d_part_key = nullptr;
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_part_key, (pkey.size() + 4) * sizeof(uint32_t)));
    CUDA_TRY(ctx, cudaMemcpy(ctx->d_part_key, pkey.data(), pkey.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    f.part_key = ctx->d_part_key;
    // every answer of this node, by a kernel (ans_table.cuh); the host keeps a copy for single decisions
    const uint32_t n_ent = ans_entries(T.n);
    const size_t ans_bytes = ((size_t)n_ent * sizeof(uint16_t) + 15) & ~(size_t)15;
    if (ctx->d_ans) cudaFree(ctx->d_ans);
    if (ctx->d_ans8) cudaFree(ctx->d_ans8);
    ctx->d_ans = nullptr;
    ctx->d_ans8 = nullptr;
    ctx->h_ans.clear();
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_ans, ans_bytes));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_ans8, table_ans_bytes(n_ent)));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_ans, 0, ans_bytes, ctx->stream));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_ans8, 0, table_ans_bytes(n_ent), ctx->stream));
    answer_table_build_kernel<<<(n_ent + 255) / 256, 256, 0, ctx->stream>>>(f, ctx->d_ans, ctx->d_ans8, n_ent, 0);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    std::vector<uint16_t> h(n_ent);
    CUDA_TRY(ctx, cudaMemcpyAsync(h.data(), ctx->d_ans, (size_t)n_ent * sizeof(uint16_t), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->h_ans.swap(h);
    f.ans = ctx->d_ans;
    f.ans8 = ctx->d_ans8;
    f.ans_entries = n_ent;
  }
  ctx->fast = f;
  return GPA_OK;
}

// Waits for the answer of a single-request kernel where it lands: the kernel's last thread writes the record into the
// mapped mailbox and then its launch number.  The stream is polled now and then so that a failed launch ends the wait.
int wait_mailbox(gpa_ctx *ctx, unsigned int want) {
  volatile unsigned int *seq = &ctx->h_res->seq;
  unsigned spins = 0;
  while (*seq != want) {
    if ((++spins & 0x3FFu) == 0) {
      cudaError_t q = cudaStreamQuery(ctx->stream);
      if (q != cudaErrorNotReady) {
        if (q != cudaSuccess) {
          ctx->last_error = std::string("search kernel: ") + cudaGetErrorString(q);
          return GPA_ERR_CUDA;
        }
        if (*seq != want) {  // stream drained without an answer: cannot happen unless the kernel was not run
          ctx->last_error = "search kernel finished without publishing its answer";
          return GPA_ERR_CUDA;
        }
      }
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  return GPA_OK;
}

unsigned int next_seq(gpa_ctx *ctx) {
  unsigned int s = ++ctx->seq;
  if (s == 0) s = ++ctx->seq;  // 0 is the mailbox's initial value
  return s;
}

}  // namespace

namespace {
template <typename F>
void parallel_ranges(long long count, int threads, F &&fn) {
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  if (count < (1 << 16)) threads = 1;
  if (threads == 1) {
    fn(0, count);
    return;
  }
  std::vector<std::thread> pool;
  const long long per = (count + threads - 1) / threads;
  for (int t = 0; t < threads; t++) {
    const long long b = per * t < count ? per * t : count, e = per * (t + 1) < count ? per * (t + 1) : count;
    if (b < e) pool.emplace_back([&fn, b, e] { fn(b, e); });
  }
  for (auto &th : pool) th.join();
}

// Host buffers in, host buffers out, for every batch format: the records ride a ring of four streams in chunks, so that
// while chunk i is in its kernel, chunk i+1 is on the H2D copy engine and chunk i-1 on the D2H one.  Pinned caller
// buffers (gpa_alloc_host / cudaHostRegister) are DMA'd directly; pageable ones are staged through the context's pinned
// buffers so that every copy stays asynchronous.  Whatever fails, the streams are drained before the call returns: no
// copy may still be writing into the caller's buffer afterwards.
// zero_copy_ok (the mask formats) and pinned buffers: nothing is staged at all, the kernel reads the records from, and
// writes the answers to, the caller's pinned memory across PCIe in one launch - measured 6-13 % faster than the staged
// pipeline for 4- and 8-byte records (and than copying the records in while the kernel writes the answers straight out:
// 10.3 vs 11.4 G decisions/s, and 10.7 with all the copies back to back on one stream and the kernels chained to it by
// events) and 4 % slower for the 32-byte ones, which therefore stay staged.  GPA_ZERO_COPY=0 / 1 in the environment
// forces it off / on for every format.
template <typename Launch>
int host_batch(gpa_ctx *ctx, const void *reqs, void *out, int count, size_t in_size, size_t out_size, int chunk, bool zero_copy_ok,
               Launch &&launch) {
  cudaPointerAttributes pa_in{}, pa_out{};
  const bool pinned = cudaPointerGetAttributes(&pa_in, reqs) == cudaSuccess && pa_in.type == cudaMemoryTypeHost &&
                      cudaPointerGetAttributes(&pa_out, out) == cudaSuccess && pa_out.type == cudaMemoryTypeHost;
  cudaGetLastError();
  if (pinned && (ctx->zero_copy == 1 || (ctx->zero_copy < 0 && zero_copy_ok))) {
    void *d_in = nullptr, *d_out = nullptr;
    if (cudaHostGetDevicePointer(&d_in, const_cast<void *>(reqs), 0) == cudaSuccess && cudaHostGetDevicePointer(&d_out, out, 0) == cudaSuccess) {
      ctx->launch_on_host_memory = true;
      int rc = launch(d_in, count, d_out, ctx->pipe[0]);
      ctx->launch_on_host_memory = false;
      cudaError_t e = cudaStreamSynchronize(ctx->pipe[0]);
      if (e != cudaSuccess && rc == GPA_OK) {
        ctx->last_error = std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e);
        rc = GPA_ERR_CUDA;
      }
      return rc;
    }
    cudaGetLastError();
  }
  // the staging buffers are sized in 32 / 16-byte record units
  const size_t units = std::max(((size_t)count * in_size + sizeof(gpa_request) - 1) / sizeof(gpa_request),
                                ((size_t)count * out_size + sizeof(gpa_result) - 1) / sizeof(gpa_result)) + 1;
  int rc = ensure_batch_capacity(ctx, units, !pinned);
  if (rc != GPA_OK) return rc;
  if (!pinned) memcpy(ctx->h_reqs, reqs, (size_t)count * in_size);
  const char *src = static_cast<const char *>(pinned ? reqs : (const void *)ctx->h_reqs);
  char *dst = static_cast<char *>(pinned ? out : (void *)ctx->h_out);
  char *d_in = reinterpret_cast<char *>(ctx->d_reqs), *d_out = reinterpret_cast<char *>(ctx->d_out);
  int c = 0;
  for (int off = 0; off < count && rc == GPA_OK; off += chunk, c++) {
    const int cnt = count - off < chunk ? count - off : chunk;
    cudaStream_t st = ctx->pipe[c & 3];
    cudaError_t e = cudaMemcpyAsync(d_in + (size_t)off * in_size, src + (size_t)off * in_size, (size_t)cnt * in_size, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
      rc = launch(d_in + (size_t)off * in_size, cnt, d_out + (size_t)off * out_size, st);
      if (rc != GPA_OK) break;
      e = cudaMemcpyAsync(dst + (size_t)off * out_size, d_out + (size_t)off * out_size, (size_t)cnt * out_size, cudaMemcpyDeviceToHost, st);
    }
    if (e != cudaSuccess) {
      ctx->last_error = std::string("cudaMemcpyAsync: ") + cudaGetErrorString(e);
      rc = GPA_ERR_CUDA;
    }
  }
  for (auto &ps : ctx->pipe) {
    cudaError_t e = cudaStreamSynchronize(ps);
    if (e != cudaSuccess && rc == GPA_OK) {
      ctx->last_error = std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e);
      rc = GPA_ERR_CUDA;
    }
  }
  if (rc != GPA_OK) return rc;
  if (!pinned) memcpy(out, ctx->h_out, (size_t)count * out_size);
  return GPA_OK;
}

}  // namespace

extern "C" {

const char *gpa_strerror(int code) {
  switch (code) {
    case GPA_OK: return "ok";
    case GPA_ERR_INVALID: return "invalid argument";
    case GPA_ERR_CUDA: return "CUDA error";
    case GPA_ERR_NO_TOPOLOGY: return "no topology installed";
    case GPA_ERR_ASYMMETRIC:
      return "Internal error in bestEffort GPU allocator: all P2PLinks between 2 GPUs should be bidirectional";
    case GPA_ERR_TOO_LARGE: return "request too large";
    case GPA_ERR_WEIGHT_RANGE: return "pair weights out of range for the score table";
    case GPA_ERR_BUFFER: return "output buffer too small";
    case GPA_ERR_NVML: return "NVML error";
    case GPA_ERR_PANIC: return "runtime error: makeslice: cap out of range";
    case GPA_ERR_NO_DEVICES_LEFT: return "no devices left to allocate";
    case GPA_ERR_MUST_MISSING: return "device in mustIncludeDeviceIDs is missing from availableDeviceIDs";
    case GPA_ERR_UNKNOWN_DEVICE: return "unknown device";
    case GPA_ERR_NOT_IMPLEMENTED: return "GetPreferredAllocation() not implemented in this case";
    default: return "unknown error";
  }
}

int gpa_create(int cuda_device, gpa_ctx **out) {
  if (!out) return GPA_ERR_INVALID;
  *out = nullptr;
  gpa_ctx *ctx = new gpa_ctx();
  ctx->device = cuda_device;
  auto fail = [&](int code) {
    // keep the text reachable for the caller's log
    fprintf(stderr, "gpa_create: %s\n", ctx->last_error.c_str());
    gpa_destroy(ctx);
    return code;
  };
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || cuda_device < 0 || cuda_device >= ndev) {
    ctx->last_error = e != cudaSuccess ? cudaGetErrorString(e) : "no such CUDA device";
    return fail(GPA_ERR_CUDA);
  }
#define CREATE_TRY(expr)                                              \
  do {                                                                \
    cudaError_t e_ = (expr);                                          \
    if (e_ != cudaSuccess) {                                          \
      ctx->last_error = std::string(#expr) + ": " + cudaGetErrorString(e_); \
      return fail(GPA_ERR_CUDA);                                      \
    }                                                                 \
  } while (0)
  CREATE_TRY(cudaSetDevice(cuda_device));
  CREATE_TRY(cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, cuda_device));
  CREATE_TRY(cudaDeviceGetAttribute(&ctx->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, cuda_device));
  CREATE_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  for (auto &ps : ctx->pipe) CREATE_TRY(cudaStreamCreateWithFlags(&ps, cudaStreamNonBlocking));
  CREATE_TRY(cudaEventCreate(&ctx->ev0));
  CREATE_TRY(cudaEventCreate(&ctx->ev1));
  CREATE_TRY(cudaMalloc(&ctx->d_key, 2 * sizeof(unsigned long long)));  // [0] reduction key, [1] item cursor
  CREATE_TRY(cudaMalloc(&ctx->d_ticket, sizeof(unsigned int)));
  CREATE_TRY(cudaHostAlloc(&ctx->h_res, sizeof(SearchResult), cudaHostAllocMapped));
  memset(ctx->h_res, 0, sizeof(SearchResult));
  CREATE_TRY(cudaHostGetDevicePointer(&ctx->d_res, ctx->h_res, 0));
  CREATE_TRY(cudaMalloc(&ctx->d_dp_any, ((size_t)1 << kMaxDevices) * sizeof(uint16_t)));
  CREATE_TRY(cudaMalloc(&ctx->d_dp_ok, ((size_t)1 << (kMaxDevices - 1)) * sizeof(uint16_t)));
  CREATE_TRY(cudaMemset(ctx->d_key, 0, 2 * sizeof(unsigned long long)));
  CREATE_TRY(cudaMemset(ctx->d_ticket, 0, sizeof(unsigned int)));
  {
    const char *zc = getenv("GPA_ZERO_COPY");
    ctx->zero_copy = (zc && *zc) ? (*zc == '1' ? 1 : 0) : -1;
    const char *walk = getenv("GPA_SUBSET_WALK");
    ctx->sub_dfs_only = walk && strcmp(walk, "dfs") == 0;
  }
  fill_binom(ctx->binom);
  for (uint32_t x = 0; x < 256; x++) ctx->h_tern[x] = (uint16_t)ans_tern(x);
  CREATE_TRY(cudaMemcpyToSymbol(c_binom, &ctx->binom, sizeof(BinomTable)));
  {
    uint32_t parts[13][13];
    memset(parts, 0, sizeof parts);
    for (int n = 2; n <= 12; n++)
      for (int k = 2; k <= n; k++) {
        uint64_t c = partition_count(make_geom(n, k, 0), ctx->binom);
        parts[n][k] = c > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)c;
      }
    CREATE_TRY(cudaMemcpyToSymbol(c_parts, parts, sizeof parts));
    // bins by descending work: the 66 real shapes (2 <= k <= n <= 12) come first, so 2 * bin + 1 < kShapeBins
    std::vector<std::pair<uint32_t, int>> shapes;
    for (int n = 0; n <= 12; n++)
      for (int k = 0; k <= 12; k++) shapes.push_back({parts[n][k], n * 13 + k});
    std::sort(shapes.begin(), shapes.end(), [](const std::pair<uint32_t, int> &a, const std::pair<uint32_t, int> &b) {
      return a.first != b.first ? a.first > b.first : a.second < b.second;
    });
    uint8_t bins[13][13];
    for (size_t r = 0; r < shapes.size(); r++) bins[shapes[r].second / 13][shapes[r].second % 13] = (uint8_t)r;
    CREATE_TRY(cudaMemcpyToSymbol(c_shape_bin, bins, sizeof bins));
  }
  CREATE_TRY(cudaDeviceSynchronize());
#undef CREATE_TRY
  *out = ctx;
  return GPA_OK;
}

void gpa_destroy(gpa_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->d_key) cudaFree(ctx->d_key);
  if (ctx->d_ticket) cudaFree(ctx->d_ticket);
  if (ctx->d_dp_any) cudaFree(ctx->d_dp_any);
  if (ctx->d_dp_ok) cudaFree(ctx->d_dp_ok);
  if (ctx->h_res) cudaFreeHost(ctx->h_res);
  if (ctx->d_wq) cudaFree(ctx->d_wq);
  if (ctx->d_asym) cudaFree(ctx->d_asym);
  if (ctx->d_sfull) cudaFree(ctx->d_sfull);
  if (ctx->d_part_words) cudaFree(ctx->d_part_words);
  if (ctx->d_part_index) cudaFree(ctx->d_part_index);
  if (ctx->d_part_key) cudaFree(ctx->d_part_key);
  if (ctx->d_ans) cudaFree(ctx->d_ans);
  if (ctx->d_ans8) cudaFree(ctx->d_ans8);
  if (ctx->d_sub_binom) cudaFree(ctx->d_sub_binom);
  if (ctx->d_sub_pair) cudaFree(ctx->d_sub_pair);
  if (ctx->d_sub_next) cudaFree(ctx->d_sub_next);
  for (auto &ev : ctx->sub_ev)
    if (ev) cudaEventDestroy(ev);
  if (ctx->d_reqs) cudaFree(ctx->d_reqs);
  if (ctx->d_out) cudaFree(ctx->d_out);
  if (ctx->h_reqs) cudaFreeHost(ctx->h_reqs);
  if (ctx->h_out) cudaFreeHost(ctx->h_out);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  for (auto &ps : ctx->pipe)
    if (ps) cudaStreamDestroy(ps);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char *gpa_last_error(gpa_ctx *ctx) { return ctx ? ctx->last_error.c_str() : ""; }

unsigned long long gpa_launch_count(gpa_ctx *ctx) {
  if (!ctx) return 0;
  std::lock_guard<std::mutex> lk(ctx->mu);
  return ctx->launches;
}

int gpa_link_type_score(int link_type) {
  // the switch of besteffort_policy.go:314-353 over nvml.P2PLinkType (nvml.go:132-154):
  // six PCIe-ancestry levels worth 10..60, then n NVLinks worth 100 each
  if (link_type >= 1 && link_type <= 6) return 10 * link_type;
  if (link_type >= 7 && link_type <= 18) return 100 * (link_type - 6);
  return 0;
}

int gpa_set_topology(gpa_ctx *ctx, int n, const int32_t *W, const uint8_t *nlinks) {
  if (!ctx || !W || !nlinks || n < 1 || n > GPA_MAX_TOPOLOGY) return GPA_ERR_INVALID;
  Topology t;
  t.n = n;
  t.W.assign(W, W + (size_t)n * n);
  t.nlinks.assign(nlinks, nlinks + (size_t)n * n);
  int g = 0, mx = 0;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (i == j) continue;  // a device is never paired with itself (:303-305)
      int w = W[i * n + j];
      if (w < 0) return GPA_ERR_WEIGHT_RANGE;
      g = gcd_int(g, w);
      mx = w > mx ? w : mx;
    }
  if (g == 0) g = 1;
  if ((long long)(mx / g) * kMaxPairs > kScoreCap) return GPA_ERR_WEIGHT_RANGE;
  bool asym_any = false;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) asym_any = asym_any || (i != j && nlinks[i * n + j] != nlinks[j * n + i]);
  // with asymmetric link counts a partition's honest score must stay below the poison weight (kernels_common.cuh)
  if (asym_any && (long long)(mx / g) * kMaxPairs >= (long long)kAsymPoison) return GPA_ERR_WEIGHT_RANGE;
  t.unit = g;
  t.wq.assign((size_t)n * n, 0);
  t.asym.assign((size_t)n * n, 0);
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (i == j) continue;
      t.wq[i * n + j] = (uint16_t)(W[i * n + j] / g);
      t.asym[i * n + j] = nlinks[i * n + j] != nlinks[j * n + i];
      t.any_asym = t.any_asym || t.asym[i * n + j];
    }
  std::lock_guard<std::mutex> lk(ctx->mu);
  if ((int)ctx->topo.order.size() == n) t.order = ctx->topo.order;  // a device order set earlier survives a re-install
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  // Host-buffer requests are serialised on ctx->mu and finish before they return; the *_device entry points only
  // enqueue on the caller's stream, so a kernel of theirs may still be reading the tables that are replaced below:
  // wait for the device before anything is freed (installing a topology is rare and rebuilds tables anyway).
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  uint16_t *d_wq = nullptr;
  uint8_t *d_asym = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&d_wq, t.wq.size() * sizeof(uint16_t)));
  CUDA_TRY(ctx, cudaMalloc(&d_asym, t.asym.size()));
  CUDA_TRY(ctx, cudaMemcpy(d_wq, t.wq.data(), t.wq.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
  CUDA_TRY(ctx, cudaMemcpy(d_asym, t.asym.data(), t.asym.size(), cudaMemcpyHostToDevice));
  if (ctx->d_wq) cudaFree(ctx->d_wq);
  if (ctx->d_asym) cudaFree(ctx->d_asym);
  ctx->d_wq = d_wq;
  ctx->d_asym = d_asym;
  ctx->topo = std::move(t);
  ctx->has_topo = true;
  return rebuild_fast(ctx);
}

int gpa_set_device_order(gpa_ctx *ctx, const int32_t *order, int n) {
  if (!ctx || !order || n < 1) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (n != ctx->topo.n) return GPA_ERR_INVALID;
  unsigned long long seen = 0;
  for (int i = 0; i < n; i++) {
    if (order[i] < 0 || order[i] >= n || (seen >> order[i] & 1ull)) return GPA_ERR_INVALID;  // must be a permutation
    seen |= 1ull << order[i];
  }
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaDeviceSynchronize());  // a batch enqueued through a *_device entry point may still read the old tables
  ctx->topo.order.assign(order, order + n);
  return rebuild_fast(ctx);
}

int gpa_set_topology_links(gpa_ctx *ctx, int n, const int32_t *links2) {
  if (!ctx || !links2 || n < 1 || n > GPA_MAX_TOPOLOGY) return GPA_ERR_INVALID;
  std::vector<int32_t> W((size_t)n * n, 0);
  std::vector<uint8_t> nl((size_t)n * n, 0);
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (i == j) continue;
      for (int s = 0; s < 2; s++) {
        int t = links2[((size_t)i * n + j) * 2 + s];
        if (t != 0) {  // device.go:56-58, :64-66 — P2PLinkUnknown is not recorded
          W[i * n + j] += gpa_link_type_score(t);
          nl[i * n + j]++;
        }
      }
    }
  return gpa_set_topology(ctx, n, W.data(), nl.data());
}

int gpa_get_topology(gpa_ctx *ctx, int *n, int32_t *W, uint8_t *nlinks, int cap) {
  if (!ctx || !n) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  *n = ctx->topo.n;
  int need = ctx->topo.n * ctx->topo.n;
  if ((W || nlinks) && cap < need) return GPA_ERR_BUFFER;
  if (W) memcpy(W, ctx->topo.W.data(), sizeof(int32_t) * (size_t)need);
  if (nlinks) memcpy(nlinks, ctx->topo.nlinks.data(), (size_t)need);
  return GPA_OK;
}

int gpa_get_answer_table(gpa_ctx *ctx, uint16_t *table, int cap, int *n_entries) {
  if (!ctx || !n_entries) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (ctx->h_ans.empty()) return GPA_ERR_TOO_LARGE;
  *n_entries = (int)ctx->h_ans.size();
  if (!table) return GPA_OK;
  if (cap < (int)ctx->h_ans.size()) return GPA_ERR_BUFFER;
  memcpy(table, ctx->h_ans.data(), ctx->h_ans.size() * sizeof(uint16_t));
  return GPA_OK;
}

int gpa_besteffort(gpa_ctx *ctx, const int32_t *avail, int n_avail, const int32_t *req, int n_req, int size,
                   int mode_flags, int32_t *out, int *n_out) {
  if (!ctx || !n_out || n_avail < 0 || n_req < 0 || (n_avail > 0 && !avail) || (n_req > 0 && !req)) return GPA_ERR_INVALID;
  *n_out = 0;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  const Topology &T = ctx->topo;
  memset(&ctx->stats, 0, sizeof ctx->stats);
  ctx->timing_pending = false;
  ctx->stats.best_score = -1;
  ctx->stats.score_unit = T.unit;

  // argument hygiene (the Go seam takes pointers, which cannot be out of range)
  unsigned long long seen = 0;
  for (int i = 0; i < n_avail; i++) {
    if (avail[i] < 0 || avail[i] >= T.n) return GPA_ERR_INVALID;
    if (seen >> avail[i] & 1ull) return GPA_ERR_INVALID;  // the same pointer twice: outside the reference's domain
    seen |= 1ull << avail[i];
  }
  bool foreign = false;
  unsigned long long rseen = 0;
  for (int i = 0; i < n_req; i++) {
    if (req[i] == GPA_FOREIGN) {
      foreign = true;
      continue;
    }
    if (req[i] < 0 || req[i] >= T.n) return GPA_ERR_INVALID;
    if (rseen >> req[i] & 1ull) return GPA_ERR_INVALID;
    rseen |= 1ull << req[i];
    if (!(seen >> req[i] & 1ull)) foreign = true;  // required but not available: contained in no set
  }

  // besteffort_policy.go:35-45
  if (size <= 0) return GPA_OK;
  if (n_avail < size) return GPA_OK;
  if (n_req > size) return GPA_OK;
  if (n_avail > GPA_MAX_AVAILABLE) return GPA_ERR_TOO_LARGE;
  if (!out) return GPA_ERR_INVALID;
  // a required pointer that is in no candidate set: the filter at :53 rejects everything
  if (foreign) return GPA_OK;

  uint32_t rmask = 0;
  for (int i = 0; i < n_req; i++)
    for (int p = 0; p < n_avail; p++)
      if (avail[p] == req[i]) rmask |= 1u << p;

  if (size == 1) {
    // :219-224 — one single-device candidate per device, in order; all score 0, so the first accepted one stays
    int p = 0;
    if (rmask) p = __builtin_ctz(rmask);
    out[0] = avail[p];
    *n_out = 1;
    return GPA_OK;
  }

  // A node of up to 8 GPUs whose devices are listed in node order (the reference's sorted-UUID order): the answer is
  // one entry of the table the GPU built when the topology was installed.  No launch, no PCIe round trip - this is
  // what a single kubelet request costs.  (The table exists only for symmetric link counts, so :307-310 cannot fire.)
  if (!ctx->h_ans.empty() && !(mode_flags & (GPA_MODE_FORCE_ENUM | GPA_MODE_FORCE_DP | GPA_MODE_FORCE_KERNEL))) {
    uint32_t mask = 0, rlab = 0;
    bool in_order = true;
    for (int p = 0; p < n_avail && in_order; p++) {
      const int lab = ctx->fast.rank[avail[p]];
      const uint32_t bit = lab < kAnsMaxTopo ? 1u << lab : 0u;
      in_order = bit > mask;  // strictly ascending labels; a device the table does not know fails the same test
      mask |= bit;
      if (rmask >> p & 1u) rlab |= bit;
    }
    if (in_order) {
      const uint16_t e = ctx->h_ans[ans_index(ctx->h_tern, mask, rlab, size)];
      int o = 0;
      for (uint32_t c = ans_chosen(e); c; c &= c - 1) out[o++] = ctx->fast.order[__builtin_ctz(c)];
      for (int q = 0; q < ans_nil(e); q++) out[o++] = GPA_NIL;
      *n_out = o;
      const Geom g = make_geom(n_avail, size, rmask);
      ctx->stats.partitions = partition_count(g, ctx->binom);
      ctx->stats.items = 1;
      ctx->stats.sets = g.m;
      ctx->stats.padding = g.P;
      ctx->stats.algorithm = 6;
      return GPA_OK;
    }
  }

  // Pairs on which calculateGPUPairScore would panic (:307-310).  Go only panics when such a pair is SCORED, so the
  // search decides: the pair gets the weight kAsymPoison and the answer is GPA_ERR_ASYMMETRIC exactly when an accepted
  // candidate - or a set the final choice scores - holds one (kernels_common.cuh).
  bool asym_avail = false;
  for (int a = 0; a < n_avail && T.any_asym; a++)
    for (int b = 0; b < n_avail; b++)
      if (a != b && T.asym[avail[a] * T.n + avail[b]]) asym_avail = true;

  SearchParams sp;
  memset(&sp, 0, sizeof sp);
  sp.g = make_geom(n_avail, size, rmask);
  sp.intended = (mode_flags & GPA_MODE_INTENDED) ? 1 : 0;
  sp.seq = next_seq(ctx);
  for (int a = 0; a < n_avail; a++)
    for (int b = a + 1; b < n_avail; b++)
      sp.wp[a * kMaxDevices + b] = T.asym[avail[a] * T.n + avail[b]] ? (uint16_t)kAsymPoison : T.wq[avail[a] * T.n + avail[b]];
  const uint64_t parts = partition_count(sp.g, ctx->binom);
  if (parts > GPA_MAX_PARTITIONS) return GPA_ERR_TOO_LARGE;

  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  // A small node with its partition table installed, devices listed in node order: one warp answers.
  bool use_small = false;
  uint32_t small_mask = 0, small_rlab = 0;
  if (ctx->d_sfull && ctx->fast.part_words && !(mode_flags & (GPA_MODE_FORCE_ENUM | GPA_MODE_FORCE_DP))) {  // incl. GPA_MODE_FORCE_KERNEL
    use_small = true;
    int prev = -1;
    for (int p = 0; p < n_avail && use_small; p++) {
      const int lab = ctx->fast.rank[avail[p]];
      if (lab == 0xFF || lab <= prev) use_small = false;
      prev = lab;
      small_mask |= 1u << (lab & 31);
      if (rmask >> p & 1u) small_rlab |= 1u << (lab & 31);
    }
  }
  // Enumeration visits every partition in one launch; the subset DP needs m launches but far less work.  The DP wins
  // once the tree is large (crossover measured around 10^5 partitions on B200).
  // Enumeration visits every partition in one launch; the subset DP needs a launch per layer but far less work.  The DP
  // wins once the tree is large (crossover measured around 10^5 partitions on B200, tools/search_sweep.py).  A one-CTA
  // DP with its tables in shared memory was tried for <= 14 devices and measured 3-12x SLOWER than the layered launches
  // (one SM cannot issue the ~10^5..10^6 stepper iterations fast enough), so it is not kept.
  // With 15-16 devices the enumeration kernel's per-CTA score table (64-128 KB) costs more than the whole DP, whatever k.
  // (the DP's 16-bit tables have no room for poisoned scores: requests over such a pair are always enumerated)
  const bool use_dp = !asym_avail && ((mode_flags & GPA_MODE_FORCE_DP) ||
                                      (!(mode_flags & GPA_MODE_FORCE_ENUM) && (parts > 150000ull || n_avail >= 15)));
  int grid = 0, block = 0;
  size_t smem = 0;
  if (use_small) {
    grid = 1;
    block = 32;
    // no event pair around this one: two cudaEventRecord calls cost as much as the kernel itself (kernel_ms stays 0)
    besteffort_small_kernel<<<1, 32, 0, ctx->stream>>>(ctx->fast, small_mask, small_rlab, size, sp.intended, sp.seq, ctx->d_res);
    ctx->launches++;
    sp.cut.d = 0;
    sp.cut.n_items = parts;
  } else if (use_dp) {
    DpParams dp;
    memset(&dp, 0, sizeof dp);
    dp.g = sp.g;
    dp.intended = sp.intended;
    dp.seq = sp.seq;
    dp.rbit = rmask ? (rmask & (0u - rmask)) : 0u;
    memcpy(dp.wp, sp.wp, sizeof dp.wp);
    const uint32_t total = 1u << n_avail;
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    {
      grid = (int)((total + kDpThreads - 1) / kDpThreads);
      if (grid > ctx->sm_count * 8) grid = ctx->sm_count * 8;
      block = kDpThreads;
      dp_scores_kernel<<<grid, kDpThreads, 0, ctx->stream>>>(dp, ctx->d_dp_any);
      ctx->launches++;
      for (int j = 2; j < dp.g.m; j++) {
        // A thread per remainder walks its first sets one after the other; a warp per remainder spreads them over its
        // lanes by literal rank.  Measured: the warp form wins once a remainder has ~100 first sets.
        if (binom_at(ctx->binom, j * dp.g.k - 1, dp.g.k - 1) >= 96u) {
          // one warp per remainder of the layer (numbered: masks of j*k and of j*k - P devices)
          const int s1 = j * dp.g.k, s2 = dp.g.P ? j * dp.g.k - dp.g.P : 0;
          const uint64_t rems = (s1 <= n_avail ? binom_at(ctx->binom, n_avail, s1) : 0u) +
                                ((s2 >= 1 && s2 <= n_avail) ? binom_at(ctx->binom, n_avail, s2) : 0u);
          int wgrid = (int)((rems + kDpThreads / 32 - 1) / (kDpThreads / 32));
          if (wgrid > ctx->sm_count * 8) wgrid = ctx->sm_count * 8;
          if (wgrid < 1) wgrid = 1;
          dp_layer_warp_kernel<<<wgrid, kDpThreads, 0, ctx->stream>>>(dp, j, ctx->d_dp_any, ctx->d_dp_ok);
        } else {
          dp_layer_kernel<<<grid, kDpThreads, 0, ctx->stream>>>(dp, j, ctx->d_dp_any, ctx->d_dp_ok);
        }
        ctx->launches++;
      }
      dp_final_kernel<<<1, 1024, 0, ctx->stream>>>(dp, ctx->d_dp_any, ctx->d_dp_ok, ctx->d_res);
      ctx->launches++;
    }
    sp.cut.d = dp.g.m >= 2 ? 1 : 0;
    sp.cut.n_items = dp.g.m >= 2 ? binom_at(ctx->binom, dp.g.m * dp.g.k - 1, dp.g.k - 1) : 1;
  } else {
  const bool large = ((size_t)1 << n_avail) * sizeof(uint16_t) > 32 * 1024;
  block = large ? kSearchBlockLarge : kSearchBlockSmall;
  smem = search_smem_bytes(n_avail, block);
  if ((int)smem > ctx->max_smem_optin - 2048) return GPA_ERR_TOO_LARGE;
  auto kern = large ? besteffort_search_kernel<kSearchBlockLarge> : besteffort_search_kernel<kSearchBlockSmall>;
  if (smem > 40 * 1024) CUDA_TRY(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 1;
  CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, block, smem));
  if (occ < 1) occ = 1;
  const uint64_t resident_threads = (uint64_t)ctx->sm_count * occ * block;
  sp.cut = choose_cut(sp.g, ctx->binom, resident_threads * 4);
  uint64_t blocks = (sp.cut.n_items + block - 1) / block;
  uint64_t cap = (uint64_t)ctx->sm_count * occ;
  grid = (int)(blocks < cap ? blocks : cap);
  if (grid < 1) grid = 1;

  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  kern<<<grid, block, smem, ctx->stream>>>(sp, ctx->d_key, ctx->d_ticket, ctx->d_res);
  ctx->launches++;
  }
  CUDA_TRY(ctx, cudaGetLastError());
  if (!use_small) CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  {
    int rc = wait_mailbox(ctx, sp.seq);
    if (rc != GPA_OK) return rc;
  }
  ctx->timing_pending = !use_small;

  ctx->stats.partitions = parts;
  ctx->stats.items = sp.cut.n_items;
  ctx->stats.cut_depth = sp.cut.d;
  ctx->stats.sets = sp.g.m;
  ctx->stats.padding = sp.g.P;
  ctx->stats.best_score = ctx->h_res->best_score;
  ctx->stats.grid = grid;
  ctx->stats.block = block;
  ctx->stats.smem_bytes = (int)smem;
  ctx->stats.kernel_ms = 0;
  ctx->stats.algorithm = use_small ? 3 : (use_dp ? 2 : 1);

  const SearchResult &r = *ctx->h_res;
  if (asym_avail) {
    // an accepted candidate held such a pair (panic while iterating, :56), or the chosen set - the best-scoring one
    // among those the final loop scores, :77-85 - does
    bool chosen_asym = false;
    for (int i = 0; i < r.count; i++)
      for (int j = 0; j < r.count; j++)
        if (i != j && r.pos[i] >= 0 && r.pos[j] >= 0 && T.asym[avail[r.pos[i]] * T.n + avail[r.pos[j]]]) chosen_asym = true;
    if (r.best_score >= (int)kAsymPoison || chosen_asym) return GPA_ERR_ASYMMETRIC;
  }
  if (r.count < 0) {
    ctx->last_error = "search epilogue could not replay the winning item";
    return GPA_ERR_CUDA;
  }
  for (int i = 0; i < r.count; i++) out[i] = r.pos[i] < 0 ? GPA_NIL : avail[r.pos[i]];
  *n_out = r.count;
  return GPA_OK;
}

int gpa_last_search_stats(gpa_ctx *ctx, gpa_search_stats *out) {
  if (!ctx || !out) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (ctx->timing_pending) {  // the device time of the last search is read here, off the decision's critical path
    float ms = 0;
    if (cudaEventSynchronize(ctx->ev1) == cudaSuccess && cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) == cudaSuccess)
      ctx->stats.kernel_ms = ms;
    ctx->timing_pending = false;
  }
  *out = ctx->stats;
  return GPA_OK;
}

int gpa_besteffort_batch_device(gpa_ctx *ctx, const gpa_request *d_reqs, int count, int max_avail, int mode_flags,
                                gpa_result *d_out, void *stream) {
  if (!ctx || count < 0 || (count > 0 && (!d_reqs || !d_out))) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return launch_batch(ctx, d_reqs, count, max_avail, mode_flags, d_out, (cudaStream_t)stream);
}

int gpa_besteffort_batch(gpa_ctx *ctx, const gpa_request *reqs, int count, int max_avail, int mode_flags, gpa_result *out) {
  if (!ctx || count < 0 || max_avail < 0 || (count > 0 && (!reqs || !out))) return GPA_ERR_INVALID;
  if (count == 0) return GPA_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (max_avail == 0) {  // caller does not know: one pass over the records
    for (int i = 0; i < count; i++) {
      int na = reqs[i].n_avail;
      if (na > max_avail && na <= GPA_BATCH_MAX_AVAILABLE) max_avail = na;
    }
  }
  return host_batch(ctx, reqs, out, count, sizeof(gpa_request), sizeof(gpa_result), 1 << 18, false,
                    [&](void *d_in, int cnt, void *d_out, cudaStream_t st) {
                      return launch_batch(ctx, static_cast<const gpa_request *>(d_in), cnt, max_avail, mode_flags,
                                          static_cast<gpa_result *>(d_out), st);
                    });
}

int gpa_besteffort_batch_masks_device(gpa_ctx *ctx, const gpa_mask_request *d_reqs, int count, int mode_flags,
                                      gpa_mask_result *d_out, void *stream) {
  if (!ctx || count < 0 || (count > 0 && (!d_reqs || !d_out))) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (!ctx->d_sfull) return GPA_ERR_TOO_LARGE;  // masks need the node-wide score table: <= 12 GPUs, symmetric link counts
  if (count == 0) return GPA_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return launch_masks(ctx, d_reqs, count, mode_flags, d_out, (cudaStream_t)stream);
}

int gpa_besteffort_batch_masks(gpa_ctx *ctx, const gpa_mask_request *reqs, int count, int mode_flags, gpa_mask_result *out) {
  if (!ctx || count < 0 || (count > 0 && (!reqs || !out))) return GPA_ERR_INVALID;
  if (count == 0) return GPA_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (!ctx->d_sfull) return GPA_ERR_TOO_LARGE;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return host_batch(ctx, reqs, out, count, sizeof(gpa_mask_request), sizeof(gpa_mask_result), 1 << 20, true,
                    [&](void *d_in, int cnt, void *d_out, cudaStream_t st) { return launch_masks(ctx, d_in, cnt, mode_flags, d_out, st); });
}

int gpa_besteffort_batch_masks8_device(gpa_ctx *ctx, const gpa_mask8_request *d_reqs, int count, gpa_mask8_result *d_out, void *stream) {
  if (!ctx || count < 0 || (count > 0 && (!d_reqs || !d_out))) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (!ctx->fast.ans) return GPA_ERR_TOO_LARGE;  // the compact format is answered from the answer table only: <= 8 GPUs
  if (count == 0) return GPA_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return launch_masks8(ctx, d_reqs, count, d_out, (cudaStream_t)stream);
}

int gpa_besteffort_batch_masks8(gpa_ctx *ctx, const gpa_mask8_request *reqs, int count, gpa_mask8_result *out) {
  if (!ctx || count < 0 || (count > 0 && (!reqs || !out))) return GPA_ERR_INVALID;
  if (count == 0) return GPA_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (!ctx->fast.ans) return GPA_ERR_TOO_LARGE;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return host_batch(ctx, reqs, out, count, sizeof(gpa_mask8_request), sizeof(gpa_mask8_result), 1 << 20, true,
                    [&](void *d_in, int cnt, void *d_out, cudaStream_t st) { return launch_masks8(ctx, d_in, cnt, d_out, st); });
}

// ---- host-side conversion between the two batch formats (plain C++, no device work) ---------------------------
int gpa_pack_mask_requests(gpa_ctx *ctx, const gpa_request *reqs, int count, gpa_mask_request *out, int threads,
                           int *n_unpackable) {
  if (!ctx || count < 0 || (count > 0 && (!reqs || !out))) return GPA_ERR_INVALID;
  uint8_t rank[GPA_MAX_TOPOLOGY];
  {
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
    if (!ctx->d_sfull) return GPA_ERR_TOO_LARGE;  // no node-wide label space: more than 12 GPUs or asymmetric link counts
    memcpy(rank, ctx->fast.rank, sizeof rank);
  }
  std::atomic<int> bad_total{0};
  parallel_ranges(count, threads, [&](long long b, long long e) {
    int bad_here = 0;
    for (long long i = b; i < e; i++) {
      const gpa_request &r = reqs[i];
      uint32_t mask = 0, rl = 0;
      bool bad = r.n_avail > GPA_MAX_AVAILABLE || (r.n_avail < 16 && (r.req_mask >> r.n_avail) != 0);
      const int n = r.n_avail > GPA_MAX_AVAILABLE ? 0 : r.n_avail;
      for (int p = 0; p < n; p++) {
        const int lab = r.avail[p] < GPA_MAX_TOPOLOGY ? rank[r.avail[p]] : 0xFF;
        const uint32_t bit = lab < 16 ? 1u << lab : 0u;
        bad = bad || bit <= mask;  // strictly ascending labels; an unknown device fails the same test
        mask |= bit;
        if (r.req_mask >> p & 1u) rl |= bit;
      }
      gpa_mask_request m;
      m.avail_mask = bad ? (uint16_t)0xFFFF : (uint16_t)mask;
      m.req_mask = bad ? (uint16_t)0 : (uint16_t)rl;
      m.size = r.size;
      m.flags = r.flags;
      m.n_req_extra = r.n_req_extra;
      m.reserved = 0;
      out[i] = m;
      bad_here += bad;
    }
    bad_total += bad_here;
  });
  if (n_unpackable) *n_unpackable = bad_total.load();
  return GPA_OK;
}

int gpa_unpack_mask_results(gpa_ctx *ctx, const gpa_mask_result *res, int count, gpa_result *out, int threads) {
  if (!ctx || count < 0 || (count > 0 && (!res || !out))) return GPA_ERR_INVALID;
  uint8_t order[16];
  {
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
    if (!ctx->d_sfull) return GPA_ERR_TOO_LARGE;
    memcpy(order, ctx->fast.order, sizeof order);
  }
  parallel_ranges(count, threads, [&](long long b, long long e) {
    for (long long i = b; i < e; i++) {
      gpa_result o;
      memset(o.dev, GPA_RES_END, sizeof o.dev);
      const gpa_mask_result &m = res[i];
      if (m.status != GPA_OK) {
        o.dev[0] = GPA_RES_ERROR;
        o.dev[1] = m.status;
      } else {
        int k = 0;
        for (uint32_t c = m.chosen_mask; c; c &= c - 1) o.dev[k++] = (int8_t)order[__builtin_ctz(c)];
        for (int q = 0; q < m.n_nil && k < GPA_MAX_AVAILABLE; q++) o.dev[k++] = GPA_NIL;
      }
      out[i] = o;
    }
  });
  return GPA_OK;
}

#include "subset_host.inc"

/* pinned host memory for batch buffers (so gpa_besteffort_batch can skip its staging copy) */
void *gpa_alloc_host(size_t bytes) {
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
  return p;
}
void gpa_free_host(void *p) {
  if (p) cudaFreeHost(p);
}

}  // extern "C"

// =======================================================================================
// Allocation-free batch entry point.
//
// Who needs it.  This library decides GPU allocations on a node whose GPUs are SHARED: the
// same device that runs these kernels also holds the memory of the pods the plugin has
// handed it to.  Two kinds of caller therefore must not allocate when a request comes in:
//   * a daemon on a GPU whose tenants have filled HBM — cudaMalloc would fail exactly when
//     the node is busiest, and an allocation decision must not depend on free device memory;
//   * a caller that may not stall its tenants — cudaMalloc / cudaFree synchronise the
//     whole device, so every pod sharing the GPU would feel each request.
// gpa_create() + gpa_besteffort_batch() allocate (context, staging buffers, topology
// tables).  This entry point takes everything from the caller and keeps the device-side
// staging that the batch path needs — topology tables, request records, result records — in
// ONE fixed array that is part of the loaded module.  The price: the array is resident on
// every GPU that loads the library until the process exits, its size caps the batch, and it
// is shared by all calls.
// =======================================================================================
namespace {

constexpr size_t kNaWqOff = 0;                                                                      // uint16[64*64]
constexpr size_t kNaAsymOff = kNaWqOff + sizeof(uint16_t) * GPA_MAX_TOPOLOGY * GPA_MAX_TOPOLOGY;    // uint8[64*64]
constexpr size_t kNaReqOff = kNaAsymOff + (size_t)GPA_MAX_TOPOLOGY * GPA_MAX_TOPOLOGY;              // gpa_request[max]
constexpr size_t kNaResOff = kNaReqOff + sizeof(gpa_request) * GPA_NOALLOC_MAX_REQUESTS;            // gpa_result[max]
constexpr size_t kNaBytes = kNaResOff + sizeof(gpa_result) * GPA_NOALLOC_MAX_REQUESTS;              // 208 896 B
static_assert(kNaReqOff % 256 == 0 && kNaResOff % 256 == 0, "sub-buffers keep the 256-byte alignment");
static_assert(kNaBytes == GPA_NOALLOC_SCRATCH_BYTES, "header and implementation agree on the footprint");

}  // namespace

// The scratch itself, at file scope.  In HOST code this name is only a handle for the
// runtime: it is not a device address and must never be passed to a kernel.  The address
// comes from cudaGetSymbolAddress (noalloc_setup below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

namespace {

struct NoallocDevice {
  bool ready = false;
  unsigned char *base = nullptr;  // device address of g_noalloc_scratch on this GPU
  int sm_count = 0;
};
NoallocDevice g_noalloc_dev[64];
std::mutex g_noalloc_mu;

void na_msg(char *errbuf, int errbuf_len, const std::string &m) {
  if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%s", m.c_str());
}

// Once per device, with that device current: resolve the array's device address and load the
// constant table the kernels read.  Neither call allocates.
int noalloc_setup(int dev, cudaStream_t st, char *errbuf, int errbuf_len) {
  NoallocDevice &d = g_noalloc_dev[dev];
  if (d.ready) return GPA_OK;
  void *p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
  if (e != cudaSuccess || p == nullptr) {
    na_msg(errbuf, errbuf_len, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
    return GPA_ERR_CUDA;
  }
  // The table goes in on the CALLER's stream, in front of the kernel that reads it, and the host waits for it (the
  // source is a local): a copy on the default stream would not be ordered with a non-blocking caller stream.
  BinomTable b;
  fill_binom(b);
  e = cudaMemcpyToSymbolAsync(c_binom, &b, sizeof b, 0, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) {
    na_msg(errbuf, errbuf_len, std::string("no-alloc set-up: ") + cudaGetErrorString(e));
    return GPA_ERR_CUDA;
  }
  d.base = static_cast<unsigned char *>(p);
  d.ready = true;
  return GPA_OK;
}

}  // namespace

extern "C" int gpa_besteffort_batch_noalloc(int cuda_device, int n_topology, const int32_t *W, const uint8_t *nlinks,
                                            const gpa_request *reqs, int count, int max_avail, int mode_flags,
                                            gpa_result *out, void *stream, char *errbuf, int errbuf_len) {
  na_msg(errbuf, errbuf_len, "");
  if (cuda_device < 0 || cuda_device >= 64 || !W || !nlinks || count < 0 || max_avail < 0 || (count > 0 && (!reqs || !out))) {
    na_msg(errbuf, errbuf_len, "invalid argument");
    return GPA_ERR_INVALID;
  }
  if (n_topology < 1 || n_topology > GPA_MAX_TOPOLOGY) {
    na_msg(errbuf, errbuf_len, "topology of " + std::to_string(n_topology) + " devices: the fixed scratch holds tables for 1.." +
                                   std::to_string(GPA_MAX_TOPOLOGY) + " devices");
    return GPA_ERR_TOO_LARGE;
  }
  if (count > GPA_NOALLOC_MAX_REQUESTS) {
    na_msg(errbuf, errbuf_len, "batch of " + std::to_string(count) + " requests exceeds the fixed scratch of this entry point (" +
                                   std::to_string(GPA_NOALLOC_MAX_REQUESTS) + " requests); split the batch or use gpa_besteffort_batch");
    return GPA_ERR_TOO_LARGE;
  }
  if (count == 0) return GPA_OK;

  // weights -> the 16-bit scaled table the kernel reads (same rule as gpa_set_topology)
  const int n = n_topology;
  uint16_t wq[GPA_MAX_TOPOLOGY * GPA_MAX_TOPOLOGY];
  uint8_t asym[GPA_MAX_TOPOLOGY * GPA_MAX_TOPOLOGY];
  int g = 0, mx = 0;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (i == j) continue;
      int w = W[i * n + j];
      if (w < 0) {
        na_msg(errbuf, errbuf_len, "negative pair weight");
        return GPA_ERR_WEIGHT_RANGE;
      }
      g = gcd_int(g, w);
      mx = w > mx ? w : mx;
    }
  if (g == 0) g = 1;
  if ((long long)(mx / g) * kMaxPairs > kScoreCap) {
    na_msg(errbuf, errbuf_len, "pair weights too large for the 16-bit score table");
    return GPA_ERR_WEIGHT_RANGE;
  }
  bool asym_any = false;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      wq[i * n + j] = i == j ? 0 : (uint16_t)(W[i * n + j] / g);
      asym[i * n + j] = i != j && nlinks[i * n + j] != nlinks[j * n + i];
      asym_any = asym_any || asym[i * n + j];
    }
  if (asym_any && (long long)(mx / g) * kMaxPairs >= (long long)kAsymPoison) {  // same rule as gpa_set_topology
    na_msg(errbuf, errbuf_len, "pair weights too large for a topology with asymmetric link counts");
    return GPA_ERR_WEIGHT_RANGE;
  }
  if (max_avail == 0)
    for (int i = 0; i < count; i++)
      if (reqs[i].n_avail > max_avail && reqs[i].n_avail <= GPA_BATCH_MAX_AVAILABLE) max_avail = reqs[i].n_avail;

  // One call at a time per process: every call uses the same array.
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  cudaError_t e = cudaSetDevice(cuda_device);
  if (e != cudaSuccess) {
    na_msg(errbuf, errbuf_len, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    return GPA_ERR_CUDA;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int rc = noalloc_setup(cuda_device, st, errbuf, errbuf_len);
  if (rc != GPA_OK) return rc;
  const NoallocDevice &d = g_noalloc_dev[cuda_device];
  uint16_t *d_wq = reinterpret_cast<uint16_t *>(d.base + kNaWqOff);
  uint8_t *d_asym = d.base + kNaAsymOff;
  gpa_request *d_reqs = reinterpret_cast<gpa_request *>(d.base + kNaReqOff);
  gpa_result *d_out = reinterpret_cast<gpa_result *>(d.base + kNaResOff);

  // On a failure the stream is drained before the lock on the array is given up: copies or the kernel queued so far
  // may still be touching it, and the next caller must not overlap them.
#define NA_TRY(expr)                                                              \
  do {                                                                            \
    cudaError_t e_ = (expr);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      na_msg(errbuf, errbuf_len, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
      cudaStreamSynchronize(st);                                                  \
      return GPA_ERR_CUDA;                                                        \
    }                                                                             \
  } while (0)
  NA_TRY(cudaMemcpyAsync(d_wq, wq, sizeof(uint16_t) * (size_t)n * n, cudaMemcpyHostToDevice, st));
  NA_TRY(cudaMemcpyAsync(d_asym, asym, (size_t)n * n, cudaMemcpyHostToDevice, st));
  NA_TRY(cudaMemcpyAsync(d_reqs, reqs, sizeof(gpa_request) * (size_t)count, cudaMemcpyHostToDevice, st));
  // The array still holds the previous call's answers.  The kernel writes every record it is
  // asked for; clearing them first - to GPA_RES_END bytes, i.e. to empty answers - means a
  // launch that fails cannot hand back stale ones.
  NA_TRY(cudaMemsetAsync(d_out, 0x80, sizeof(gpa_result) * (size_t)count, st));

  int n_cap = max_avail < 2 ? 2 : max_avail;
  if (n_cap > GPA_BATCH_MAX_AVAILABLE) n_cap = GPA_BATCH_MAX_AVAILABLE;
  size_t smem = (size_t)kBatchWarps * (((size_t)1 << n_cap) + kMaxDevices * kMaxDevices) * sizeof(uint16_t);
  if (smem > 48 * 1024)
    NA_TRY(cudaFuncSetAttribute(besteffort_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int blocks = (count + kBatchWarps - 1) / kBatchWarps;
  int cap = d.sm_count * 4;
  int grid = blocks < cap ? blocks : cap;
  BatchTopo t{d_wq, d_asym, n};
  besteffort_batch_kernel<<<grid, kBatchWarps * 32, smem, st>>>(d_reqs, count, n_cap, mode_flags & GPA_MODE_INTENDED, t, d_out, 0);
  NA_TRY(cudaGetLastError());
  NA_TRY(cudaMemcpyAsync(out, d_out, sizeof(gpa_result) * (size_t)count, cudaMemcpyDeviceToHost, st));
  NA_TRY(cudaStreamSynchronize(st));
#undef NA_TRY
  return GPA_OK;
}
