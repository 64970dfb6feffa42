// Replica-set search kernels: which `size` of the available shared replicas are best connected (subset_core.cuh has
// the definition and says why there is no reference counterpart).
//   ONE request per launch, answered into the mailbox:
//     subset_counts_kernel    per-GPU count vectors in descending lexicographic order, one rank per thread at a time
//     subset_enum_kernel<K>   the literal walk over C(R, K) combinations: a thread per (K-1)-prefix in colex order, runs
//                             over the last pick
//   MANY requests per launch (count-form records in, "take this many more from each GPU" out):
//     subset_batch_kernel<G>  one warp per request, depth-first walk of the count vectors with carried partial scores
// Part of the single translation unit gpualloc.cu; not compiled on its own.
#pragma once
#include "kernels_common.cuh"
#include "subset_core.cuh"

namespace {

constexpr int kSubThreads = 256;

struct SubsetCountsParams {
  SubsetCounts q;
  unsigned long long total;  // count vectors = cnt[0][K]
  int32_t size;
  unsigned int seq;
};

struct SubsetEnumParams {
  SubsetEnum q;
  unsigned long long items;  // (K-1)-prefixes = C(R-1, K-1); a thread walks the last pick of one prefix at a time
  int32_t size;
  unsigned int seq;
};

// Block-wide maximum of the keys, then the grid-wide one through g_scratch[0]; the last CTA to arrive publishes it.
// The host turns the winning rank back into positions (it holds the same tables).
__device__ __forceinline__ void subset_reduce_publish(unsigned long long best, int size, unsigned int seq, unsigned long long *g_scratch,
                                                      unsigned int *g_ticket, SearchResult *g_res) {
  __shared__ unsigned long long wkey[kSubThreads / 32];
  __shared__ int is_last;
  const int tid = threadIdx.x;
  best = warp_max_u64(best);
  if ((tid & 31) == 0) wkey[tid >> 5] = best;
  __syncthreads();
  if (tid == 0) {
    unsigned long long b = wkey[0];
#pragma unroll
    for (int w = 1; w < kSubThreads / 32; w++) b = wkey[w] > b ? wkey[w] : b;
    if (b) atomicMax(g_scratch, b);
    __threadfence();
    const unsigned int t = atomicAdd(g_ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last && tid == 0) {
    const unsigned long long key = atomicMax(g_scratch, 0ull);
    SearchResult r;
    r.key = key;
    r.best_score = key ? key_score(key) : -1;
    r.count = key ? size : -1;
#pragma unroll
    for (int i = 0; i < 32; i++) r.pos[i] = 0;
    *g_scratch = 0ull;  // scratch is left clean for the next launch on this stream
    *g_ticket = 0u;
    publish_result(g_res, r, seq);
  }
}

__global__ void __launch_bounds__(kSubThreads) subset_counts_kernel(const SubsetCountsParams p, unsigned long long *g_scratch,
                                                                     unsigned int *g_ticket, SearchResult *g_res) {
  __shared__ uint64_t cnt[(kSubGroups + 1) * kSubCntStride];
  __shared__ uint16_t w[kSubGroups * kSubGroups];
  const int tid = threadIdx.x;
  for (int i = tid; i < kSubGroups * kSubGroups; i += kSubThreads) w[i] = p.q.w[i];
  // class counts, bottom level first; a level needs only the one below
  if (tid <= p.q.K) cnt[p.q.G * kSubCntStride + tid] = tid == 0 ? 1ull : 0ull;
  __syncthreads();
  for (int g = p.q.G - 1; g >= 0; g--) {
    if (tid <= p.q.K) cnt[g * kSubCntStride + tid] = sub_cnt_entry(p.q, cnt, g, tid);
    __syncthreads();
  }
  unsigned long long best = 0;
  const unsigned long long stride = (unsigned long long)gridDim.x * kSubThreads;
  for (unsigned long long rank = (unsigned long long)blockIdx.x * kSubThreads + tid; rank < p.total; rank += stride) {
    const uint32_t score = sub_unrank_score(p.q, cnt, w, rank, nullptr);
    const unsigned long long key = pack_key((int)score, rank);
    best = key > best ? key : best;
  }
  subset_reduce_publish(best, p.size, p.seq, g_scratch, g_ticket, g_res);
}

template <int K>
__global__ void __launch_bounds__(kSubThreads) subset_enum_kernel(const SubsetEnumParams p, const uint64_t *__restrict__ binom,
                                                                   unsigned long long *g_scratch, unsigned int *g_ticket,
                                                                   SearchResult *g_res) {
  __shared__ uint8_t grp[kSubEnumMaxFree];
  __shared__ uint32_t wreq[kSubEnumMaxFree];
  __shared__ uint16_t w[kSubGroups * kSubGroups];
  __shared__ uint64_t s_binom[kSubBinomRows * kSubBinomCols];  // 18.5 KB: every lane probes its own entries while unranking
  const int tid = threadIdx.x;
  for (int i = tid; i < kSubBinomRows * kSubBinomCols; i += kSubThreads) s_binom[i] = binom[i];
  for (int i = tid; i < kSubEnumMaxFree; i += kSubThreads) {
    grp[i] = p.q.grp[i];
    wreq[i] = p.q.wreq[i];
  }
  for (int i = tid; i < kSubGroups * kSubGroups; i += kSubThreads) w[i] = p.q.w[i];
  __syncthreads();
  const int R = p.q.R;
  unsigned long long best = 0;
  const unsigned long long stride = (unsigned long long)gridDim.x * kSubThreads;
  for (unsigned long long item = (unsigned long long)blockIdx.x * kSubThreads + tid; item < p.items; item += stride)
    best = sub_walk_prefix<K>(s_binom, grp, wreq, w, R, p.q.base, item, best);
  subset_reduce_publish(best, p.size, p.seq, g_scratch, g_ticket, g_res);
}

// --------------------------------------------------------------------------------------------------------------------
// MANY replica-set requests in count form (gpa_subset_request): one warp per request.
// Labels of the node order stand for the groups (a GPU with no replica in the request is a group that can take
// nothing), so the class order and the tie-break are those of the single-request kernel.
//
// The count vectors are walked depth first, one GPU per level, in descending order of the count at every level
// (= descending lexicographic order of the vectors).  The score sum_{g<h} c_g c_h w[g][h] is carried down the walk: P =
// score of the levels above, L[h] = sum_{g above} c_g w[g][h], so a step at level g costs G-g-1 additions and a leaf two
// operations, instead of the G^2/2 of scoring a vector from scratch.  The lanes of the warp split the walk by the counts
// of the first kSubPrefix GPUs (an "item"); bests are merged on (score, lexicographically largest vector).

constexpr int kSubBatchMaxExtra = GPA_SUBSET_BATCH_MAX_EXTRA;  // replicas to choose beyond the required ones
constexpr int kSubBatchWarps = kSubThreads / 32;
constexpr int kSubPrefix = 4;
constexpr int kSubDraw = 4;  // requests a warp takes from the launch's counter at a time
static_assert(kSubBatchMaxExtra <= 15, "a count beyond the required ones is packed in 4 bits");

struct SubsetBatchTopo {
  uint16_t w[kSubGroups * kSubGroups];  // scaled pair weight in LABEL space, [a * 16 + b] for labels a < b
  uint16_t asym_row[kSubGroups];        // bit b of row a: link counts of labels a and b differ (calculateGPUPairScore panics)
  int32_t nt;                           // GPUs of the node
};

static_assert(sizeof(gpa_subset_request) == 48 && sizeof(gpa_subset_result) == 24, "record layouts of gpualloc.h");

// What a warp knows about its request after the guards (besteffort_policy.go:35-45 and the limits of gpualloc.h)
struct SubBatchDecoded {
  int my_f, my_r;  // lane g < 16: free / must-include replicas of GPU label g
  int size, K;     // K = replicas to choose beyond the must-include ones
  int n_req;
  int status;
  bool search;
};

__device__ __forceinline__ SubBatchDecoded sub_batch_decode(const gpa_subset_request *reqs, long long i, int nt, const uint16_t *asym_row,
                                                            const BinomTable &s_binom, int lane) {
  SubBatchDecoded d;
  // 48-byte record = 12 words: 0-3 free[], 4-7 must[], 8 size
  const uint32_t *rec = reinterpret_cast<const uint32_t *>(reqs + i);
  const uint32_t word = lane < 12 ? __ldg(rec + lane) : 0u;
  d.size = (int)__shfl_sync(0xffffffffu, word, 8);
  const uint32_t fw = __shfl_sync(0xffffffffu, word, (lane >> 2) & 3);
  const uint32_t rw = __shfl_sync(0xffffffffu, word, 4 + ((lane >> 2) & 3));
  d.my_f = lane < kSubGroups ? (int)((fw >> (8 * (lane & 3))) & 0xFFu) : 0;
  d.my_r = lane < kSubGroups ? (int)((rw >> (8 * (lane & 3))) & 0xFFu) : 0;
  const unsigned present = __ballot_sync(0xffffffffu, d.my_f + d.my_r > 0);
  const unsigned has_free = __ballot_sync(0xffffffffu, d.my_f > 0);
  const int n_free = (int)__reduce_add_sync(0xffffffffu, (unsigned)d.my_f);  // lanes 16.. hold zeros
  const int n_req = (int)__reduce_add_sync(0xffffffffu, (unsigned)d.my_r);
  d.n_req = n_req;
  d.K = d.size - n_req;
  d.status = GPA_OK;
  d.search = true;
  if (present >> nt) {
    d.status = GPA_ERR_INVALID;  // replicas of a GPU the node does not have
    d.search = false;
  } else if (d.size <= 0 || n_free + n_req < d.size || n_req > d.size) {
    d.search = false;  // besteffort_policy.go:35-45: the empty answer
  } else if (d.size > GPA_SUBSET_MAX_SIZE || d.K > kSubBatchMaxExtra ||
             binom_at(s_binom, d.K + __popc(has_free) - 1, __popc(has_free) - 1) > (uint32_t)GPA_SUBSET_BATCH_MAX_CLASSES) {
    d.status = GPA_ERR_TOO_LARGE;  // (K > 0 implies a GPU with free replicas here, so the binomial is well formed)
    d.search = false;
  } else if (d.size >= 2) {
    const bool bad = lane < kSubGroups && (present >> lane & 1u) && (asym_row[lane] & present);
    if (__any_sync(0xffffffffu, bad)) {
      d.status = GPA_ERR_ASYMMETRIC;
      d.search = false;
    }
  }
  return d;
}

__device__ __forceinline__ void sub_batch_write_empty(gpa_subset_result *out, long long i, int status, int lane) {
  if (lane == 0) {
    uint2 *dst = reinterpret_cast<uint2 *>(out + i);
    dst[0] = make_uint2(0u, 0u);
    dst[1] = make_uint2(0u, 0u);
    dst[2] = make_uint2(0xFFFFFFFFu, (uint32_t)(uint16_t)0 | ((uint32_t)(uint16_t)(int16_t)status << 16));
  }
}

// The depth-first walk for one request, whole warp: f / r / cap / exact are the warp's scratch in shared memory.
template <int G>
__device__ __forceinline__ void sub_batch_dfs(const SubBatchDecoded &d, uint8_t *f, uint8_t *r, uint16_t *cap, uint32_t *exact,
                                              const uint16_t *w, gpa_subset_result *out, long long i, int lane) {
  const int K = d.K, my_f = d.my_f, my_r = d.my_r, size = d.size;
  // free counts clamped to K, required counts, and what the GPUs from g on can still absorb
  const int fc = my_f < K ? my_f : K;
  int suffix = lane < kSubGroups ? fc : 0;
#pragma unroll
  for (int o = 1; o < kSubGroups; o <<= 1) {  // inclusive suffix sum over lanes 0..15
    const int up = __shfl_down_sync(0xffffffffu, suffix, o);
    if (lane + o < kSubGroups) suffix += up;
  }
  if (lane < kSubGroups) {
    f[lane] = (uint8_t)fc;
    r[lane] = (uint8_t)my_r;
    cap[lane] = (uint16_t)suffix;
  }
  if (lane == 0) cap[kSubGroups] = 0;
  __syncwarp();

  // Items = count prefixes of the first kSubPrefix GPUs that respect the free counts and leave no more than the GPUs
  // below can absorb.  Items that leave the SAME number of replicas walk identical subtrees below the prefix (only P
  // and L differ), so they are numbered class by class - most left first - and a warp's 32 lanes run in lockstep
  // except where a round straddles two classes.  exact[p][s] = ways for prefix GPUs p.. to take exactly s.
  if (lane <= K) exact[kSubPrefix * 16 + lane] = lane == 0 ? 1u : 0u;
  __syncwarp();
#pragma unroll
  for (int p = kSubPrefix - 1; p >= 0; p--) {
    if (lane <= K) {
      uint32_t sum = 0;
      const int hi = (int)f[p] < lane ? (int)f[p] : lane;
      for (int x = 0; x <= hi; x++) sum += exact[(p + 1) * 16 + lane - x];
      exact[p * 16 + lane] = sum;
    }
    __syncwarp();
  }
  const int s_lo = K > (int)cap[kSubPrefix] ? K - (int)cap[kSubPrefix] : 0;  // the prefix must take at least this much
  uint32_t n_items = 0;
  for (int sp = s_lo; sp <= K; sp++) n_items += exact[sp];
  int bs = -1;
  unsigned long long bp = 0;
  uint32_t blex = 0;
  for (uint32_t item = lane; item < n_items; item += 32) {
    sub_batch_item<G, kSubPrefix>(f, r, cap, w, exact, K, s_lo, item, bs, bp, blex);
  }
  const unsigned long long mine = bs >= 0 ? ((unsigned long long)(unsigned)(bs + 1) << 32) | blex : 0ull;
  const unsigned long long best = warp_max_u64(mine);
  if (mine == best && best) {  // exactly one lane: prefixes are distinct
    uint2 *dst = reinterpret_cast<uint2 *>(out + i);
    uint32_t take[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int g = 0; g < G; g++) take[g >> 2] |= (uint32_t)((bp >> (4 * g)) & 0xFull) << (8 * (g & 3));
    dst[0] = make_uint2(take[0], take[1]);
    dst[1] = make_uint2(take[2], take[3]);
    dst[2] = make_uint2((uint32_t)bs, (uint32_t)(uint16_t)size | ((uint32_t)(uint16_t)GPA_OK << 16));
  }
  __syncwarp();  // f / r / cap are rewritten for the warp's next request
}

template <int G>
__global__ void __launch_bounds__(kSubThreads, G <= 8 ? 3 : 1) subset_batch_kernel(const gpa_subset_request *__restrict__ reqs, int count,
                                                                    const SubsetBatchTopo t,
                                                                    const gpa_subset_result *__restrict__ g_free_node,
                                                                    unsigned int *__restrict__ g_next,
                                                                    gpa_subset_result *__restrict__ out) {
  __shared__ uint16_t w[kSubGroups * kSubGroups];
  __shared__ uint8_t fr_all[kSubBatchWarps][2 * kSubGroups];
  __shared__ uint16_t cap_all[kSubBatchWarps][kSubGroups + 1];
  __shared__ uint32_t exact_all[kSubBatchWarps][(kSubPrefix + 1) * 16];
  __shared__ uint16_t asym_row[kSubGroups];
  __shared__ BinomTable s_binom;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  stage_binom(s_binom, tid, kSubThreads);
  for (int i = tid; i < kSubGroups * kSubGroups; i += kSubThreads) w[i] = t.w[i];
  if (tid < kSubGroups) asym_row[tid] = t.asym_row[tid];
  __syncthreads();
  // requests drawn from a counter and the empty node's answers: as in subset_batch_pair_kernel below (both optional)
  for (unsigned int turn = 0;; turn++) {
    unsigned int first = ((unsigned int)(blockIdx.x * kSubBatchWarps + warp) + turn * (unsigned int)(gridDim.x * kSubBatchWarps)) * kSubDraw;
    if (g_next) {
      if (lane == 0) first = atomicAdd(g_next, (unsigned int)kSubDraw);
      first = __shfl_sync(0xffffffffu, first, 0);
    }
    if (first >= (unsigned int)count) break;
    const unsigned int last = first + kSubDraw < (unsigned int)count ? first + kSubDraw : (unsigned int)count;
    for (long long i = first; i < last; i++) {
      const SubBatchDecoded d = sub_batch_decode(reqs, i, t.nt, asym_row, s_binom, lane);
      if (!d.search) {
        sub_batch_write_empty(out, i, d.status, lane);
      } else if (g_free_node != nullptr && d.n_req == 0 && __all_sync(0xffffffffu, lane >= t.nt || d.my_f >= d.K)) {
        if (lane < 3) reinterpret_cast<uint2 *>(out + i)[lane] = __ldg(reinterpret_cast<const uint2 *>(g_free_node + d.K) + lane);
      } else {
        sub_batch_dfs<G>(d, fr_all[warp], fr_all[warp] + kSubGroups, cap_all[warp], exact_all[warp], w, out, i, lane);
      }
    }
  }
}

// --------------------------------------------------------------------------------------------------------------------
// Nodes of up to 8 GPUs: the two-halves walk (subset_core.cuh "count form, two halves").  The node's table of
// half-vectors (16 KB, built by sub_pair_fill when the topology or the order is installed) is staged in shared memory
// once per CTA; a warp filters it down to what its request's free counts allow - two ballots per 32 compositions - and
// then visits (prefix, suffix) pairs at one shared-memory load and five integer operations each.  Requests that choose
// more than kSubHalfMaxK further replicas take the depth-first walk above.
constexpr int kSubListLen = (kSubHalfEntries + 1 + 7) & ~7;

template <int G>
__device__ __noinline__ void sub_batch_dfs_call(const SubBatchDecoded &d, uint8_t *f, uint8_t *r, uint16_t *cap, uint32_t *exact,
                                                const uint16_t *w, gpa_subset_result *out, long long i, int lane) {
  sub_batch_dfs<G>(d, f, r, cap, exact, w, out, i, lane);
}

__global__ void __launch_bounds__(kSubThreads, 5) subset_batch_pair_kernel(const gpa_subset_request *__restrict__ reqs, int count,
                                                                           const SubsetBatchTopo t, const uint4 *__restrict__ g_blob,
                                                                           const uint32_t *__restrict__ g_flat,
                                                                           const gpa_subset_result *__restrict__ g_free_node,
                                                                           unsigned int *__restrict__ g_next,
                                                                           gpa_subset_result *__restrict__ out) {
  __shared__ __align__(16) uint32_t blob[kSubPairBlobBytes / 4];
  __shared__ uint32_t split_tab[32];
  __shared__ uint16_t ident_list[kSubListLen], start_tab[kSubHalfMaxK + 2];  // the lists of a request no free count binds
  __shared__ uint16_t plist_all[kSubBatchWarps][kSubListLen], slist_all[kSubBatchWarps][kSubListLen];
  __shared__ uint16_t off_all[kSubBatchWarps][2 * (kSubHalfMaxK + 2)];
  __shared__ uint32_t lin_all[kSubBatchWarps][8];
  __shared__ uint16_t w[kSubGroups * kSubGroups];
  __shared__ uint8_t fr_all[kSubBatchWarps][2 * kSubGroups];
  __shared__ uint16_t cap_all[kSubBatchWarps][kSubGroups + 1];
  __shared__ uint32_t exact_all[kSubBatchWarps][(kSubPrefix + 1) * 16];
  __shared__ uint16_t asym_row[kSubGroups];
  __shared__ BinomTable s_binom;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  stage_binom(s_binom, tid, kSubThreads);
  for (int i = tid; i < kSubPairBlobBytes / 16; i += kSubThreads) reinterpret_cast<uint4 *>(blob)[i] = __ldg(g_blob + i);
  for (int i = tid; i < kSubGroups * kSubGroups; i += kSubThreads) w[i] = t.w[i];
  if (tid < kSubGroups) asym_row[tid] = t.asym_row[tid];
  if (tid < 32) split_tab[tid] = sub_pair_split_entry(tid);
  for (int i = tid; i < kSubListLen; i += kSubThreads) ident_list[i] = (uint16_t)i;
  if (tid < kSubHalfMaxK + 2) start_tab[tid] = (uint16_t)sub_half_start(tid);
  __syncthreads();
  uint16_t *const my_plist = plist_all[warp], *const my_slist = slist_all[warp];
  uint16_t *const my_poff = off_all[warp], *const my_soff = off_all[warp] + kSubHalfMaxK + 2;
  uint32_t *lin = lin_all[warp];
  const int my_start = sub_half_start(lane < kSubHalfMaxK + 2 ? lane : 0);  // where class `lane` starts in the table
  const uint32_t lt = (1u << lane) - 1u;

  const uint32_t wmax = blob[kSubPairMetaOff / 4];
  // Requests cost between a few hundred and several thousand instructions (1 to 6 435 count vectors), so the warps
  // draw them kSubDraw at a time from a counter (*g_next, zeroed by the host before the launch) instead of striding:
  // no SM waits for its slowest warp's fixed share.  Without a counter (g_next == nullptr) they stride.
  for (unsigned int turn = 0;; turn++) {
    unsigned int first = ((unsigned int)(blockIdx.x * kSubBatchWarps + warp) + turn * (unsigned int)(gridDim.x * kSubBatchWarps)) * kSubDraw;
    if (g_next) {
      if (lane == 0) first = atomicAdd(g_next, (unsigned int)kSubDraw);
      first = __shfl_sync(0xffffffffu, first, 0);
    }
    if (first >= (unsigned int)count) break;
    const unsigned int last = first + kSubDraw < (unsigned int)count ? first + kSubDraw : (unsigned int)count;
  for (long long i = first; i < last; i++) {
    const SubBatchDecoded d = sub_batch_decode(reqs, i, t.nt, asym_row, s_binom, lane);
    if (!d.search) {
      sub_batch_write_empty(out, i, d.status, lane);
      continue;
    }
    // A request that no free count binds - every GPU of the node has K replicas free, none is must-include - has the
    // answer of "K replicas on the empty node", which this kernel worked out itself when the topology was installed
    // (g_free_node[K], rebuild_subset_tables(); nullptr while it does).  On a lightly loaded node that is most requests.
    if (g_free_node != nullptr && d.n_req == 0 && __all_sync(0xffffffffu, lane >= t.nt || d.my_f >= d.K)) {
      if (lane < 3) reinterpret_cast<uint2 *>(out + i)[lane] = __ldg(reinterpret_cast<const uint2 *>(g_free_node + d.K) + lane);
      continue;
    }
    if (d.K > kSubHalfMaxK) {
      sub_batch_dfs_call<8>(d, fr_all[warp], fr_all[warp] + kSubGroups, cap_all[warp], exact_all[warp], w, out, i, lane);
      continue;
    }
    const int K = d.K;
    // free counts clamped to K, one byte per GPU, label 0 / 4 in the highest byte of its half
    uint32_t fsh = lane < 8 ? (uint32_t)(d.my_f < K ? d.my_f : K) << (8 * (3 - (lane & 3))) : 0u;
    fsh |= __shfl_xor_sync(0xffffffffu, fsh, 1);
    fsh |= __shfl_xor_sync(0xffffffffu, fsh, 2);
    const uint32_t fpack_p = __shfl_sync(0xffffffffu, fsh, 0), fpack_s = __shfl_sync(0xffffffffu, fsh, 4);
    // the must-include replicas: among themselves (r0) and against one more replica of GPU `lane` (lin)
    uint32_t r0 = 0, my_lin = 0;
    if (d.n_req > 0) {
      uint32_t above = 0;
#pragma unroll
      for (int g = 0; g < 8; g++) {
        const uint32_t rg = (uint32_t)__shfl_sync(0xffffffffu, d.my_r, g);
        if (lane < 8 && g != lane) {
          const uint32_t wgl = w[(g < lane ? g : lane) * kSubGroups + (g < lane ? lane : g)];
          my_lin += rg * wgl;
          if (g > lane) above += rg * wgl;
        }
      }
      r0 = (uint32_t)d.my_r * above;
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) r0 += __shfl_xor_sync(0xffffffffu, r0, o);  // lanes 8.. hold zeros
      r0 = __shfl_sync(0xffffffffu, r0, 0);
    }
    if (lane < 8) lin[lane] = my_lin;
    uint32_t lin_max = lane >= 4 && lane < 8 ? my_lin : 0u;
    if (d.n_req > 0) {
      lin_max = max(lin_max, __shfl_xor_sync(0xffffffffu, lin_max, 1));
      lin_max = max(lin_max, __shfl_xor_sync(0xffffffffu, lin_max, 2));
      lin_max = __shfl_sync(0xffffffffu, lin_max, 4);
    }
    if (K >= 1 && K <= kSubFlatMaxK) {
      // a few further replicas: one pass over the class in the node's table of all count vectors (sub_flat_key)
      __syncwarp();  // lin[] is read by every lane
      const unsigned long long best = warp_max_u64(sub_flat_lane(g_flat, K, fpack_p, fpack_s, r0, lin, d.n_req > 0, lane));
      if (best == 0) {
        sub_batch_write_empty(out, i, GPA_OK, lane);
      } else if (lane == 0) {
        const int e = (int)~(uint32_t)best;
        uint2 *dst = reinterpret_cast<uint2 *>(out + i);
        dst[0] = make_uint2(__byte_perm(g_flat[2 * e], 0u, 0x0123), __byte_perm(g_flat[2 * e + 1], 0u, 0x0123));
        dst[1] = make_uint2(0u, 0u);
        dst[2] = make_uint2((uint32_t)(best >> 32) - 1u, (uint32_t)(uint16_t)d.size | ((uint32_t)(uint16_t)GPA_OK << 16));
      }
      __syncwarp();  // lin[] is rewritten for the warp's next request
      continue;
    }
    const int mode = sub_pair_mode(K, wmax, lin_max, d.n_req);
    // the lists: compositions of classes 0..K that fit, for each half, and where every class starts in them.  Where every
    // GPU has K replicas free (deep replica sets, most requests) nothing is dropped and the table itself is the list.
    const int n_static = sub_half_start(K + 1);
    const uint32_t all_k = (uint32_t)K * 0x01010101u;
    const bool whole = fpack_p == all_k && fpack_s == all_k;
    const uint16_t *plist = ident_list, *slist = ident_list, *poff = start_tab, *soff = start_tab;
    int base_s = n_static;
    if (!whole) {
      plist = my_plist;
      slist = my_slist;
      poff = my_poff;
      soff = my_soff;
      int base_p = 0;
      base_s = 0;
      for (int i0 = 0; i0 < n_static; i0 += 32) {
        const int e = i0 + lane;
        const bool in = e < n_static;
        const uint32_t x = in ? blob[kSubPairCompOff / 4 + e] : 0u;
        const bool okp = in && sub_half_fits(x, fpack_p), oks = in && sub_half_fits(x, fpack_s);
        const unsigned bp = __ballot_sync(0xffffffffu, okp), bsf = __ballot_sync(0xffffffffu, oks);
        if (okp) my_plist[base_p + __popc(bp & lt)] = (uint16_t)e;
        if (oks) my_slist[base_s + __popc(bsf & lt)] = (uint16_t)e;
        if (lane <= K && my_start >= i0 && my_start < i0 + 32) {
          const uint32_t before = (1u << (my_start - i0)) - 1u;
          my_poff[lane] = (uint16_t)(base_p + __popc(bp & before));
          my_soff[lane] = (uint16_t)(base_s + __popc(bsf & before));
        }
        base_p += __popc(bp);
        base_s += __popc(bsf);
      }
      if (lane == 0) {
        my_poff[K + 1] = (uint16_t)base_p;
        my_soff[K + 1] = (uint16_t)base_s;
      }
    }
    __syncwarp();
    SubPairBest best{0ull, 0};
    sub_pair_lane(blob, plist, poff, slist, soff, K, r0, lin, split_tab, base_s == n_static, mode, lane, best);
    // the warp's best (score, prefix word); lanes that share it hold the same prefix.  Then the first suffix of that
    // prefix's class that reaches the score: the whole warp walks the class once more.
    const unsigned long long b1 = warp_max_u64(best.key);
    if (b1 == 0) {
      sub_batch_write_empty(out, i, GPA_OK, lane);
    } else {
      const unsigned holders = __ballot_sync(0xffffffffu, best.key == b1);
      const int pe = __shfl_sync(0xffffffffu, best.pe, __ffs(holders) - 1);
      const uint32_t xp = (uint32_t)b1;
      const int score = (int)(b1 >> 32) - 1;
      const int tp = (int)__dp4a(xp, 0x01010101u, 0u);  // sum of the prefix counts
      const uint32_t first = __reduce_min_sync(0xffffffffu, sub_pair_find_suffix(blob, pe, slist, soff[K - tp], soff[K - tp + 1], r0, lin, score, lane));
      if (first == 0xFFFFFFFFu) {  // cannot happen: some lane measured `score` on this very class
        sub_batch_write_empty(out, i, GPA_ERR_CUDA, lane);
      } else if (lane == 0) {
        const uint32_t xs = blob[kSubPairCompOff / 4 + slist[first]];
        uint2 *dst = reinterpret_cast<uint2 *>(out + i);
        dst[0] = make_uint2(__byte_perm(xp, 0u, 0x0123), __byte_perm(xs, 0u, 0x0123));
        dst[1] = make_uint2(0u, 0u);
        dst[2] = make_uint2((uint32_t)score, (uint32_t)(uint16_t)d.size | ((uint32_t)(uint16_t)GPA_OK << 16));
      }
    }
    __syncwarp();  // the lists are rewritten for the warp's next request
  }
  }
}

}  // namespace
