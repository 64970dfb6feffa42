// Best-effort partition search on the GPU: kernels, context and the C ABI around them.
//
// Replaces, behind include/gpualloc.h, the reference's
//   bestEffortPolicy.Allocate   vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:34-89
//   iterateGPUPartitions        ... :209-290
//   calculateGPU{Pair,Set,Partition}Score ... :298-380
// and the per-request weight source gpuallocator.NewDevices (gpuallocator/device.go:33-72),
// which becomes a device-resident matrix installed once by gpa_set_topology().
//
// Two kernels, both built from the enumeration core in gpa_enum.cuh:
//   besteffort_search_kernel : ONE request, the whole machine.  The partition tree is cut
//       into order-numbered work items; every CTA builds the set-score table in shared
//       memory, threads walk items grid-stride, a 64-bit (score, ~item) key is max-reduced
//       warp -> CTA -> one global atomicMax, and the last CTA to finish replays the winning
//       item and writes the answer.  No second launch, no host round trip in between.
//   besteffort_batch_kernel  : MANY requests (the ContainerRequests loop), one warp each.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/gpualloc.h"
#include "gpa_enum.cuh"

namespace {

using namespace gpa;

__constant__ BinomTable c_binom;
// candidate partitions of (n available, size k), n,k <= 12, saturated at 2^32-1: work estimate of the fast batch path
__constant__ uint32_t c_parts[13][13];
// shape (n, k) -> bin of the fast batch kernel's counting sort: bins are numbered by DESCENDING work, so the sorted
// order of a tile starts with its most expensive searches (longest-processing-time-first scheduling)
__constant__ uint8_t c_shape_bin[13][13];

constexpr int kSearchBlockSmall = 256;   // score table <= 32 KB: several CTAs per SM
constexpr int kSearchBlockLarge = 1024;  // 64-128 KB table: one or two CTAs per SM, so make them wide
constexpr int kBatchWarps = 8;
constexpr int kScoreCap = 65535;                        // uint16 set-score table
constexpr int kMaxPairs = (kMaxDevices * (kMaxDevices - 1)) / 2;  // 120

// The binomial table is read with a DIFFERENT index in every lane while item ranks are decoded; constant memory
// serialises such reads one address at a time, shared memory does not, so kernels that decode stage it first.
__device__ __forceinline__ void stage_binom(BinomTable &dst, int tid, int nthreads) {
  const uint32_t *src = &c_binom.c[0][0];
  uint32_t *d = &dst.c[0][0];
  for (int i = tid; i < 32 * 17; i += nthreads) d[i] = src[i];
}

struct STab {
  const uint16_t *s;
  __device__ __forceinline__ uint32_t operator()(uint32_t mask) const { return s[mask]; }
};

struct SearchParams {
  Geom g;
  Cut cut;
  int intended;
  uint16_t wp[kMaxDevices * kMaxDevices];  // scaled weight of (earlier position a, later position b) at [a*16+b]
};

struct SearchResult {
  int32_t count;  // k, 0 = empty answer, -1 = internal inconsistency
  int32_t best_score;
  int8_t pos[32];  // positions, -1 = nil pad
  unsigned long long key;
};

// Score table of every mask a set can be: k reals, or k-P reals next to the pads.
__device__ __forceinline__ void fill_score_table(uint16_t *S, const uint16_t *wp, const Geom &g, int tid, int nthreads) {
  const uint32_t total = 1u << g.n;
  const int kp = g.P ? g.k - g.P : -1;
  for (uint32_t mask = tid; mask < total; mask += nthreads) {
    int pc = __popc(mask);
    if (pc != g.k && pc != kp) continue;
    uint32_t sum = 0;
    uint32_t ma = mask;
    while (ma) {
      int a = __ffs(ma) - 1;
      ma &= ma - 1;
      uint32_t mb = ma;
      while (mb) {
        int b = __ffs(mb) - 1;
        mb &= mb - 1;
        sum += wp[a * kMaxDevices + b];
      }
    }
    S[mask] = (uint16_t)sum;
  }
}

__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long u = __shfl_xor_sync(0xffffffffu, v, o);
    v = u > v ? u : v;
  }
  return v;
}

// One decoded work item waiting for its walk.
struct QueuedItem {
  unsigned long long item;
  uint32_t rem;
  int32_t acc_ok;  // acc | ok << 30
};
constexpr int kQueueSlots = 64;  // per warp: at most 31 left over + 32 new
constexpr int kGrabChunks = 8;   // a warp draws 8 x 32 consecutive items per visit to the global cursor

constexpr size_t search_smem_bytes(int n, int block) {
  return ((((size_t)1 << n) * sizeof(uint16_t) + 15) & ~(size_t)15) + (size_t)(block / 32) * kQueueSlots * sizeof(QueuedItem);
}

// g_scratch[0] = reduction key, g_scratch[1] = item cursor; both zero between launches.
template <int kSearchBlock>
__global__ void __launch_bounds__(kSearchBlock, kSearchBlock >= 1024 ? 1 : 4) besteffort_search_kernel(const SearchParams p, unsigned long long *g_scratch,
                                                                          unsigned int *g_ticket, SearchResult *g_res) {
  extern __shared__ __align__(16) uint16_t S[];
  __shared__ uint16_t wp[kMaxDevices * kMaxDevices];
  __shared__ unsigned long long wkey[kSearchBlock / 32];
  __shared__ int is_last;
  __shared__ BinomTable s_binom;
  stage_binom(s_binom, threadIdx.x, kSearchBlock);
  unsigned long long *g_key = g_scratch;
  unsigned long long *g_cursor = g_scratch + 1;

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  for (int i = tid; i < kMaxDevices * kMaxDevices; i += kSearchBlock) wp[i] = p.wp[i];
  __syncthreads();
  fill_score_table(S, wp, p.g, tid, kSearchBlock);
  __syncthreads();

  const STab stab{S};
  QueuedItem *q = reinterpret_cast<QueuedItem *>(reinterpret_cast<unsigned char *>(S) +
                                                 ((((size_t)1 << p.g.n) * sizeof(uint16_t) + 15) & ~(size_t)15)) +
                  (tid >> 5) * kQueueSlots;
  const int suffix = p.g.m - p.cut.d;
  unsigned long long best = 0;
  unsigned qhead = 0, qtail = 0;  // warp-uniform

  // Items are decoded one per lane, but about half of them die in the padding / `required` filters and the rest
  // walk a few hundred leaves each.  Survivors are therefore compacted into a per-warp queue and walked 32 at a
  // time, so the expensive part runs with full warps; warps draw their items from a global cursor, so a warp
  // whose items happen to be cheap simply comes back for more.
  // small searches take one chunk per visit so that every warp gets a share; big ones amortise the atomic
  const unsigned long long per_warp = p.cut.n_items / ((unsigned long long)gridDim.x * (kSearchBlock / 32) * 32ull);
  const int grab = per_warp >= (unsigned long long)kGrabChunks ? kGrabChunks : (per_warp < 1 ? 1 : (int)per_warp);
  unsigned long long base = 0;
  int chunk_left = 0;
  bool exhausted = false;
  for (;;) {
    // refill: decode chunks until a full warp of survivors is waiting (or the items ran out)
    while (qtail - qhead < 32u && !exhausted) {
      if (chunk_left == 0) {
        if (lane == 0) base = atomicAdd(g_cursor, (unsigned long long)(32 * grab));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= p.cut.n_items) {
          exhausted = true;
          break;
        }
        chunk_left = grab;
      }
      const unsigned long long item = base + (unsigned long long)(grab - chunk_left) * 32 + lane;
      chunk_left--;
      ItemState st;
      st.valid = false;
      if (item < p.cut.n_items) st = decode_item(p.g, p.cut, s_binom, stab, item, nullptr);
      const unsigned vm = __ballot_sync(0xffffffffu, st.valid);
      if (st.valid) {
        QueuedItem e;
        e.item = item;
        e.rem = st.rem;
        e.acc_ok = st.acc | (st.ok ? (1 << 30) : 0);
        q[(qtail + __popc(vm & ((1u << lane) - 1u))) & (kQueueSlots - 1)] = e;
      }
      qtail += __popc(vm);
      __syncwarp();
    }
    const unsigned waiting = qtail - qhead;
    if (waiting == 0) break;
    const unsigned take = waiting < 32u ? waiting : 32u;
    QueuedItem e;
    e.item = 0;
    e.rem = 0;
    e.acc_ok = 0;
    if (lane < take) e = q[(qhead + lane) & (kQueueSlots - 1)];
    qhead += take;
    __syncwarp();
    if (lane < take) {
      int b = -1;
      walk_dyn(suffix, p.g, stab, e.rem, e.acc_ok & 0x3FFFFFFF, (e.acc_ok >> 30) != 0, b, nullptr);
      if (b >= 0) {
        const unsigned long long kk = pack_key(b, e.item);
        best = kk > best ? kk : best;
      }
    }
  }

  best = warp_max_u64(best);
  if ((tid & 31) == 0) wkey[tid >> 5] = best;
  __syncthreads();
  if (tid == 0) {
    unsigned long long b = wkey[0];
#pragma unroll
    for (int w = 1; w < kSearchBlock / 32; w++) b = wkey[w] > b ? wkey[w] : b;
    if (b) atomicMax(g_key, b);
    __threadfence();
    unsigned int t = atomicAdd(g_ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last && tid == 0) {
    // every CTA's atomicMax is visible: its fence preceded its ticket
    unsigned long long key = atomicMax(g_key, 0ull);
    SearchResult r;
    r.key = key;
    r.best_score = key ? key_score(key) : -1;
    r.count = finalize(p.g, p.cut, s_binom, stab, key, p.intended != 0, r.pos);
    *g_res = r;
    *g_key = 0ull;  // leave the scratch clean for the next launch on this stream
    *g_cursor = 0ull;
    *g_ticket = 0u;
  }
}

// ---------------------------------------------------------------------------------------
// Subset dynamic programme for ONE large request (gpa_enum.cuh "Subset dynamic programme"): instead of visiting
// every partition, fill best[R] for every reachable remainder R, one data-parallel pass per layer, then rebuild the
// first best partition.  n = 16, k = 3: ~1 M (mask, set) pairs instead of 22.4 M partitions.  Tables live in global
// memory (128 KB + 64 KB, L2-resident); each layer is its own launch, which is the grid-wide barrier between layers.
struct DpParams {
  Geom g;
  int intended;
  uint32_t rbit;  // lowest required position as a bit, 0 if none
  uint16_t wp[kMaxDevices * kMaxDevices];
};
struct TabAny {
  const uint16_t *t;
  __device__ __forceinline__ uint32_t operator()(uint32_t mask) const { return t[mask]; }
};
struct TabOk {
  const uint16_t *t;
  uint32_t rbit;
  __device__ __forceinline__ uint32_t operator()(uint32_t mask) const { return t[dp_ok_index(mask, rbit)]; }
};
constexpr int kDpThreads = 256;

// layer 1: the score of every mask a set can be
__global__ void __launch_bounds__(kDpThreads) dp_scores_kernel(const DpParams p, uint16_t *__restrict__ T) {
  __shared__ uint16_t wp[kMaxDevices * kMaxDevices];
  for (int i = threadIdx.x; i < kMaxDevices * kMaxDevices; i += kDpThreads) wp[i] = p.wp[i];
  __syncthreads();
  fill_score_table(T, wp, p.g, blockIdx.x * kDpThreads + threadIdx.x, gridDim.x * kDpThreads);
}

// layer j in 2..m-1: best completion of every remainder of j sets, both tables
__global__ void __launch_bounds__(kDpThreads) dp_layer_kernel(const DpParams p, int j, uint16_t *T, uint16_t *Tok) {
  const uint32_t total = 1u << p.g.n;
  const TabAny tany{T};
  const TabOk tok{Tok, p.rbit};
  for (uint32_t R = blockIdx.x * kDpThreads + threadIdx.x; R < total; R += gridDim.x * kDpThreads) {
    if (dp_layer_of(p.g, R) != j) continue;
    // reads touch layers 1 and j-1 only, writes layer j only: no hazard inside the launch
    const uint32_t va = dp_value(p.g, tany, tok, R, false);
    T[R] = (uint16_t)va;
    if (p.g.rmask != 0 && (R & p.g.rmask) == p.g.rmask) Tok[dp_ok_index(R, p.rbit)] = (uint16_t)dp_value(p.g, tany, tok, R, true);
  }
}

// One level of the recursion for remainder R, spread over threads: every first set iterate() would try on R, by its
// literal rank (decode_item on the sub-request "R with j sets to go"), valued with the tables of the layer below.
// Returns the thread's best (value, ~rank) key; rank order is visiting order, so the max key is the FIRST best set.
__device__ __forceinline__ unsigned long long dp_level_key(const Geom &g, const BinomTable &binom, uint32_t R, bool need_ok,
                                                           const TabAny &tany, const TabOk &tok, int first, int stride,
                                                           Cut &cut_out) {
  const Geom sg = make_geom_sparse(R, g.k, need_ok ? g.rmask : 0u, g.tok_pos);
  Cut cut;
  cut.d = 1;
#pragma unroll
  for (int q = 0; q < kMaxCutDepth; q++) cut.radix[q] = 1;
  cut.radix[0] = binom_at(binom, sg.m * sg.k - 1, sg.k - 1);
  cut.n_items = cut.radix[0];
  cut_out = cut;
  unsigned long long best = 0;
  for (unsigned long long item = first; item < cut.n_items; item += stride) {
    const ItemState st = decode_item(sg, cut, binom, tany, item, nullptr);
    if (!st.valid) continue;
    const uint32_t rest = st.rem & g.real_mask;
    const bool need2 = need_ok && !st.ok;
    const uint32_t t = !need2 ? tany(rest) : (sg.m == 2 ? dp_ok_base(g, tany(rest), rest) : tok(rest));
    if (t == kDpInvalid) continue;
    const unsigned long long kk = pack_key(st.acc + (int)t, item);
    best = kk > best ? kk : best;
  }
  return best;
}

// layer j, ONE WARP per remainder: for layers whose remainders have many first sets (large k) a single thread
// walking them one after the other would be the tail of the whole search
__global__ void __launch_bounds__(kDpThreads) dp_layer_warp_kernel(const DpParams p, int j, uint16_t *T, uint16_t *Tok) {
  __shared__ BinomTable s_binom;
  stage_binom(s_binom, threadIdx.x, kDpThreads);
  __syncthreads();
  const uint32_t total = 1u << p.g.n;
  const TabAny tany{T};
  const TabOk tok{Tok, p.rbit};
  const int lane = threadIdx.x & 31;
  const uint32_t warps = gridDim.x * (kDpThreads / 32);
  for (uint32_t R = blockIdx.x * (kDpThreads / 32) + (threadIdx.x >> 5); R < total; R += warps) {
    if (dp_layer_of(p.g, R) != j) continue;
    Cut cut;
    unsigned long long ka = warp_max_u64(dp_level_key(p.g, s_binom, R, false, tany, tok, lane, 32, cut));
    unsigned long long ko = 0;
    const bool has_req = p.g.rmask != 0 && (R & p.g.rmask) == p.g.rmask;
    if (has_req) ko = warp_max_u64(dp_level_key(p.g, s_binom, R, true, tany, tok, lane, 32, cut));
    if (lane == 0) {
      T[R] = (uint16_t)(ka ? (uint32_t)key_score(ka) : kDpInvalid);
      if (has_req) Tok[dp_ok_index(R, p.rbit)] = (uint16_t)(ko ? (uint32_t)key_score(ko) : kDpInvalid);
    }
  }
}

// top layer + reconstruction in one CTA: level by level, all threads value the first sets of the current remainder,
// the block-wide max key names the first best one, and the remainder shrinks by it.  No serial replay.
__global__ void __launch_bounds__(1024) dp_final_kernel(const DpParams p, const uint16_t *T, const uint16_t *Tok, SearchResult *g_res) {
  __shared__ unsigned long long wkey[32];
  __shared__ unsigned long long level_key;
  __shared__ uint32_t sets[kMaxSets];
  __shared__ BinomTable s_binom;
  stage_binom(s_binom, threadIdx.x, 1024);
  __syncthreads();
  const TabAny tany{T};
  const TabOk tok{Tok, p.rbit};
  const Geom &g = p.g;
  const int tid = threadIdx.x;
  uint32_t R = g.real_mask;
  bool need = g.rmask != 0;
  bool have = true;
  int top_score = -1;
  for (int level = 0; level < g.m; level++) {
    const int left = g.m - level;
    const bool has_tok = g.P != 0 && (__popc(R) % g.k) != 0;
    if (left == 1) {
      if (need && !((R & g.rmask) == g.rmask && !has_tok)) have = false;
      if (tid == 0) sets[level] = R | (has_tok ? g.tok_bit : 0u);
      if (level == 0 && have) top_score = (int)tany(R);
      break;
    }
    Cut cut;
    unsigned long long best = warp_max_u64(dp_level_key(g, s_binom, R, need, tany, tok, tid, 1024, cut));
    if ((tid & 31) == 0) wkey[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < 32; w++) best = wkey[w] > best ? wkey[w] : best;
      level_key = best;
    }
    __syncthreads();
    const unsigned long long key = level_key;
    if (key == 0) {
      have = false;
      break;
    }
    if (level == 0) top_score = key_score(key);
    // every thread decodes the winner, so R and `need` stay uniform without another broadcast
    const Geom sg = make_geom_sparse(R, g.k, need ? g.rmask : 0u, g.tok_pos);
    uint32_t set = 0;
    const ItemState st = decode_item(sg, cut, s_binom, tany, key_item(key), &set);
    if (tid == 0) sets[level] = set;
    R = st.rem & g.real_mask;
    need = need && !st.ok;
    __syncthreads();  // wkey / level_key are reused by the next level
  }
  __syncthreads();
  if (tid == 0) {
    SearchResult r;
    r.key = 0;
    r.count = 0;
    r.best_score = have ? top_score : -1;
    if (have) {
      uint32_t local_sets[kMaxSets];
      for (int q = 0; q < g.m; q++) local_sets[q] = sets[q];
      uint32_t best_set = 0;
      if (choose_from_partition(g, tany, local_sets, p.intended != 0, best_set) > 0) {
        int o = 0;
        uint32_t reals = best_set & g.real_mask;
        while (reals) {
          r.pos[o++] = (int8_t)(__ffs(reals) - 1);
          reals &= reals - 1;
        }
        if (best_set & g.tok_bit)
          for (int q = 0; q < g.P; q++) r.pos[o++] = -1;
        r.count = o;
      }
    }
    *g_res = r;
  }
}

// ---------------------------------------------------------------------------------------
// Batch: one warp per container request.

constexpr int8_t kResDefer = -126;  // internal: "left to the general kernel"; never leaves the library

struct BatchTopo {
  const uint16_t *wq;   // nt*nt scaled weights
  const uint8_t *asym;  // nt*nt, 1 where the link counts of (i,j) and (j,i) differ
  int nt;
};

__device__ __forceinline__ void write_result(gpa_result *out, const int8_t *dev) {
  const int4 *src = reinterpret_cast<const int4 *>(dev);
  *reinterpret_cast<int4 *>(out->dev) = *src;
}

// General batch kernel: any `available` order, per-request score table, one warp per request.
// With deferred_only it answers just the records the fast kernel marked GPA_RES_DEFER: each
// warp scans 32 result records at a time (one coalesced load) and works on the marked ones.
__global__ void __launch_bounds__(kBatchWarps * 32) besteffort_batch_kernel(const gpa_request *__restrict__ reqs, int count,
                                                                            int n_cap, int intended, BatchTopo topo,
                                                                            gpa_result *out, int deferred_only) {
  extern __shared__ __align__(16) uint16_t smem[];
  __shared__ BinomTable s_binom;
  stage_binom(s_binom, threadIdx.x, kBatchWarps * 32);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  uint16_t *S = smem + (size_t)warp * ((1u << n_cap) + kMaxDevices * kMaxDevices);
  uint16_t *wp = S + (1u << n_cap);
  const STab stab{S};

  auto process = [&](int r) {
    // 32-byte record, same address in every lane: one broadcast transaction each
    const uint4 lo = __ldg(reinterpret_cast<const uint4 *>(reqs + r));
    const uint4 hi = __ldg(reinterpret_cast<const uint4 *>(reqs + r) + 1);
    const uint32_t availw[4] = {lo.x, lo.y, lo.z, lo.w};
    const uint32_t rmask = hi.x & 0xFFFFu;
    const int n = (hi.x >> 16) & 0xFF;
    const int size = (hi.x >> 24) & 0xFF;
    const int flags = hi.y & 0xFF;
    const int n_req = __popc(rmask) + ((hi.y >> 8) & 0xFF);
    auto dev_at = [&](int pidx) -> int { return (availw[pidx >> 2] >> ((pidx & 3) * 8)) & 0xFF; };

    __align__(16) int8_t res[GPA_MAX_AVAILABLE];
#pragma unroll
    for (int i = 0; i < GPA_MAX_AVAILABLE; i++) res[i] = GPA_RES_END;
    int err = 0;
    bool done = false;

    // guards of besteffort_policy.go:35-45, in the reference's order
    if (size <= 0 || n < size || n_req > size) done = true;
    if (!done) {
      if (n > n_cap || n > GPA_BATCH_MAX_AVAILABLE) {
        err = GPA_ERR_TOO_LARGE;
      } else {
        unsigned long long seen = 0;
        for (int a = 0; a < n; a++) {
          int d = dev_at(a);
          if (d >= topo.nt) {
            err = GPA_ERR_INVALID;
          } else {
            if (seen >> d & 1ull) err = GPA_ERR_INVALID;  // the same device twice
            seen |= 1ull << d;
          }
        }
        if ((rmask >> n) != 0) err = GPA_ERR_INVALID;
      }
      if (err) done = true;
    }
    if (!done && (flags & GPA_REQ_FOREIGN)) done = true;  // no candidate can contain a foreign pointer (:53)
    if (!done && size == 1) {
      // :219-224 — single-device candidates in `available` order, first accepted wins (score 0 never beats 0)
      res[0] = (int8_t)dev_at(rmask ? __ffs(rmask) - 1 : 0);
      done = true;
    }
    if (!done) {
      // link-count symmetry over the available devices (where calculateGPUPairScore would panic, :307-310)
      int bad = 0;
      for (int i = lane; i < n * n; i += 32) {
        int a = i / n, b = i - a * n;
        int da = dev_at(a), db = dev_at(b);
        bad |= topo.asym[da * topo.nt + db];
        if (a < b) wp[a * kMaxDevices + b] = topo.wq[da * topo.nt + db];
      }
      bad = __any_sync(0xffffffffu, bad);
      if (bad) {
        err = GPA_ERR_ASYMMETRIC;
        done = true;
      }
    }
    if (!done) {
      __syncwarp();
      const Geom g = make_geom(n, size, rmask);
      fill_score_table(S, wp, g, lane, 32);
      __syncwarp();
      const Cut cut = choose_cut(g, s_binom, 64);
      unsigned long long best = 0;
      for (unsigned long long item = lane; item < cut.n_items; item += 32) {
        int sc = search_item(g, cut, s_binom, stab, item);
        if (sc >= 0) {
          unsigned long long kk = pack_key(sc, item);
          best = kk > best ? kk : best;
        }
      }
      best = warp_max_u64(best);
      if (lane == 0) {
        int8_t pos[32];
        int cnt = finalize(g, cut, s_binom, stab, best, intended != 0, pos);
        if (cnt < 0) {
          err = GPA_ERR_INVALID;
        } else {
          for (int i = 0; i < cnt; i++) res[i] = pos[i] < 0 ? (int8_t)GPA_NIL : (int8_t)dev_at(pos[i]);
        }
      }
      __syncwarp();  // S / wp are reused by the next request of this warp
    }
    if (lane == 0) {
      if (err) {
        res[0] = GPA_RES_ERROR;
        res[1] = (int8_t)err;
      }
      write_result(out + r, res);
    }
  };

  if (!deferred_only) {
    for (int r = blockIdx.x * kBatchWarps + warp; r < count; r += gridDim.x * kBatchWarps) process(r);
  } else {
    for (long long base = ((long long)blockIdx.x * kBatchWarps + warp) * 32; base < count;
         base += (long long)gridDim.x * kBatchWarps * 32) {
      const long long r = base + lane;
      const bool marked = r < count && out[r].dev[0] == kResDefer;
      unsigned todo = __ballot_sync(0xffffffffu, marked);
      while (todo) {
        const int b = __ffs(todo) - 1;
        todo &= todo - 1;
        process((int)(base + b));
        __syncwarp();
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// Fast batch kernel for requests in CANONICAL order: the devices of every request are listed
// in one node-wide order (the reference always does: stripReplicas sorts the UUIDs,
// cmd/nvidia-device-plugin/replica.go:43).  Devices are relabelled by their rank in that
// order, so a request is just a label MASK, positions are labels, and one score table over
// all 2^nt label masks serves every request of the node: no per-request table, no gather.
//
// One LANE answers one request.  Requests of different shapes (n available, size k) walk
// differently shaped trees, and lanes of a warp that take different paths run one after the
// other, so each CTA first BINS its tile of kTile requests by shape in shared memory
// (counting sort on n*13+k) and then hands the sorted order to its threads: the 32 lanes of
// a warp get neighbours in that order, i.e. (almost always) the same shape.
//   pass 1  coalesced load, parse, guards; trivial answers are stored at once; searches leave
//           a compact record {label mask, required mask, size} and take a rank in their bin;
//   pass 2  one warp turns the bin counts into offsets;
//   pass 3  records are threaded into the sorted index;
//   pass 4  lane p takes sorted record p: search, epilogue, 16-byte store.
// Anything unusual (order foreign to the node, unknown device, more than 12 devices) is marked
// GPA_RES_DEFER for the general kernel.
constexpr int kFastMaxTopo = 12;
constexpr int kTile = 4096;
constexpr int kFastThreads = 256;
constexpr int kShapeBins = 13 * 13 + 7;  // n*13 + k for n,k <= 12, padded to 176

struct FastTopo {
  const uint16_t *s_full;  // [1 << nt] scaled set scores over label masks
  int nt;
  uint8_t order[16];  // label -> NVML index
  uint8_t rank[GPA_MAX_TOPOLOGY];  // NVML index -> label, 0xFF = not on this node
};

__device__ __noinline__ int fast_search_item(const Geom &g, const Cut &cut, const uint16_t *S, unsigned long long item) {
  return search_item(g, cut, c_binom, STab{S}, item);
}
__device__ __noinline__ int fast_finalize(const Geom &g, const Cut &cut, const uint16_t *S, unsigned long long key, bool intended,
                                          uint32_t &best_set) {
  return finalize_set(g, cut, c_binom, STab{S}, key, intended, best_set);
}

// Whole-tree search of a shallow request (m <= 4 sets) in one lane; returns like finalize_set.
__device__ __forceinline__ int fast_tracked_search(const Geom &g, const uint16_t *S, bool intended, uint32_t &best_set) {
  const STab stab{S};
  Track tr;
  tr.best = -1;
  uint32_t path[kMaxSuffix];
  const uint32_t all = g.real_mask | g.tok_bit;
  switch (g.m) {
    case 1: walk_track<1, 1>(g, stab, all, 0, false, path, tr); break;
    case 2: walk_track<2, 2>(g, stab, all, 0, false, path, tr); break;
    case 3: walk_track<3, 3>(g, stab, all, 0, false, path, tr); break;
    default: walk_track<4, 4>(g, stab, all, 0, false, path, tr); break;
  }
  best_set = 0;
  if (tr.best < 0) return 0;  // no accepted partition: bestPartition == nil (:72-74)
  return choose_from_partition(g, stab, tr.sets, intended, best_set);
}

__device__ __forceinline__ void res_set_byte(unsigned long long &r0, unsigned long long &r1, int i, int v) {
  unsigned long long b = (unsigned long long)(v & 0xFF);
  if (i < 8) r0 = (r0 & ~(0xFFull << (8 * i))) | (b << (8 * i));
  else r1 = (r1 & ~(0xFFull << (8 * (i - 8)))) | (b << (8 * (i - 8)));
}
__device__ __forceinline__ void res_store(gpa_result *dst, unsigned long long r0, unsigned long long r1) {
  uint4 v;
  v.x = (uint32_t)r0;
  v.y = (uint32_t)(r0 >> 32);
  v.z = (uint32_t)r1;
  v.w = (uint32_t)(r1 >> 32);
  *reinterpret_cast<uint4 *>(dst->dev) = v;
}

constexpr size_t fast_smem_bytes(int nt) {
  return ((size_t)1 << nt) * sizeof(uint16_t)  // S
         + (size_t)kTile * sizeof(uint32_t)    // w_mr
         + (size_t)kTile * sizeof(uint16_t) * 2  // w_rank, sorted_idx
         + (size_t)kTile * 2                     // w_k, w_key
         + (size_t)(kShapeBins + 1) * sizeof(uint32_t) * 2;  // hist, offs
}

// kMasks = false: 32-byte gpa_request in, 16-byte gpa_result out.
// kMasks = true : 8-byte gpa_mask_request in (the caller already holds label masks), 4-byte gpa_mask_result out.
template <bool kMasks>
__global__ void __launch_bounds__(kFastThreads) besteffort_batch_fast_kernel(const void *__restrict__ reqs_v, int count,
                                                                             int intended, const FastTopo ft,
                                                                             void *__restrict__ out_v) {
  const gpa_request *__restrict__ reqs = static_cast<const gpa_request *>(reqs_v);
  gpa_result *__restrict__ out = static_cast<gpa_result *>(out_v);
  const uint2 *__restrict__ mreqs = static_cast<const uint2 *>(reqs_v);
  uint32_t *__restrict__ mout = static_cast<uint32_t *>(out_v);
  extern __shared__ __align__(16) unsigned char fsmem[];
  uint32_t *w_mr = reinterpret_cast<uint32_t *>(fsmem);
  uint32_t *hist = w_mr + kTile;
  uint32_t *offs = hist + (kShapeBins + 1);
  uint16_t *S = reinterpret_cast<uint16_t *>(offs + (kShapeBins + 1));
  uint16_t *w_rank = S + ((size_t)1 << ft.nt);
  uint16_t *sorted_idx = w_rank + kTile;
  uint8_t *w_k = reinterpret_cast<uint8_t *>(sorted_idx + kTile);
  uint8_t *w_key = w_k + kTile;

  const int tid = threadIdx.x;
  for (uint32_t i = tid; i < (1u << ft.nt); i += kFastThreads) S[i] = ft.s_full[i];

  const long long n_tiles = ((long long)count + kTile - 1) / kTile;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long tile_base = tile * kTile;
    for (int b = tid; b <= kShapeBins; b += kFastThreads) hist[b] = 0;
    __syncthreads();

    // ---- pass 1: parse; answer the trivial ones; bin the searches by shape
#pragma unroll 1
    for (int i = tid; i < kTile; i += kFastThreads) {
      const long long r = tile_base + i;
      int bin = 0xFF;
      if (r < count) {
        if constexpr (kMasks) {
          // {avail_mask:16, req_mask:16} {size:8, flags:8, n_req_extra:8, reserved:8}
          const uint2 rec = __ldg(mreqs + r);
          const uint32_t mask = rec.x & 0xFFFFu, rlab_in = rec.x >> 16;
          const int size = rec.y & 0xFF, flags = (rec.y >> 8) & 0xFF;
          const int n = __popc(mask);
          const int n_req = __popc(rlab_in) + ((rec.y >> 16) & 0xFF);
          uint32_t result = 0;  // {chosen_mask:16, n_nil:8, status:8}: the empty answer
          bool search = false;
          if ((mask >> ft.nt) != 0 || (rlab_in >> ft.nt) != 0) {
            result = (uint32_t)(GPA_ERR_INVALID & 0xFF) << 24;
          } else if (!(size <= 0 || n < size || n_req > size)) {  // guards of besteffort_policy.go:35-45
            if ((flags & GPA_REQ_FOREIGN) || (rlab_in & ~mask) != 0) {
              // a required device that is not available is in no set (:53) -> empty answer
            } else if (size == 1) {
              result = rlab_in ? (rlab_in & (0u - rlab_in)) : (mask & (0u - mask));  // :219-224
            } else {
              search = true;
              w_mr[i] = mask | (rlab_in << 16);
              w_k[i] = (uint8_t)size;
              bin = c_shape_bin[n][size];
              w_rank[i] = (uint16_t)atomicAdd(&hist[bin], 1u);
            }
          }
          if (!search) mout[r] = result;
        } else {
        const uint4 lo = __ldg(reinterpret_cast<const uint4 *>(reqs + r));
        const uint4 hi = __ldg(reinterpret_cast<const uint4 *>(reqs + r) + 1);
        const uint32_t rmask = hi.x & 0xFFFFu;
        const int n = (hi.x >> 16) & 0xFF;
        const int size = (hi.x >> 24) & 0xFF;
        const int flags = hi.y & 0xFF;
        const int n_req = __popc(rmask) + ((hi.y >> 8) & 0xFF);
        unsigned long long r0 = 0x8080808080808080ull, r1 = 0x8080808080808080ull;  // 16 x GPA_RES_END
        bool search = false;
        // guards of besteffort_policy.go:35-45
        if (!(size <= 0 || n < size || n_req > size)) {
          if (n > kFastMaxTopo || (rmask >> n) != 0) {
            res_set_byte(r0, r1, 0, kResDefer);
          } else {
            // relabel: device -> rank in the node order; the list must be strictly ascending
            const uint32_t w[4] = {lo.x, lo.y, lo.z, lo.w};
            uint32_t mask = 0, rlab = 0;
            int prev = -1, first_dev = 0, first_req_dev = -1;
            bool defer = false;
#pragma unroll
            for (int p = 0; p < kFastMaxTopo; p++) {
              if (p < n) {
                const int d = (w[p >> 2] >> ((p & 3) * 8)) & 0xFF;
                const int lab = d < GPA_MAX_TOPOLOGY ? ft.rank[d] : 0xFF;
                defer = defer || lab == 0xFF || lab <= prev;
                prev = lab;
                mask |= 1u << (lab & 31);
                if (rmask >> p & 1u) {
                  rlab |= 1u << (lab & 31);
                  if (first_req_dev < 0) first_req_dev = d;
                }
                if (p == 0) first_dev = d;
              }
            }
            if (defer) {
              res_set_byte(r0, r1, 0, kResDefer);
            } else if (flags & GPA_REQ_FOREIGN) {
              // a required pointer that is in no set: every candidate is rejected (:53) -> empty answer
            } else if (size == 1) {
              // :219-224 — first accepted single-device candidate
              res_set_byte(r0, r1, 0, first_req_dev >= 0 ? first_req_dev : first_dev);
            } else {
              search = true;
              w_mr[i] = (mask & 0xFFFFu) | (rlab << 16);
              w_k[i] = (uint8_t)size;
              bin = c_shape_bin[n][size];
              w_rank[i] = (uint16_t)atomicAdd(&hist[bin], 1u);
            }
          }
        }
        if (!search) res_store(out + r, r0, r1);
        }
      }
      w_key[i] = (uint8_t)bin;
    }
    __syncthreads();

    // ---- pass 2: bin counts -> offsets (one warp, 6 bins per lane)
    if (tid < 32) {
      constexpr int kPer = (kShapeBins + 31) / 32;
      uint32_t local[kPer];
      uint32_t sum = 0;
#pragma unroll
      for (int q = 0; q < kPer; q++) {
        const int b = tid * kPer + q;
        local[q] = b < kShapeBins ? hist[b] : 0u;
        sum += local[q];
      }
      uint32_t incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (tid >= o) incl += v;
      }
      uint32_t run = incl - sum;
#pragma unroll
      for (int q = 0; q < kPer; q++) {
        const int b = tid * kPer + q;
        if (b < kShapeBins) offs[b] = run;
        run += local[q];
      }
      if (tid == 31) offs[kShapeBins] = incl;
    }
    __syncthreads();

    // ---- pass 3: thread the records into shape order
#pragma unroll 1
    for (int i = tid; i < kTile; i += kFastThreads) {
      const int bin = w_key[i];
      if (bin != 0xFF) sorted_idx[offs[bin] + w_rank[i]] = (uint16_t)i;
    }
    __syncthreads();

    // ---- pass 4: one lane per search, neighbours in shape order.  Warps draw chunks of 32 sorted records from
    // a shared cursor (hist[kShapeBins], zero after pass 1), heaviest shapes first, so no warp is left holding
    // a long tail while the others wait at the barrier.
    const int n_search = (int)offs[kShapeBins];
    const int lane = tid & 31;
#pragma unroll 1
    for (;;) {
      int p0 = 0;
      if (lane == 0) p0 = (int)atomicAdd(&hist[kShapeBins], 32u);
      p0 = __shfl_sync(0xffffffffu, p0, 0);
      if (p0 >= n_search) break;
      const int p = p0 + lane;
      if (p >= n_search) continue;
      const int i = sorted_idx[p];
      const uint32_t mr = w_mr[i];
      const Geom g = make_geom_sparse(mr & 0xFFFFu, w_k[i], mr >> 16, ft.nt);
      unsigned long long r0 = 0x8080808080808080ull, r1 = 0x8080808080808080ull;
      uint32_t best_set = 0;
      int rc;
      if (g.m <= kMaxSuffix) {
        // the whole tree in this lane, best partition tracked in registers
        rc = fast_tracked_search(g, S, intended != 0, best_set);
      } else {
        // deep trees (k = 2 on 9+ devices): the first m - 4 levels come from an item index, then replay
        const Cut cut = choose_cut(g, c_binom, 1);
        unsigned long long key = 0;
        for (unsigned long long item = 0; item < cut.n_items; item++) {
          const int sc = fast_search_item(g, cut, S, item);
          if (sc >= 0) {
            const unsigned long long kk = pack_key(sc, item);
            key = kk > key ? kk : key;
          }
        }
        rc = fast_finalize(g, cut, S, key, intended != 0, best_set);
      }
      if constexpr (kMasks) {
        uint32_t result = 0;
        if (rc < 0) result = (uint32_t)(GPA_ERR_INVALID & 0xFF) << 24;
        else if (rc > 0) result = (best_set & g.real_mask) | ((best_set & g.tok_bit) ? ((uint32_t)g.P << 16) : 0u);
        mout[tile_base + i] = result;
      } else {
        if (rc < 0) {
          res_set_byte(r0, r1, 0, GPA_RES_ERROR);
          res_set_byte(r0, r1, 1, GPA_ERR_INVALID);
        } else if (rc > 0) {
          int o = 0;
          uint32_t reals = best_set & g.real_mask;
          while (reals) {
            const int lab = __ffs(reals) - 1;
            reals &= reals - 1;
            res_set_byte(r0, r1, o++, ft.order[lab]);
          }
          if (best_set & g.tok_bit)
            for (int q = 0; q < g.P; q++) res_set_byte(r0, r1, o++, GPA_NIL);
        }
        res_store(out + tile_base + i, r0, r1);
      }
    }
    __syncthreads();  // the tile's shared arrays are reused by the next tile
  }
}

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
  SearchResult *d_res = nullptr;
  SearchResult *h_res = nullptr;  // pinned
  // topology (host copy + device tables)
  Topology topo;
  bool has_topo = false;
  uint16_t *d_wq = nullptr;
  uint8_t *d_asym = nullptr;
  uint16_t *d_dp_any = nullptr;  // subset DP tables of gpa_besteffort(): 2^16 and 2^15 uint16
  uint16_t *d_dp_ok = nullptr;
  uint16_t *d_sfull = nullptr;  // fast batch path: set scores over label masks, nullptr = path unavailable
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

int launch_batch(gpa_ctx *ctx, const gpa_request *d_reqs, int count, int max_avail, int mode_flags, gpa_result *d_out,
                 cudaStream_t stream) {
  if (count == 0) return GPA_OK;
  if (!ctx->d_sfull) return launch_general(ctx, d_reqs, count, max_avail, mode_flags, d_out, stream, 0);
  // canonical-order fast path, then the general kernel mops up whatever was deferred
  int rc = launch_fast(ctx, d_reqs, count, mode_flags, d_out, stream, false);
  if (rc != GPA_OK) return rc;
  return launch_general(ctx, d_reqs, count, max_avail, mode_flags, d_out, stream, 1);
}

// (Re)builds what the fast batch path needs from the installed topology and device order: the
// label <-> NVML index maps and the score table over all label masks.  The path is available
// for nodes of up to kFastMaxTopo devices whose link counts are symmetric.
int rebuild_fast(gpa_ctx *ctx) {
  Topology &T = ctx->topo;
  if (ctx->d_sfull) {
    cudaFree(ctx->d_sfull);
    ctx->d_sfull = nullptr;
  }
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
  ctx->fast = f;
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
  CREATE_TRY(cudaMalloc(&ctx->d_res, sizeof(SearchResult)));
  CREATE_TRY(cudaMalloc(&ctx->d_dp_any, ((size_t)1 << kMaxDevices) * sizeof(uint16_t)));
  CREATE_TRY(cudaMalloc(&ctx->d_dp_ok, ((size_t)1 << (kMaxDevices - 1)) * sizeof(uint16_t)));
  CREATE_TRY(cudaMallocHost(&ctx->h_res, sizeof(SearchResult)));
  CREATE_TRY(cudaMemset(ctx->d_key, 0, 2 * sizeof(unsigned long long)));
  CREATE_TRY(cudaMemset(ctx->d_ticket, 0, sizeof(unsigned int)));
  fill_binom(ctx->binom);
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
    // bins by descending work; at most 13*13 = 169 shapes < kShapeBins
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
  if (ctx->d_res) cudaFree(ctx->d_res);
  if (ctx->d_dp_any) cudaFree(ctx->d_dp_any);
  if (ctx->d_dp_ok) cudaFree(ctx->d_dp_ok);
  if (ctx->h_res) cudaFreeHost(ctx->h_res);
  if (ctx->d_wq) cudaFree(ctx->d_wq);
  if (ctx->d_asym) cudaFree(ctx->d_asym);
  if (ctx->d_sfull) cudaFree(ctx->d_sfull);
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
  // Requests are serialised on ctx->mu and each one finishes before it returns, so no
  // kernel can still be reading the old tables here.
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

int gpa_besteffort(gpa_ctx *ctx, const int32_t *avail, int n_avail, const int32_t *req, int n_req, int size,
                   int mode_flags, int32_t *out, int *n_out) {
  if (!ctx || !n_out || n_avail < 0 || n_req < 0 || (n_avail > 0 && !avail) || (n_req > 0 && !req)) return GPA_ERR_INVALID;
  *n_out = 0;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  const Topology &T = ctx->topo;
  memset(&ctx->stats, 0, sizeof ctx->stats);
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

  // where calculateGPUPairScore panics (:307-310); see DESIGN.md for the (conservative) scope of this check
  for (int a = 0; a < n_avail; a++)
    for (int b = 0; b < n_avail; b++)
      if (a != b && T.asym[avail[a] * T.n + avail[b]]) return GPA_ERR_ASYMMETRIC;

  SearchParams sp;
  memset(&sp, 0, sizeof sp);
  sp.g = make_geom(n_avail, size, rmask);
  sp.intended = (mode_flags & GPA_MODE_INTENDED) ? 1 : 0;
  for (int a = 0; a < n_avail; a++)
    for (int b = a + 1; b < n_avail; b++) sp.wp[a * kMaxDevices + b] = T.wq[avail[a] * T.n + avail[b]];
  const uint64_t parts = partition_count(sp.g, ctx->binom);
  if (parts > GPA_MAX_PARTITIONS) return GPA_ERR_TOO_LARGE;

  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  // Enumeration visits every partition in one launch; the subset DP needs m launches but far less work.  The DP wins
  // once the tree is large (crossover measured around 10^5 partitions on B200).
  const bool use_dp = (mode_flags & GPA_MODE_FORCE_DP) || (!(mode_flags & GPA_MODE_FORCE_ENUM) && parts > 100000ull);
  int grid = 0, block = 0;
  size_t smem = 0;
  if (use_dp) {
    DpParams dp;
    memset(&dp, 0, sizeof dp);
    dp.g = sp.g;
    dp.intended = sp.intended;
    dp.rbit = rmask ? (rmask & (0u - rmask)) : 0u;
    memcpy(dp.wp, sp.wp, sizeof dp.wp);
    const uint32_t total = 1u << n_avail;
    grid = (int)((total + kDpThreads - 1) / kDpThreads);
    if (grid > ctx->sm_count * 8) grid = ctx->sm_count * 8;
    block = kDpThreads;
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    dp_scores_kernel<<<grid, kDpThreads, 0, ctx->stream>>>(dp, ctx->d_dp_any);
    ctx->launches++;
    for (int j = 2; j < dp.g.m; j++) {
      // A thread per remainder walks its first sets one after the other; a warp per remainder spreads them over its lanes
      // by literal rank.  Measured on B200 (tools/search_sweep.py): the warp form wins once a remainder has ~100 first sets.
      if (binom_at(ctx->binom, j * dp.g.k - 1, dp.g.k - 1) >= 96u) {
        int wgrid = (int)((total + kDpThreads / 32 - 1) / (kDpThreads / 32));
        if (wgrid > ctx->sm_count * 8) wgrid = ctx->sm_count * 8;
        dp_layer_warp_kernel<<<wgrid, kDpThreads, 0, ctx->stream>>>(dp, j, ctx->d_dp_any, ctx->d_dp_ok);
      } else {
        dp_layer_kernel<<<grid, kDpThreads, 0, ctx->stream>>>(dp, j, ctx->d_dp_any, ctx->d_dp_ok);
      }
      ctx->launches++;
    }
    dp_final_kernel<<<1, 1024, 0, ctx->stream>>>(dp, ctx->d_dp_any, ctx->d_dp_ok, ctx->d_res);
    ctx->launches++;
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
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_res, ctx->d_res, sizeof(SearchResult), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);

  ctx->stats.partitions = parts;
  ctx->stats.items = sp.cut.n_items;
  ctx->stats.cut_depth = sp.cut.d;
  ctx->stats.sets = sp.g.m;
  ctx->stats.padding = sp.g.P;
  ctx->stats.best_score = ctx->h_res->best_score;
  ctx->stats.grid = grid;
  ctx->stats.block = block;
  ctx->stats.smem_bytes = (int)smem;
  ctx->stats.kernel_ms = ms;
  ctx->stats.algorithm = use_dp ? 2 : 1;

  const SearchResult &r = *ctx->h_res;
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
  // Pinned caller buffers (gpa_alloc_host / cudaHostRegister) are DMA'd directly; pageable
  // ones are staged through the context's pinned mailbox so that every copy stays async.
  cudaPointerAttributes pa_in{}, pa_out{};
  bool pinned = cudaPointerGetAttributes(&pa_in, reqs) == cudaSuccess && pa_in.type == cudaMemoryTypeHost &&
                cudaPointerGetAttributes(&pa_out, out) == cudaSuccess && pa_out.type == cudaMemoryTypeHost;
  cudaGetLastError();
  int rc = ensure_batch_capacity(ctx, (size_t)count, !pinned);
  if (rc != GPA_OK) return rc;
  if (max_avail == 0) {  // caller does not know: one pass over the records
    for (int i = 0; i < count; i++) {
      int na = reqs[i].n_avail;
      if (na > max_avail && na <= GPA_BATCH_MAX_AVAILABLE) max_avail = na;
    }
  }
  if (!pinned) memcpy(ctx->h_reqs, reqs, (size_t)count * sizeof(gpa_request));
  const gpa_request *src = pinned ? reqs : ctx->h_reqs;
  gpa_result *dst = pinned ? out : ctx->h_out;
  // Chunks ride a ring of streams: while chunk i is in the kernel, chunk i+1 is on the H2D
  // copy engine and chunk i-1 on the D2H one.
  const int kChunk = 1 << 18;
  int c = 0;
  for (int off = 0; off < count; off += kChunk, c++) {
    int cnt = count - off < kChunk ? count - off : kChunk;
    cudaStream_t st = ctx->pipe[c & 3];
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->d_reqs + off, src + off, (size_t)cnt * sizeof(gpa_request), cudaMemcpyHostToDevice, st));
    rc = launch_batch(ctx, ctx->d_reqs + off, cnt, max_avail, mode_flags, ctx->d_out + off, st);
    if (rc != GPA_OK) break;
    CUDA_TRY(ctx, cudaMemcpyAsync(dst + off, ctx->d_out + off, (size_t)cnt * sizeof(gpa_result), cudaMemcpyDeviceToHost, st));
  }
  for (auto &ps : ctx->pipe) {
    cudaError_t e = cudaStreamSynchronize(ps);
    if (e != cudaSuccess && rc == GPA_OK) {
      ctx->last_error = std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e);
      rc = GPA_ERR_CUDA;
    }
  }
  if (rc != GPA_OK) return rc;
  if (!pinned) memcpy(out, ctx->h_out, (size_t)count * sizeof(gpa_result));
  return GPA_OK;
}

int gpa_besteffort_batch_masks_device(gpa_ctx *ctx, const gpa_mask_request *d_reqs, int count, int mode_flags,
                                      gpa_mask_result *d_out, void *stream) {
  if (!ctx || count < 0 || (count > 0 && (!d_reqs || !d_out))) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (!ctx->d_sfull) return GPA_ERR_TOO_LARGE;  // masks need the node-wide score table: <= 12 GPUs, symmetric link counts
  if (count == 0) return GPA_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return launch_fast(ctx, d_reqs, count, mode_flags, d_out, (cudaStream_t)stream, true);
}

int gpa_besteffort_batch_masks(gpa_ctx *ctx, const gpa_mask_request *reqs, int count, int mode_flags, gpa_mask_result *out) {
  if (!ctx || count < 0 || (count > 0 && (!reqs || !out))) return GPA_ERR_INVALID;
  if (count == 0) return GPA_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!ctx->has_topo) return GPA_ERR_NO_TOPOLOGY;
  if (!ctx->d_sfull) return GPA_ERR_TOO_LARGE;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  cudaPointerAttributes pa_in{}, pa_out{};
  bool pinned = cudaPointerGetAttributes(&pa_in, reqs) == cudaSuccess && pa_in.type == cudaMemoryTypeHost &&
                cudaPointerGetAttributes(&pa_out, out) == cudaSuccess && pa_out.type == cudaMemoryTypeHost;
  cudaGetLastError();
  // the staging buffers are sized for the 32/16-byte records, so count/4 of them hold count mask records
  int rc = ensure_batch_capacity(ctx, (size_t)count / 4 + 1, !pinned);
  if (rc != GPA_OK) return rc;
  gpa_mask_request *d_in = reinterpret_cast<gpa_mask_request *>(ctx->d_reqs);
  gpa_mask_result *d_res = reinterpret_cast<gpa_mask_result *>(ctx->d_out);
  gpa_mask_request *h_in = reinterpret_cast<gpa_mask_request *>(ctx->h_reqs);
  gpa_mask_result *h_res = reinterpret_cast<gpa_mask_result *>(ctx->h_out);
  if (!pinned) memcpy(h_in, reqs, (size_t)count * sizeof(gpa_mask_request));
  const gpa_mask_request *src = pinned ? reqs : h_in;
  gpa_mask_result *dst = pinned ? out : h_res;
  const int kChunk = 1 << 20;
  int c = 0;
  for (int off = 0; off < count; off += kChunk, c++) {
    int cnt = count - off < kChunk ? count - off : kChunk;
    cudaStream_t st = ctx->pipe[c & 3];
    CUDA_TRY(ctx, cudaMemcpyAsync(d_in + off, src + off, (size_t)cnt * sizeof(gpa_mask_request), cudaMemcpyHostToDevice, st));
    rc = launch_fast(ctx, d_in + off, cnt, mode_flags, d_res + off, st, true);
    if (rc != GPA_OK) break;
    CUDA_TRY(ctx, cudaMemcpyAsync(dst + off, d_res + off, (size_t)cnt * sizeof(gpa_mask_result), cudaMemcpyDeviceToHost, st));
  }
  for (auto &ps : ctx->pipe) {
    cudaError_t e = cudaStreamSynchronize(ps);
    if (e != cudaSuccess && rc == GPA_OK) {
      ctx->last_error = std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e);
      rc = GPA_ERR_CUDA;
    }
  }
  if (rc != GPA_OK) return rc;
  if (!pinned) memcpy(out, h_res, (size_t)count * sizeof(gpa_mask_result));
  return GPA_OK;
}

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
int noalloc_setup(int dev, char *errbuf, int errbuf_len) {
  NoallocDevice &d = g_noalloc_dev[dev];
  if (d.ready) return GPA_OK;
  void *p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
  if (e != cudaSuccess || p == nullptr) {
    na_msg(errbuf, errbuf_len, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
    return GPA_ERR_CUDA;
  }
  BinomTable b;
  fill_binom(b);
  e = cudaMemcpyToSymbol(c_binom, &b, sizeof b);
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
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      wq[i * n + j] = i == j ? 0 : (uint16_t)(W[i * n + j] / g);
      asym[i * n + j] = i != j && nlinks[i * n + j] != nlinks[j * n + i];
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
  int rc = noalloc_setup(cuda_device, errbuf, errbuf_len);
  if (rc != GPA_OK) return rc;
  const NoallocDevice &d = g_noalloc_dev[cuda_device];
  cudaStream_t st = (cudaStream_t)stream;
  uint16_t *d_wq = reinterpret_cast<uint16_t *>(d.base + kNaWqOff);
  uint8_t *d_asym = d.base + kNaAsymOff;
  gpa_request *d_reqs = reinterpret_cast<gpa_request *>(d.base + kNaReqOff);
  gpa_result *d_out = reinterpret_cast<gpa_result *>(d.base + kNaResOff);

#define NA_TRY(expr)                                                              \
  do {                                                                            \
    cudaError_t e_ = (expr);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      na_msg(errbuf, errbuf_len, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
      return GPA_ERR_CUDA;                                                        \
    }                                                                             \
  } while (0)
  NA_TRY(cudaMemcpyAsync(d_wq, wq, sizeof(uint16_t) * (size_t)n * n, cudaMemcpyHostToDevice, st));
  NA_TRY(cudaMemcpyAsync(d_asym, asym, (size_t)n * n, cudaMemcpyHostToDevice, st));
  NA_TRY(cudaMemcpyAsync(d_reqs, reqs, sizeof(gpa_request) * (size_t)count, cudaMemcpyHostToDevice, st));
  // The array still holds the previous call's answers.  The kernel writes every record it is
  // asked for; clearing them first means a launch that fails cannot hand back stale answers.
  NA_TRY(cudaMemsetAsync(d_out, 0, sizeof(gpa_result) * (size_t)count, st));

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
