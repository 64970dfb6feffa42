// Subset dynamic programme kernels: ONE large request without visiting every partition.
// Part of the single translation unit gpualloc.cu (the kernels share __constant__ tables and the
// enumeration core); not compiled on its own.
#pragma once
#include "kernels_common.cuh"

namespace {

// ---------------------------------------------------------------------------------------
// Subset dynamic programme for ONE large request (gpa_enum.cuh "Subset dynamic programme"): instead of visiting
// every partition, fill best[R] for every reachable remainder R, one data-parallel pass per layer, then rebuild the
// first best partition.  n = 16, k = 3: ~1 M (mask, set) pairs instead of 22.4 M partitions.  Tables live in global
// memory (128 KB + 64 KB, L2-resident); each layer is its own launch, which is the grid-wide barrier between layers.
struct DpParams {
  Geom g;
  int intended;
  unsigned int seq;
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

// the idx-th mask with s of the n lowest bits set (colex order; any bijection onto the C(n, s) masks would do)
__device__ __forceinline__ uint32_t nth_mask(const BinomTable &B, int n, int s, uint32_t idx) {
  uint32_t m = 0;
  for (int pos = n - 1; pos >= 0 && s > 0; pos--) {
    const uint32_t c = binom_at(B, pos, s);  // masks of s bits below pos
    if (idx >= c) {
      m |= 1u << pos;
      idx -= c;
      s--;
    }
  }
  return m;
}

// layer j, ONE WARP per remainder: for layers whose remainders have many first sets (large k) a single thread
// walking them one after the other would be the tail of the whole search.  The remainders of a layer are the masks of
// j*k devices and, with pads, of j*k - P devices; they are NUMBERED (nth_mask) so that every warp gets its share of them -
// striding over all 2^n masks instead left most warps without a remainder and a few with several (n = 16, k = 5: the
// third layer has 4 384 remainders among 65 536 masks and took 224 us; numbered, they fit one round of the grid).
__global__ void __launch_bounds__(kDpThreads) dp_layer_warp_kernel(const DpParams p, int j, uint16_t *T, uint16_t *Tok) {
  __shared__ BinomTable s_binom;
  stage_binom(s_binom, threadIdx.x, kDpThreads);
  __syncthreads();
  const TabAny tany{T};
  const TabOk tok{Tok, p.rbit};
  const int lane = threadIdx.x & 31;
  const uint32_t warps = gridDim.x * (kDpThreads / 32);
  const int s1 = j * p.g.k, s2 = p.g.P ? j * p.g.k - p.g.P : 0;
  const uint32_t c1 = s1 <= p.g.n ? binom_at(s_binom, p.g.n, s1) : 0u;
  const uint32_t c2 = (s2 >= 1 && s2 <= p.g.n) ? binom_at(s_binom, p.g.n, s2) : 0u;
  for (uint32_t idx = blockIdx.x * (kDpThreads / 32) + (threadIdx.x >> 5); idx < c1 + c2; idx += warps) {
    const uint32_t R = idx < c1 ? nth_mask(s_binom, p.g.n, s1, idx) : nth_mask(s_binom, p.g.n, s2, idx - c1);
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
__device__ __forceinline__ void dp_final_stage(const DpParams &p, const uint16_t *T, const uint16_t *Tok, SearchResult *g_res) {
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
#pragma unroll
    for (int q = 0; q < 32; q++) r.pos[q] = 0;
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
    publish_result(g_res, r, p.seq);
  }
}

__global__ void __launch_bounds__(1024) dp_final_kernel(const DpParams p, const uint16_t *T, const uint16_t *Tok, SearchResult *g_res) {
  dp_final_stage(p, T, Tok, g_res);
}

}  // namespace
