// Replica-set search kernels: which `size` of the available shared replicas are best connected (subset_core.cuh has
// the definition and says why there is no reference counterpart).  ONE request per launch, answered into the mailbox.
//   subset_counts_kernel   per-GPU count vectors in descending lexicographic order, one rank per thread at a time
//   subset_enum_kernel<K>  the literal walk over C(R, K) combinations, a run of consecutive ranks per thread
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
  unsigned long long total;  // C(R, K)
  unsigned int chunk;        // consecutive ranks per thread visit
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
  const int tid = threadIdx.x;
  for (int i = tid; i < kSubEnumMaxFree; i += kSubThreads) {
    grp[i] = p.q.grp[i];
    wreq[i] = p.q.wreq[i];
  }
  for (int i = tid; i < kSubGroups * kSubGroups; i += kSubThreads) w[i] = p.q.w[i];
  __syncthreads();
  const int R = p.q.R;
  unsigned long long best = 0;
  const unsigned long long stride = (unsigned long long)gridDim.x * kSubThreads * p.chunk;
  for (unsigned long long start = ((unsigned long long)blockIdx.x * kSubThreads + tid) * p.chunk; start < p.total; start += stride) {
    int e[K], ge[K];
    uint32_t ps[K + 1];
    sub_comb_unrank<K>(binom, R, start, e);
    ps[0] = p.q.base;
    sub_comb_rescore<K>(grp, wreq, w, e, ge, ps, 0);
    const unsigned long long end = start + p.chunk < p.total ? start + p.chunk : p.total;
    for (unsigned long long rank = start;;) {
      const unsigned long long key = pack_key((int)ps[K], rank);
      best = key > best ? key : best;
      if (++rank >= end) break;
      const int j = sub_comb_next<K>(R, e);
      sub_comb_rescore<K>(grp, wreq, w, e, ge, ps, j);
    }
  }
  subset_reduce_publish(best, p.size, p.seq, g_scratch, g_ticket, g_res);
}

}  // namespace
