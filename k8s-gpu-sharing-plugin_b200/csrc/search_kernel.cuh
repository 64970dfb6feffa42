// Enumeration search kernel: ONE request, every partition visited, the whole machine.
// Part of the single translation unit gpualloc.cu (the kernels share __constant__ tables and the
// enumeration core); not compiled on its own.
#pragma once
#include "kernels_common.cuh"

namespace {

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
#pragma unroll
    for (int i = 0; i < 32; i++) r.pos[i] = 0;
    r.count = finalize(p.g, p.cut, s_binom, stab, key, p.intended != 0, r.pos);
    *g_key = 0ull;  // leave the scratch clean for the next launch on this stream
    *g_cursor = 0ull;
    *g_ticket = 0u;
    publish_result(g_res, r, p.seq);
  }
}

}  // namespace
