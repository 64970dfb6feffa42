// Shared device-side pieces of the allocation kernels: constant tables, score-table fill, reductions.
// Part of the single translation unit gpualloc.cu (the kernels share __constant__ tables and the
// enumeration core); not compiled on its own.
#pragma once
#include <cstddef>

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
// calculateGPUPairScore panics on a pair whose two link lists differ in length (besteffort_policy.go:307-310), but only
// when such a pair is actually SCORED: inside a candidate partition that passed the `required` filter (:53-56), or inside
// a set of the saved bestPartition that holds every required device (:77-80).  The kernels reproduce that by giving such a
// pair the weight kAsymPoison: on a topology with asymmetric link counts the scaled weights are kept small enough
// (gpa_set_topology) that a whole partition without such a pair scores below kAsymPoison, so "the best accepted
// partition scores >= kAsymPoison" says exactly "some accepted partition holds such a pair", and the same for the
// final set choice.
constexpr uint32_t kAsymPoison = 0x8000u;

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
  unsigned int seq;
  uint16_t wp[kMaxDevices * kMaxDevices];  // scaled weight of (earlier position a, later position b) at [a*16+b]
};

// Answer of a single-request search.  It lives in PINNED, DEVICE-MAPPED host memory: the last thread of the search
// writes the record straight across PCIe, fences, and then publishes `seq`; the host spins on `seq` instead of paying
// for a D2H copy and a stream synchronisation (about 10 us of a 30 us call).
struct __align__(16) SearchResult {
  int32_t count;  // k, 0 = empty answer, -1 = internal inconsistency
  int32_t best_score;
  int8_t pos[32];  // positions, -1 = nil pad
  unsigned long long key;
  unsigned int seq;  // launch number this record belongs to; written last
  unsigned int reserved;
};

static_assert(sizeof(SearchResult) == 64 && offsetof(SearchResult, seq) == 48, "three 16-byte stores cover everything before seq");

__device__ __forceinline__ void publish_result(SearchResult *dst, const SearchResult &r, unsigned int seq) {
  // 48 bytes of payload as three 16-byte stores (one PCIe write each), then the fence, then the flag
  const uint4 *src = reinterpret_cast<const uint4 *>(&r);
  uint4 *out = reinterpret_cast<uint4 *>(dst);
  out[0] = src[0];
  out[1] = src[1];
  out[2] = src[2];
  __threadfence_system();
  *reinterpret_cast<volatile unsigned int *>(&dst->seq) = seq;
}

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
    S[mask] = (uint16_t)(sum > 0xFFFFu ? 0xFFFFu : sum);  // only poisoned sets (kAsymPoison) can get there
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

}  // namespace
