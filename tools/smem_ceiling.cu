// The two on-chip ceilings the scoring kernels run against ("measure this ceiling with a SMEM-LUT-gather +
// IADD microbenchmark on the box"): how many warp instructions the SMs issue, and how many shared-memory table lookups at
// lane-random addresses they complete, per second.
//   issue    eight independent chains of integer adds per thread, nothing else: warp instructions / s against 4 schedulers
//            x SM clock (the same loop with multiply-adds runs at half of it: IMAD issues every other clock)
//   gather1  one-byte lookups in a 51 KB table at lane-random indices, four independent chains per thread, index update =
//            one multiply-add + one mask (the answer-table lookup of the batch kernels)
//   gather8  the same with 8-byte entries of a 2 KB table (their per-mask device lists)
//   dot      8-byte loads at a warp-uniform address + two 2-way dot products + max (the pair loop of the replica-set batch
//            kernel), pairs / s
// Build: make -C tools      Usage: smem_ceiling [reps=5]     prints one JSON line
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess) {                                                               \
      fprintf(stderr, "%s: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                             \
    }                                                                                      \
  } while (0)

constexpr int kThreads = 1024;
constexpr int kIters = 4096;
constexpr int kTable = 52488;  // bytes: the answer table of an 8-GPU node

__global__ void __launch_bounds__(kThreads) issue_kernel(uint32_t *sink, uint32_t seed) {
  uint32_t a[8];
#pragma unroll
  for (int j = 0; j < 8; j++) a[j] = seed + threadIdx.x * 8 + j;
#pragma unroll 8
  for (int i = 0; i < kIters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] += a[(j + 1) & 7];  // one IADD each; the chains feed each other, so nothing folds
  }
  uint32_t x = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) x ^= a[j];
  if (x == 0x12345u) *sink = x;
}

template <typename T, int kEntries>
__global__ void __launch_bounds__(kThreads) gather_kernel(uint32_t *sink, uint32_t seed) {
  extern __shared__ __align__(16) unsigned char smem[];
  T *tab = reinterpret_cast<T *>(smem);
  for (int i = threadIdx.x; i < kEntries; i += kThreads) tab[i] = (T)(i * 2654435761u);
  __syncthreads();
  uint32_t idx[4], acc[4] = {0, 0, 0, 0};
#pragma unroll
  for (int j = 0; j < 4; j++) idx[j] = (seed + threadIdx.x * 4 + j) * 2654435761u;
  for (int i = 0; i < kIters; i++) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t e = (uint32_t)(((unsigned long long)idx[j] * kEntries) >> 32);  // lane-random entry
      acc[j] += (uint32_t)tab[e];
      idx[j] = idx[j] * 1664525u + 1013904223u;
    }
  }
  if ((acc[0] ^ acc[1] ^ acc[2] ^ acc[3]) == 0x12345u) *sink = acc[0];
}

__global__ void __launch_bounds__(256) dot_kernel(uint32_t *sink, uint32_t seed) {
  __shared__ uint2 suf[512];
  for (int i = threadIdx.x; i < 512; i += 256) suf[i] = make_uint2(i * 7u, (i * 2654435761u) & 0x07070707u);
  __syncthreads();
  const uint32_t q0 = seed + threadIdx.x, q1 = seed * 3 + threadIdx.x;
  int best = -1;
  for (int i = 0; i < kIters / 4; i++) {
#pragma unroll 4
    for (int s = 0; s < 496; s += 4) {
      const uint2 a = suf[s], b = suf[s + 1], c = suf[s + 2], d = suf[s + 3];
      const int ga = (int)__dp2a_lo(q1, a.y, __dp2a_hi(q0, a.y, a.x)), gb = (int)__dp2a_lo(q1, b.y, __dp2a_hi(q0, b.y, b.x));
      const int gc = (int)__dp2a_lo(q1, c.y, __dp2a_hi(q0, c.y, c.x)), gd = (int)__dp2a_lo(q1, d.y, __dp2a_hi(q0, d.y, d.x));
      best = max(best, max(max(ga, gb), max(gc, gd)));
    }
  }
  if (best == 0x12345) *sink = best;
}

int main(int argc, char **argv) {
  const int reps = argc > 1 ? atoi(argv[1]) : 5;
  int sms = 0, clock_khz = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  CK(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0));
  uint32_t *sink;
  CK(cudaMalloc(&sink, 4));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  auto g1 = gather_kernel<uint8_t, kTable>;
  auto g8 = gather_kernel<unsigned long long, 256>;
  CK(cudaFuncSetAttribute(g1, cudaFuncAttributeMaxDynamicSharedMemorySize, kTable + 16));
  auto time_ms = [&](auto &&launch) {
    std::vector<float> ms;
    for (int r = 0; r < reps + 2; r++) {
      CK(cudaEventRecord(e0));
      launch();
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float t;
      CK(cudaEventElapsedTime(&t, e0, e1));
      if (r >= 2) ms.push_back(t);
    }
    std::sort(ms.begin(), ms.end());
    return (double)ms[ms.size() / 2];
  };
  const double t_issue = time_ms([&] { issue_kernel<<<sms * 2, kThreads>>>(sink, 1u); });
  const double t_g1 = time_ms([&] { g1<<<sms, kThreads, kTable + 16>>>(sink, 1u); });
  const double t_g8 = time_ms([&] { g8<<<sms, kThreads, 2048>>>(sink, 1u); });
  const double t_dot = time_ms([&] { dot_kernel<<<sms * 8, 256>>>(sink, 1u); });
  CK(cudaGetLastError());
  const double warp_instr = (double)sms * 2 * (kThreads / 32) * (double)kIters * 8;
  const double gathers = (double)sms * kThreads * (double)kIters * 4;
  const double pairs = (double)sms * 8 * 256 * (double)(kIters / 4) * 496;
  printf("{\"sms\": %d, \"sm_clock_mhz_max\": %d, \"issue\": {\"warp_instructions_per_s\": %.4g, \"nominal_4_per_clock_per_sm\": %.4g}, "
         "\"gather_1_byte_51KB\": {\"lookups_per_s\": %.4g, \"per_sm_per_clock\": %.3f}, "
         "\"gather_8_byte_2KB\": {\"lookups_per_s\": %.4g, \"per_sm_per_clock\": %.3f}, "
         "\"pair_loop\": {\"pairs_per_s\": %.4g, \"per_sm_per_clock\": %.3f}}\n",
         sms, clock_khz / 1000, warp_instr / (t_issue * 1e-3), (double)sms * 4 * clock_khz * 1e3, gathers / (t_g1 * 1e-3),
         gathers / (t_g1 * 1e-3) / sms / (clock_khz * 1e3), gathers / (t_g8 * 1e-3), gathers / (t_g8 * 1e-3) / sms / (clock_khz * 1e3),
         pairs / (t_dot * 1e-3), pairs / (t_dot * 1e-3) / sms / (clock_khz * 1e3));
  return 0;
}
