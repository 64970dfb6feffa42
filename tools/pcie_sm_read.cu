// How fast can the SMs themselves pull records out of pinned host memory?  The compact-mask batch call reads its
// records across PCIe from inside the kernel; this probe times the read side alone, three ways, beside
// the copy engine:
//   ldg      16-byte loads, one per thread per step, grid-stride (what the mask kernels do)
//   ldg4     the same with four loads in flight per thread
//   bulk     cp.async.bulk tiles into shared memory behind an mbarrier ring (local host memory, no multicast)
//   ldg_st   the 16-byte loads, with 8 bytes per load stored to a second pinned buffer (the compact-mask call's traffic:
//            4 bytes in, 2 bytes out per request); GB/s counts the bytes read
//   memcpy   cudaMemcpyAsync H2D of the same bytes
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/bin/pcie_sm_read tools/pcie_sm_read.cu
// Usage: pcie_sm_read [MiB=16] [reps=20]     prints one JSON line
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) {                                                                    \
      fprintf(stderr, "%s: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__);      \
      exit(1);                                                                                  \
    }                                                                                           \
  } while (0)

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(512) ldg_kernel(const uint4 *__restrict__ src, size_t n16, uint32_t *sink) {
  uint32_t acc = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = __ldcs(src + i);
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

__global__ void __launch_bounds__(512) ldg4_kernel(const uint4 *__restrict__ src, size_t n16, uint32_t *sink) {
  uint32_t acc = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n16; i += 4 * stride) {
    const uint4 a = __ldcs(src + i), b = __ldcs(src + i + stride), c = __ldcs(src + i + 2 * stride), d = __ldcs(src + i + 3 * stride);
    acc ^= a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w ^ c.x ^ c.y ^ c.z ^ c.w ^ d.x ^ d.y ^ d.z ^ d.w;
  }
  for (; i < n16; i += stride) {
    const uint4 v = __ldcs(src + i);
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

__global__ void __launch_bounds__(512) ldg_st_kernel(const uint4 *__restrict__ src, size_t n16, uint2 *__restrict__ dst) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = __ldcs(src + i);
    dst[i] = make_uint2(v.x ^ v.y, v.z ^ v.w);
  }
}

constexpr int kStages = 4;
template <int kTile>
__global__ void __launch_bounds__(512) bulk_kernel(const char *__restrict__ src, size_t bytes, uint32_t *sink) {
  extern __shared__ __align__(128) char ring[];
  __shared__ unsigned long long full[kStages];
  const int tid = threadIdx.x;
  const size_t tiles = (bytes + kTile - 1) / kTile;
  const size_t my_tiles = tiles > blockIdx.x ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto issue = [&](size_t i, int s) {
    const size_t off = (blockIdx.x + i * gridDim.x) * (size_t)kTile;
    const uint32_t b = (uint32_t)(bytes - off < (size_t)kTile ? bytes - off : (size_t)kTile);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&full[s])), "r"(b) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(ring + (size_t)s * kTile)),
                 "l"(src + off), "r"(b), "r"(smem_addr(&full[s]))
                 : "memory");
  };
  if (tid == 0) {
    for (int s = 0; s < kStages; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(&full[s])), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (size_t i = 0; i < kStages && i < my_tiles; i++) issue(i, (int)i);
  }
  __syncthreads();
  uint32_t acc = 0;
  for (size_t i = 0; i < my_tiles; i++) {
    const int s = (int)(i % kStages);
    const uint32_t parity = (uint32_t)((i / kStages) & 1);
    asm volatile(
        "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(
            smem_addr(&full[s])),
        "r"(parity)
        : "memory");
    const uint4 *t = reinterpret_cast<const uint4 *>(ring + (size_t)s * kTile);
    for (int q = tid; q < kTile / 16; q += 512) {
      const uint4 v = t[q];
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    __syncthreads();
    if (tid == 0 && i + kStages < my_tiles) issue(i + kStages, s);
  }
  if (acc == 0x12345678u) *sink = acc;
}

int main(int argc, char **argv) {
  const size_t mib = argc > 1 ? (size_t)atol(argv[1]) : 16;
  const int reps = argc > 2 ? atoi(argv[2]) : 20;
  const size_t bytes = mib << 20;
  char *h = nullptr, *d = nullptr;
  uint32_t *sink = nullptr;
  CK(cudaHostAlloc(&h, bytes, cudaHostAllocMapped));
  for (size_t i = 0; i < bytes; i++) h[i] = (char)(i * 131u >> 3);
  CK(cudaMalloc(&d, bytes));
  CK(cudaMalloc(&sink, 4));
  char *hd = nullptr, *h_out = nullptr, *hd_out = nullptr;
  CK(cudaHostGetDevicePointer(&hd, h, 0));
  CK(cudaHostAlloc(&h_out, bytes / 2, cudaHostAllocMapped));
  CK(cudaHostGetDevicePointer(&hd_out, h_out, 0));
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  constexpr int kTileA = 16384, kTileB = 4096;
  CK(cudaFuncSetAttribute(bulk_kernel<kTileA>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStages * kTileA));
  auto time_it = [&](auto &&fn) {
    std::vector<float> ms;
    for (int r = 0; r < reps + 3; r++) {
      CK(cudaEventRecord(e0, st));
      fn();
      CK(cudaEventRecord(e1, st));
      CK(cudaStreamSynchronize(st));
      float t;
      CK(cudaEventElapsedTime(&t, e0, e1));
      if (r >= 3) ms.push_back(t);
    }
    std::sort(ms.begin(), ms.end());
    return (double)bytes / (ms[ms.size() / 2] * 1e-3) / 1e9;
  };
  const size_t n16 = bytes / 16;
  printf("{\"mib\": %zu, \"reps\": %d, \"sms\": %d, \"unit\": \"GB/s, median\"", mib, reps, sms);
  printf(", \"memcpy_h2d\": %.1f", time_it([&] { CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, st)); }));
  for (int per_sm : {1, 2, 4})
    printf(", \"ldg_%dcta\": %.1f", per_sm, time_it([&] { ldg_kernel<<<sms * per_sm, 512, 0, st>>>((const uint4 *)hd, n16, sink); }));
  for (int per_sm : {1, 2})
    printf(", \"ldg4_%dcta\": %.1f", per_sm, time_it([&] { ldg4_kernel<<<sms * per_sm, 512, 0, st>>>((const uint4 *)hd, n16, sink); }));
  for (int per_sm : {1, 2})
    printf(", \"ldg_st_%dcta\": %.1f", per_sm,
           time_it([&] { ldg_st_kernel<<<sms * per_sm, 512, 0, st>>>((const uint4 *)hd, n16, (uint2 *)hd_out); }));
  printf(", \"bulk16k_1cta\": %.1f", time_it([&] { bulk_kernel<kTileA><<<sms, 512, kStages * kTileA, st>>>(hd, bytes, sink); }));
  printf(", \"bulk16k_quarter_grid\": %.1f", time_it([&] { bulk_kernel<kTileA><<<sms / 4, 512, kStages * kTileA, st>>>(hd, bytes, sink); }));
  printf(", \"bulk4k_1cta\": %.1f", time_it([&] { bulk_kernel<kTileB><<<sms, 512, kStages * kTileB, st>>>(hd, bytes, sink); }));
  printf(", \"bulk4k_2cta\": %.1f", time_it([&] { bulk_kernel<kTileB><<<sms * 2, 512, kStages * kTileB, st>>>(hd, bytes, sink); }));
  CK(cudaGetLastError());
  printf("}\n");
  return 0;
}
