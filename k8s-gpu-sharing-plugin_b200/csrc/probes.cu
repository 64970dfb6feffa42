// Topology validation probes (BASELINE config 5): measured bandwidths that check what NVML
// reports.  No reference counterpart — the reference never touches device memory; these are
// the two probe kernels `north_star` asks for:
//   hbm_stream_*  : per-GPU HBM streaming (read-only sum, and copy), reported against the
//                   measured HBM peak — bytes = read + written;
//   p2p_read      : pairwise peer load over NVLink 5 / NVSwitch (the consumer GPU pulls from
//                   the producer's HBM with plain LDG.128 on a peer-mapped pointer) next to
//                   cudaMemcpyPeerAsync — bytes = moved one way.
// Both are HBM-/link-bound streaming kernels: coalesced 128-bit accesses, 8 independent
// loads in flight per thread, streaming cache hints, persistent grid = SMs x resident CTAs.
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/gpualloc.h"

namespace {

constexpr int kThreads = 512;
constexpr int kUnroll = 8;

__device__ __forceinline__ uint4 ld_stream(const uint4 *p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream(uint4 *p, const uint4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w));
}

// dst == nullptr: read-only (xor-reduce into sink so the loads cannot be dropped)
__global__ void __launch_bounds__(kThreads) stream_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t n_vec,
                                                           unsigned int *__restrict__ sink) {
  const size_t stride = (size_t)gridDim.x * kThreads;
  size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (; i + (kUnroll - 1) * stride < n_vec; i += kUnroll * stride) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; u++) v[u] = ld_stream(src + i + u * stride);
    if (dst) {
#pragma unroll
      for (int u = 0; u < kUnroll; u++) st_stream(dst + i + u * stride, v[u]);
    } else {
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        acc.x ^= v[u].x;
        acc.y ^= v[u].y;
        acc.z ^= v[u].z;
        acc.w ^= v[u].w;
      }
    }
  }
  for (; i < n_vec; i += stride) {
    uint4 v = ld_stream(src + i);
    if (dst) st_stream(dst + i, v);
    else {
      acc.x ^= v.x;
      acc.y ^= v.y;
      acc.z ^= v.z;
      acc.w ^= v.w;
    }
  }
  if (!dst) {
    unsigned int r = acc.x ^ acc.y ^ acc.z ^ acc.w;
    if (r == 0x9E3779B9u) atomicAdd(sink, 1u);  // practically never; keeps the loads alive
  }
}

__global__ void fill_kernel(uint4 *p, size_t n_vec, unsigned int seed) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n_vec; i += stride) {
    unsigned int x = (unsigned int)i * 2654435761u + seed;
    p[i] = make_uint4(x, x ^ 0x55555555u, x + 7u, ~x);
  }
}

// Ceiling of what the allocation kernels do at their core: a dependent-address gather from a shared-memory table
// plus an integer add (SURVEY.md 8d asks for this figure).  Every lane chases its own index through a 2^log2-entry
// uint16 table; 4 independent chains per lane hide the LDS latency.
__global__ void __launch_bounds__(256) lut_kernel(int log2_entries, int iters, unsigned int *__restrict__ sink) {
  extern __shared__ uint16_t lut[];
  const uint32_t n = 1u << log2_entries;
  for (uint32_t i = threadIdx.x; i < n; i += 256) lut[i] = (uint16_t)((i * 40503u + 7u) & 0xFFFFu);
  __syncthreads();
  const uint32_t m = n - 1u;
  uint32_t a = threadIdx.x & m, b = (threadIdx.x * 7u + 1u) & m, c = (threadIdx.x * 13u + 2u) & m, d = (threadIdx.x * 29u + 3u) & m;
  uint32_t acc = blockIdx.x;
  for (int i = 0; i < iters; i++) {
    const uint32_t va = lut[a], vb = lut[b], vc = lut[c], vd = lut[d];
    acc += va + vb + vc + vd;
    a = (a + va) & m;
    b = (b + vb) & m;
    c = (c + vc) & m;
    d = (d + vd) & m;
  }
  if (acc == 0x9E3779B9u) atomicAdd(sink, 1u);
}

#define PROBE_TRY(expr)                                            \
  do {                                                             \
    cudaError_t e_ = (expr);                                       \
    if (e_ != cudaSuccess) {                                       \
      if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%s: %s", #expr, cudaGetErrorString(e_)); \
      goto fail;                                                   \
    }                                                              \
  } while (0)

int grid_for(int device) {
  int sms = 148, occ = 1;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, stream_kernel, kThreads, 0);
  return sms * (occ < 1 ? 1 : occ);
}

}  // namespace

extern "C" {

// mode 0: read-only sum, bytes counted = bytes read; mode 1: copy, bytes counted = read + written.
int gpa_probe_hbm(int cuda_device, unsigned long long bytes, int reps, int mode, double *best_gbs, double *median_gbs,
                  char *errbuf, int errbuf_len) {
  if (bytes < (1ull << 20) || reps < 1 || (mode != 0 && mode != 1) || !best_gbs) return GPA_ERR_INVALID;
  uint4 *a = nullptr, *b = nullptr;
  unsigned int *sink = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  std::vector<float> ms;
  const size_t n_vec = bytes / sizeof(uint4);
  int grid;
  PROBE_TRY(cudaSetDevice(cuda_device));
  grid = grid_for(cuda_device);
  PROBE_TRY(cudaMalloc(&a, n_vec * sizeof(uint4)));
  if (mode == 1) PROBE_TRY(cudaMalloc(&b, n_vec * sizeof(uint4)));
  PROBE_TRY(cudaMalloc(&sink, sizeof(unsigned int)));
  PROBE_TRY(cudaMemset(sink, 0, sizeof(unsigned int)));
  PROBE_TRY(cudaEventCreate(&e0));
  PROBE_TRY(cudaEventCreate(&e1));
  fill_kernel<<<grid, 256>>>(a, n_vec, 1u);
  for (int w = 0; w < 3; w++) stream_kernel<<<grid, kThreads>>>(a, b, n_vec, sink);
  PROBE_TRY(cudaDeviceSynchronize());
  for (int r = 0; r < reps; r++) {
    PROBE_TRY(cudaEventRecord(e0));
    stream_kernel<<<grid, kThreads>>>(a, b, n_vec, sink);
    PROBE_TRY(cudaEventRecord(e1));
    PROBE_TRY(cudaEventSynchronize(e1));
    float t = 0;
    PROBE_TRY(cudaEventElapsedTime(&t, e0, e1));
    ms.push_back(t);
  }
  PROBE_TRY(cudaGetLastError());
  {
    std::sort(ms.begin(), ms.end());
    const double moved = (double)(n_vec * sizeof(uint4)) * (mode == 1 ? 2.0 : 1.0);
    *best_gbs = moved / (ms.front() * 1e-3) / 1e9;
    if (median_gbs) *median_gbs = moved / (ms[ms.size() / 2] * 1e-3) / 1e9;
  }
  cudaFree(a);
  cudaFree(b);
  cudaFree(sink);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return GPA_OK;
fail:
  if (a) cudaFree(a);
  if (b) cudaFree(b);
  if (sink) cudaFree(sink);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  return GPA_ERR_CUDA;
}

// Shared-memory table gather + integer add rate, in 10^9 lookups per second over the whole GPU.
int gpa_probe_lut(int cuda_device, int log2_entries, int reps, double *best_glookups_per_s, char *errbuf, int errbuf_len) {
  if (log2_entries < 4 || log2_entries > 15 || reps < 1 || !best_glookups_per_s) return GPA_ERR_INVALID;
  unsigned int *sink = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  const int iters = 4096;
  const size_t smem = ((size_t)1 << log2_entries) * sizeof(uint16_t);
  int sms = 148, occ = 1, grid;
  double best = 0;
  PROBE_TRY(cudaSetDevice(cuda_device));
  PROBE_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cuda_device));
  if (smem > 48 * 1024) PROBE_TRY(cudaFuncSetAttribute(lut_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  PROBE_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lut_kernel, 256, smem));
  grid = sms * (occ < 1 ? 1 : occ);
  PROBE_TRY(cudaMalloc(&sink, sizeof(unsigned int)));
  PROBE_TRY(cudaMemset(sink, 0, sizeof(unsigned int)));
  PROBE_TRY(cudaEventCreate(&e0));
  PROBE_TRY(cudaEventCreate(&e1));
  for (int r = 0; r < reps + 2; r++) {
    PROBE_TRY(cudaEventRecord(e0));
    lut_kernel<<<grid, 256, smem>>>(log2_entries, iters, sink);
    PROBE_TRY(cudaEventRecord(e1));
    PROBE_TRY(cudaEventSynchronize(e1));
    float t = 0;
    PROBE_TRY(cudaEventElapsedTime(&t, e0, e1));
    const double g = (double)grid * 256.0 * iters * 4.0 / (t * 1e-3) / 1e9;
    if (r >= 2 && g > best) best = g;
  }
  PROBE_TRY(cudaGetLastError());
  *best_glookups_per_s = best;
  cudaFree(sink);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return GPA_OK;
fail:
  if (sink) cudaFree(sink);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  return GPA_ERR_CUDA;
}

// n x n matrix (row = consumer GPU that pulls, column = producer GPU whose HBM is read), GB/s one way.
// mode 0: peer-load kernel on the consumer; mode 1: cudaMemcpyPeerAsync.  The diagonal is the local copy.
// can_access (optional, n x n) receives cudaDeviceCanAccessPeer.
int gpa_probe_p2p(int n_devices, unsigned long long bytes, int reps, int mode, double *matrix_gbs, int *can_access, char *errbuf,
                  int errbuf_len) {
  if (n_devices < 1 || n_devices > 16 || bytes < (1ull << 20) || reps < 1 || !matrix_gbs || mode < 0 || mode > 3)
    return GPA_ERR_INVALID;
  const int n = n_devices;
  const size_t n_vec = bytes / sizeof(uint4);
  std::vector<uint4 *> src(n, nullptr), dst(n, nullptr);
  std::vector<unsigned int *> sink(n, nullptr);
  std::vector<cudaEvent_t> e0(n, nullptr), e1(n, nullptr);
  int ndev = 0;
  PROBE_TRY(cudaGetDeviceCount(&ndev));
  if (ndev < n) {
    if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%d CUDA devices visible, %d asked for", ndev, n);
    return GPA_ERR_INVALID;
  }
  for (int d = 0; d < n; d++) {
    PROBE_TRY(cudaSetDevice(d));
    PROBE_TRY(cudaMalloc(&src[d], n_vec * sizeof(uint4)));
    PROBE_TRY(cudaMalloc(&dst[d], n_vec * sizeof(uint4)));
    PROBE_TRY(cudaMalloc(&sink[d], sizeof(unsigned int)));
    PROBE_TRY(cudaEventCreate(&e0[d]));
    PROBE_TRY(cudaEventCreate(&e1[d]));
    fill_kernel<<<grid_for(d), 256>>>(src[d], n_vec, 17u * d + 1u);
    PROBE_TRY(cudaDeviceSynchronize());
  }
  for (int c = 0; c < n; c++) {
    PROBE_TRY(cudaSetDevice(c));
    for (int p = 0; p < n; p++) {
      int ok = 1;
      if (p != c) {
        PROBE_TRY(cudaDeviceCanAccessPeer(&ok, c, p));
        if (ok) {
          cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
          if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = 0;
          cudaGetLastError();
        }
      }
      if (can_access) can_access[c * n + p] = ok;
    }
  }
  if (mode >= 2) {
    // ALL GPUs at once: in round s every GPU c pulls from GPU (c + s) mod n at the same time, so every NVLink port is
    // busy in both directions and the switch carries n streams.  Round 0 is the local copy on every GPU.
    const bool use_kernel = mode == 2;
    for (int s = 0; s < n; s++) {
      std::vector<double> best(n, 0.0);
      for (int r = 0; r < reps + 2; r++) {
        for (int c = 0; c < n; c++) {
          const int p = (c + s) % n;
          PROBE_TRY(cudaSetDevice(c));
          if (use_kernel && can_access && !can_access[c * n + p]) continue;
          PROBE_TRY(cudaEventRecord(e0[c]));
          if (use_kernel) stream_kernel<<<grid_for(c), kThreads>>>(src[p], dst[c], n_vec, sink[c]);
          else PROBE_TRY(cudaMemcpyPeerAsync(dst[c], c, src[p], p, n_vec * sizeof(uint4)));
          PROBE_TRY(cudaEventRecord(e1[c]));
        }
        for (int c = 0; c < n; c++) {
          const int p = (c + s) % n;
          if (use_kernel && can_access && !can_access[c * n + p]) continue;
          PROBE_TRY(cudaSetDevice(c));
          PROBE_TRY(cudaEventSynchronize(e1[c]));
          float t = 0;
          PROBE_TRY(cudaEventElapsedTime(&t, e0[c], e1[c]));
          const double g = (double)(n_vec * sizeof(uint4)) / (t * 1e-3) / 1e9;
          if (r >= 2 && g > best[c]) best[c] = g;
        }
      }
      for (int c = 0; c < n; c++) matrix_gbs[c * n + (c + s) % n] = best[c];
    }
  } else
  for (int c = 0; c < n; c++) {
    PROBE_TRY(cudaSetDevice(c));
    const int grid = grid_for(c);
    for (int p = 0; p < n; p++) {
      double best = 0;
      if (can_access && !can_access[c * n + p] && mode == 0) {
        matrix_gbs[c * n + p] = 0;
        continue;
      }
      for (int r = 0; r < reps + 2; r++) {
        PROBE_TRY(cudaEventRecord(e0[c]));
        if (mode == 0) stream_kernel<<<grid, kThreads>>>(src[p], dst[c], n_vec, sink[c]);
        else PROBE_TRY(cudaMemcpyPeerAsync(dst[c], c, src[p], p, n_vec * sizeof(uint4)));
        PROBE_TRY(cudaEventRecord(e1[c]));
        PROBE_TRY(cudaEventSynchronize(e1[c]));
        float t = 0;
        PROBE_TRY(cudaEventElapsedTime(&t, e0[c], e1[c]));
        double g = (double)(n_vec * sizeof(uint4)) / (t * 1e-3) / 1e9;
        if (r >= 2 && g > best) best = g;
      }
      PROBE_TRY(cudaGetLastError());
      matrix_gbs[c * n + p] = best;
    }
  }
  for (int d = 0; d < n; d++) {
    cudaSetDevice(d);
    cudaFree(src[d]);
    cudaFree(dst[d]);
    cudaFree(sink[d]);
    cudaEventDestroy(e0[d]);
    cudaEventDestroy(e1[d]);
  }
  return GPA_OK;
fail:
  for (int d = 0; d < n; d++) {
    cudaSetDevice(d);
    if (src[d]) cudaFree(src[d]);
    if (dst[d]) cudaFree(dst[d]);
    if (sink[d]) cudaFree(sink[d]);
    if (e0[d]) cudaEventDestroy(e0[d]);
    if (e1[d]) cudaEventDestroy(e1[d]);
  }
  return GPA_ERR_CUDA;
}

}  // extern "C"
