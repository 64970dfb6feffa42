// Answer-table kernels for nodes of up to 8 GPUs (ans_table.cuh): the builder, and the batch kernels that answer a
// request with one lookup.  Part of the single translation unit gpualloc.cu; not compiled on its own.
//
//   answer_table_build_kernel        one thread per (mask, required, size): the partition-table search of
//                                    part_table.cuh, run once when a topology is installed (52 488 entries for 8 GPUs)
//   besteffort_batch_table_kernel    MANY requests in node order, one LANE per request, requests and answers in RECORD
//                                    order: 256-bit load of the record, relabel to a device mask, one table lookup from
//                                    shared memory, 128-bit store.  No sort, no divergence: the kernel moves 48 bytes per
//                                    decision and its ceiling is HBM bandwidth.
// Replaces, for such nodes, bestEffortPolicy.Allocate
// (vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:34-89) per request.
#pragma once
#include "ans_table.cuh"
#include "batch_kernels.cuh"

namespace {

constexpr int kTabThreads = 1024;
constexpr int kTabBlocksPerSm = 2;  // 2 x (105 KB answers + 3 KB of small tables) of shared memory per SM

__global__ void __launch_bounds__(256) answer_table_build_kernel(const FastTopo ft, uint16_t *__restrict__ table, uint32_t n_entries,
                                                                 int intended) {
  const GlobalScores S{ft.s_full};
  for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n_entries; e += gridDim.x * blockDim.x) {
    uint32_t mask, rl;
    ans_decode(e / (uint32_t)kAnsSizes, mask, rl);
    const int size = (int)(e % (uint32_t)kAnsSizes) + 1;
    table[e] = ans_compute_entry(ft.nt, mask, rl, size, ft.part_words, ft.part_index, ft.part_key, S, intended != 0);
  }
}

constexpr size_t table_smem_bytes(uint32_t n_entries) {
  return (size_t)256 * sizeof(unsigned long long)   // s_dev: chosen mask -> its devices' NVML indices, one per byte
         + (((size_t)n_entries * sizeof(uint16_t) + 15) & ~(size_t)15)  // answers
         + 256 * sizeof(uint16_t)                   // tern
         + 256;                                     // s_bit: NVML index -> label bit
}

__device__ __forceinline__ void ld_record_256(const gpa_request *p, uint32_t (&w)[8]) {
  // the whole 32-byte record in one load: a warp reads 1 KiB of consecutive records per instruction; streamed, no reuse
  asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
               : "l"(p));
}

// kMasks = false: 32-byte gpa_request in, 16-byte gpa_result out; a record the table cannot answer (devices not in
//                 node order, unknown device) is marked kResDefer for the general kernel and its warp's byte in
//                 `defer_flags` (one per 32 records, optional) is set so that the general kernel finds it without
//                 reading every result back.
// kMasks = true : 8-byte gpa_mask_request in, 4-byte gpa_mask_result out; nothing is ever deferred.
template <bool kMasks>
__global__ void __launch_bounds__(kTabThreads, kTabBlocksPerSm) besteffort_batch_table_kernel(const void *__restrict__ reqs_v, int count,
                                                                                const FastTopo ft, void *__restrict__ out_v,
                                                                                uint8_t *__restrict__ defer_flags) {
  extern __shared__ __align__(16) unsigned char tsmem[];
  const uint32_t n_entries = ft.ans_entries;
  unsigned long long *s_dev = reinterpret_cast<unsigned long long *>(tsmem);
  uint16_t *s_ans = reinterpret_cast<uint16_t *>(s_dev + 256);
  uint16_t *s_tern = reinterpret_cast<uint16_t *>(reinterpret_cast<unsigned char *>(s_ans) +
                                                  (((size_t)n_entries * sizeof(uint16_t) + 15) & ~(size_t)15));
  uint8_t *s_bit = reinterpret_cast<uint8_t *>(s_tern + 256);

  const int tid = threadIdx.x;
  {
    // answers: 16 bytes per thread per step (the table is padded to a multiple of 16 bytes on the device)
    const uint4 *src = reinterpret_cast<const uint4 *>(ft.ans);
    uint4 *dst = reinterpret_cast<uint4 *>(s_ans);
    const uint32_t n16 = (n_entries * (uint32_t)sizeof(uint16_t) + 15u) / 16u;
    for (uint32_t i = tid; i < n16; i += kTabThreads) dst[i] = __ldg(src + i);
  }
  if (tid < 256) {
    s_tern[tid] = (uint16_t)ans_tern((uint32_t)tid);
    const int lab = tid < GPA_MAX_TOPOLOGY ? ft.rank[tid] : 0xFF;
    s_bit[tid] = lab < 8 ? (uint8_t)(1u << lab) : (uint8_t)0;
    unsigned long long dev = 0x8080808080808080ull;  // 8 x GPA_RES_END
    int o = 0;
    for (int lab2 = 0; lab2 < 8; lab2++)
      if (tid >> lab2 & 1) {
        dev ^= (unsigned long long)((ft.order[lab2] ^ 0x80) & 0xFF) << (8 * o);
        o++;
      }
    s_dev[tid] = dev;
  }
  __syncthreads();

  if constexpr (kMasks) {
    const uint2 *__restrict__ mreqs = static_cast<const uint2 *>(reqs_v);
    uint32_t *__restrict__ mout = static_cast<uint32_t *>(out_v);
    const uint32_t node_mask = (1u << ft.nt) - 1u;
    auto answer = [&](uint32_t x, uint32_t y) -> uint32_t {
      // {avail_mask:16, req_mask:16} {size:8, flags:8, n_req_extra:8, reserved:8} -> {chosen_mask:16, n_nil:8, status:8}
      const uint32_t mask = x & 0xFFFFu, rl = x >> 16;
      const int size = (int)(y & 0xFFu), flags = (int)((y >> 8) & 0xFFu);
      const int n = __popc(mask);
      const int n_req = __popc(rl) + (int)((y >> 16) & 0xFFu);
      if (((mask | rl) & ~node_mask) != 0) return (uint32_t)(GPA_ERR_INVALID & 0xFF) << 24;
      // guards of besteffort_policy.go:35-45; a required device that is not available is in no set (:53)
      if (size <= 0 || n < size || n_req > size || (flags & GPA_REQ_FOREIGN) || (rl & ~mask) != 0) return 0u;
      const uint32_t e = s_ans[ans_index(s_tern, mask, rl, size)];
      return (e & 0xFFu) | ((e >> 8) << 16);
    };
    if (((reinterpret_cast<uintptr_t>(reqs_v) & 15u) | (reinterpret_cast<uintptr_t>(out_v) & 7u)) == 0) {
      // two records per lane per step: 16 bytes in, 8 bytes out
      const long long n_pairs = (long long)count >> 1;
      for (long long p = (long long)blockIdx.x * kTabThreads + tid; p < n_pairs; p += (long long)gridDim.x * kTabThreads) {
        const uint4 rec = __ldg(reinterpret_cast<const uint4 *>(mreqs) + p);
        uint2 res;
        res.x = answer(rec.x, rec.y);
        res.y = answer(rec.z, rec.w);
        reinterpret_cast<uint2 *>(mout)[p] = res;
      }
      if ((count & 1) && blockIdx.x == 0 && tid == 0) {
        const uint2 rec = __ldg(mreqs + (count - 1));
        mout[count - 1] = answer(rec.x, rec.y);
      }
    } else {  // buffers that only have the records' own 8 / 4-byte alignment
      for (long long r = (long long)blockIdx.x * kTabThreads + tid; r < count; r += (long long)gridDim.x * kTabThreads) {
        const uint2 rec = __ldg(mreqs + r);
        mout[r] = answer(rec.x, rec.y);
      }
    }
  } else {
    const gpa_request *__restrict__ reqs = static_cast<const gpa_request *>(reqs_v);
    gpa_result *__restrict__ out = static_cast<gpa_result *>(out_v);
    const long long n_blocks = ((long long)count + kTabThreads - 1) / kTabThreads;
    for (long long blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
      const long long r = blk * kTabThreads + tid;
      bool defer = false;
      if (r < count) {
        uint32_t w[8];
        ld_record_256(reqs + r, w);
        const uint32_t rmask = w[4] & 0xFFFFu;
        const int n = (int)((w[4] >> 16) & 0xFFu);
        const int size = (int)(w[4] >> 24);
        const int flags = (int)(w[5] & 0xFFu);
        const int n_req = __popc(rmask) + (int)((w[5] >> 8) & 0xFFu);
        unsigned long long r0 = 0x8080808080808080ull;  // 8 x GPA_RES_END: the empty answer
        // guards of besteffort_policy.go:35-45
        if (!(size <= 0 || n < size || n_req > size)) {
          // relabel: device -> its bit in the node order.  The list must be strictly ascending in that order, i.e. every
          // bit must lie above all bits so far; a device the node does not have maps to 0 and fails the same test.
          uint32_t mask = 0, rl = 0;
          bool bad = n > 8 || (rmask >> n) != 0;
#pragma unroll
          for (int p = 0; p < 8; p++) {
            const uint32_t d = (w[p >> 2] >> (8 * (p & 3))) & 0xFFu;
            const uint32_t bit = p < n ? (uint32_t)s_bit[d] : 0x100u;
            bad = bad | (bit <= mask);
            mask |= bit & 0xFFu;
            rl |= (rmask >> p & 1u) ? (bit & 0xFFu) : 0u;
          }
          if (bad) {
            defer = true;
            r0 ^= (unsigned long long)((kResDefer ^ 0x80) & 0xFF);
          } else if (!(flags & GPA_REQ_FOREIGN)) {  // a required pointer that is in no set: every candidate is rejected (:53)
            const uint32_t e = s_ans[ans_index(s_tern, mask, rl, size)];
            const uint32_t chosen = e & 0xFFu;
            const int n_nil = (int)(e >> 8);
            r0 = s_dev[chosen];
            // nil pads follow the devices (D5): GPA_RES_END (0x80) -> GPA_NIL (0xFF) in n_nil bytes
            if (n_nil) r0 ^= (0x7F7F7F7F7F7F7F7Full >> (8 * (8 - n_nil))) << (8 * __popc(chosen));
          }
        }
        uint4 v;
        v.x = (uint32_t)r0;
        v.y = (uint32_t)(r0 >> 32);
        v.z = 0x80808080u;
        v.w = 0x80808080u;
        __stcs(reinterpret_cast<uint4 *>(out + r), v);
      }
      if (defer_flags != nullptr) {
        const unsigned any = __ballot_sync(0xffffffffu, defer);
        if ((tid & 31) == 0 && r < count) defer_flags[r >> 5] = any ? 1 : 0;
      }
    }
  }
}

}  // namespace
