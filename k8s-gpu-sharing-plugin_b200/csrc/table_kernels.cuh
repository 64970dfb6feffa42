// Answer-table kernels for nodes of up to 8 GPUs (ans_table.cuh): the builder, and the batch kernels that answer a
// request with one lookup.  Part of the single translation unit gpualloc.cu; not compiled on its own.
//
//   answer_table_build_kernel        one thread per (mask, required, size): the partition-table search of
//                                    part_table.cuh, run once when a topology is installed (52 488 entries for 8 GPUs)
//   besteffort_batch_table_kernel    MANY 32-byte requests in node order, one LANE per request, requests and answers in
//                                    RECORD order.  One CTA per SM keeps the answer table in shared memory and streams
//                                    64 KB request tiles through a ring of shared-memory stages filled by bulk
//                                    asynchronous copies (cp.async.bulk + mbarrier), so that several tiles are in flight
//                                    per SM while the lanes relabel, look up and store.  No sort, no divergence: the
//                                    kernel moves 48 bytes per decision and its ceiling is HBM bandwidth.
//   besteffort_masks_table_kernel    the same for 8-byte mask records (4-byte answers): nothing to relabel.
// Replaces, for such nodes, bestEffortPolicy.Allocate
// (vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:34-89) per request.
#pragma once
#include "ans_table.cuh"
#include "batch_kernels.cuh"

namespace {

// The record kernel is a template over (threads per CTA, stages of the request ring, records per thread and tile); which
// instantiation runs is chosen on the host (launch_table).

// table: the 2-byte entries of ans_table.cuh (what gpa_get_answer_table() hands out); table8: the chosen masks alone, for
// the batch kernels' shared memory - the nil pads of an answer follow from it: a non-empty answer of `size` entries that
// names fewer devices is padded with size - popc(chosen) nils.
__global__ void __launch_bounds__(256) answer_table_build_kernel(const FastTopo ft, uint16_t *__restrict__ table,
                                                                 uint8_t *__restrict__ table8, uint32_t n_entries, int intended) {
  const GlobalScores S{ft.s_full};
  for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n_entries; e += gridDim.x * blockDim.x) {
    uint32_t mask, rl;
    ans_decode(e / (uint32_t)kAnsSizes, mask, rl);
    const int size = (int)(e % (uint32_t)kAnsSizes) + 1;
    const uint16_t a = ans_compute_entry(ft.nt, mask, rl, size, ft.part_words, ft.part_index, ft.part_key, S, intended != 0);
    table[e] = a;
    table8[e] = (uint8_t)ans_chosen(a);
  }
}

// ---- shared-memory layout common to the batch kernels ---------------------------------------------------------
__host__ __device__ constexpr size_t table_ans_bytes(uint32_t n_entries) { return ((size_t)n_entries + 15) & ~(size_t)15; }  // one byte each
constexpr size_t kTabSmallTables = 256 * sizeof(unsigned long long) + 256 * sizeof(uint16_t);  // s_dev + tern
constexpr size_t masks_smem_bytes(uint32_t n_entries) { return kTabSmallTables + table_ans_bytes(n_entries); }
constexpr size_t table_smem_bytes(uint32_t n_entries, int tile_records, int stages) {
  return (size_t)stages * tile_records * sizeof(gpa_request)  // request ring (first: 128-byte aligned)
         + kTabSmallTables + table_ans_bytes(n_entries) + (size_t)2 * stages * sizeof(unsigned long long);  // + the stages' mbarriers (full, empty)
}

__device__ __forceinline__ void stage_tables(const FastTopo &ft, unsigned long long *s_dev, uint16_t *s_tern, uint8_t *s_ans, int tid,
                                             int nthreads) {
  // answers: 16 bytes per thread per step (the table is padded to a multiple of 16 bytes on the device)
  const uint4 *src = reinterpret_cast<const uint4 *>(ft.ans8);
  uint4 *dst = reinterpret_cast<uint4 *>(s_ans);
  const uint32_t n16 = (uint32_t)(table_ans_bytes(ft.ans_entries) / 16);
  for (uint32_t i = tid; i < n16; i += nthreads) dst[i] = __ldg(src + i);
  if (tid < 256) {
    s_tern[tid] = (uint16_t)ans_tern((uint32_t)tid);
    // chosen label mask -> the NVML indices of its devices, one per byte in label order, then GPA_RES_END
    unsigned long long dev = 0x8080808080808080ull;
    int o = 0;
#pragma unroll
    for (int lab = 0; lab < 8; lab++)
      if (tid >> lab & 1) {
        dev ^= (unsigned long long)((ft.order[lab] ^ 0x80) & 0xFF) << (8 * o);
        o++;
      }
    s_dev[tid] = dev;
  }
}

// ---- mbarrier / bulk-copy primitives (sm_90+ PTX; the copy reads this GPU's own memory, no multicast) -------------
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, uint32_t arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(arrivals));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(parity)
      : "memory");
}
// arrive, and say how many arrivals the phase was still waiting for BEFORE this one (1 = this one completed it)
__device__ __forceinline__ uint32_t mbar_arrive_pending(unsigned long long *bar) {
  unsigned long long state;
  uint32_t pending;
  asm volatile("mbarrier.arrive.shared::cta.b64 %0, [%1];" : "=l"(state) : "r"(smem_addr(bar)) : "memory");
  asm volatile("mbarrier.pending_count.b64 %0, %1;" : "=r"(pending) : "l"(state));
  return pending;
}
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gmem_src, uint32_t bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}

// Byte tricks of the relabel step.  nibbles_of_bytes: bytes 0x0a 0x0b 0x0c 0x0d (low to high) -> 0xdcba, the selector
// of a byte permute; or_of_bytes: OR of the four bytes; bytes_of_bits: bits 0..3 -> bytes 0x00 / 0xFF.
__device__ __forceinline__ uint32_t nibbles_of_bytes(uint32_t w) {
  const uint32_t t = (w | (w >> 4)) & 0x00FF00FFu;
  return (t | (t >> 8)) & 0xFFFFu;
}
__device__ __forceinline__ uint32_t or_of_bytes(uint32_t x) {
  x |= x >> 16;
  return (x | (x >> 8)) & 0xFFu;
}
__device__ __forceinline__ uint32_t bytes_of_bits(uint32_t b4) { return ((b4 * 0x00204081u) & 0x01010101u) * 0xFFu; }

// 32-byte gpa_request in, 16-byte gpa_result out.  A record the table cannot answer (devices not in node order, unknown
// device) is marked kResDefer while the tiles stream through; a CTA that marked any goes over those tiles once more when
// it is done and answers them itself, one warp per record, through general_request() - with the request ring, idle by
// then, as the warps' scratch.  A batch in node order (every batch the reference's call site can produce) never enters
// that part, and the whole step is ONE launch.
// reqs must be 16-byte aligned (bulk copies), out 16-byte aligned.
constexpr int kTabFlagWords = 64;  // tiles of one CTA whose "has a marked record" bit is kept; later tiles share the last bit
template <int kTabThreads, int kTabStages, int kRec>
__global__ void __launch_bounds__(kTabThreads, kTabThreads >= 1024 ? 1 : 2) besteffort_batch_table_kernel(const gpa_request *__restrict__ reqs, int count,
                                                                                const FastTopo ft, gpa_result *__restrict__ out,
                                                                                const BatchTopo topo, int n_cap, int intended) {
  constexpr int kTabTile = kTabThreads * kRec;  // records per tile: kRec per thread, kTabThreads apart
  extern __shared__ __align__(128) unsigned char tsmem[];
  gpa_request *ring = reinterpret_cast<gpa_request *>(tsmem);
  unsigned long long *s_dev = reinterpret_cast<unsigned long long *>(ring + (size_t)kTabStages * kTabTile);
  uint16_t *s_tern = reinterpret_cast<uint16_t *>(s_dev + 256);
  uint8_t *s_ans = reinterpret_cast<uint8_t *>(s_tern + 256);
  unsigned long long *full = reinterpret_cast<unsigned long long *>(s_ans + table_ans_bytes(ft.ans_entries));
  unsigned long long *empty = full + kTabStages;  // one arrival per warp: "my lanes hold their records of this stage"
  __shared__ uint32_t s_marked[kTabFlagWords];
  __shared__ uint2 s_keep[9];  // s_keep[n] = byte mask of the first n of 8 bytes, as two words

  const int tid = threadIdx.x;
  const int n_tiles = (int)(((long long)count + kTabTile - 1) / kTabTile);  // <= 2^21
  // tile i of this CTA is global tile blockIdx.x + i * gridDim.x
  const int my_tiles = (int)blockIdx.x < n_tiles ? (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto issue = [&](int i, int s) {  // one thread: fill stage s with this CTA's tile i
    const long long base = ((long long)blockIdx.x + (long long)i * gridDim.x) * kTabTile;
    const long long left = (long long)count - base;
    const uint32_t bytes = (uint32_t)(left < kTabTile ? left : kTabTile) * (uint32_t)sizeof(gpa_request);
    mbar_expect_tx(&full[s], bytes);
    bulk_load(ring + (size_t)s * kTabTile, reqs + base, bytes, &full[s]);
  };
  if (tid < kTabFlagWords) s_marked[tid] = 0;
  if (tid < 9) s_keep[tid] = make_uint2(tid >= 4 ? 0xFFFFFFFFu : (1u << (8 * tid)) - 1u, tid <= 4 ? 0u : (tid >= 8 ? 0xFFFFFFFFu : (1u << (8 * (tid - 4))) - 1u));
  if (tid == 0) {
    for (int s = 0; s < kTabStages; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kTabThreads / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // the first tiles are on their way while the CTA stages its tables
    for (int i = 0; i < kTabStages && i < my_tiles; i++) issue(i, i);
  }
  stage_tables(ft, s_dev, s_tern, s_ans, tid, kTabThreads);
  // device NVML index -> its bit in the node order, an 8-byte table in two registers (a node of up to 8 GPUs has NVML
  // indices 0..7)
  uint32_t bit_lo = 0, bit_hi = 0;
#pragma unroll
  for (int d = 0; d < 8; d++) {
    const int lab = ft.rank[d];
    const uint32_t b = (lab < 8 ? (1u << lab) : 0u) << (8 * (d & 3));
    if (d < 4) bit_lo |= b;
    else bit_hi |= b;
  }
  __syncthreads();

  int s = 0;
  uint32_t parity = 0;
  long long r = (long long)blockIdx.x * kTabTile + tid;
  for (int i = 0; i < my_tiles; i++, r += (long long)gridDim.x * kTabTile) {
    mbar_wait(&full[s], parity);
    // this lane's records: device bytes 0..7 and the 8 bytes of masks / sizes / flags (devices 8..15 cannot belong to a
    // request this table answers)
    uint2 dv[kRec], hd[kRec];
#pragma unroll
    for (int q = 0; q < kRec; q++) {
      dv[q] = make_uint2(0, 0);
      hd[q] = make_uint2(0, 0);
      if (r + q * kTabThreads < count) {
        const uint2 *rec = reinterpret_cast<const uint2 *>(ring + (size_t)s * kTabTile + q * kTabThreads + tid);
        dv[q] = rec[0];
        hd[q] = rec[2];
      }
    }
    // The warps do not wait for each other here: each says "my lanes hold their records" on the stage's second barrier,
    // and the warp whose arrival is the last one refills the stage.  A block-wide barrier in this place makes all 32 warps
    // start every tile together, and then queue for the same pipes together.
    __syncwarp();
    if ((tid & 31) == 0 && mbar_arrive_pending(&empty[s]) == 1 && i + kTabStages < my_tiles) {
      mbar_wait(&empty[s], parity);  // complete already: taken for its acquire of the other warps' arrivals
      issue(i + kTabStages, s);
    }
    if (++s == kTabStages) {
      s = 0;
      parity ^= 1u;
    }

    bool defer = false;
#pragma unroll
    for (int q = 0; q < kRec; q++) {
      const long long rq = r + q * kTabThreads;
      if (rq >= count) continue;
      const uint2 dvq = dv[q], hdq = hd[q];
      // Straight-line code: every lane computes a lookup (with indices made safe) and the guards select afterwards - the
      // lanes of a warp hold requests of every kind, so branches would only serialise them.
      const uint32_t rmask = hdq.x & 0xFFFFu;
      const int n = (int)((hdq.x >> 16) & 0xFFu);
      const int size = (int)(hdq.x >> 24);
      const int n_req = __popc(rmask) + (int)((hdq.y >> 8) & 0xFFu);
      // guards of besteffort_policy.go:35-45: the empty answer, whatever the device list holds
      const bool empty = size <= 0 || n < size || n_req > size;
      // relabel: device -> its bit in the node order, four devices per PRMT: the 8-byte "device -> label bit" table sits in
      // two registers, a byte-permute selector holds four table indices, one per nibble.  Bytes beyond the n listed
      // devices are cleared afterwards (s_keep[n] = byte mask of the first n bytes); a listed byte that is no device of
      // the node (>= 8) makes the lookup meaningless, and is caught by the comparison below.
      const int nc = n < 8 ? n : 8;
      const uint2 keep = s_keep[nc];
      const uint32_t bits_lo = __byte_perm(bit_lo, bit_hi, nibbles_of_bytes(dvq.x)) & keep.x;
      const uint32_t bits_hi = __byte_perm(bit_lo, bit_hi, nibbles_of_bytes(dvq.y)) & keep.y;
      const uint32_t mask = or_of_bytes(bits_lo | bits_hi);
      // The list must be strictly ascending in the node order.  A set of devices has exactly one such list, the one the
      // result path keeps in s_dev[mask] (devices, then GPA_RES_END bytes): n distinct devices whose bytes equal it are
      // known, and in order.
      const unsigned long long want = s_dev[mask];
      const uint32_t act_lo = (dvq.x & keep.x) | (0x80808080u & ~keep.x), act_hi = (dvq.y & keep.y) | (0x80808080u & ~keep.y);
      const bool bad = n > 8 || (rmask >> nc) != 0 || __popc(mask) != n || (uint32_t)want != act_lo || (uint32_t)(want >> 32) != act_hi;
      // required positions -> label bits: spread the 8 position bits into byte masks and select
      const uint32_t rl = or_of_bytes((bits_lo & bytes_of_bits(rmask & 0xFu)) | (bits_hi & bytes_of_bits((rmask >> 4) & 0xFu)));
      const int sz = size < 1 ? 1 : (size > 8 ? 8 : size);
      uint32_t chosen = s_ans[ans_index(s_tern, mask, rl, sz)];
      // no answer to look up: guards, a list the table does not know, or a required pointer that is in no set (:53)
      if (empty || bad || (hdq.y & GPA_REQ_FOREIGN)) chosen = 0;
      const int pc = __popc(chosen);
      // nil pads follow the devices (D5): bytes pc .. size-1 turn from GPA_RES_END (0x80) into GPA_NIL (0xFF)
      const uint2 k_sz = s_keep[chosen ? sz : 0], k_pc = s_keep[pc];
      const unsigned long long devs = s_dev[chosen];
      uint4 v;
      v.x = (uint32_t)devs ^ (k_sz.x & ~k_pc.x & 0x7F7F7F7Fu);
      v.y = (uint32_t)(devs >> 32) ^ (k_sz.y & ~k_pc.y & 0x7F7F7F7Fu);
      const bool defer_q = bad && !empty;
      if (defer_q) v.x = 0x80808000u | (uint32_t)(uint8_t)kResDefer;
      defer = defer || defer_q;
      v.z = 0x80808080u;
      v.w = 0x80808080u;
      __stcs(reinterpret_cast<uint4 *>(out + rq), v);
    }
    if (__any_sync(0xffffffffu, defer) && (tid & 31) == 0) {
      const int bit = i < kTabFlagWords * 32 ? i : kTabFlagWords * 32 - 1;
      atomicOr(&s_marked[bit >> 5], 1u << (bit & 31));
    }
  }

  // ---- the records this CTA marked (none for a batch in node order) ------------------------------------------------
  __syncthreads();  // every lane's marks and results are visible to the CTA; the ring is idle
  bool any_marked = false;
  for (int w = 0; w < kTabFlagWords; w++) any_marked = any_marked || s_marked[w] != 0;
  if (!any_marked) return;
  // scratch in the ring: the binomial table, then per warp a 256-entry score table and the position weights
  BinomTable *s_binom = reinterpret_cast<BinomTable *>(ring);
  uint16_t *scratch = reinterpret_cast<uint16_t *>(reinterpret_cast<unsigned char *>(ring) + ((sizeof(BinomTable) + 127) & ~(size_t)127));
  stage_binom(*s_binom, tid, kTabThreads);
  __syncthreads();
  const int lane = tid & 31, warp = tid >> 5;
  uint16_t *S = scratch + (size_t)warp * (256 + kMaxDevices * kMaxDevices);
  uint16_t *wp = S + 256;
  for (int i = 0; i < my_tiles; i++) {
    const int bit = i < kTabFlagWords * 32 ? i : kTabFlagWords * 32 - 1;
    if (!(s_marked[bit >> 5] >> (bit & 31) & 1u)) continue;
    for (int q = 0; q < kRec; q++) {
      const long long base = ((long long)blockIdx.x + (long long)i * gridDim.x) * kTabTile + q * kTabThreads + warp * 32;
      const long long r = base + lane;
      const bool marked = r < count && out[r].dev[0] == kResDefer;
      unsigned todo = __ballot_sync(0xffffffffu, marked);
      while (todo) {
        const int b = __ffs(todo) - 1;
        todo &= todo - 1;
        general_request(reqs, base + b, n_cap, 8, intended, topo, out, S, wp, *s_binom, lane);
        __syncwarp();
      }
    }
  }
}

// 8-byte gpa_mask_request in, 4-byte gpa_mask_result out; nothing is ever deferred.
constexpr int kMaskThreads = 512;
__global__ void __launch_bounds__(kMaskThreads) besteffort_masks_table_kernel(const uint2 *__restrict__ mreqs, int count,
                                                                              const FastTopo ft, uint32_t *__restrict__ mout) {
  extern __shared__ __align__(128) unsigned char tsmem[];
  unsigned long long *s_dev = reinterpret_cast<unsigned long long *>(tsmem);
  uint16_t *s_tern = reinterpret_cast<uint16_t *>(s_dev + 256);
  uint8_t *s_ans = reinterpret_cast<uint8_t *>(s_tern + 256);
  const int tid = threadIdx.x;
  // the first records are asked for before the tables are staged: across PCIe (pinned caller buffers) they take
  // longer to arrive than the staging does
  const bool wide = ((reinterpret_cast<uintptr_t>(mreqs) & 15u) | (reinterpret_cast<uintptr_t>(mout) & 7u)) == 0;
  const long long n_pairs = (long long)count >> 1, stride = (long long)gridDim.x * kMaskThreads;
  long long p = (long long)blockIdx.x * kMaskThreads + tid;
  uint4 rec = make_uint4(0, 0, 0, 0);
  if (wide && p < n_pairs) rec = __ldg(reinterpret_cast<const uint4 *>(mreqs) + p);
  stage_tables(ft, s_dev, s_tern, s_ans, tid, kMaskThreads);
  __syncthreads();

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
    const uint32_t chosen = s_ans[ans_index(s_tern, mask, rl, size)];
    return chosen | ((chosen ? (uint32_t)(size - __popc(chosen)) : 0u) << 16);
  };
  if (wide) {
    // two records per lane per step: 16 bytes in, 8 bytes out; the next step's records are in flight during this one
    while (p < n_pairs) {
      const long long nxt = p + stride;
      uint4 rec_next = make_uint4(0, 0, 0, 0);
      if (nxt < n_pairs) rec_next = __ldg(reinterpret_cast<const uint4 *>(mreqs) + nxt);
      uint2 res;
      res.x = answer(rec.x, rec.y);
      res.y = answer(rec.z, rec.w);
      reinterpret_cast<uint2 *>(mout)[p] = res;
      rec = rec_next;
      p = nxt;
    }
    if ((count & 1) && blockIdx.x == 0 && tid == 0) {
      const uint2 rec = __ldg(mreqs + (count - 1));
      mout[count - 1] = answer(rec.x, rec.y);
    }
  } else {  // buffers that only have the records' own 8 / 4-byte alignment
    for (long long r = (long long)blockIdx.x * kMaskThreads + tid; r < count; r += (long long)gridDim.x * kMaskThreads) {
      const uint2 rec = __ldg(mreqs + r);
      mout[r] = answer(rec.x, rec.y);
    }
  }
}

// 4-byte gpa_mask8_request in, 2-byte gpa_mask8_result out: the compact format of nodes of up to 8 GPUs.  Four records
// per lane per step (16 bytes in, 8 bytes out) when the buffers are aligned for it.
__global__ void __launch_bounds__(kMaskThreads) besteffort_masks8_table_kernel(const uint32_t *__restrict__ reqs, int count,
                                                                               const FastTopo ft, uint16_t *__restrict__ out) {
  extern __shared__ __align__(128) unsigned char tsmem[];
  unsigned long long *s_dev = reinterpret_cast<unsigned long long *>(tsmem);
  uint16_t *s_tern = reinterpret_cast<uint16_t *>(s_dev + 256);
  uint8_t *s_ans = reinterpret_cast<uint8_t *>(s_tern + 256);
  const int tid = threadIdx.x;
  // as in the 8-byte kernel: first records requested before the tables are staged
  const bool wide = ((reinterpret_cast<uintptr_t>(reqs) & 15u) | (reinterpret_cast<uintptr_t>(out) & 7u)) == 0;
  const long long n_quads = (long long)count >> 2, stride = (long long)gridDim.x * kMaskThreads;
  long long q = (long long)blockIdx.x * kMaskThreads + tid;
  uint4 rec = make_uint4(0, 0, 0, 0);
  if (wide && q < n_quads) rec = __ldg(reinterpret_cast<const uint4 *>(reqs) + q);
  stage_tables(ft, s_dev, s_tern, s_ans, tid, kMaskThreads);
  __syncthreads();

  const uint32_t node_mask = (1u << ft.nt) - 1u;
  auto answer = [&](uint32_t x) -> uint32_t {
    // {avail_mask:8, req_mask:8, size:8, flags:8} -> {chosen_mask:8, n_nil:8}; n_nil 0x80 = a device the node does not have
    const uint32_t mask = x & 0xFFu, rl = (x >> 8) & 0xFFu;
    const int size = (int)((x >> 16) & 0xFFu);
    if (((mask | rl) & ~node_mask) != 0) return 0x8000u;
    // guards of besteffort_policy.go:35-45; a required device that is not available is in no set (:53)
    if (size <= 0 || __popc(mask) < size || __popc(rl) > size || ((x >> 24) & GPA_REQ_FOREIGN) || (rl & ~mask) != 0) return 0u;
    const uint32_t chosen = s_ans[ans_index(s_tern, mask, rl, size)];
    return chosen | ((chosen ? (uint32_t)(size - __popc(chosen)) : 0u) << 8);  // the 2-byte entry of ans_table.cuh
  };
  if (wide) {
    while (q < n_quads) {
      const long long nxt = q + stride;
      uint4 rec_next = make_uint4(0, 0, 0, 0);
      if (nxt < n_quads) rec_next = __ldg(reinterpret_cast<const uint4 *>(reqs) + nxt);
      uint2 res;
      res.x = answer(rec.x) | (answer(rec.y) << 16);
      res.y = answer(rec.z) | (answer(rec.w) << 16);
      reinterpret_cast<uint2 *>(out)[q] = res;
      rec = rec_next;
      q = nxt;
    }
    if (blockIdx.x == 0 && tid < (count & 3)) {
      const long long r = (n_quads << 2) + tid;
      out[r] = (uint16_t)answer(__ldg(reqs + r));
    }
  } else {
    for (long long r = (long long)blockIdx.x * kMaskThreads + tid; r < count; r += (long long)gridDim.x * kMaskThreads)
      out[r] = (uint16_t)answer(__ldg(reqs + r));
  }
}

}  // namespace
