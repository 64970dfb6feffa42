// Batch kernels: MANY requests per launch (general warp-per-request kernel, fast tile-binned lane-per-request kernel).
// Part of the single translation unit gpualloc.cu (the kernels share __constant__ tables and the
// enumeration core); not compiled on its own.
#pragma once
#include "kernels_common.cuh"
#include "part_table.cuh"

namespace {

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

// One request answered by one WARP: any `available` order, per-request score table S (2^s_log2 uint16) and weights wp
// (kMaxDevices^2 uint16) in this warp's shared memory.  The general path of the batch kernels.
__device__ __noinline__ void general_request(const gpa_request *__restrict__ reqs, long long r, int n_cap, int s_log2, int intended,
                                             const BatchTopo topo, gpa_result *out, uint16_t *S, uint16_t *wp,
                                             const BinomTable &s_binom, int lane) {
  const STab stab{S};
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
      // the caller's S holds 2^s_log2 masks (the table kernel passes 8: more distinct devices than its node has cannot
      // be listed, so a longer list has failed the checks above already)
      if (!err && n > s_log2) err = GPA_ERR_INVALID;
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
    // the request's weights by position; a pair whose link counts differ is poisoned, so that the search itself finds
    // out whether Go would ever have scored it (kAsymPoison, kernels_common.cuh)
    for (int i = lane; i < n * n; i += 32) {
      int a = i / n, b = i - a * n;
      int da = dev_at(a), db = dev_at(b);
      if (a < b) wp[a * kMaxDevices + b] = topo.asym[da * topo.nt + db] ? (uint16_t)kAsymPoison : topo.wq[da * topo.nt + db];
    }
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
      if (best != 0 && key_score(best) >= (int)kAsymPoison) {
        err = GPA_ERR_ASYMMETRIC;  // an accepted candidate holds such a pair: the panic of :307-310 while iterating
      } else if (cnt < 0) {
        err = GPA_ERR_INVALID;
      } else {
        // the chosen set is the best-scoring set of the saved partition that holds the required devices: if any of
        // those sets held such a pair it is this one (the panic inside the loop of :77-85)
        uint32_t chosen = 0;
        for (int i = 0; i < cnt; i++)
          if (pos[i] >= 0) chosen |= 1u << pos[i];
        if (cnt > 0 && S[chosen] >= kAsymPoison) {
          err = GPA_ERR_ASYMMETRIC;
        } else {
          for (int i = 0; i < cnt; i++) res[i] = pos[i] < 0 ? (int8_t)GPA_NIL : (int8_t)dev_at(pos[i]);
        }
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
}

// General batch kernel: one warp per request through general_request().
// With deferred_only it answers just the records the tile-binned kernel marked kResDefer: each warp scans 32 result
// records at a time (one coalesced load) and works on the marked ones.
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

  if (!deferred_only) {
    for (long long r = (long long)blockIdx.x * kBatchWarps + warp; r < count; r += (long long)gridDim.x * kBatchWarps)
      general_request(reqs, r, n_cap, n_cap, intended, topo, out, S, wp, s_binom, lane);
  } else {
    for (long long base = ((long long)blockIdx.x * kBatchWarps + warp) * 32; base < count;
         base += (long long)gridDim.x * kBatchWarps * 32) {
      const long long r = base + lane;
      const bool marked = r < count && out[r].dev[0] == kResDefer;
      unsigned todo = __ballot_sync(0xffffffffu, marked);
      while (todo) {
        const int b = __ffs(todo) - 1;
        todo &= todo - 1;
        general_request(reqs, base + b, n_cap, n_cap, intended, topo, out, S, wp, s_binom, lane);
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
constexpr int kShapeBins = 176;  // 2 x (shape bin of c_shape_bin, 66 shapes with 2 <= k <= n <= 12) + (has required devices)

// Partition lists for small nodes.  Which partitions iterate() visits, and in which order, depends only on the
// availability mask and the size, not on the weights; for a node of up to 8 GPUs all of them (6 289 for 8 GPUs) fit a
// 25 KB table: one 32-bit word per partition, byte j = device mask of its j-th set (a padded set is the one with
// fewer than k devices).  A lane then scores a partition with one table word and m score lookups instead of
// stepping through combinations.  Built once per node size on the host with the same stepper (gpualloc.cu).

struct FastTopo {
  const uint16_t *s_full;  // [1 << nt] scaled set scores over label masks
  const uint32_t *part_words;  // partition words, grouped by (mask, size); nullptr: no table (more than 8 GPUs)
  const uint32_t *part_index;  // [256 * 9]: offset | count << 16 of the group of (mask, size)
  const uint32_t *part_key;    // per partition word: its score under the installed weights << 16 | (0xFFFF - index in its group)
  const uint16_t *ans;         // answer table (ans_table.cuh), padded to 16 bytes; nullptr: none (more than 8 GPUs)
  const uint8_t *ans8;         // its chosen masks alone, one byte per entry, padded to 16 bytes
  uint32_t ans_entries;
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

// Table-driven search for nodes of up to 8 GPUs (m <= 4 sets): same result as fast_tracked_search (part_table.cuh).
__device__ __forceinline__ int fast_table_search(const Geom &g, const FastTopo &ft, const uint16_t *S, bool intended,
                                                 uint32_t &best_set) {
  return part_table_search(g, ft.part_words, ft.part_index, ft.part_key, STab{S}, intended, best_set);
}

// ONE small request on a small node, one warp: the partitions of (mask, size) come from the table, lanes stride over
// them, a shuffle max of (score, ~index) keys names the first best one, lane 0 applies the epilogue and publishes the
// answer into the mapped mailbox.  A few hundred instructions end to end: this is what a single kubelet request costs.
struct GlobalScores {
  const uint16_t *s;
  __device__ __forceinline__ uint32_t operator()(uint32_t mask) const { return __ldg(s + mask); }
};

__global__ void __launch_bounds__(32) besteffort_small_kernel(const FastTopo ft, uint32_t mask, uint32_t rlab, int k, int intended,
                                                              unsigned int seq, SearchResult *g_res) {
  const int lane = threadIdx.x;
  const Geom g = make_geom_sparse(mask, k, rlab, ft.nt);
  const GlobalScores S{ft.s_full};
  const uint32_t entry = __ldg(ft.part_index + mask * 9 + k);
  const uint32_t *w = ft.part_words + (entry & 0xFFFFu);
  const int count = (int)(entry >> 16);
  unsigned long long best = 0;
  for (int i = lane; i < count; i += 32) {
    const uint32_t word = __ldg(w + i);
    const uint32_t s0 = word & 0xFFu, s1 = (word >> 8) & 0xFFu, s2 = (word >> 16) & 0xFFu, s3 = word >> 24;
    const int sc = (int)S(s0) + (int)S(s1) + (int)S(s2) + (int)S(s3);
    bool ok = true;
    if (rlab != 0)
      ok = ((s0 & rlab) == rlab && __popc(s0) == k) || ((s1 & rlab) == rlab && __popc(s1) == k) ||
           ((s2 & rlab) == rlab && __popc(s2) == k) || ((s3 & rlab) == rlab && __popc(s3) == k);
    if (ok) {
      const unsigned long long kk = pack_key(sc, (unsigned long long)i);
      best = kk > best ? kk : best;
    }
  }
  best = warp_max_u64(best);
  if (lane == 0) {
    SearchResult r;
    r.key = best;
    r.count = 0;
    r.best_score = best ? key_score(best) : -1;
#pragma unroll
    for (int q = 0; q < 32; q++) r.pos[q] = 0;
    if (best != 0) {
      const uint32_t word = __ldg(w + (int)key_item(best));
      uint32_t sets[kMaxSuffix];
#pragma unroll
      for (int j = 0; j < kMaxSuffix; j++) {
        const uint32_t sj = (word >> (8 * j)) & 0xFFu;
        sets[j] = sj | ((j < g.m && __popc(sj) < g.k) ? g.tok_bit : 0u);
      }
      uint32_t best_set = 0;
      if (choose_from_partition(g, S, sets, intended != 0, best_set) > 0) {
        int o = 0;
        uint32_t reals = best_set & g.real_mask;
        while (reals) {
          const uint32_t b = reals & (0u - reals);
          reals ^= b;
          r.pos[o++] = (int8_t)__popc(mask & (b - 1u));  // position of that device in the caller's `available`
        }
        if (best_set & g.tok_bit)
          for (int q = 0; q < g.P; q++) r.pos[o++] = -1;
        r.count = o;
      }
    }
    publish_result(g_res, r, seq);
  }
}

__device__ __forceinline__ void res_set_byte(unsigned long long &r0, unsigned long long &r1, int i, int v) {
  // every byte starts as GPA_RES_END (0x80) and is written at most once, so a XOR sets it
  const unsigned long long b = (unsigned long long)((v ^ 0x80) & 0xFF);
  if (i < 8) r0 ^= b << (8 * i);
  else r1 ^= b << (8 * (i - 8));
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
         + (size_t)(kShapeBins + 1) * sizeof(uint32_t) * 2  // hist, offs
         + GPA_MAX_TOPOLOGY + 16;                            // rank, order
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
  // the relabelling tables are read with a different index in every lane: shared memory, not the constant bank
  uint8_t *s_rank = w_key + kTile;
  uint8_t *s_order = s_rank + GPA_MAX_TOPOLOGY;

  const int tid = threadIdx.x;
  for (uint32_t i = tid; i < (1u << ft.nt); i += kFastThreads) S[i] = ft.s_full[i];
  if (tid < GPA_MAX_TOPOLOGY) s_rank[tid] = ft.rank[tid];
  if (tid < 16) s_order[tid] = ft.order[tid];

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
              bin = 2 * c_shape_bin[n][size] + (rlab_in != 0);
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
            uint32_t mask = 0, rlab = 0;
            int prev = -1;
            bool defer = false;
            const int first_dev = (int)(lo.x & 0xFFu);
            // one step per listed device, unrolled and predicated: the lanes of a warp hold requests of every length
            auto relabel = [&](int p, uint32_t wd) {
              if (p < n) {
                const int d = (int)((wd >> (8 * (p & 3))) & 0xFFu);
                const int lab = d < GPA_MAX_TOPOLOGY ? s_rank[d] : 0xFF;
                defer = defer | (lab == 0xFF) | (lab <= prev);
                prev = lab;
                const uint32_t bit = 1u << (lab & 31);
                mask |= bit;
                rlab |= (rmask >> p & 1u) ? bit : 0u;
              }
            };
#pragma unroll
            for (int p = 0; p < 4; p++) relabel(p, lo.x);
#pragma unroll
            for (int p = 4; p < 8; p++) relabel(p, lo.y);
            if (n > 8) {
#pragma unroll
              for (int p = 8; p < 12; p++) relabel(p, lo.z);
            }
            int first_req_dev = -1;
            if (rmask != 0) {
              const int q = __ffs(rmask) - 1;
              const uint32_t wd = q < 4 ? lo.x : (q < 8 ? lo.y : (q < 12 ? lo.z : lo.w));
              first_req_dev = (int)((wd >> (8 * (q & 3))) & 0xFFu);
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
              bin = 2 * c_shape_bin[n][size] + (rlab != 0);
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
      if (ft.part_words != nullptr) {
        // small node: the partitions of (mask, size) come from the table
        rc = fast_table_search(g, ft, S, intended != 0, best_set);
      } else if (g.m <= kMaxSuffix) {
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
            res_set_byte(r0, r1, o++, s_order[lab]);
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

}  // namespace
