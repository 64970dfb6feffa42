// Partition tables of small nodes (up to 8 GPUs): builders (host) and the search that reads them (host + device).
//
// Which partitions iterateGPUPartitions visits for (availability mask, size), and in which order, depends on neither
// the weights nor the request's required devices (besteffort_policy.go:209-290), so for a node of up to 8 GPUs all of
// them fit a table: one 32-bit WORD per partition (byte j = device mask of its j-th set; the padded set is the one with
// fewer than k devices), grouped by (mask, size), 6 289 words for 8 GPUs.  A partition's score does depend on the
// weights, so it is tabulated when they are installed, packed with the partition's place in its group:
//     KEY = score << 16 | (0xFFFF - index in group)
// and the first best partition in visiting order (strict `>`, :57) is the LARGEST key among the partitions that pass
// the `required` filter (:53).  Compiles under nvcc and under g++ (tests/csrc/enum_check.cpp drives it on the CPU).
#pragma once
#include <stdint.h>

#include <vector>

#include "gpa_enum.cuh"

#if defined(__CUDA_ARCH__)
#define GPA_LDG(p) __ldg(p)
#else
#define GPA_LDG(p) (*(p))
#endif

namespace gpa {

constexpr int kPartTableMaxTopo = 8;

// Every partition iterate() visits for one geometry, in visiting order, generated with the stepper the kernels walk with.
inline void gen_partitions(const Geom &g, uint32_t rem, int level, uint32_t word, std::vector<uint32_t> &out) {
  if (level == g.m - 1) {  // base case :245-248 — the remainder is the last set
    out.push_back(word | ((rem & g.real_mask) << (8 * level)));
    return;
  }
  uint32_t anchor;
  Stepper st = make_stepper(g, rem, anchor);
  uint32_t chosen;
  if (!st.first(chosen)) return;
  do {
    const uint32_t set = anchor | chosen;
    gen_partitions(g, rem & ~set, level + 1, word | ((set & g.real_mask) << (8 * level)), out);
  } while (st.next(chosen));
}

// words + index[mask * 9 + k] = offset | count << 16 for a node of nt <= 8 GPUs.  false = does not fit (cannot happen).
inline bool build_partition_words(int nt, std::vector<uint32_t> &words, std::vector<uint32_t> &index) {
  words.clear();
  index.assign(256 * 9, 0);
  for (uint32_t mask = 1; mask < (1u << nt); mask++) {
    const int n = __builtin_popcount(mask);
    for (int k = 2; k <= n; k++) {
      const Geom g = make_geom_sparse(mask, k, 0, nt);
      const size_t off = words.size();
      gen_partitions(g, g.real_mask | g.tok_bit, 0, 0, words);
      const size_t cnt = words.size() - off;
      if (off > 0xFFFF || cnt > 0xFFFF) return false;  // 6 289 words for nt = 8
      index[mask * 9 + k] = (uint32_t)off | ((uint32_t)cnt << 16);
    }
  }
  if (words.empty()) words.push_back(0);
  return true;
}

// keys under the installed weights; sfull[mask] = scaled score of the device set `mask` (label space).  A partition's
// sets are disjoint, so its pairs are among the node's 120 at most and the sum stays within 16 bits (gpa_set_topology).
inline void build_partition_keys(int nt, const std::vector<uint32_t> &words, const std::vector<uint32_t> &index, const uint16_t *sfull,
                                 std::vector<uint32_t> &keys) {
  keys.assign(words.size(), 0);
  for (uint32_t mask = 1; mask < (1u << nt); mask++)
    for (int k = 2; k <= 8; k++) {
      const uint32_t e = index[mask * 9 + k];
      const uint32_t off = e & 0xFFFFu, cnt = e >> 16;
      for (uint32_t i = 0; i < cnt; i++) {
        const uint32_t wd = words[off + i];
        const uint32_t sc = (uint32_t)sfull[wd & 0xFFu] + sfull[(wd >> 8) & 0xFFu] + sfull[(wd >> 16) & 0xFFu] + sfull[wd >> 24];
        keys[off + i] = ((sc > 0xFFFFu ? 0xFFFFu : sc) << 16) | (0xFFFFu - i);
      }
    }
}

GPA_HD int part_first_bit(uint32_t x) {  // x != 0
#if defined(__CUDA_ARCH__)
  return __ffs(x) - 1;
#else
  return __builtin_ctz(x);
#endif
}

GPA_HD uint32_t part_umax(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return max(a, b);
#else
  return a > b ? a : b;
#endif
}

// Largest key of the group that passes the `required` filter; 0 = none.
GPA_HD uint32_t part_best_key(const uint32_t *w, const uint32_t *pk, int count, uint32_t rl, int k) {
  uint32_t best = 0;
  if (rl == 0) {
    int i = 0;
    for (; i < count && ((reinterpret_cast<uintptr_t>(pk + i) & 15u) != 0); i++) best = part_umax(best, GPA_LDG(pk + i));
#pragma unroll 2
    for (; i + 4 <= count; i += 4) {
#if defined(__CUDA_ARCH__)
      const uint4 q = __ldg(reinterpret_cast<const uint4 *>(pk + i));
      best = max(best, max(max(q.x, q.y), max(q.z, q.w)));
#else
      best = part_umax(best, part_umax(part_umax(pk[i], pk[i + 1]), part_umax(pk[i + 2], pk[i + 3])));
#endif
    }
    for (; i < count; i++) best = part_umax(best, GPA_LDG(pk + i));
  } else {
    // :53 / :112-119 — some set holds every required device and is not the padded one.  Only the set that holds the
    // LOWEST required device can do that, and every device is in exactly one byte of the word, so that set is picked out
    // without a branch: no lane of the warp waits for another one's short-circuit.
    const uint32_t low4 = (rl & (0u - rl)) * 0x01010101u;
#pragma unroll 4
    for (int i = 0; i < count; i++) {
      const uint32_t word = GPA_LDG(w + i);
      const uint32_t key = GPA_LDG(pk + i);
      const int sh = part_first_bit(word & low4) & 24;
      const uint32_t set = (word >> sh) & 0xFFu;
      const bool ok = ((set & rl) == rl) & (popc32(set) == k);
      best = part_umax(best, ok ? key : 0u);
    }
  }
  return best;
}

// Table-driven search for nodes of up to 8 GPUs (m <= 4 sets): the answer of bestEffortPolicy.Allocate for geometry g.
// returns like finalize_set: 1 and the set, 0 for the empty answer.
template <typename STabT>
GPA_HD int part_table_search(const Geom &g, const uint32_t *part_words, const uint32_t *part_index, const uint32_t *part_key,
                             const STabT &S, bool intended, uint32_t &best_set) {
  const uint32_t entry = GPA_LDG(part_index + g.real_mask * 9 + g.k);
  const uint32_t *w = part_words + (entry & 0xFFFFu);
  const int count = (int)(entry >> 16);
  const uint32_t *pk = part_key + (entry & 0xFFFFu);
  const uint32_t best = part_best_key(w, pk, count, g.rmask, g.k);
  best_set = 0;
  if (best == 0) return 0;  // no accepted partition (a key is never 0: index < 0xFFFF)
  const uint32_t best_word = GPA_LDG(w + (0xFFFFu - (best & 0xFFFFu)));
  uint32_t sets[kMaxSuffix];
#pragma unroll
  for (int j = 0; j < kMaxSuffix; j++) {
    const uint32_t sj = (best_word >> (8 * j)) & 0xFFu;
    sets[j] = sj | ((j < g.m && popc32(sj) < g.k) ? g.tok_bit : 0u);
  }
  return choose_from_partition(g, S, sets, intended, best_set);
}

}  // namespace gpa
