// Answer table of small nodes (up to 8 GPUs): the COMPLETE function bestEffortPolicy.Allocate computes on such a node.
//
// With the weights installed, the answer of
//   vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:34-89
// depends on three things only: which devices are available (a mask over the node order), which of them are required
// (a sub-mask) and the size.  A device is in one of three states - absent, available, available and required - so a
// node of nt GPUs has 3^nt (mask, required) pairs; times 8 sizes that is 52 488 answers for 8 GPUs, two bytes each:
//     entry = chosen device mask | nil pads in the returned set << 8          (0 = the empty answer)
//     index = (tern[mask] + tern[required]) * 8 + size - 1,   tern[x] = sum of 3^i over the bits i of x
// One entry is computed from the partition tables (part_table.cuh) with the same search the batch kernel used to run
// per request; a kernel fills the whole table when the topology is installed (table_kernels.cuh) and the batch
// kernels - and, through a host copy, gpa_besteffort - then answer a request with ONE lookup.
//
// Both alias modes share the table: with at most 8 devices and k >= 2 a partition has m <= 4 sets, and the slice
// aliasing of the Go code only bites from m = 7; tests/test_enum_core.py checks that on every entry.
// Compiles under nvcc and under g++ (tests/csrc/enum_check.cpp drives it on the CPU).
#pragma once
#include <stdint.h>

#include "part_table.cuh"

namespace gpa {

constexpr int kAnsMaxTopo = kPartTableMaxTopo;  // 8
constexpr int kAnsSizes = 8;                    // sizes 1..8

GPA_HD uint32_t ans_pow3(int nt) {
  uint32_t p = 1;
  for (int i = 0; i < nt; i++) p *= 3u;
  return p;
}
GPA_HD uint32_t ans_entries(int nt) { return ans_pow3(nt) * (uint32_t)kAnsSizes; }

// tern[x], x < 256: the digit of device i is 1 where x has bit i
GPA_HD uint32_t ans_tern(uint32_t x) {
  uint32_t t = 0, p = 1;
  for (int i = 0; i < 8; i++, p *= 3u) t += (x >> i & 1u) ? p : 0u;
  return t;
}

// index of (mask, required subset of mask, size 1..8); tern = table of ans_tern values
GPA_HD uint32_t ans_index(const uint16_t *tern, uint32_t mask, uint32_t rl, int size) {
  return ((uint32_t)tern[mask] + (uint32_t)tern[rl]) * (uint32_t)kAnsSizes + (uint32_t)(size - 1);
}

// the (mask, required) pair of a ternary number: digit 0 absent, 1 available, 2 available and required
GPA_HD void ans_decode(uint32_t idx3, uint32_t &mask, uint32_t &rl) {
  mask = 0;
  rl = 0;
  for (int i = 0; i < 8; i++) {
    const uint32_t d = idx3 % 3u;
    idx3 /= 3u;
    mask |= (d >= 1u ? 1u : 0u) << i;
    rl |= (d == 2u ? 1u : 0u) << i;
  }
}

// One entry: what Allocate returns for (mask, rl, size) on a node of nt GPUs whose partition tables and set scores
// (S: label mask -> scaled score) are given.
template <typename STabT>
GPA_HD uint16_t ans_compute_entry(int nt, uint32_t mask, uint32_t rl, int size, const uint32_t *part_words,
                                  const uint32_t *part_index, const uint32_t *part_key, const STabT &S, bool intended) {
  const int n = popc32(mask);
  // besteffort_policy.go:35-45 (size <= 0 cannot be indexed; len(required) > size; len(available) < size)
  if (size < 1 || n < size || popc32(rl) > size) return 0;
  if (size == 1) return (uint16_t)low_bit(rl ? rl : mask);  // :219-224 - first accepted single-device candidate
  const Geom g = make_geom_sparse(mask, size, rl, nt);
  uint32_t best_set = 0;
  if (part_table_search(g, part_words, part_index, part_key, S, intended, best_set) <= 0) return 0;
  return (uint16_t)((best_set & g.real_mask) | ((best_set & g.tok_bit) ? ((uint32_t)g.P << 8) : 0u));
}

GPA_HD uint32_t ans_chosen(uint16_t entry) { return entry & 0xFFu; }
GPA_HD int ans_nil(uint16_t entry) { return (entry >> 8) & 7; }

}  // namespace gpa
