// Host-side planning of one replica-set search (gpa_subset_besteffort): argument checks, guards, grouping of the
// positions by GPU, the parameter blocks of the two kernels, and the way back from the winning rank to positions.
// Pure C++ (no CUDA calls) so that tests/csrc/subset_check.cpp can drive the same code on the CPU with a loop in place
// of the kernels.  Definition and reference citations: subset_core.cuh.
#pragma once
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../include/gpualloc.h"
#include "subset_core.cuh"

namespace gpa {

struct SubsetPlan {
  bool done = false;     // answered without a search (guards, or the required positions fill the request)
  bool literal = false;  // literal combinations (SubsetEnum) instead of count vectors (SubsetCounts)
  int size = 0, K = 0, G = 0, R = 0;
  unsigned long long total = 0;  // candidates of the form used
  std::vector<uint8_t> is_req, gid;
  int gpu_of[kSubGroups];
  std::vector<int32_t> answer;    // required positions (ascending); the decode step adds the chosen ones
  std::vector<int32_t> free_pos;  // non-required positions, ascending
  SubsetCounts qc;
  std::vector<uint64_t> cnt;
  SubsetEnum qe;
};

// -> GPA_OK (plan.done tells whether a search is needed) or a GPA_ERR_* code.
// wq / asym: the context's scaled weights and link-count asymmetry flags, n_topo x n_topo; binom: sub_fill_binom table.
inline int subset_plan(int n_topo, const uint16_t *wq, const uint8_t *asym, const uint64_t *binom, const int32_t *phys, int n_avail,
                       const int32_t *req_pos, int n_req, int size, int mode_flags, SubsetPlan &pl) {
  pl = SubsetPlan();
  if (n_avail > GPA_SUBSET_MAX_AVAILABLE) return GPA_ERR_TOO_LARGE;
  for (int p = 0; p < n_avail; p++)
    if (phys[p] < 0 || phys[p] >= n_topo) return GPA_ERR_INVALID;
  pl.is_req.assign((size_t)n_avail + 1, 0);
  for (int i = 0; i < n_req; i++) {
    if (req_pos[i] < 0 || req_pos[i] >= n_avail || pl.is_req[req_pos[i]]) return GPA_ERR_INVALID;
    pl.is_req[req_pos[i]] = 1;
  }
  pl.done = true;
  // besteffort_policy.go:35-45
  if (size <= 0) return GPA_OK;
  if (n_avail < size) return GPA_OK;
  if (n_req > size) return GPA_OK;
  if (size > GPA_SUBSET_MAX_SIZE) return GPA_ERR_TOO_LARGE;
  pl.size = size;

  // groups = GPUs in order of their first position; `grouped` = every GPU's positions are contiguous
  int gid_of[GPA_MAX_TOPOLOGY];
  for (int i = 0; i < GPA_MAX_TOPOLOGY; i++) gid_of[i] = -1;
  bool grouped = true;
  pl.gid.assign((size_t)n_avail, 0);
  for (int p = 0; p < n_avail; p++) {
    int g = gid_of[phys[p]];
    if (g < 0) {
      if (pl.G == kSubGroups) return GPA_ERR_TOO_LARGE;
      g = pl.G++;
      gid_of[phys[p]] = g;
      pl.gpu_of[g] = phys[p];
    } else if (phys[p - 1] != phys[p]) {
      grouped = false;
    }
    pl.gid[p] = (uint8_t)g;
  }
  const int G = pl.G;
  if (size >= 2)
    for (int a = 0; a < G; a++)
      for (int b = 0; b < G; b++)
        if (a != b && asym[pl.gpu_of[a] * n_topo + pl.gpu_of[b]]) return GPA_ERR_ASYMMETRIC;

  const int K = pl.K = size - n_req;
  for (int p = 0; p < n_avail; p++) (pl.is_req[p] ? pl.answer : pl.free_pos).push_back(p);
  pl.R = (int)pl.free_pos.size();
  if (K == 0) return GPA_OK;  // the required positions are the only candidate

  pl.literal = (mode_flags & GPA_MODE_FORCE_ENUM) || !grouped;
  if (!pl.literal) {
    if (K > kSubMaxExtra) return GPA_ERR_TOO_LARGE;
    SubsetCounts &q = pl.qc;
    memset(&q, 0, sizeof q);
    q.G = G;
    q.K = K;
    int free_of[kSubGroups] = {0};
    for (int p = 0; p < n_avail; p++) {
      if (pl.is_req[p])
        q.r[pl.gid[p]]++;
      else
        free_of[pl.gid[p]]++;
    }
    for (int g = 0; g < G; g++) q.f[g] = (uint8_t)(free_of[g] < K ? free_of[g] : K);
    for (int g = 0; g < G; g++)
      for (int h = g + 1; h < G; h++) q.w[g * kSubGroups + h] = wq[pl.gpu_of[g] * n_topo + pl.gpu_of[h]];
    pl.cnt.assign((size_t)(kSubGroups + 1) * kSubCntStride, 0);
    pl.cnt[(size_t)G * kSubCntStride] = 1;
    for (int g = G - 1; g >= 0; g--)
      for (int s = 0; s <= K; s++) pl.cnt[(size_t)g * kSubCntStride + s] = sub_cnt_entry(q, pl.cnt.data(), g, s);
    pl.total = pl.cnt[K];
    if (pl.total > GPA_SUBSET_MAX_CLASSES) return GPA_ERR_TOO_LARGE;
  } else {
    const int R = pl.R;
    if (R > kSubEnumMaxFree || K > kSubEnumMaxExtra) return GPA_ERR_TOO_LARGE;
    pl.total = binom[(size_t)R * kSubBinomCols + K];
    if (pl.total > GPA_SUBSET_MAX_CANDIDATES) return GPA_ERR_TOO_LARGE;
    SubsetEnum &q = pl.qe;
    memset(&q, 0, sizeof q);
    q.R = R;
    q.K = K;
    auto pw = [&](int pa, int pb) -> uint32_t {  // pa < pb: the earlier position's links (:314), 0 on one GPU (:303-305)
      return pl.gid[pa] == pl.gid[pb] ? 0u : (uint32_t)wq[pl.gpu_of[pl.gid[pa]] * n_topo + pl.gpu_of[pl.gid[pb]]];
    };
    for (int g = 0; g < G; g++)
      for (int h = 0; h < G; h++)
        if (g != h) q.w[g * kSubGroups + h] = wq[pl.gpu_of[g] * n_topo + pl.gpu_of[h]];
    for (int i = 0; i < R; i++) {
      const int p = pl.free_pos[i];
      q.grp[i] = pl.gid[p];
      uint32_t s = 0;
      for (int32_t r : pl.answer) s += r < p ? pw(r, p) : pw(p, r);
      q.wreq[i] = s;
    }
    for (size_t a = 0; a < pl.answer.size(); a++)
      for (size_t b = a + 1; b < pl.answer.size(); b++) q.base += pw(pl.answer[a], pl.answer[b]);
  }
  if (pl.total == 0) return GPA_OK;  // cannot happen: n_avail >= size
  pl.done = false;
  return GPA_OK;
}

// Winning key of a search -> pl.answer (all `size` positions, ascending).  false = the rank does not replay to the score.
inline bool subset_decode(SubsetPlan &pl, const uint64_t *binom, uint64_t key) {
  uint64_t rank = key_item(key);
  if (rank >= pl.total) return false;
  if (!pl.literal) {
    uint8_t x[kSubGroups] = {0};
    const uint32_t score = sub_unrank_score(pl.qc, pl.cnt.data(), pl.qc.w, rank, x);
    if ((int)score != key_score(key)) return false;
    // the first member of the winning class: the first x[g] free replicas of every GPU next to the required ones
    int taken[kSubGroups] = {0};
    for (int32_t p : pl.free_pos)
      if (taken[pl.gid[p]] < x[pl.gid[p]]) {
        taken[pl.gid[p]]++;
        pl.answer.push_back(p);
      }
  } else {
    int v = 0;
    for (int i = 0; i < pl.K; i++) {  // lexicographic unrank: the i-th pick is the first v with rank < C(R-1-v, K-1-i)
      for (;; v++) {
        const uint64_t with_v = binom[(size_t)(pl.R - 1 - v) * kSubBinomCols + (pl.K - 1 - i)];
        if (rank < with_v) break;
        rank -= with_v;
      }
      pl.answer.push_back(pl.free_pos[v++]);
    }
  }
  std::sort(pl.answer.begin(), pl.answer.end());
  return (int)pl.answer.size() == pl.size;
}

}  // namespace gpa
