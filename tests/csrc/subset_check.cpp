// CPU harness for the replica-set search: the product's own plan / rank / score / decode code
// (csrc/subset_plan.h, csrc/subset_core.cuh) with a sequential loop standing where the CUDA kernels are.
// Test infrastructure; built by tests/test_subset_core.py with g++.
#include <stdint.h>

#include <algorithm>
#include <vector>

#include "../../k8s-gpu-sharing-plugin_b200/csrc/subset_plan.h"

namespace {

template <int K>
uint64_t walk_literal(const gpa::SubsetPlan &pl, const uint64_t *binom, unsigned chunk) {
  using namespace gpa;
  uint64_t best = 0;
  // the kernel's loop: a work item per (K-1)-prefix, taken here in a scrambled order (`chunk` picks the stride) because
  // the kernel's threads take them in no particular order either
  const uint64_t items = sub_prefix_count(binom, pl.qe.R, K);
  const uint64_t step = chunk | 1;
  for (uint64_t s0 = 0; s0 < step && s0 < items; s0++)
    for (uint64_t item = s0; item < items; item += step)
      best = sub_walk_prefix<K>(binom, pl.qe.grp, pl.qe.wreq, pl.qe.w, pl.qe.R, pl.qe.base, item, best);
  return best;
}

}  // namespace

// W: n_topo x n_topo weights (already scaled), asym: n_topo x n_topo flags.  -> GPA_* code; *form = 0 none, 4 counts, 5 literal
extern "C" int subset_check(int n_topo, const uint16_t *wq, const uint8_t *asym, const int32_t *phys, int n_avail,
                            const int32_t *req_pos, int n_req, int size, int mode_flags, unsigned chunk, int32_t *out_pos, int *n_out,
                            int *form, unsigned long long *total, int *best_score) {
  using namespace gpa;
  static std::vector<uint64_t> binom;
  if (binom.empty()) {
    binom.resize((size_t)kSubBinomRows * kSubBinomCols);
    sub_fill_binom(binom.data());
  }
  *n_out = 0;
  *form = 0;
  *total = 0;
  *best_score = -1;
  SubsetPlan pl;
  int rc = subset_plan(n_topo, wq, asym, binom.data(), phys, n_avail, req_pos, n_req, size, mode_flags, pl);
  if (rc != GPA_OK) return rc;
  if (!pl.done) {
    uint64_t best = 0;
    if (!pl.literal) {
      for (uint64_t rank = 0; rank < pl.total; rank++) {
        const uint32_t s = sub_unrank_score(pl.qc, pl.cnt.data(), pl.qc.w, rank, nullptr);
        const uint64_t key = pack_key((int)s, rank);
        best = key > best ? key : best;
      }
    } else {
      if (chunk < 1) chunk = 1;
      switch (pl.K) {
        case 1: best = walk_literal<1>(pl, binom.data(), chunk); break;
        case 2: best = walk_literal<2>(pl, binom.data(), chunk); break;
        case 3: best = walk_literal<3>(pl, binom.data(), chunk); break;
        case 4: best = walk_literal<4>(pl, binom.data(), chunk); break;
        case 5: best = walk_literal<5>(pl, binom.data(), chunk); break;
        case 6: best = walk_literal<6>(pl, binom.data(), chunk); break;
        case 7: best = walk_literal<7>(pl, binom.data(), chunk); break;
        case 8: best = walk_literal<8>(pl, binom.data(), chunk); break;
        default: return -100;
      }
    }
    if (!best || !subset_decode(pl, binom.data(), best)) return -101;
    *form = pl.literal ? 5 : 4;
    *total = pl.total;
    *best_score = key_score(best);
  }
  for (size_t i = 0; i < pl.answer.size(); i++) out_pos[i] = pl.answer[i];
  *n_out = (int)pl.answer.size();
  return GPA_OK;
}

// The batch kernel's search for ONE count-form request, lanes emulated one after the other: free / must counts per GPU
// label (G = 8 or 16 levels), label-space weights w[a * 16 + b] (a < b).  Guards and statuses are the kernel's own
// (warp-parallel) code and are not restated here: the caller passes searchable requests only (0 < K <= 15 or K == 0 with
// size == sum(must)).  -> take[16], *score
template <int G>
static void batch_search(const uint16_t *w, const uint8_t *free_in, const uint8_t *must, int size, uint8_t *take, int *score) {
  using namespace gpa;
  constexpr int D = 4;
  int n_req = 0;
  for (int g = 0; g < 16; g++) n_req += must[g];
  const int K = size - n_req;
  uint8_t f[16], r[16];
  uint16_t cap[17];
  for (int g = 0; g < 16; g++) {
    f[g] = (uint8_t)(free_in[g] < K ? free_in[g] : K);
    r[g] = must[g];
  }
  cap[16] = 0;
  for (int g = 15; g >= 0; g--) cap[g] = (uint16_t)(cap[g + 1] + f[g]);
  uint32_t exact[(D + 1) * 16] = {0};
  exact[D * 16] = 1;
  for (int p = D - 1; p >= 0; p--)
    for (int s = 0; s <= K; s++) exact[p * 16 + s] = sub_exact_entry(f, exact, p, s);
  const int s_lo = K > (int)cap[D] ? K - (int)cap[D] : 0;
  uint32_t n_items = 0;
  for (int sp = s_lo; sp <= K; sp++) n_items += exact[sp];
  unsigned long long best = 0, best_path = 0;
  int best_score = -1;
  for (int lane = 0; lane < 32; lane++) {
    int bs = -1;
    unsigned long long bp = 0;
    uint32_t blex = 0;
    for (uint32_t item = (uint32_t)lane; item < n_items; item += 32) sub_batch_item<G, D>(f, r, cap, w, exact, K, s_lo, item, bs, bp, blex);
    const unsigned long long mine = bs >= 0 ? ((unsigned long long)(unsigned)(bs + 1) << 32) | blex : 0ull;
    if (mine > best) {  // the warp's shuffle max
      best = mine;
      best_path = bp;
      best_score = bs;
    }
  }
  for (int g = 0; g < 16; g++) take[g] = g < G ? (uint8_t)((best_path >> (4 * g)) & 0xF) : 0;
  *score = best_score;
}

// The two-halves walk of the batch kernel on nodes of up to 8 GPUs (subset_core.cuh "count form, two halves"): the
// per-node blob, the per-request lists (built here one composition after the other; the kernel compacts them with
// ballots), then the product's own lane routine for lanes 0..31 and the warp's two-step maximum.
static void batch_pair_search(const uint16_t *w, const uint8_t *free_in, const uint8_t *must, int size, uint8_t *take, int *score) {
  using namespace gpa;
  static std::vector<uint32_t> blob;
  static std::vector<uint16_t> blob_w;
  if (blob.empty() || !std::equal(blob_w.begin(), blob_w.end(), w)) {
    blob.assign(kSubPairBlobBytes / 4, 0u);
    blob_w.assign(w, w + kSubGroups * kSubGroups);
    sub_pair_fill(w, blob.data());
  }
  int n_req = 0;
  for (int g = 0; g < 8; g++) n_req += must[g];
  const int K = size - n_req;
  uint32_t fpack[2] = {0, 0}, lin[8], r0 = 0;
  for (int g = 0; g < 8; g++) {
    const uint32_t fc = (uint32_t)(free_in[g] < K ? free_in[g] : K);
    fpack[g >> 2] |= fc << (8 * (3 - (g & 3)));
    lin[g] = 0;
    for (int h = 0; h < 8; h++)
      if (h != g) lin[g] += (uint32_t)must[h] * w[(g < h ? g : h) * kSubGroups + (g < h ? h : g)];
    for (int h = g + 1; h < 8; h++) r0 += (uint32_t)must[g] * must[h] * w[g * kSubGroups + h];
  }
  if (K >= 1 && K <= kSubFlatMaxK) {  // the kernel's short cut: one pass over the class in the node's table of all vectors
    static std::vector<uint32_t> flat;
    static std::vector<uint16_t> flat_w;
    if (flat.empty() || !std::equal(flat_w.begin(), flat_w.end(), w)) {
      flat.assign(kSubFlatBytes / 4, 0u);
      flat_w.assign(w, w + kSubGroups * kSubGroups);
      sub_flat_fill(w, flat.data());
    }
    unsigned long long best = 0;
    for (int lane = 0; lane < 32; lane++) {
      const unsigned long long k = sub_flat_lane(flat.data(), K, fpack[0], fpack[1], r0, lin, n_req > 0, lane);
      best = k > best ? k : best;
    }
    for (int g = 0; g < 16; g++) take[g] = 0;
    *score = -1;
    if (best) {
      const int e = (int)~(uint32_t)best;
      for (int g = 0; g < 4; g++) {
        take[g] = (uint8_t)(flat[2 * e] >> (8 * (3 - g)));
        take[4 + g] = (uint8_t)(flat[2 * e + 1] >> (8 * (3 - g)));
      }
      *score = (int)(best >> 32) - 1;
    }
    return;
  }
  uint16_t plist[kSubHalfEntries], slist[kSubHalfEntries], poff[kSubHalfMaxK + 2], soff[kSubHalfMaxK + 2];
  int np = 0, ns = 0;
  for (int t = 0; t <= K; t++) {
    poff[t] = (uint16_t)np;
    soff[t] = (uint16_t)ns;
    for (int e = sub_half_start(t); e < sub_half_start(t + 1); e++) {
      if (sub_half_fits(blob[e], fpack[0])) plist[np++] = (uint16_t)e;
      if (sub_half_fits(blob[e], fpack[1])) slist[ns++] = (uint16_t)e;
    }
  }
  poff[K + 1] = (uint16_t)np;
  soff[K + 1] = (uint16_t)ns;
  uint32_t split_tab[32];
  for (int i = 0; i < 32; i++) split_tab[i] = sub_pair_split_entry(i);
  const bool ident = ns == sub_half_start(K + 1);
  uint32_t lin_max = 0;
  for (int g = 4; g < 8; g++) lin_max = lin[g] > lin_max ? lin[g] : lin_max;
  const int mode = sub_pair_mode(K, blob[kSubPairMetaOff / 4], lin_max, n_req);
  SubPairBest lanes[32];
  int win = -1;
  for (int lane = 0; lane < 32; lane++) {
    lanes[lane] = SubPairBest{0ull, 0};
    sub_pair_lane(blob.data(), plist, poff, slist, soff, K, r0, lin, split_tab, ident, mode, lane, lanes[lane]);
    if (lanes[lane].key > (win < 0 ? 0ull : lanes[win].key)) win = lane;  // the warp's shuffle max; equal keys hold the same prefix
  }
  for (int g = 0; g < 16; g++) take[g] = 0;
  *score = -1;
  if (win >= 0) {
    const uint32_t xp = sub_pair_prefix(lanes[win]);
    const int t = (int)((xp >> 24) + ((xp >> 16) & 0xFF) + ((xp >> 8) & 0xFF) + (xp & 0xFF));
    uint32_t first = 0xFFFFFFFFu;
    for (int lane = 0; lane < 32; lane++) {
      const uint32_t f = sub_pair_find_suffix(blob.data(), lanes[win].pe, slist, soff[K - t], soff[K - t + 1], r0, lin, sub_pair_score(lanes[win]), lane);
      first = f < first ? f : first;
    }
    const uint32_t xs = blob[kSubPairCompOff / 4 + slist[first]];
    for (int g = 0; g < 4; g++) {
      take[g] = (uint8_t)(xp >> (8 * (3 - g)));
      take[4 + g] = (uint8_t)(xs >> (8 * (3 - g)));
    }
    *score = sub_pair_score(lanes[win]);
  }
}

// form: 0 = what the kernel would pick, 1 = the depth-first walk, 2 = the two-halves walk (nt <= 8, K <= 8 only)
extern "C" void subset_batch_check_form(int nt, const uint16_t *w, const uint8_t *free_in, const uint8_t *must, int size, int form,
                                        uint8_t *take, int *score) {
  int n_req = 0;
  for (int g = 0; g < 16; g++) n_req += must[g];
  const bool pair_ok = nt <= 8 && size - n_req <= gpa::kSubHalfMaxK;
  if (form == 2 || (form == 0 && pair_ok))
    batch_pair_search(w, free_in, must, size, take, score);
  else if (nt <= 8)
    batch_search<8>(w, free_in, must, size, take, score);
  else
    batch_search<16>(w, free_in, must, size, take, score);
}

extern "C" void subset_batch_check(int nt, const uint16_t *w, const uint8_t *free_in, const uint8_t *must, int size, uint8_t *take,
                                   int *score) {
  if (nt <= 8)
    batch_search<8>(w, free_in, must, size, take, score);
  else
    batch_search<16>(w, free_in, must, size, take, score);
}
