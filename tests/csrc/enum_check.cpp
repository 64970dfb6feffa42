// Host-side unit harness for k8s-gpu-sharing-plugin_b200/csrc/gpa_enum.cuh.
//
// TEST CODE, not a product path: it runs the enumeration core that the CUDA kernels are
// built from as a plain sequential loop over work items, so that the ordering / tie-break /
// padding / aliasing logic can be checked against the oracle on a machine without a GPU.
// The product library never contains this loop (it fails loudly without CUDA).
#include <cstdint>
#include <vector>

#include "../../k8s-gpu-sharing-plugin_b200/csrc/gpa_enum.cuh"
#include "../../k8s-gpu-sharing-plugin_b200/csrc/part_table.cuh"
#include "../../k8s-gpu-sharing-plugin_b200/csrc/ans_table.cuh"

extern "C" int enum_check_search(int nt, const int32_t *W, const int32_t *avail, int n_avail, uint32_t rmask, int size,
                                 int intended, unsigned long long want_items, int32_t *out, int *n_out,
                                 unsigned long long *n_items, int *cut_depth) {
  using namespace gpa;
  *n_out = 0;
  if (size < 2 || n_avail < size || n_avail > kMaxDevices) return -1;
  BinomTable B;
  fill_binom(B);
  int n = n_avail;
  std::vector<uint16_t> S((size_t)1 << n, 0);
  for (uint32_t mask = 1; mask < (1u << n); mask++) {
    int s = 0;
    for (int a = 0; a < n; a++)
      if (mask >> a & 1)
        for (int b = a + 1; b < n; b++)
          if (mask >> b & 1) s += W[avail[a] * nt + avail[b]];
    S[mask] = (uint16_t)s;
  }
  auto Sf = [&](uint32_t mask) { return S[mask]; };
  Geom g = make_geom(n, size, rmask);
  Cut cut = choose_cut(g, B, want_items);
  if (n_items) *n_items = cut.n_items;
  if (cut_depth) *cut_depth = cut.d;
  uint64_t key = 0;
  for (uint64_t item = 0; item < cut.n_items; item++) {
    int sc = search_item(g, cut, B, Sf, item);
    if (sc >= 0) {
      uint64_t kk = pack_key(sc, item);
      if (kk > key) key = kk;
    }
  }
  int8_t pos[32];
  int cnt = finalize(g, cut, B, Sf, key, intended != 0, pos);
  if (cnt < 0) return -2;
  for (int i = 0; i < cnt; i++) out[i] = pos[i] < 0 ? -1 : avail[pos[i]];
  *n_out = cnt;
  return 0;
}

// exposes the stepper alone: all sets of one level, as masks (anchor included)
extern "C" int enum_check_level_sets(int n, int k, int tok_present, uint32_t rem_reals, uint32_t *out, int cap) {
  using namespace gpa;
  Geom g = make_geom(n, k, 0);
  uint32_t rem = rem_reals | (tok_present ? g.tok_bit : 0u);
  uint32_t anchor;
  Stepper st = make_stepper(g, rem, anchor);
  uint32_t chosen;
  int cnt = 0;
  if (!st.first(chosen)) return 0;
  do {
    if (cnt < cap) out[cnt] = anchor | chosen;
    cnt++;
  } while (st.next(chosen));
  return cnt;
}

// Sparse variant: devices are the set bits of real_mask (labels 0..nt-1, ascending label order =
// `available` order), the pads sit at position nt, and set scores come from ONE table over all
// 2^nt label masks — the layout of the canonical-order fast path of the batch kernel.
extern "C" int enum_check_search_sparse(int nt, const int32_t *W, uint32_t real_mask, uint32_t rmask, int size, int intended,
                                        unsigned long long want_items, int32_t *out, int *n_out) {
  using namespace gpa;
  *n_out = 0;
  BinomTable B;
  fill_binom(B);
  std::vector<uint16_t> S((size_t)1 << nt, 0);
  for (uint32_t mask = 1; mask < (1u << nt); mask++) {
    int s = 0;
    for (int a = 0; a < nt; a++)
      if (mask >> a & 1)
        for (int b = a + 1; b < nt; b++)
          if (mask >> b & 1) s += W[a * nt + b];
    S[mask] = (uint16_t)s;
  }
  auto Sf = [&](uint32_t mask) { return S[mask]; };
  Geom g = make_geom_sparse(real_mask, size, rmask, nt);
  if (size < 2 || g.n < size) return -1;
  Cut cut = choose_cut(g, B, want_items);
  uint64_t key = 0;
  for (uint64_t item = 0; item < cut.n_items; item++) {
    int sc = search_item(g, cut, B, Sf, item);
    if (sc >= 0) {
      uint64_t kk = pack_key(sc, item);
      if (kk > key) key = kk;
    }
  }
  int8_t pos[32];
  int cnt = finalize(g, cut, B, Sf, key, intended != 0, pos);
  if (cnt < 0) return -2;
  for (int i = 0; i < cnt; i++) out[i] = pos[i];
  *n_out = cnt;
  return 0;
}

// Tracked whole-tree walk (the lane-per-request path of the fast batch kernel), sparse layout.
extern "C" int enum_check_tracked_sparse(int nt, const int32_t *W, uint32_t real_mask, uint32_t rmask, int size, int intended,
                                         int32_t *out, int *n_out) {
  using namespace gpa;
  *n_out = 0;
  std::vector<uint16_t> S((size_t)1 << nt, 0);
  for (uint32_t mask = 1; mask < (1u << nt); mask++) {
    int s = 0;
    for (int a = 0; a < nt; a++)
      if (mask >> a & 1)
        for (int b = a + 1; b < nt; b++)
          if (mask >> b & 1) s += W[a * nt + b];
    S[mask] = (uint16_t)s;
  }
  auto Sf = [&](uint32_t mask) { return S[mask]; };
  Geom g = make_geom_sparse(real_mask, size, rmask, nt);
  if (size < 2 || g.n < size || g.m > kMaxSuffix) return -1;
  Track tr;
  tr.best = -1;
  uint32_t path[kMaxSuffix];
  const uint32_t all = g.real_mask | g.tok_bit;
  switch (g.m) {
    case 1: walk_track<1, 1>(g, Sf, all, 0, false, path, tr); break;
    case 2: walk_track<2, 2>(g, Sf, all, 0, false, path, tr); break;
    case 3: walk_track<3, 3>(g, Sf, all, 0, false, path, tr); break;
    default: walk_track<4, 4>(g, Sf, all, 0, false, path, tr); break;
  }
  if (tr.best < 0) return 0;
  uint32_t best_set = 0;
  if (choose_from_partition(g, Sf, tr.sets, intended != 0, best_set) > 0) {
    int o = 0;
    for (int b = 0; b < nt; b++)
      if (best_set >> b & 1) out[o++] = b;
    if (best_set & g.tok_bit)
      for (int i = 0; i < g.P; i++) out[o++] = -1;
    *n_out = o;
  }
  return 0;
}

// Subset dynamic programme (gpa_enum.cuh dp_*), run layer by layer on the CPU: dense positions 0..n-1.
extern "C" int enum_check_dp(int nt, const int32_t *W, const int32_t *avail, int n_avail, uint32_t rmask, int size, int intended,
                             int32_t *out, int *n_out) {
  using namespace gpa;
  *n_out = 0;
  if (size < 2 || n_avail < size || n_avail > kMaxDevices) return -1;
  const int n = n_avail;
  Geom g = make_geom(n, size, rmask);
  std::vector<uint16_t> T((size_t)1 << n, 0), Tok((size_t)1 << (n > 0 ? n - 1 : 0), (uint16_t)kDpInvalid);
  const uint32_t rbit = rmask ? (rmask & (0u - rmask)) : 0u;
  auto Tany = [&](uint32_t m) -> uint32_t { return T[m]; };
  auto TokF = [&](uint32_t m) -> uint32_t { return Tok[dp_ok_index(m, rbit)]; };
  for (uint32_t mask = 1; mask < (1u << n); mask++) {  // layer 1 = set scores
    if (dp_layer_of(g, mask) != 1) continue;
    int s = 0;
    for (int a = 0; a < n; a++)
      if (mask >> a & 1)
        for (int b = a + 1; b < n; b++)
          if (mask >> b & 1) s += W[avail[a] * nt + avail[b]];
    T[mask] = (uint16_t)s;
  }
  for (int j = 2; j <= g.m; j++)
    for (uint32_t mask = 1; mask < (1u << n); mask++) {
      if (dp_layer_of(g, mask) != j) continue;
      // both tables of layer j are computed from layer j-1 before either is stored
      uint32_t va = dp_value(g, Tany, TokF, mask, false);
      uint32_t vo = (rmask && (mask & rmask) == rmask) ? dp_value(g, Tany, TokF, mask, true) : kDpInvalid;
      T[mask] = (uint16_t)va;
      if (rmask && (mask & rmask) == rmask) Tok[dp_ok_index(mask, rbit)] = (uint16_t)vo;
    }
  uint32_t sets[kMaxSets];
  if (!dp_reconstruct(g, Tany, TokF, g.real_mask, rmask != 0, 0, sets)) return 0;
  uint32_t best_set = 0;
  if (choose_from_partition(g, Tany, sets, intended != 0, best_set) > 0) {
    int o = 0;
    for (int b = 0; b < n; b++)
      if (best_set >> b & 1) out[o++] = avail[b];
    if (best_set & g.tok_bit)
      for (int i = 0; i < g.P; i++) out[o++] = -1;
    *n_out = o;
  }
  return 0;
}

// Partition tables of a small node (part_table.cuh, the fast batch kernel's path on up to 8 GPUs): words, index and keys
// are built as the library builds them, then ONE request is answered by part_table_search.  `shift` moves the key
// table to every 16-byte phase so that the aligned-vector body of the key scan and its scalar head / tail are all hit.
struct EnumTables {
  int nt = 0;
  std::vector<uint16_t> S;
  std::vector<uint32_t> words, index, keys;
};

extern "C" void *enum_check_tables_new(int nt, const int32_t *W) {
  using namespace gpa;
  if (nt < 1 || nt > kPartTableMaxTopo) return nullptr;
  EnumTables *t = new EnumTables();
  t->nt = nt;
  t->S.assign((size_t)1 << nt, 0);
  for (uint32_t mask = 1; mask < (1u << nt); mask++) {
    int s = 0;
    for (int a = 0; a < nt; a++)
      if (mask >> a & 1)
        for (int b = a + 1; b < nt; b++)
          if (mask >> b & 1) s += W[a * nt + b];  // the earlier label's view (:314)
    t->S[mask] = (uint16_t)s;
  }
  if (!build_partition_words(nt, t->words, t->index)) {
    delete t;
    return nullptr;
  }
  build_partition_keys(nt, t->words, t->index, t->S.data(), t->keys);
  return t;
}

extern "C" void enum_check_tables_free(void *p) { delete static_cast<EnumTables *>(p); }

extern "C" int enum_check_tables_words(void *p) { return (int)static_cast<EnumTables *>(p)->words.size(); }

extern "C" int enum_check_table_search(void *p, uint32_t real_mask, uint32_t rmask, int size, int intended, int shift, int32_t *out,
                                       int *n_out) {
  using namespace gpa;
  EnumTables *t = static_cast<EnumTables *>(p);
  *n_out = 0;
  Geom g = make_geom_sparse(real_mask, size, rmask, t->nt);
  if (size < 2 || g.n < size || g.m > kMaxSuffix) return -1;
  // a copy of the keys at a chosen 16-byte phase
  std::vector<uint32_t> buf(t->keys.size() + 8);
  uint32_t *base = buf.data();
  while ((reinterpret_cast<uintptr_t>(base) & 15u) != 0) base++;
  base += shift & 3;
  for (size_t i = 0; i < t->keys.size(); i++) base[i] = t->keys[i];
  const uint16_t *S = t->S.data();
  auto Sf = [S](uint32_t mask) { return S[mask]; };
  uint32_t best_set = 0;
  if (part_table_search(g, t->words.data(), t->index.data(), base, Sf, intended != 0, best_set) > 0) {
    int o = 0;
    for (int b = 0; b < t->nt; b++)
      if (best_set >> b & 1) out[o++] = b;
    if (best_set & g.tok_bit)
      for (int i = 0; i < g.P; i++) out[o++] = -1;
    *n_out = o;
  }
  return 0;
}

// ans_table.cuh: every entry of the answer table as the build kernel computes it (one ans_compute_entry per index),
// and the index arithmetic the batch kernels and gpa_besteffort() look an answer up with.
extern "C" int enum_check_answer_table(void *p, int intended, uint16_t *out, int cap) {
  using namespace gpa;
  EnumTables *t = static_cast<EnumTables *>(p);
  const uint32_t n = ans_entries(t->nt);
  if ((uint32_t)cap < n) return -1;
  const uint16_t *S = t->S.data();
  auto Sf = [S](uint32_t mask) { return S[mask]; };
  for (uint32_t e = 0; e < n; e++) {
    uint32_t mask, rl;
    ans_decode(e / (uint32_t)kAnsSizes, mask, rl);
    const int size = (int)(e % (uint32_t)kAnsSizes) + 1;
    out[e] = ans_compute_entry(t->nt, mask, rl, size, t->words.data(), t->index.data(), t->keys.data(), Sf, intended != 0);
  }
  return (int)n;
}

extern "C" uint32_t enum_check_answer_index(uint32_t mask, uint32_t rl, int size) {
  using namespace gpa;
  uint16_t tern[256];
  for (uint32_t x = 0; x < 256; x++) tern[x] = (uint16_t)ans_tern(x);
  return ans_index(tern, mask, rl, size);
}
