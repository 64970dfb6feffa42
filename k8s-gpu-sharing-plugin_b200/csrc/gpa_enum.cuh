// Partition enumeration core of the best-effort search, written once for device and host.
//
// What it reproduces: the visiting order and the accept rules of
//   vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go
// ("alloc/besteffort_policy.go" below): iterateGPUPartitions (:209-290) over the padded
// device list, the padding filter (:273-276), the `required` filter (:53, :112-119), the
// strict-`>` / first-valid accept rule (:57) and the final set choice (:65-88).
//
// How it differs from the Go code (this is a re-design, not a translation):
//   * devices are POSITIONS 0..n-1 in the caller's `available` order, sets and remainders
//     are 32-bit masks, the P nil pads are ONE token bit at position n (they always move
//     together: :273-276 admits only 0 or P of them per set, :278-283 drops them together);
//   * a set's score is one lookup in a table S[mask] (all 2^n masks, built once per
//     request in shared memory) instead of C(k,2) map lookups;
//   * the tree is cut at depth d into independent, ORDER-NUMBERED work items: item index
//     = mixed-radix number of the lexicographic ranks of the first d sets, each rank taken
//     among the (k-1)-combinations of the *literal* (nil-expanded) remainder, exactly the
//     order iterateGPUSets (:155-204) emits them.  A larger item index is a later visit,
//     so "max score, then smallest item index, then first leaf inside the item" is the
//     reference's first-wins rule;
//   * the levels below the cut are walked by a register-only combination stepper
//     (next_set) instead of heap slices.
//
// The file compiles under nvcc (device + host) and under g++ (host only, for the unit
// harness tests/csrc/enum_check.cpp that checks this logic against the oracle on the CPU).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GPA_HD __host__ __device__ __forceinline__
#else
#define GPA_HD inline
#endif

namespace gpa {

constexpr int kMaxDevices = 16;  // available devices per request (positions)
constexpr int kMaxSets = 16;     // sets per partition (m <= n for k >= 1)
constexpr int kMaxCutDepth = 8;  // levels resolved from the item index
constexpr int kMaxSuffix = 4;    // levels walked per item by the templated DFS

GPA_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
GPA_HD uint32_t top_bit(uint32_t x) {  // x != 0
#if defined(__CUDA_ARCH__)
  return 0x80000000u >> __clz(x);
#else
  return 0x80000000u >> __builtin_clz(x);
#endif
}
GPA_HD uint32_t low_bit(uint32_t x) { return x & (0u - x); }
GPA_HD int bit_pos(uint32_t onehot) {
#if defined(__CUDA_ARCH__)
  return __ffs(onehot) - 1;
#else
  return __builtin_ctz(onehot);
#endif
}

// Shape of one request after the guards of alloc/besteffort_policy.go:35-45.
struct Geom {
  int n;               // available devices
  int k;               // allocation size (>= 2 here; size 1 is answered without a search)
  int P;               // nil pads: N - n with N = ceil(n/k)*k   (:232-233)
  int m;               // sets per partition, N / k
  int t;               // picks per set besides the anchor, k - 1
  uint32_t real_mask;  // (1 << n) - 1
  uint32_t tok_bit;    // P ? 1 << tok_pos : 0 — stands for all P nil pads
  int tok_pos;         // position of the pads: n for dense positions, above every device for sparse ones
  uint32_t rmask;      // positions of the `required` devices
};

GPA_HD Geom make_geom(int n, int k, uint32_t rmask) {
  Geom g;
  g.n = n;
  g.k = k;
  int N = ((n + k - 1) / k) * k;
  g.P = N - n;
  g.m = N / k;
  g.t = k - 1;
  g.real_mask = (n >= 32) ? 0xFFFFFFFFu : ((1u << n) - 1u);
  g.tok_bit = g.P ? (1u << n) : 0u;
  g.tok_pos = n;
  g.rmask = rmask;
  return g;
}

// Same request over SPARSE positions: the devices sit at the set bits of real_mask (their
// ascending bit order is the caller's `available` order) and the pads at tok_pos, above all
// of them.  Everything downstream only ever asks for "lowest bit", "bits above" and "count".
GPA_HD Geom make_geom_sparse(uint32_t real_mask, int k, uint32_t rmask, int tok_pos) {
  Geom g;
  g.n = popc32(real_mask);
  g.k = k;
  int N = ((g.n + k - 1) / k) * k;
  g.P = N - g.n;
  g.m = N / k;
  g.t = k - 1;
  g.real_mask = real_mask;
  g.tok_bit = g.P ? (1u << tok_pos) : 0u;
  g.tok_pos = tok_pos;
  g.rmask = rmask;
  return g;
}

// ---------------------------------------------------------------------------------------
// Combination stepper: the sets {anchor} + chosen that one iterate() level tries
// (:266-276), in the order iterateGPUSets emits them, skipping the ones the padding filter
// rejects without visiting them.
//
//   cand : real candidates (remainder minus its lowest element, the anchor)
//   tok  : g.tok_bit if the nil pads are still in the remainder, else 0
// A set is complete with t real picks, or with t-P real picks plus the token.  Because the
// token is the highest position, "x then every extension of x by P reals, then x + token"
// is lexicographic order.
struct Stepper {
  uint32_t cand;
  uint32_t tok;
  int t;
  int tp;  // t - P: number of real picks that must precede the token

  GPA_HD uint32_t above(uint32_t chosen) const {
    if (chosen == 0) return cand;
    uint32_t hp = top_bit(chosen);
    return cand & ~((hp << 1) - 1u);
  }
  // after making pick number j+1 (a real) with `left` real candidates above it: can the set
  // still be completed?
  GPA_HD bool feasible_after(int j, int left) const {
    int j1 = j + 1;
    if (left >= t - j1) return true;
    return tok != 0 && j1 <= tp && left >= tp - j1;
  }
  // complete `chosen` (all real so far) with the lowest admissible picks
  GPA_HD bool extend(uint32_t &chosen) const {
    int j = popc32(chosen);
    for (;;) {
      if (j == t) return true;
      uint32_t A = above(chosen);
      int cnt = popc32(A);
      if (cnt > 0 && feasible_after(j, cnt - 1)) {
        chosen |= low_bit(A);
        j++;
        continue;
      }
      if (tok != 0 && j == tp) {
        chosen |= tok;
        return true;
      }
      return false;
    }
  }
  GPA_HD bool first(uint32_t &chosen) const {
    chosen = 0;
    return extend(chosen);
  }
  GPA_HD bool next(uint32_t &chosen) const {
    chosen &= ~tok;  // the token set is the last one of its real prefix
    for (;;) {
      if (chosen == 0) return false;
      uint32_t hp = top_bit(chosen);
      chosen ^= hp;
      int j = popc32(chosen);  // index of the pick being replaced
      uint32_t A = cand & ~((hp << 1) - 1u);
      int cnt = popc32(A);
      if (cnt > 0 && feasible_after(j, cnt - 1)) {
        chosen |= low_bit(A);
        if (extend(chosen)) return true;
        // unreachable when feasible_after said yes; fall through defensively
        chosen &= ~tok;
        continue;
      }
      if (tok != 0 && j == tp) {
        chosen |= tok;
        return true;
      }
    }
  }
};

GPA_HD Stepper make_stepper(const Geom &g, uint32_t rem, uint32_t &anchor) {
  uint32_t reals = rem & g.real_mask;
  anchor = low_bit(reals);
  Stepper s;
  s.cand = reals ^ anchor;
  s.tok = rem & g.tok_bit;
  s.t = g.t;
  s.tp = g.t - g.P;
  return s;
}

// `required` bookkeeping for one chosen set (:53, :112-119).
//   returns false when no partition below can pass the filter any more (the required
//   devices were split, or sit in the padded set) — pruning those changes nothing, the
//   callback would have returned at :54.
GPA_HD bool account_required(const Geom &g, uint32_t set, bool &ok) {
  uint32_t hit = set & g.rmask;
  bool all = hit == g.rmask;
  bool unpadded = (set & g.tok_bit) == 0;
  if (g.rmask != 0 && hit != 0 && !(all && unpadded)) return false;
  ok = ok || (all && unpadded);
  return true;
}

// ---------------------------------------------------------------------------------------
// Suffix walk: the last L levels of one work item.  STab is any callable mask -> score.
// best: highest score of an accepted leaf so far, -1 = none; strict `>` keeps the first.
// If rec != nullptr the walk is a REPLAY: it stops at the first accepted leaf whose score
// equals rec->target and leaves that leaf's sets in rec->sets[rec->base ...].
struct Replay {
  int target;
  int base;  // index of the first suffix set in sets[]
  bool found;
  uint32_t sets[kMaxSets];
};

template <int L, typename STab>
GPA_HD void walk(const Geom &g, const STab &S, uint32_t rem, int acc, bool ok, int &best, Replay *rec, int depth) {
  if constexpr (L == 1) {
    // base case :245-248 — the remainder is the last set, no filter applies
    bool okk = ok;
    account_required(g, rem, okk);
    // a padded last set never makes the partition valid, and a split `required` cannot
    // happen here without having been pruned above, so okk is the :53 test
    int sc = acc + (int)S(rem & g.real_mask);
    if (rec) {
      if (!rec->found && okk && sc == rec->target) {
        rec->sets[rec->base + depth] = rem;
        rec->found = true;
      }
      return;
    }
    if (okk && sc > best) best = sc;
  } else {
    uint32_t anchor;
    Stepper st = make_stepper(g, rem, anchor);
    uint32_t chosen;
    if (!st.first(chosen)) return;
    do {
      uint32_t set = anchor | chosen;
      bool okk = ok;
      if (!account_required(g, set, okk)) continue;
      if (rec) rec->sets[rec->base + depth] = set;
      walk<L - 1>(g, S, rem & ~set, acc + (int)S(set & g.real_mask), okk, best, rec, depth + 1);
      if (rec && rec->found) return;
    } while (st.next(chosen));
  }
}

template <typename STab>
GPA_HD void walk_dyn(int L, const Geom &g, const STab &S, uint32_t rem, int acc, bool ok, int &best, Replay *rec) {
  switch (L) {
    case 1: walk<1>(g, S, rem, acc, ok, best, rec, 0); break;
    case 2: walk<2>(g, S, rem, acc, ok, best, rec, 0); break;
    case 3: walk<3>(g, S, rem, acc, ok, best, rec, 0); break;
    case 4: walk<4>(g, S, rem, acc, ok, best, rec, 0); break;
    default: break;
  }
}

// Tracked walk, for callers that own a whole (shallow) tree: the sets of the first best leaf are kept
// in registers as the walk goes, so no replay is needed afterwards.  L0 = depth of the tree (all m sets),
// L = levels still to choose; every index into path[] / tr.sets[] is a compile-time constant.
struct Track {
  int best;  // score of the first best accepted leaf, -1 = none
  uint32_t sets[kMaxSuffix];
};

template <int L0, int L, typename STab>
GPA_HD void walk_track(const Geom &g, const STab &S, uint32_t rem, int acc, bool ok, uint32_t (&path)[kMaxSuffix], Track &tr) {
  if constexpr (L == 1) {
    bool okk = ok;
    account_required(g, rem, okk);
    int sc = acc + (int)S(rem & g.real_mask);
    if (okk && sc > tr.best) {  // strict `>`: the first best leaf in reference order stays (:57)
      tr.best = sc;
#if defined(__CUDACC__)
#pragma unroll
#endif
      for (int i = 0; i < L0 - 1; i++) tr.sets[i] = path[i];
      tr.sets[L0 - 1] = rem;
    }
  } else {
    uint32_t anchor;
    Stepper st = make_stepper(g, rem, anchor);
    uint32_t chosen;
    if (!st.first(chosen)) return;
    do {
      uint32_t set = anchor | chosen;
      bool okk = ok;
      if (!account_required(g, set, okk)) continue;
      path[L0 - L] = set;
      walk_track<L0, L - 1>(g, S, rem & ~set, acc + (int)S(set & g.real_mask), okk, path, tr);
    } while (st.next(chosen));
  }
}

// ---------------------------------------------------------------------------------------
// Work items: the first d levels, decoded from a mixed-radix index over LITERAL positions
// (the P nil pads as positions n..N-1), where every node at depth j has exactly
// C(N - j*k - 1, k-1) children — a regular tree, so the index is a closed form.  Children
// the padding filter rejects (:273-276) are decoded and dropped: they cost one idle item.
struct Cut {
  int d;                         // levels in the item index, 0 <= d <= m-1
  uint32_t radix[kMaxCutDepth];  // C(N - j*k - 1, k - 1)
  uint64_t n_items;              // product of the radices
};

// binom(a, b) for a < 32 — small exact table filled by the host, passed by pointer/constant
struct BinomTable {
  uint32_t c[32][17];
};

GPA_HD uint32_t binom_at(const BinomTable &B, int a, int b) {
  if (b < 0 || a < 0 || b > a) return 0;
  if (b > a - b) b = a - b;
  if (b > 16) return 0;  // never needed: k-1 <= 15 and the symmetric half is used
  return B.c[a][b];
}

struct ItemState {
  uint32_t rem;  // token form
  int acc;       // score of the sets fixed by the item
  bool ok;       // :53 already satisfied by one of them
  bool valid;    // false: rejected by the padding filter or by `required` pruning
};

// Decode item -> state; when sets_out != nullptr also stores the d token-form sets.
template <typename STab>
GPA_HD ItemState decode_item(const Geom &g, const Cut &cut, const BinomTable &B, const STab &S, uint64_t item,
                             uint32_t *sets_out) {
  ItemState st;
  // literal remainder: the devices plus P separate pad positions tok_pos .. tok_pos+P-1
  uint32_t lit = g.real_mask | (g.P ? (((1u << g.P) - 1u) << g.tok_pos) : 0u);
  const uint32_t nil_lit = lit & ~g.real_mask;
  st.acc = 0;
  st.ok = false;
  st.valid = true;
  // digits, most significant first.  64-bit division is emulated on the GPU (~100 instructions), so it
  // is only used for the rare searches with more than 2^32 items.
  uint32_t digit[kMaxCutDepth];
  if (cut.n_items <= 0xFFFFFFFFull) {
    uint32_t it = (uint32_t)item;
    for (int j = cut.d - 1; j >= 0; j--) {
      const uint32_t q = it / cut.radix[j];
      digit[j] = it - q * cut.radix[j];
      it = q;
    }
  } else {
    for (int j = cut.d - 1; j >= 0; j--) {
      digit[j] = (uint32_t)(item % cut.radix[j]);
      item /= cut.radix[j];
    }
  }
  for (int j = 0; j < cut.d; j++) {
    uint32_t anchor = low_bit(lit);
    uint32_t bits = lit ^ anchor;
    int left = popc32(bits);
    int need = g.t;
    uint32_t c = digit[j];
    uint32_t set_lit = anchor;
    while (need > 0) {
      uint32_t p = low_bit(bits);
      bits ^= p;
      left--;
      uint32_t with_p = binom_at(B, left, need - 1);  // combinations whose next pick is p
      if (c < with_p) {
        set_lit |= p;
        need--;
      } else {
        c -= with_p;
      }
    }
    int pads = popc32(set_lit & nil_lit);
    if (!(pads == 0 || pads == g.P)) {  // :273-276
      st.valid = false;
      break;
    }
    uint32_t set = (set_lit & g.real_mask) | (pads ? g.tok_bit : 0u);
    if (!account_required(g, set, st.ok)) {
      st.valid = false;
      break;
    }
    if (sets_out) sets_out[j] = set;
    st.acc += (int)S(set & g.real_mask);
    lit &= ~set_lit;
    if (pads) lit &= ~nil_lit;  // :278-283 — a nil in the set removes every nil
  }
  st.rem = (lit & g.real_mask) | ((lit & nil_lit) ? g.tok_bit : 0u);
  return st;
}

// 64-bit reduction key: larger is better.  0 = "no accepted leaf".
//   [63:40] score + 1        [39:0] ~item index  (earlier item wins a score tie)
constexpr uint64_t kItemBits = 40;
constexpr uint64_t kItemMask = (1ull << kItemBits) - 1ull;
GPA_HD uint64_t pack_key(int score, uint64_t item) { return ((uint64_t)(score + 1) << kItemBits) | (kItemMask - item); }
GPA_HD int key_score(uint64_t key) { return (int)(key >> kItemBits) - 1; }
GPA_HD uint64_t key_item(uint64_t key) { return kItemMask - (key & kItemMask); }

// Score of the best accepted leaf of one item, -1 if none.
template <typename STab>
GPA_HD int search_item(const Geom &g, const Cut &cut, const BinomTable &B, const STab &S, uint64_t item) {
  ItemState st = decode_item(g, cut, B, S, item, nullptr);
  if (!st.valid) return -1;
  int best = -1;
  walk_dyn(g.m - cut.d, g, S, st.rem, st.acc, st.ok, best, nullptr);
  return best;
}

// number of leading candidate slots that live in an array private to their prefix when Go
// grows `accum` by append (cap 0,1,2,4,8,16: a new array is made when len == cap); see
// oracle/ref_besteffort.c fast_fresh_slots and SURVEY.md §8a row A3 (defect D6).
GPA_HD int fresh_slots(int m) {
  int L = 0;
  if (m > 1) {
    L = 1;
    while (L * 2 < m) L *= 2;
  }
  return L + 1;
}

// What Allocate() does with the saved bestPartition (:65-88): the partition `sets` (m sets in visiting order)
// first suffers what Go's aliased append did to it (go_faithful), then the sets containing all `required`
// devices compete on their own score, first best wins.  returns 1 and the set, or 0 for the empty answer.
template <typename STab>
GPA_HD int choose_from_partition(const Geom &g, const STab &S, uint32_t *sets, bool intended, uint32_t &best_set) {
  if (!intended) {
    // D6: slots fresh..m-1 of the saved candidate are shared with every later candidate
    // under the same prefix and end up holding the LAST one: anchor + the highest picks.
    int f = fresh_slots(g.m);
    if (f < g.m) {
      uint32_t rem = g.real_mask | g.tok_bit;
      for (int s = 0; s < f; s++) rem &= ~sets[s];
      for (int s = f; s < g.m; s++) {
        uint32_t reals = rem & g.real_mask;
        uint32_t set = low_bit(reals);
        uint32_t cand = reals ^ set;
        int need = g.t;
        if (rem & g.tok_bit) {  // the last k-1 literal elements include all P nils
          set |= g.tok_bit;
          need -= g.P;
        }
        while (need > 0) {
          uint32_t hp = top_bit(cand);
          set |= hp;
          cand ^= hp;
          need--;
        }
        sets[s] = set;
        rem &= ~set;
      }
    }
  }
  // :65-85 — no padding test here: a padded set can win (defect D5)
  int best = -1;
  best_set = 0;
  for (int s = 0; s < g.m; s++) {
    uint32_t set = sets[s];
    if ((set & g.rmask) != g.rmask) continue;
    int sc = (int)S(set & g.real_mask);
    if (sc > best) {
      best = sc;
      best_set = set;
    }
  }
  return best < 0 ? 0 : 1;
}

// Epilogue (:65-88) for the winning item: replay it to recover the first best partition,
// apply what Go's slice aliasing does to the saved bestPartition (go_faithful), filter
// the sets by `required` and return the first best-scoring one.
//   returns 1 and the winning set (positions + token bit) in best_set, 0 for the empty
//   answer, -1 if the key cannot be replayed (cannot happen for a key from this request).
template <typename STab>
GPA_HD int finalize_set(const Geom &g, const Cut &cut, const BinomTable &B, const STab &S, uint64_t key, bool intended,
                        uint32_t &best_set) {
  best_set = 0;
  if (key == 0) return 0;  // bestPartition == nil -> empty answer (:72-74)
  Replay rec;
  rec.target = key_score(key);
  rec.base = cut.d;
  rec.found = false;
  ItemState st = decode_item(g, cut, B, S, key_item(key), rec.sets);
  int dummy = -1;
  walk_dyn(g.m - cut.d, g, S, st.rem, st.acc, st.ok, dummy, &rec);
  if (!st.valid || !rec.found) return -1;

  return choose_from_partition(g, S, rec.sets, intended, best_set);
}

//   out_pos[0..k): positions in ascending order, -1 for each nil pad.  returns k, 0, or -1.
template <typename STab>
GPA_HD int finalize(const Geom &g, const Cut &cut, const BinomTable &B, const STab &S, uint64_t key, bool intended,
                    int8_t *out_pos) {
  uint32_t best_set;
  int rc = finalize_set(g, cut, B, S, key, intended, best_set);
  if (rc <= 0) return rc;
  int o = 0;
  uint32_t reals = best_set & g.real_mask;
  while (reals) {
    uint32_t b = low_bit(reals);
    out_pos[o++] = (int8_t)bit_pos(b);
    reals ^= b;
  }
  if (best_set & g.tok_bit)
    for (int i = 0; i < g.P; i++) out_pos[o++] = -1;
  return o;
}

// ---------------------------------------------------------------------------------------
// Subset dynamic programme: the same answer without visiting every partition.
//
//   best[R] = max over the sets s that iterate() would try first on remainder R of  S[s] + best[R \ s]
//
// R is a mask of real devices; whether the pads are still in it follows from its size (|R| = j*k without pads,
// j*k - P with them), so one table indexed by the real mask holds every layer j = 1..m, and layer 1 IS the set
// score table S.  Filling layer j needs only layers 1 and j-1, so the layers are m-1 data-parallel passes.
// The reference's answer is the FIRST best partition in visiting order (strict `>` at :57); that one is rebuilt
// level by level afterwards: at each level the first set, in stepper order, whose value reaches best[R].
// `required` devices add a second table for "a set holding all of them, unpadded, is still to come" (:53, :112-119),
// defined on the masks that contain them all and indexed with the lowest required bit squeezed out.
constexpr uint32_t kDpInvalid = 0xFFFFu;  // no acceptable completion (scores stay below: 120 pairs x 546)

GPA_HD uint32_t dp_ok_index(uint32_t R, uint32_t rbit) {
  return ((R >> 1) & ~(rbit - 1u)) | (R & (rbit - 1u));
}
// layer of a real mask, 0 if its size belongs to none (sizes j*k, and j*k - P when there are pads)
GPA_HD int dp_layer_of(const Geom &g, uint32_t R) {
  int pc = popc32(R);
  if (pc == 0) return 0;
  if (pc % g.k == 0) return pc / g.k;
  if (g.P && (pc + g.P) % g.k == 0) return (pc + g.P) / g.k;
  return 0;
}

// base of the required-still-to-come recursion: R is the last set, acceptable iff it holds them all unpadded
GPA_HD uint32_t dp_ok_base(const Geom &g, uint32_t S_of_R, uint32_t R) {
  return (popc32(R) == g.k && (R & g.rmask) == g.rmask) ? S_of_R : kDpInvalid;
}

// value of ONE first set of remainder R with the tables of the layers below filled; -1 = not acceptable
template <typename TAny, typename TOk>
GPA_HD int dp_set_value(const Geom &g, const TAny &Tany, const TOk &Tok, uint32_t R, uint32_t set, bool want_ok, bool last_pair) {
  const uint32_t sr = set & g.real_mask;
  const uint32_t rest = R & ~sr;
  if (!want_ok) return (int)Tany(sr) + (int)Tany(rest);
  const uint32_t hit = sr & g.rmask;
  if (hit == 0) {
    const uint32_t t = last_pair ? dp_ok_base(g, Tany(rest), rest) : Tok(rest);
    if (t == kDpInvalid) return -1;
    return (int)Tany(sr) + (int)t;
  }
  if (hit == g.rmask && (set & g.tok_bit) == 0) return (int)Tany(sr) + (int)Tany(rest);
  return -1;  // the required devices are split, or sit in the padded set: never acceptable (:112-119)
}

// value of best[R] in the plain (want_ok = false) or the required-still-to-come (true) table; layers below filled
template <typename TAny, typename TOk>
GPA_HD uint32_t dp_value(const Geom &g, const TAny &Tany, const TOk &Tok, uint32_t R, bool want_ok) {
  const bool last_pair = dp_layer_of(g, R) == 2;  // what is left after one more set is a single, final set
  const bool has_tok = g.P != 0 && (popc32(R) % g.k) != 0;
  const uint32_t rem = R | (has_tok ? g.tok_bit : 0u);
  uint32_t anchor;
  Stepper st = make_stepper(g, rem, anchor);
  uint32_t chosen;
  int best = -1;
  if (!st.first(chosen)) return kDpInvalid;
  do {
    const int v = dp_set_value(g, Tany, Tok, R, anchor | chosen, want_ok, last_pair);
    best = v > best ? v : best;
  } while (st.next(chosen));
  return best < 0 ? kDpInvalid : (uint32_t)best;
}

// Rebuilds the first best partition from filled tables, starting below `level0` sets already fixed in sets[]
// (R = what is left, need_ok = no fixed set holds the required devices yet).  Returns false if none is acceptable.
template <typename TAny, typename TOk>
GPA_HD bool dp_reconstruct(const Geom &g, const TAny &Tany, const TOk &Tok, uint32_t R, bool need_ok, int level0, uint32_t *sets) {
  for (int level = level0; level < g.m; level++) {
    const int left = g.m - level;
    const bool has_tok = g.P != 0 && (popc32(R) % g.k) != 0;
    if (left == 1) {
      if (need_ok && !((R & g.rmask) == g.rmask && !has_tok)) return false;
      sets[level] = R | (has_tok ? g.tok_bit : 0u);
      return true;
    }
    const uint32_t target = need_ok ? Tok(R) : Tany(R);
    if (target == kDpInvalid) return false;
    const uint32_t rem = R | (has_tok ? g.tok_bit : 0u);
    uint32_t anchor;
    Stepper st = make_stepper(g, rem, anchor);
    uint32_t chosen;
    bool found = false;
    if (!st.first(chosen)) return false;
    do {
      const uint32_t set = anchor | chosen;
      const uint32_t sr = set & g.real_mask;
      const uint32_t rest = R & ~sr;
      bool next_need = need_ok;
      uint32_t v;
      if (!need_ok) {
        v = Tany(sr) + Tany(rest);
      } else {
        const uint32_t hit = sr & g.rmask;
        if (hit == 0) {
          const uint32_t t = left == 2 ? dp_ok_base(g, Tany(rest), rest) : Tok(rest);
          if (t == kDpInvalid) continue;
          v = Tany(sr) + t;
        } else if (hit == g.rmask && (set & g.tok_bit) == 0) {
          v = Tany(sr) + Tany(rest);
          next_need = false;
        } else {
          continue;
        }
      }
      if (v == target) {
        sets[level] = set;
        R = rest;
        need_ok = next_need;
        found = true;
        break;
      }
    } while (st.next(chosen));
    if (!found) return false;
  }
  return true;
}

// Number of candidate partitions the reference visits (closed form of SURVEY.md §8a A3):
// P = 0: n!/((k!)^m m!);  P > 0: C(n, k-P) * ((m-1)k)!/((k!)^(m-1) (m-1)!).
GPA_HD uint64_t partition_count(const Geom &g, const BinomTable &B) {
  uint64_t full = 1;
  int mm = g.P == 0 ? g.m : g.m - 1;
  for (int r = mm * g.k; r > 0; r -= g.k) full *= binom_at(B, r - 1, g.k - 1);
  if (g.P == 0) return full;
  return (uint64_t)binom_at(B, g.n, g.k - g.P) * full;
}

// Choice of the cut depth: deep enough that the suffix fits the templated walk and that
// there are about `want_items` items to spread over the machine — but never so deep that
// the LITERAL item count (which includes the children the padding filter drops) dwarfs the
// number of real partitions: with many pads (k close to n) almost every literal child is
// dropped, and the stepper inside one item skips them for free.
GPA_HD Cut choose_cut(const Geom &g, const BinomTable &B, uint64_t want_items) {
  Cut c;
  c.d = 0;
  c.n_items = 1;
  const int N = g.m * g.k;
  const uint64_t parts = partition_count(g, B);
  int min_d = g.m - kMaxSuffix;
  if (min_d < 0) min_d = 0;
  while (c.d < g.m - 1 && c.d < kMaxCutDepth) {
    uint32_t r = binom_at(B, N - c.d * g.k - 1, g.k - 1);
    if (c.d >= min_d && (c.n_items >= want_items || c.n_items * r > 8 * parts)) break;
    c.radix[c.d] = r;
    c.n_items *= r;
    c.d++;
  }
  for (int j = c.d; j < kMaxCutDepth; j++) c.radix[j] = 1;
  return c;
}

inline void fill_binom(BinomTable &B) {
  for (int a = 0; a < 32; a++)
    for (int b = 0; b <= 16; b++) {
      uint64_t v;
      if (b > a) v = 0;
      else if (b == 0) v = 1;
      else v = (uint64_t)B.c[a - 1][b - 1] + (b <= a - 1 ? (uint64_t)B.c[a - 1][b] : 0);
      B.c[a][b] = (uint32_t)v;  // C(31,15) = 300540195 fits
    }
}

}  // namespace gpa
