// Core of the replica-set search: which `size` of the available SHARED replicas are best connected.
//
// NO reference counterpart as a whole.  The reference answers a request over replicas with prioritizeDevices
// (cmd/replica.go:96-198) and never looks at links between replicas; BASELINE.json's north_star and SURVEY.md 8d
// ("config 3c") ask for the search the best-effort policy WOULD do if it were handed the replicas: every k-subset of the
// available virtual devices, scored against the link-weight matrix.  It is assembled from reference pieces only:
//   candidates   every `size`-combination of the available positions in lexicographic index order
//                (iterateGPUSets, vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:155-204)
//                that contains all `required` positions (gpuSetContainsAll, :102-110)
//   score        sum over pairs a < b of calculateGPUPairScore(dev(a), dev(b)) (:298-356, :360-368): the links of the
//                EARLIER position's GPU towards the later one, and 0 for two replicas of one GPU (same *Device, :303-305)
//   choice       the first candidate with the highest score (strict `>`, as at :55 and :81)
//
// Two equivalent forms, both here for host and device:
//   literal      walk the combinations (SubsetEnum): C(R, K) candidates
//   counts       replicas of one GPU are interchangeable for the score, so a candidate class is a vector of per-GPU
//                COUNTS c_g with r_g <= c_g <= r_g + f_g (required / free replicas of GPU g) and sum = size; its score is
//                sum_{g<h} c_g c_h w[g][h], and the lexicographically first member of a class takes the first free
//                replicas of every GPU.  When the positions of every GPU are contiguous (the plugin's sorted id order,
//                cmd/replica.go:41) a class with a LARGER count in the first differing group starts earlier, so
//                enumerating count vectors in descending lexicographic order with first-wins gives the literal answer.
//                C(K+G-1, G-1) classes at most instead of C(R, K): 330 instead of 174 792 640 for 4 of 256 replicas
//                on 8 GPUs.
#pragma once
#include <stdint.h>

#include "gpa_enum.cuh"

namespace gpa {

constexpr int kSubGroups = 16;          // GPUs behind one request
constexpr int kSubMaxExtra = 64;        // replicas to choose beyond the required ones (counts form)
constexpr int kSubCntStride = kSubMaxExtra + 1;
constexpr uint64_t kSubSat = 1ull << 62;  // class counts saturate here; ranks stay far below
constexpr int kSubEnumMaxFree = 256;    // literal form: free positions
constexpr int kSubEnumMaxExtra = 8;     // literal form: replicas to choose beyond the required ones

// --------------------------------------------------------------------------------------------------------------------
// counts form

struct SubsetCounts {
  int32_t G;                           // groups (GPUs) in order of their first position
  int32_t K;                           // replicas to choose beyond the required ones, 1..kSubMaxExtra
  uint8_t f[kSubGroups];               // free non-required replicas of the group, clamped to K
  uint8_t r[kSubGroups];               // required replicas of the group
  uint16_t w[kSubGroups * kSubGroups];  // scaled pair weight, [g * 16 + h] for group ranks g < h
};

GPA_HD uint64_t sub_sat_add(uint64_t a, uint64_t b) {
  const uint64_t s = a + b;
  return s > kSubSat ? kSubSat : s;
}

// One level of the class-count table: cnt[g][s] = number of ways groups g.. can take s more replicas.
// Level G is {1, 0, 0, ...}; callers fill levels G-1 .. 0, entry s of a level depends on the level below only.
GPA_HD uint64_t sub_cnt_entry(const SubsetCounts &q, const uint64_t *cnt, int g, int s) {
  uint64_t sum = 0;
  const int hi = q.f[g] < s ? q.f[g] : s;
  for (int x = 0; x <= hi; x++) sum = sub_sat_add(sum, cnt[(g + 1) * kSubCntStride + s - x]);
  return sum;
}

// Class `rank` in descending lexicographic order of the count vectors -> its score; x_out (optional) gets the counts
// beyond the required ones.  rank < cnt[0][K].  f / r: free and required replicas per group; cnt rows are kStride apart.
template <typename CntT, int kStride>
GPA_HD uint32_t sub_unrank_score_t(int G, int K, const uint8_t *f, const uint8_t *r, const CntT *cnt, const uint16_t *w, uint64_t rank,
                                   uint8_t *x_out) {
  uint32_t c[kSubGroups];
  uint32_t score = 0;
  int left = K;
#pragma unroll
  for (int g = 0; g < kSubGroups; g++) {
    if (g < G) {
      const int fg = f[g];
      int x = fg < left ? fg : left;
      while (x > 0) {
        const uint64_t below = cnt[(g + 1) * kStride + left - x];
        if (rank < below) break;
        rank -= below;
        x--;
      }
      left -= x;
      if (x_out) x_out[g] = (uint8_t)x;
      c[g] = (uint32_t)r[g] + (uint32_t)x;
      uint32_t lin = 0;
#pragma unroll
      for (int i = 0; i < kSubGroups; i++)
        if (i < g) lin += c[i] * w[i * kSubGroups + g];
      score += c[g] * lin;
    }
  }
  return score;
}

GPA_HD uint32_t sub_unrank_score(const SubsetCounts &q, const uint64_t *cnt, const uint16_t *w, uint64_t rank, uint8_t *x_out) {
  return sub_unrank_score_t<uint64_t, kSubCntStride>(q.G, q.K, q.f, q.r, cnt, w, rank, x_out);
}

// --------------------------------------------------------------------------------------------------------------------
// literal form: K-combinations of the R free positions, required positions folded into `wreq` and `base`

struct SubsetEnum {
  int32_t R;                               // free (non-required) positions, in position order
  int32_t K;                               // how many of them to choose, 1..kSubEnumMaxExtra
  uint32_t base;                           // score of the required positions among themselves
  uint8_t grp[kSubEnumMaxFree];            // group of free position i
  uint32_t wreq[kSubEnumMaxFree];          // score of free position i against all required positions
  uint16_t w[kSubGroups * kSubGroups];     // scaled pair weight [g * 16 + h]: earlier position in g, later in h; 0 for g == h
};

// C(n, j) for n <= 256, j <= 8, saturated: row-major [n * 9 + j]
constexpr int kSubBinomRows = kSubEnumMaxFree + 1;
constexpr int kSubBinomCols = kSubEnumMaxExtra + 1;

inline void sub_fill_binom(uint64_t *b) {
  for (int n = 0; n < kSubBinomRows; n++)
    for (int j = 0; j < kSubBinomCols; j++) {
      uint64_t v;
      if (j == 0)
        v = 1;
      else if (n == 0)
        v = 0;
      else
        v = sub_sat_add(b[(n - 1) * kSubBinomCols + j - 1], b[(n - 1) * kSubBinomCols + j]);
      b[n * kSubBinomCols + j] = v;
    }
}

// prefix scores: ps[t + 1] = score of (required + e[0..t]) for t < N; ge[t] = 16 * group of e[t]
template <int N>
GPA_HD void sub_prefix_score(const uint8_t *grp, const uint32_t *wreq, const uint16_t *w, const int *e, int *ge, uint32_t *ps) {
#pragma unroll
  for (int t = 0; t < N; t++) {
    const int g = grp[e[t]];
    uint32_t s = ps[t] + wreq[e[t]];
#pragma unroll
    for (int u = 0; u < N; u++)
      if (u < t) s += w[ge[u] + g];
    ge[t] = g * kSubGroups;
    ps[t + 1] = s;
  }
}

// Number of work items of the literal walk: the (K-1)-prefixes, i.e. combinations of the first R-1 free positions.
GPA_HD uint64_t sub_prefix_count(const uint64_t *binom, int R, int K) { return binom[(R - 1) * kSubBinomCols + (K - 1)]; }

// One work item of the literal walk: all combinations that share their first K-1 picks, i.e. a run over the last pick.
// Prefixes are numbered in COLEXICOGRAPHIC order (by their largest pick first), so neighbouring items - the lanes of a
// warp - have runs of the same length and step through them together; the reference's visiting order survives in the
// key, which carries each candidate's LEXICOGRAPHIC rank: C(R,K) - 1 - sum_i C(R-1-e_i, K-i).
// Returns the larger of `best` and the run's first best candidate (score, earliest rank).
template <int K>
GPA_HD unsigned long long sub_walk_prefix(const uint64_t *binom, const uint8_t *grp, const uint32_t *wreq, const uint16_t *w, int R,
                                          uint32_t base, uint64_t item, unsigned long long best) {
  int e[K], ge[K];  // [K-1] unused
  uint32_t ps[K + 1];
  uint64_t rest = item;
  uint64_t after = 0;  // sum_i C(R-1-e_i, K-i) over the prefix
#pragma unroll
  for (int i = K - 2; i >= 0; i--) {  // colex unrank: the largest c with C(c, i+1) <= rest, binary search
    int lo = i, hi = R - 2;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (binom[mid * kSubBinomCols + i + 1] <= rest)
        lo = mid;
      else
        hi = mid - 1;
    }
    e[i] = lo;
    rest -= binom[lo * kSubBinomCols + i + 1];
    after += binom[(R - 1 - lo) * kSubBinomCols + (K - i)];
  }
  ps[0] = base;
  sub_prefix_score<K - 1>(grp, wreq, w, e, ge, ps);
  const int x0 = K > 1 ? e[K > 1 ? K - 2 : 0] + 1 : 0;
  const uint32_t pb = ps[K - 1];
  int run_best = -1, run_x = x0;
  for (int x = x0; x < R; x++) {
    const int g = grp[x];
    uint32_t sc = pb + wreq[x];
#pragma unroll
    for (int u = 0; u < K - 1; u++) sc += w[ge[u] + g];
    if ((int)sc > run_best) {  // strict: the first best of the run
      run_best = (int)sc;
      run_x = x;
    }
  }
  const uint64_t rank = binom[R * kSubBinomCols + K] - 1 - after - (uint64_t)(R - 1 - run_x);
  const unsigned long long key = pack_key(run_best, rank);
  return key > best ? key : best;
}

// --------------------------------------------------------------------------------------------------------------------
// count form, depth first (the batch kernel's walk; also driven by the CPU harness)
//
// The count vectors are walked one GPU per level, in descending order of the count at every level (= descending
// lexicographic order of the vectors).  P = score of the levels above, L[h] = sum_{g above} c_g w[g][h]: a step at level g
// costs G-g-1 additions and a leaf two operations.  f / r: free (clamped to K) and required replicas per GPU;
// cap[g] = what the GPUs from g on can still absorb.

template <int G, int g>
struct SubDfs {
  static GPA_HD void run(const uint8_t *f, const uint8_t *r, const uint16_t *cap, const uint16_t *w, int left, int P,
                                             int (&L)[G], int (&xs)[G], int &bs, unsigned long long &bp) {
    const int rg = r[g];
    const int fg = f[g];
    const int hi = fg < left ? fg : left;
    const int rest = left - (int)cap[g + 1];  // what the GPUs below cannot absorb
    const int lo = rest > 0 ? rest : 0;
#pragma unroll
    for (int h = g + 1; h < G; h++) L[h] += (rg + hi) * (int)w[g * kSubGroups + h];
    for (int x = hi; x >= lo; x--) {
      xs[g] = x;
      SubDfs<G, g + 1>::run(f, r, cap, w, left - x, P + (rg + x) * L[g], L, xs, bs, bp);
#pragma unroll
      for (int h = g + 1; h < G; h++) L[h] -= (int)w[g * kSubGroups + h];
    }
#pragma unroll
    for (int h = g + 1; h < G; h++) L[h] -= (rg + lo - 1) * (int)w[g * kSubGroups + h];  // back to the caller's L
  }
};

template <int G>
struct SubDfs<G, G - 1> {  // the last GPU takes what is left (the level above made sure it can)
  static GPA_HD void run(const uint8_t *, const uint8_t *r, const uint16_t *, const uint16_t *, int left, int P,
                                             int (&L)[G], int (&xs)[G], int &bs, unsigned long long &bp) {
    const int score = P + ((int)r[G - 1] + left) * L[G - 1];
    if (score > bs) {  // strict: the first best vector of the walk stays
      bs = score;
      xs[G - 1] = left;
      unsigned long long pk = 0;
#pragma unroll
      for (int i = 0; i < G; i++) pk |= (unsigned long long)(unsigned)xs[i] << (4 * i);
      bp = pk;
    }
  }
};

// exact[p * 16 + s] = ways for prefix GPUs p..D-1 to take exactly s (level D is {1, 0, ...}); one entry of level p
GPA_HD uint32_t sub_exact_entry(const uint8_t *f, const uint32_t *exact, int p, int s) {
  uint32_t sum = 0;
  const int hi = (int)f[p] < s ? (int)f[p] : s;
  for (int x = 0; x <= hi; x++) sum += exact[(p + 1) * 16 + s - x];
  return sum;
}

// One work item = one valid count prefix of the first D GPUs (numbered class by class, most replicas left first; inside
// a class in descending lexicographic order) and the whole walk below it; merges the item's first best vector into
// (bs, bp, blex) - ties go to the lexicographically larger prefix, because items do not come in vector order.
template <int G, int D>
GPA_HD void sub_batch_item(const uint8_t *f, const uint8_t *r, const uint16_t *cap, const uint16_t *w, const uint32_t *exact, int K,
                         int s_lo, uint32_t item, int &bs, unsigned long long &bp, uint32_t &blex) {
  uint32_t rest = item;
  int sp = s_lo;
  while (rest >= exact[sp]) rest -= exact[sp++];
  int xs[G];
  int todo = sp;
  uint32_t lex = 0;
#pragma unroll
  for (int p = 0; p < D; p++) {  // prefix number `rest` of its class, descending lexicographic order
    int v = (int)f[p] < todo ? (int)f[p] : todo;
    while (v > 0) {
      const uint32_t with_v = exact[(p + 1) * 16 + todo - v];
      if (rest < with_v) break;
      rest -= with_v;
      v--;
    }
    xs[p] = v;
    todo -= v;
    lex = (lex << 4) | (uint32_t)v;
  }
  int L[G];
#pragma unroll
  for (int h = 0; h < G; h++) L[h] = 0;
  int P = 0;
#pragma unroll
  for (int p = 0; p < D; p++) {
    const int c = (int)r[p] + xs[p];
    P += c * L[p];
#pragma unroll
    for (int h = p + 1; h < G; h++) L[h] += c * (int)w[p * kSubGroups + h];
  }
  int is = -1;  // this item's first best
  unsigned long long ip = 0;
  SubDfs<G, D>::run(f, r, cap, w, K - sp, P, L, xs, is, ip);
  if (is > bs || (is == bs && lex > blex)) {  // items do not come in vector order: ties go to the larger prefix
    bs = is;
    bp = ip;
    blex = lex;
  }
}

// --------------------------------------------------------------------------------------------------------------------
// count form, two halves (the batch kernel's walk on nodes of up to 8 GPUs, K <= kSubHalfMaxK)
//
// With c_g = r_g + x_g the score of a count vector splits over the GPU labels 0..3 ("prefix") and 4..7 ("suffix"):
//   sum_{g<h} c_g c_h w[g][h] = R0 + [P(xp) + xp . lin_p] + [S(xs) + xs . (lin_s + L(xp))]
//   R0      = sum_{g<h} r_g r_h w[g][h]                       the must-include replicas among themselves
//   lin_h   = sum_{g != h} r_g w[min(g,h)][max(g,h)]          what one more replica of GPU h adds against them
//   P / S   = sum_{g<h} x_g x_h w[g][h] inside a half         a property of the half-vector and the node alone
//   L_h(xp) = sum_{g<4} x_g w[g][h], h = 4..7                 likewise
// so every 4-part composition of 0..kSubHalfMaxK gets its P, L and S ONCE per node (sub_pair_fill, 495 entries), and a
// request only has to (1) drop the compositions its free counts do not allow and (2) visit the pairs (xp, xs) with
// sum(xp) + sum(xs) = K at one multiply-add chain each.  The compositions are stored class by class (sum ascending) and
// in descending lexicographic order inside a class, GPU 0 / 4 of the half in the HIGHEST byte, so that "first in
// descending order of the count vector" (subset_core's tie-break) is: larger xp word first, then larger xs word, and a
// walk through a suffix class with strict `>` keeps the right one.
constexpr int kSubHalfMaxK = 8;
constexpr int kSubHalfEntries = 495;  // C(8 + 4, 4)
GPA_HD int sub_half_start(int t) { return t * (t + 1) * (t + 2) * (t + 3) / 24; }  // compositions with sum < t

// blob layout (bytes): comp u32[495] | P u32[495] | suffix {S, x} u32x2[495] | L u32x4[495] | Q u32x2[495] | meta u32[2]
// Q = the low 16 bits of L, paired for the 2-way dot products: {GPU 5 | GPU 4 << 16, GPU 7 | GPU 6 << 16}
constexpr int kSubPairCompOff = 0;
constexpr int kSubPairPOff = 4 * kSubHalfEntries;
constexpr int kSubPairSufOff = 8 * kSubHalfEntries;
constexpr int kSubPairLOff = 16 * kSubHalfEntries;
constexpr int kSubPairQOff = 32 * kSubHalfEntries;
constexpr int kSubPairMetaOff = 40 * kSubHalfEntries;  // [0] = largest weight between a prefix and a suffix GPU
constexpr int kSubPairBlobBytes = 40 * kSubHalfEntries + 8;
static_assert(kSubPairBlobBytes % 16 == 0 && kSubPairLOff % 16 == 0 && kSubPairQOff % 8 == 0, "staged and read with vector accesses");

struct alignas(8) SubWord2 {
  uint32_t x, y;
};
struct alignas(16) SubWord4 {
  uint32_t x, y, z, w;
};

// w: label-space pair weights [a * 16 + b], a < b (labels 0..7 used)
inline void sub_pair_fill(const uint16_t *w, uint32_t *blob) {
  uint32_t *comp = blob + kSubPairCompOff / 4, *P = blob + kSubPairPOff / 4, *suf = blob + kSubPairSufOff / 4, *L = blob + kSubPairLOff / 4;
  uint32_t *meta = blob + kSubPairMetaOff / 4, *Q = blob + kSubPairQOff / 4;
  meta[0] = meta[1] = 0;
  for (int g = 0; g < 4; g++)
    for (int h = 4; h < 8; h++)
      if (w[g * kSubGroups + h] > meta[0]) meta[0] = w[g * kSubGroups + h];
  int e = 0;
  for (int t = 0; t <= kSubHalfMaxK; t++)
    for (int a = t; a >= 0; a--)
      for (int b = t - a; b >= 0; b--)
        for (int c = t - a - b; c >= 0; c--) {
          const int x[4] = {a, b, c, t - a - b - c};
          comp[e] = (uint32_t)x[0] << 24 | (uint32_t)x[1] << 16 | (uint32_t)x[2] << 8 | (uint32_t)x[3];
          uint32_t p = 0, s = 0;
          for (int g = 0; g < 4; g++)
            for (int h = g + 1; h < 4; h++) {
              p += (uint32_t)(x[g] * x[h]) * w[g * kSubGroups + h];
              s += (uint32_t)(x[g] * x[h]) * w[(4 + g) * kSubGroups + 4 + h];
            }
          P[e] = p;
          suf[2 * e] = s;
          suf[2 * e + 1] = comp[e];
          for (int h = 0; h < 4; h++) {
            uint32_t l = 0;
            for (int g = 0; g < 4; g++) l += (uint32_t)x[g] * w[g * kSubGroups + 4 + h];
            L[4 * e + h] = l;
          }
          Q[2 * e] = (L[4 * e + 1] & 0xFFFFu) | (L[4 * e + 0] << 16);
          Q[2 * e + 1] = (L[4 * e + 3] & 0xFFFFu) | (L[4 * e + 2] << 16);
          e++;
        }
}

// does composition x respect the free counts of the half?  Both words hold one count per byte, all below 128.
GPA_HD bool sub_half_fits(uint32_t x, uint32_t fpack) { return (((fpack | 0x80808080u) - x) & 0x80808080u) == 0x80808080u; }

GPA_HD uint32_t sub_dp2a_lo(uint32_t a, uint32_t b, uint32_t c) {  // c + a.lo16 * b.byte0 + a.hi16 * b.byte1
#ifdef __CUDA_ARCH__
  return __dp2a_lo(a, b, c);
#else
  return c + (a & 0xFFFFu) * (b & 0xFFu) + (a >> 16) * ((b >> 8) & 0xFFu);
#endif
}
GPA_HD uint32_t sub_dp2a_hi(uint32_t a, uint32_t b, uint32_t c) {  // c + a.lo16 * b.byte2 + a.hi16 * b.byte3
#ifdef __CUDA_ARCH__
  return __dp2a_hi(a, b, c);
#else
  return c + (a & 0xFFFFu) * ((b >> 16) & 0xFFu) + (a >> 16) * (b >> 24);
#endif
}

// Up to kSubFlatMaxK further replicas (what most containers ask for): no lists, no pairs.  All count vectors over 8 GPUs
// with sum <= 4 are 495; the node's table holds each one (prefix word, suffix word) and its score without must-include
// replicas, class by class (sum) and in descending order inside a class, so a request is one pass over its class: does
// the vector fit the free counts, table score + x . lin, maximum on (score, earliest entry).
constexpr int kSubFlatMaxK = 4;
constexpr int kSubFlatEntries = 495;  // C(4 + 8, 8)
GPA_HD int sub_flat_start(int t) {    // vectors with sum < t = C(t + 7, 8)
  return t <= 0 ? 0 : (t == 1 ? 1 : (t == 2 ? 9 : (t == 3 ? 45 : (t == 4 ? 165 : 495))));
}
// table layout (bytes): x {prefix word, suffix word} u32x2[495] | score u32[495] (+ 12 bytes of padding)
constexpr int kSubFlatScoreOff = 8 * kSubFlatEntries;
constexpr int kSubFlatBytes = 12 * kSubFlatEntries + 12;
static_assert(kSubFlatBytes % 16 == 0, "uploaded and addressed in 16-byte units");

inline void sub_flat_fill(const uint16_t *w, uint32_t *tab) {
  int e = 0;
  int x[8];
  for (int t = 0; t <= kSubFlatMaxK; t++) {
    // descending lexicographic order: odometer over x[0..6], x[7] takes the rest
    for (int i = 0; i < 8; i++) x[i] = 0;
    x[0] = t;
    for (;;) {
      uint32_t xp = 0, xs = 0, sc = 0;
      for (int g = 0; g < 4; g++) {
        xp |= (uint32_t)x[g] << (8 * (3 - g));
        xs |= (uint32_t)x[4 + g] << (8 * (3 - g));
      }
      for (int g = 0; g < 8; g++)
        for (int h = g + 1; h < 8; h++) sc += (uint32_t)(x[g] * x[h]) * w[g * kSubGroups + h];
      tab[2 * e] = xp;
      tab[2 * e + 1] = xs;
      tab[kSubFlatScoreOff / 4 + e] = sc;
      e++;
      // next vector: move one unit from the last non-zero position before x[7] to its right neighbour, and pull
      // everything right of that neighbour back to it
      int p = 6;
      while (p >= 0 && x[p] == 0) p--;
      if (p < 0) break;  // everything sits in x[7]: the class is done
      x[p]--;
      const int carry = x[7] + 1;
      x[7] = 0;
      x[p + 1] += carry;
    }
  }
  for (int i = 3 * kSubFlatEntries; i < kSubFlatBytes / 4; i++) tab[i] = 0;
}

// Entry e for a request: 0 if the free counts do not allow it, else (score + 1) << 32 | ~e (the maximum is the first best).
// kPlain: every GPU has K replicas free and none is must-include - the table score is the answer's score.
template <bool kPlain>
GPA_HD unsigned long long sub_flat_key(const uint32_t *tab, int e, uint32_t fpack_p, uint32_t fpack_s, uint32_t r0, const uint32_t *lin,
                                       bool has_req) {
  uint32_t sc = tab[kSubFlatScoreOff / 4 + e];
  if (!kPlain) {
    const SubWord2 x = *reinterpret_cast<const SubWord2 *>(tab + 2 * e);
    if (!sub_half_fits(x.x, fpack_p) || !sub_half_fits(x.y, fpack_s)) return 0ull;
    if (has_req) {
      sc += r0 + (x.x >> 24) * lin[0] + ((x.x >> 16) & 0xFFu) * lin[1] + ((x.x >> 8) & 0xFFu) * lin[2] + (x.x & 0xFFu) * lin[3] +
            (x.y >> 24) * lin[4] + ((x.y >> 16) & 0xFFu) * lin[5] + ((x.y >> 8) & 0xFFu) * lin[6] + (x.y & 0xFFu) * lin[7];
    }
  }
  return ((unsigned long long)(sc + 1u) << 32) | (uint32_t)~(uint32_t)e;
}
// the best key of lane `lane` over class K
GPA_HD unsigned long long sub_flat_lane(const uint32_t *tab, int K, uint32_t fpack_p, uint32_t fpack_s, uint32_t r0, const uint32_t *lin,
                                        bool has_req, int lane) {
  const uint32_t all_k = (uint32_t)K * 0x01010101u;
  const int e_end = sub_flat_start(K + 1);
  unsigned long long best = 0;
  if (!has_req && fpack_p == all_k && fpack_s == all_k) {
    for (int e = sub_flat_start(K) + lane; e < e_end; e += 32) {
      const unsigned long long k = sub_flat_key<true>(tab, e, fpack_p, fpack_s, r0, lin, false);
      best = k > best ? k : best;
    }
  } else {
    for (int e = sub_flat_start(K) + lane; e < e_end; e += 32) {
      const unsigned long long k = sub_flat_key<false>(tab, e, fpack_p, fpack_s, r0, lin, has_req);
      best = k > best ? k : best;
    }
  }
  return best;
}

// A lane's running best: (score + 1) << 32 | prefix word, and which composition that prefix is.  key 0 = nothing yet.
// The suffix is not tracked: the warp finds it again for the one winning prefix (sub_pair_find_suffix).
struct SubPairBest {
  unsigned long long key;
  int pe;
};
GPA_HD int sub_pair_score(const SubPairBest &b) { return (int)(b.key >> 32) - 1; }
GPA_HD uint32_t sub_pair_prefix(const SubPairBest &b) { return (uint32_t)b.key; }

// What a prefix contributes to every pair it is part of: the score without the suffix, and its L (+ lin) in 16-bit halves
struct SubPairPrefix {
  uint32_t base, q45lo, q67lo, q45hi, q67hi;
};
// kMode 0: no must-include replicas (r0 = 0, lin = 0): straight from the table, low halves only (8 * 546 < 65 536)
//       1: every L + lin of this request stays below 65 536 (sub_pair_small): low halves only
//       2: the general case
template <int kMode>
GPA_HD SubPairPrefix sub_pair_prefix_terms(const uint32_t *blob, int pe, uint32_t r0, const uint32_t *lin) {
  const uint32_t *comp = blob + kSubPairCompOff / 4, *P = blob + kSubPairPOff / 4, *L = blob + kSubPairLOff / 4;
  SubPairPrefix t;
  t.q45hi = t.q67hi = 0;
  if (kMode == 0) {
    const SubWord2 q = *reinterpret_cast<const SubWord2 *>(blob + kSubPairQOff / 4 + 2 * pe);
    t.base = P[pe];
    t.q45lo = q.x;
    t.q67lo = q.y;
    return t;
  }
  const uint32_t xp = comp[pe];
  t.base = r0 + P[pe] + (xp >> 24) * lin[0] + ((xp >> 16) & 0xFFu) * lin[1] + ((xp >> 8) & 0xFFu) * lin[2] + (xp & 0xFFu) * lin[3];
  const SubWord4 l = *reinterpret_cast<const SubWord4 *>(L + 4 * pe);
  const uint32_t l0 = l.x + lin[4], l1 = l.y + lin[5], l2 = l.z + lin[6], l3 = l.w + lin[7];
  // {lo16: GPU 7, hi16: GPU 6} and {lo16: GPU 5, hi16: GPU 4}
  t.q67lo = (l3 & 0xFFFFu) | (l2 << 16);
  t.q45lo = (l1 & 0xFFFFu) | (l0 << 16);
  if (kMode == 2) {
    t.q67hi = (l3 >> 16) | (l2 & 0xFFFF0000u);
    t.q45hi = (l1 >> 16) | (l0 & 0xFFFF0000u);
  }
  return t;
}

// what the suffix adds to the prefix's base: S + xs . (L + lin)
template <bool kFull>
GPA_HD int sub_pair_gain(const SubPairPrefix &t, const SubWord2 &rec) {
  uint32_t sc = sub_dp2a_lo(t.q67lo, rec.y, sub_dp2a_hi(t.q45lo, rec.y, rec.x));
  if (kFull) sc += sub_dp2a_lo(t.q67hi, rec.y, sub_dp2a_hi(t.q45hi, rec.y, 0u)) << 16;
  return (int)sc;
}

// One work item: the best score of prefix composition `pe` with suffixes number s_first, s_first + step, ... below s_end
// of the list (one suffix class); -1 when there is none.
//   kIdent  the suffix half dropped nothing, so the list is the table itself and is not read
template <bool kIdent, int kMode>
GPA_HD int sub_pair_item(const uint32_t *blob, int pe, const uint16_t *slist, int s_first, int s_end, int step, uint32_t r0,
                         const uint32_t *lin) {
  const SubWord2 *suf = reinterpret_cast<const SubWord2 *>(blob + kSubPairSufOff / 4);
  const SubPairPrefix t = sub_pair_prefix_terms<kMode>(blob, pe, r0, lin);
  int gain = -1;
  int s = s_first;
  for (; s + 3 * step < s_end; s += 4 * step) {  // four a turn: a quarter of the loop control, four loads in flight
    const SubWord2 a = suf[kIdent ? s : (int)slist[s]], b = suf[kIdent ? s + step : (int)slist[s + step]];
    const SubWord2 c = suf[kIdent ? s + 2 * step : (int)slist[s + 2 * step]], d = suf[kIdent ? s + 3 * step : (int)slist[s + 3 * step]];
    const int ga = sub_pair_gain<kMode == 2>(t, a), gb = sub_pair_gain<kMode == 2>(t, b);
    const int gc = sub_pair_gain<kMode == 2>(t, c), gd = sub_pair_gain<kMode == 2>(t, d);
    const int m1 = ga > gb ? ga : gb, m2 = gc > gd ? gc : gd;
    const int m3 = m1 > m2 ? m1 : m2;
    gain = m3 > gain ? m3 : gain;
  }
  for (; s < s_end; s += step) {
    const int ga = sub_pair_gain<kMode == 2>(t, suf[kIdent ? s : (int)slist[s]]);
    gain = ga > gain ? ga : gain;
  }
  return gain < 0 ? -1 : (int)t.base + gain;  // the base is the same for every pair of the item
}

// The winning prefix once more, by the whole warp: the first suffix of its class (list order = descending, so the one the
// tie-break wants) that reaches `target`.  Returns this lane's first list position that does, or 0xFFFFFFFF; the warp
// takes the minimum.
GPA_HD uint32_t sub_pair_find_suffix(const uint32_t *blob, int pe, const uint16_t *slist, int s_begin, int s_end, uint32_t r0,
                                     const uint32_t *lin, int target, int lane) {
  const SubWord2 *suf = reinterpret_cast<const SubWord2 *>(blob + kSubPairSufOff / 4);
  const SubPairPrefix t = sub_pair_prefix_terms<2>(blob, pe, r0, lin);
  for (int s = s_begin + lane; s < s_end; s += 32)
    if ((int)t.base + sub_pair_gain<true>(t, suf[slist[s]]) == target) return (uint32_t)s;
  return 0xFFFFFFFFu;
}

// can the high halves be skipped?  wmax = largest weight between a prefix and a suffix GPU; lin_max over the suffix GPUs
GPA_HD bool sub_pair_small(int K, uint32_t wmax, uint32_t lin_max) { return (uint32_t)K * wmax + lin_max < 65536u && lin_max < 65536u; }
// kMode of sub_pair_prefix_terms for a request
GPA_HD int sub_pair_mode(int K, uint32_t wmax, uint32_t lin_max, int n_req) {
  return !sub_pair_small(K, wmax, lin_max) ? 2 : (n_req == 0 ? 0 : 1);
}

// When fewer than 32 prefixes of a class are left, each gets m = 32 / rem lanes and the suffix class is dealt to them
// round-robin.  Table entry for rem = 1..31: m | ceil(1024 / m) << 8; lane / m = lane * that >> 10, exact for lane < 32.
GPA_HD uint32_t sub_pair_split_entry(int rem) {
  const uint32_t m = rem >= 1 && rem < 32 ? 32u / (uint32_t)rem : 1u;
  return m | ((1024u + m - 1u) / m) << 8;
}

// Everything lane `lane` of the warp does for one request once the lists stand: plist / slist = the compositions each half
// allows (indices into the blob, class by class), poff / soff[t] = where class t starts, [K + 1] = the end.
template <bool kIdent, int kMode>
GPA_HD void sub_pair_lane_t(const uint32_t *blob, const uint16_t *plist, const uint16_t *poff, const uint16_t *slist, const uint16_t *soff,
                            int K, uint32_t r0, const uint32_t *lin_in, const uint32_t *split_tab, int lane, SubPairBest &best) {
  uint32_t lin[8];
#pragma unroll
  for (int g = 0; g < 8; g++) lin[g] = kMode == 0 ? 0u : lin_in[g];
  for (int t = 0; t <= K; t++) {
    const int pb = poff[t], pe = poff[t + 1], sb = soff[K - t], se = soff[K - t + 1];
    if (pe == pb || se == sb) continue;
    for (int base = pb; base < pe; base += 32) {
      const int rem = pe - base;
      int j = lane, sub = 0, m = 1;
      if (rem < 32) {
        const uint32_t e = split_tab[rem];
        m = (int)(e & 0xFFu);
        j = (int)(((uint32_t)lane * (e >> 8)) >> 10);
        sub = lane - j * m;
      }
      if (j < rem) {
        const int entry = plist[base + j];
        const int is = sub_pair_item<kIdent, kMode>(blob, entry, slist, sb + sub, se, m, r0, lin);
        const unsigned long long key = ((unsigned long long)(unsigned)(is + 1) << 32) | blob[kSubPairCompOff / 4 + entry];
        if (is >= 0 && key > best.key) {
          best.key = key;
          best.pe = entry;
        }
      }
    }
  }
}

// ident / mode: the same for every lane of the warp
GPA_HD void sub_pair_lane(const uint32_t *blob, const uint16_t *plist, const uint16_t *poff, const uint16_t *slist, const uint16_t *soff,
                          int K, uint32_t r0, const uint32_t *lin, const uint32_t *split_tab, bool ident, int mode, int lane,
                          SubPairBest &best) {
#define GPA_SUB_LANE(I, M) sub_pair_lane_t<I, M>(blob, plist, poff, slist, soff, K, r0, lin, split_tab, lane, best)
  if (ident) {
    if (mode == 0)
      GPA_SUB_LANE(true, 0);
    else if (mode == 1)
      GPA_SUB_LANE(true, 1);
    else
      GPA_SUB_LANE(true, 2);
  } else {
    if (mode == 0)
      GPA_SUB_LANE(false, 0);
    else if (mode == 1)
      GPA_SUB_LANE(false, 1);
    else
      GPA_SUB_LANE(false, 2);
  }
#undef GPA_SUB_LANE
}

}  // namespace gpa
