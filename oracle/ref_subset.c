/* TEST INFRASTRUCTURE ONLY - CPU oracle of the replica-set search (gpa_subset_besteffort).
 *
 * PARITY UNPINNED, and more than that: the reference has NO function that does this as a whole.  It answers requests
 * over shared replicas with prioritizeDevices (cmd/replica.go:96-198, oracle: ref_replica.c).  What is restated here is
 * BASELINE.json north_star's "all k-subsets of the available replica set scored against the link-weight matrix"
 * (SURVEY.md 8d, config 3c), composed from reference pieces, each followed literally:
 *
 *   iterateGPUSets              vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:155-204
 *                               every `size`-combination in lexicographic index order (iterative odometer)
 *   gpuSetContainsAll           :102-110   the candidate must hold every required device
 *   calculateGPUPairScore       :298-356   0 for the same *Device (two replicas of one GPU, :303-305); panic when the
 *                               two link lists differ in length (:307-310); else the EARLIER device's links
 *   calculateGPUSetScore        :360-368   sum over the C(size, 2) pairs
 *   first best wins             strict `>` as at :55 / :81
 *   guards                      :35-45     size <= 0, fewer available than size, more required than size => empty
 *
 * Deliberately the slow, literal form (no interchangeable-replica shortcut): it is the checker. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { SUB_OK = 0, SUB_PANIC_ASYMMETRIC = 1, SUB_ERR_ARG = 4 };

typedef struct {
  int nt;
  const int32_t *W;
  const uint8_t *nl;
  const int32_t *phys;
  int panicked;
} sub_env;

/* besteffort_policy.go:298-356 with the weight sum precomputed per ordered pair (ref_links_to_weights) */
static long long pair_score(sub_env *e, int a, int b) {
  const int pa = e->phys[a], pb = e->phys[b];
  if (pa == pb) return 0; /* :303-305 */
  if (e->nl[pa * e->nt + pb] != e->nl[pb * e->nt + pa]) { /* :307-310 */
    e->panicked = 1;
    return 0;
  }
  return e->W[pa * e->nt + pb]; /* :314 - g0's links to g1 */
}

/* -> SUB_OK with *n_out = 0 (empty answer) or size; positions ascending in out_pos */
int ref_subset_besteffort(int nt, const int32_t *W, const uint8_t *nlinks, const int32_t *phys, int n_avail,
                          const int32_t *req_pos, int n_req, int size, int32_t *out_pos, int *n_out, long long *best_score,
                          unsigned long long *visited) {
  *n_out = 0;
  if (best_score) *best_score = -1;
  if (visited) *visited = 0;
  for (int i = 0; i < n_avail; i++)
    if (phys[i] < 0 || phys[i] >= nt) return SUB_ERR_ARG;
  for (int i = 0; i < n_req; i++)
    if (req_pos[i] < 0 || req_pos[i] >= n_avail) return SUB_ERR_ARG;
  if (size <= 0) return SUB_OK;     /* :35-37 */
  if (n_avail < size) return SUB_OK; /* :39-41 */
  if (n_req > size) return SUB_OK;   /* :43-45 */

  sub_env env = {nt, W, nlinks, phys, 0};
  int *idx = (int *)malloc(sizeof(int) * (size_t)size);
  int *best = (int *)malloc(sizeof(int) * (size_t)size);
  uint8_t *need = (uint8_t *)calloc((size_t)n_avail + 1, 1);
  if (!idx || !best || !need) {
    free(idx);
    free(best);
    free(need);
    return SUB_ERR_ARG;
  }
  for (int i = 0; i < n_req; i++) need[req_pos[i]] = 1;
  int n_need = 0;
  for (int i = 0; i < n_avail; i++) n_need += need[i];

  /* :163-204 - the odometer: indices start at 0..size-1, the rightmost one that can still move is advanced */
  for (int i = 0; i < size; i++) idx[i] = i;
  long long top = -1;
  unsigned long long seen = 0;
  for (;;) {
    int have = 0;
    for (int i = 0; i < size; i++) have += need[idx[i]];
    if (have == n_need) { /* gpuSetContainsAll :102-110 */
      long long s = 0;
      for (int i = 0; i < size; i++)
        for (int j = i + 1; j < size; j++) s += pair_score(&env, idx[i], idx[j]); /* :360-368 */
      seen++;
      if (s > top) { /* strict: the first best candidate stays */
        top = s;
        memcpy(best, idx, sizeof(int) * (size_t)size);
      }
    }
    int i = size - 1;
    while (i >= 0 && idx[i] == n_avail - size + i) i--;
    if (i < 0) break;
    idx[i]++;
    for (int j = i + 1; j < size; j++) idx[j] = idx[j - 1] + 1;
  }
  int rc = SUB_OK;
  if (env.panicked) {
    rc = SUB_PANIC_ASYMMETRIC;
  } else if (top >= 0) {
    for (int i = 0; i < size; i++) out_pos[i] = best[i];
    *n_out = size;
    if (best_score) *best_score = top;
  }
  if (visited) *visited = seen;
  free(idx);
  free(best);
  free(need);
  return rc;
}
