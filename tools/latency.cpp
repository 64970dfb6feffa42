// Native (no Python) latency of ONE allocation decision through the C ABI, next to the oracle: the honest small-request
// comparison.  n = 8 devices (T3 topology), every size k = 1..8, no required devices.
//   g++ -O2 -std=c++17 tools/latency.cpp -o /tmp/latency -I include -L k8s-gpu-sharing-plugin_b200 -lgpualloc -L oracle -loracle
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <vector>

#include "gpualloc.h"

extern "C" int ref_besteffort_fast(int nt, const int32_t *W, const uint8_t *nlinks, const int32_t *avail, int n_avail, const int32_t *req,
                                   int n_req, int size, int mode, int32_t *out, int *n_out, unsigned long long *visited);

extern "C" int ref_subset_besteffort(int nt, const int32_t *W, const uint8_t *nlinks, const int32_t *phys, int n_avail, const int32_t *req_pos,
                                     int n_req, int size, int32_t *out_pos, int *n_out, long long *best_score, unsigned long long *visited);

int main() {
  const int n = 8;
  // synthetic hybrid cube mesh (synthetic.topo_dgx1v_like): PCIe + NVLink weights
  int32_t W[64];
  uint8_t nl[64];
  auto pcie = [](int i, int j) {
    if ((i ^ j) == 1 && (i / 2) == (j / 2)) return 50;
    return (i < 4) == (j < 4) ? 20 : 10;
  };
  const int two[8][2] = {{0, 3}, {0, 4}, {1, 2}, {1, 5}, {2, 3}, {4, 7}, {5, 6}, {6, 7}};
  const int one[8][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 6}, {3, 7}, {4, 5}, {4, 6}, {5, 7}};
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      W[i * n + j] = i == j ? 0 : pcie(i, j);
      nl[i * n + j] = i == j ? 0 : 1;
    }
  for (auto &p : two) {
    W[p[0] * n + p[1]] += 200; W[p[1] * n + p[0]] += 200;
    nl[p[0] * n + p[1]] = nl[p[1] * n + p[0]] = 2;
  }
  for (auto &p : one) {
    W[p[0] * n + p[1]] += 100; W[p[1] * n + p[0]] += 100;
    nl[p[0] * n + p[1]] = nl[p[1] * n + p[0]] = 2;
  }
  gpa_ctx *ctx = nullptr;
  if (gpa_create(0, &ctx) != GPA_OK) return 2;
  if (gpa_set_topology(ctx, n, W, nl) != GPA_OK) return 3;
  int32_t avail[8] = {0, 1, 2, 3, 4, 5, 6, 7}, out[16], out2[16];
  printf("{\"what\": \"one decision on an 8-GPU node, native C ABI vs oracle (one host thread), microseconds; "
         "table = gpa_besteffort default (host copy of the GPU-built answer table, no launch), kernel = GPA_MODE_FORCE_KERNEL "
         "(one-warp kernel + mapped mailbox)\", \"cases\": [");
  bool first = true;
  for (int with_req = 0; with_req < 2; with_req++)
    for (int k = 1; k <= 8; k++) {
      const int32_t req[1] = {3};
      const int n_req = with_req;
      std::vector<double> tb, g, c;
      int n_out = 0, n_out2 = 0, n_out3 = 0;
      int32_t out3[16];
      for (int i = 0; i < 4200; i++) {
        auto t0 = std::chrono::steady_clock::now();
        gpa_besteffort(ctx, avail, n, req, n_req, k, GPA_MODE_GO_FAITHFUL, out3, &n_out3);
        auto t1 = std::chrono::steady_clock::now();
        ref_besteffort_fast(n, W, nl, avail, n, req, n_req, k, 0, out2, &n_out2, nullptr);
        auto t2 = std::chrono::steady_clock::now();
        if (i >= 200) {
          tb.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
          c.push_back(std::chrono::duration<double, std::micro>(t2 - t1).count());
        }
        if (i % 2 == 0) {
          auto t3 = std::chrono::steady_clock::now();
          gpa_besteffort(ctx, avail, n, req, n_req, k, GPA_MODE_GO_FAITHFUL | GPA_MODE_FORCE_KERNEL, out, &n_out);
          auto t4 = std::chrono::steady_clock::now();
          if (i >= 200) g.push_back(std::chrono::duration<double, std::micro>(t4 - t3).count());
        }
      }
      bool same = n_out == n_out2 && std::equal(out, out + n_out, out2) && n_out3 == n_out2 && std::equal(out3, out3 + n_out3, out2);
      std::sort(tb.begin(), tb.end());
      std::sort(g.begin(), g.end());
      std::sort(c.begin(), c.end());
      printf("%s{\"k\": %d, \"required\": %d, \"table_p50\": %.3f, \"table_p99\": %.3f, \"kernel_p50\": %.1f, \"kernel_p99\": %.1f, "
             "\"oracle_p50\": %.3f, \"oracle_p99\": %.3f, \"same\": %s}",
             first ? "" : ", ", k, n_req, tb[tb.size() / 2], tb[(size_t)(tb.size() * 0.99)], g[g.size() / 2], g[(size_t)(g.size() * 0.99)],
             c[c.size() / 2], c[(size_t)(c.size() * 0.99)], same ? "true" : "false");
      first = false;
    }
  printf("],\n \"replica_set\": [");
  // replica-set search (gpa_subset_besteffort, count-vector kernel): 4 of the free replicas, every second replica free
  const int reps_of[3] = {4, 16, 32};
  for (int c = 0; c < 3; c++) {
    std::vector<int32_t> phys;
    for (int g = 0; g < n; g++)
      for (int r = 0; r < reps_of[c]; r += 2) phys.push_back(g);
    int32_t pos[8], pos2[8];
    std::vector<double> g, cpu;
    int n_out = 0, n_out2 = 0;
    const int cpu_every = c == 0 ? 1 : (c == 1 ? 20 : 200);  // the literal oracle walks C(free, 4) combinations
    for (int i = 0; i < 2200; i++) {
      auto t0 = std::chrono::steady_clock::now();
      gpa_subset_besteffort(ctx, phys.data(), (int)phys.size(), nullptr, 0, 4, 0, pos, &n_out);
      auto t1 = std::chrono::steady_clock::now();
      if (i >= 200) g.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
      if (i % cpu_every == 0) {
        auto t2 = std::chrono::steady_clock::now();
        ref_subset_besteffort(n, W, nl, phys.data(), (int)phys.size(), nullptr, 0, 4, pos2, &n_out2, nullptr, nullptr);
        auto t3 = std::chrono::steady_clock::now();
        cpu.push_back(std::chrono::duration<double, std::micro>(t3 - t2).count());
      }
    }
    bool same = n_out == n_out2 && std::equal(pos, pos + n_out, pos2);
    std::sort(g.begin(), g.end());
    std::sort(cpu.begin(), cpu.end());
    printf("%s{\"virtual_devices\": %d, \"free\": %d, \"size\": 4, \"gpu_p50\": %.1f, \"gpu_p99\": %.1f, \"oracle_p50\": %.1f, \"same\": %s}",
           c ? ", " : "", n * reps_of[c], (int)phys.size(), g[g.size() / 2], g[(size_t)(g.size() * 0.99)], cpu[cpu.size() / 2],
           same ? "true" : "false");
  }
  printf("]}\n");
  gpa_destroy(ctx);
  return 0;
}
