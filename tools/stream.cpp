// BASELINE config 2, natively: 8 GPUs x 4 replicas = 32 virtual devices, a STATEFUL stream of 100 000
// steps, seed 2: with p = 0.6 GetPreferredAllocation(size in {1,2,4,8}, available = the current free set in shuffled
// order, must = one free replica with p = 0.1) followed by Allocate of the answer; with p = 0.4 an earlier allocation is
// freed.  With replicas > 1 the reference answers through prioritizeDevices (server.go:285-296), host code here too, so no
// GPU is involved.  Every preferred allocation is compared with the oracle (ref_prioritize_devices).  One JSON line:
// decisions/s, call latency percentiles of GetPreferredAllocation and Allocate, mismatches.
//   g++ -O2 -std=c++17 tools/stream.cpp -o /tmp/stream -I include -L k8s-gpu-sharing-plugin_b200 -lgpualloc -L oracle -loracle
//   /tmp/stream [replicas] [steps]
// replicas = 1 (8 virtual devices) takes the OTHER branch of the handler (server.go:297-301): the best-effort policy on the
// GPU context - answered from the host copy of the answer table the GPU built - compared with ref_besteffort_fast.  That
// branch only closes its loop with two of the reference's defects switched off (GPA_FIX_REPLICA_SUFFIX: the answer carries the
// "-replica-0" suffix Allocate knows; GPA_FIX_REQUIRED_IDENTITY: a must-include device is honoured, D4), so
// the stream runs with those two fixes and says so.  Needs a CUDA device.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "gpualloc.h"

extern "C" int ref_besteffort_fast(int nt, const int32_t *W, const uint8_t *nlinks, const int32_t *avail, int n_avail, const int32_t *req,
                                   int n_req, int size, int mode, int32_t *out, int *n_out, unsigned long long *visited);
extern "C" int ref_prioritize_devices(const char *const *availableDeviceIDs, int n_avail, const char *const *mustIncludeDeviceIDs,
                                      int n_must, int allocationSize, char *outbuf, int outbuf_len, int *n_out, char *errbuf,
                                      int errbuf_len);

static std::vector<std::string> unpack(const char *buf, int n) {
  std::vector<std::string> out;
  const char *p = buf;
  for (int i = 0; i < n; i++) {
    out.emplace_back(p);
    p += out.back().size() + 1;
  }
  return out;
}

static double pct(std::vector<float> &v, double q) {
  if (v.empty()) return 0;
  std::sort(v.begin(), v.end());
  return v[(size_t)(q * (v.size() - 1))];
}

int main(int argc, char **argv) {
  const int n = 8;
  const int replicas = argc > 1 ? atoi(argv[1]) : 4;
  const long steps = argc > 2 ? atol(argv[2]) : 100000;
  std::vector<std::string> uuids, paths(n), index(n);
  std::vector<const char *> path_ptrs(n);
  std::vector<gpa_plugin_device> devs(n);
  for (int i = 0; i < n; i++) {
    char b[64];
    snprintf(b, sizeof b, "GPU-%08x-4820-abfc-e83e-94318197576e", 0x1000u * (unsigned)(i * 7919 + 13));
    uuids.push_back(b);
  }
  for (int i = 0; i < n; i++) {
    paths[i] = "/dev/nvidia" + std::to_string(i);
    path_ptrs[i] = paths[i].c_str();
    index[i] = std::to_string(i);
    devs[i] = {uuids[i].c_str(), index[i].c_str(), &path_ptrs[i], 1, 183359ull, i / 4, i};
  }
  const bool best_effort = replicas == 1;
  gpa_ctx *ctx = nullptr;
  int32_t W[64];
  uint8_t nl[64];
  if (best_effort) {
    // synthetic hybrid cube mesh (synthetic.topo_dgx1v_like): PCIe + NVLink weights, as tools/latency.cpp
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
    if (gpa_create(0, &ctx) != GPA_OK) return 2;
    if (gpa_set_topology(ctx, n, W, nl) != GPA_OK) return 2;
  }
  gpa_plugin_config cfg = {"nvidia.com/sharedgpu", "NVIDIA_VISIBLE_DEVICES", "/", (uint32_t)replicas, 0, GPA_LIST_STRATEGY_ENVVAR,
                           GPA_ID_STRATEGY_UUID, 0, 1, best_effort ? (GPA_FIX_REPLICA_SUFFIX | GPA_FIX_REQUIRED_IDENTITY) : 0};
  gpa_plugin *plugin = nullptr;
  if (gpa_plugin_create(ctx, &cfg, devs.data(), n, &plugin) != GPA_OK) return 3;
  const unsigned long long launches0 = ctx ? gpa_launch_count(ctx) : 0;
  // the oracle of the best-effort branch: devices in sorted-uuid order (stripReplicas, replica.go:43), answer as ids
  std::vector<int> by_uuid(n);
  for (int i = 0; i < n; i++) by_uuid[i] = i;
  std::sort(by_uuid.begin(), by_uuid.end(), [&](int a, int b) { return uuids[a] < uuids[b]; });
  auto oracle_besteffort = [&](const std::vector<const char *> &avail, const std::vector<const char *> &must, int size, char *outbuf,
                               int *n_out) -> int {
    auto strip = [](const char *id) { std::string s(id); const size_t c = s.find("-replica-"); return c == std::string::npos ? s : s.substr(0, c); };
    std::vector<std::string> au, mu;
    for (auto *a : avail) au.push_back(strip(a));
    for (auto *m : must) mu.push_back(strip(m));
    int32_t av[8], rq[8], out[16];
    int na = 0, nr = 0, cnt = 0;
    for (int d : by_uuid)
      if (std::find(au.begin(), au.end(), uuids[d]) != au.end()) av[na++] = d;
    for (int d : by_uuid)
      if (std::find(mu.begin(), mu.end(), uuids[d]) != mu.end()) rq[nr++] = d;
    const int rc = ref_besteffort_fast(n, W, nl, av, na, rq, nr, size, 0, out, &cnt, nullptr);
    char *o = outbuf;
    for (int i = 0; i < cnt; i++) {
      if (out[i] < 0) return 99;  // a nil in the answer: the Go handler dies on device.UUID (server.go:299-301)
      o += sprintf(o, "%s-replica-0", uuids[out[i]].c_str()) + 1;
    }
    *n_out = cnt;
    return rc;
  };
  std::vector<std::string> free_ids;
  for (auto &u : uuids)
    for (int r = 0; r < replicas; r++) free_ids.push_back(u + "-replica-" + std::to_string(r));
  const size_t n_virtual = free_ids.size();

  std::mt19937 rng(2);
  std::uniform_real_distribution<double> uni(0.0, 1.0);
  std::vector<std::vector<std::string>> held;
  std::vector<float> lat_pref, lat_alloc;
  lat_pref.reserve((size_t)steps);
  lat_alloc.reserve((size_t)steps);
  std::vector<char> out((size_t)n_virtual * 96), out2((size_t)n_virtual * 96), json(1 << 16);
  char err[512], err2[512];
  long decisions = 0, allocs = 0, frees = 0, mismatches = 0, errors = 0;
  double t_pref = 0, t_ref = 0;
  for (long step = 0; step < steps; step++) {
    if (uni(rng) < 0.6 && !free_ids.empty()) {
      const int size = 1 << (rng() % 4);
      std::vector<const char *> avail;
      for (auto &s : free_ids) avail.push_back(s.c_str());
      std::shuffle(avail.begin(), avail.end(), rng);
      std::vector<const char *> must;
      if (uni(rng) < 0.1) must.push_back(avail[0]);
      int n_out = 0, n_out2 = 0;
      auto t0 = std::chrono::steady_clock::now();
      const int rc = gpa_plugin_get_preferred_allocation(plugin, avail.data(), (int)avail.size(), must.data(), (int)must.size(), size,
                                                         out.data(), (int)out.size(), &n_out, err, sizeof err);
      auto t1 = std::chrono::steady_clock::now();
      const int rc2 = best_effort ? oracle_besteffort(avail, must, size, out2.data(), &n_out2)
                                  : ref_prioritize_devices(avail.data(), (int)avail.size(), must.data(), (int)must.size(), size, out2.data(),
                                                           (int)out2.size(), &n_out2, err2, sizeof err2);
      auto t2 = std::chrono::steady_clock::now();
      t_pref += std::chrono::duration<double>(t1 - t0).count();
      t_ref += std::chrono::duration<double>(t2 - t1).count();
      lat_pref.push_back((float)std::chrono::duration<double, std::micro>(t1 - t0).count());
      decisions++;
      // oracle kinds: 0 none, 1 non-unique (ids still returned), 2 no devices left, 3 must missing, 4 panic
      const bool ok = rc == GPA_OK, ok2 = rc2 == 0 || rc2 == 1;
      std::vector<std::string> got = ok ? unpack(out.data(), n_out) : std::vector<std::string>();
      std::vector<std::string> want = ok2 ? unpack(out2.data(), n_out2) : std::vector<std::string>();
      if (ok != ok2 || got != want) mismatches++;
      if (!ok) {
        errors++;  // "no devices left to allocate" when fewer than `size` replicas are free: kubelet would not ask
        continue;
      }
      if (got.empty()) continue;
      std::vector<const char *> ids;
      for (auto &s : got) ids.push_back(s.c_str());
      auto t3 = std::chrono::steady_clock::now();
      const int rca = gpa_plugin_allocate(plugin, ids.data(), (int)ids.size(), json.data(), (int)json.size(), err, sizeof err);
      auto t4 = std::chrono::steady_clock::now();
      lat_alloc.push_back((float)std::chrono::duration<double, std::micro>(t4 - t3).count());
      if (rca != GPA_OK) errors++;
      allocs++;
      for (auto &s : got) free_ids.erase(std::find(free_ids.begin(), free_ids.end(), s));
      held.push_back(got);
    } else if (!held.empty()) {
      const size_t at = rng() % held.size();
      for (auto &s : held[at]) free_ids.push_back(s);
      held.erase(held.begin() + (long)at);
      frees++;
    }
  }
  printf("{\"what\": \"config 2: stateful GetPreferredAllocation -> Allocate / free stream, seed 2, host C++ behind the C ABI, one "
         "thread\", \"branch\": \"%s\", \"kernel_launches_during_the_stream\": %llu, \"virtual_devices\": %zu, \"steps\": %ld, \"decisions\": %ld, \"allocations\": %ld, \"frees\": %ld, "
         "\"request_errors_no_devices_left\": %ld, \"mismatches_vs_oracle\": %ld, \"decisions_per_s\": %.0f, "
         "\"oracle_decision_only_per_s\": %.0f, \"preferred_us\": {\"p50\": %.2f, \"p99\": %.2f}, \"allocate_us\": {\"p50\": %.2f, "
         "\"p99\": %.2f}}\n",
         best_effort ? "best-effort policy (replicas = 1), answers from the host copy of the GPU-built answer table; fixes D3 + D4 on, "
                       "oracle = ref_besteffort_fast"
                     : "prioritizeDevices (replicas > 1), oracle = ref_prioritize_devices",
         ctx ? gpa_launch_count(ctx) - launches0 : 0ull, n_virtual, steps, decisions, allocs, frees, errors, mismatches, decisions / t_pref, decisions / t_ref, pct(lat_pref, 0.5),
         pct(lat_pref, 0.99), pct(lat_alloc, 0.5), pct(lat_alloc, 0.99));
  gpa_plugin_destroy(plugin);
  if (ctx) gpa_destroy(ctx);
  return mismatches == 0 ? 0 : 1;
}
