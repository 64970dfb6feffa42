// BASELINE config 4, natively: health-check churn under concurrent Allocate load.
// 8 client threads call gpa_plugin_allocate back to back while an injector posts XID events at 50 Hz into a scripted
// NVML (tests/csrc/fake_nvml.c) and the watcher thread turns them into ListAndWatch updates.  Prints one JSON line:
// Allocate latency percentiles, error count, events seen / skipped, devices marked unhealthy, re-sent list entries.
//   g++ -O2 -std=c++17 -pthread tools/churn.cpp -o /tmp/churn -I include -L k8s-gpu-sharing-plugin_b200 -lgpualloc -ldl
//   /tmp/churn tests/csrc/libfake-nvidia-ml.so.1 [seconds]
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "gpualloc.h"

int main(int argc, char **argv) {
  if (argc < 2) {
    fprintf(stderr, "usage: churn <fake nvml .so> [seconds] [parity|fixed]\n");
    return 2;
  }
  const double seconds = argc > 2 ? atof(argv[2]) : 5.0;
  // parity: the reference's behaviour (the re-sent list never changes); fixed: GPA_FIX_HEALTH_PROPAGATION
  const bool parity = argc > 3 && std::string(argv[3]) == "parity";
  void *fake = dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL);
  if (!fake) {
    fprintf(stderr, "%s\n", dlerror());
    return 2;
  }
  auto reset = (void (*)(int))dlsym(fake, "fake_nvml_reset");
  auto set_device = (void (*)(int, const char *, const char *))dlsym(fake, "fake_nvml_set_device");
  auto inject = (void (*)(int, unsigned long long, unsigned long long))dlsym(fake, "fake_nvml_inject");
  const int n = 8, replicas = 4;
  reset(n);
  std::vector<std::string> uuids;
  for (int i = 0; i < n; i++) {
    char b[64];
    snprintf(b, sizeof b, "GPU-%08x-4820-abfc-e83e-94318197576e", 0x1000u * (unsigned)(i * 7919 + 13));
    uuids.push_back(b);
    set_device(i, b, nullptr);
  }
  gpa_nvml *nv = nullptr;
  if (gpa_nvml_open(argv[1], &nv) != GPA_OK) return 3;
  std::vector<const char *> ids;
  for (auto &u : uuids) ids.push_back(u.c_str());
  gpa_health *health = nullptr;
  if (gpa_health_open(nv, ids.data(), n, "", 20, &health) != GPA_OK) return 4;

  std::vector<std::string> paths(n);
  std::vector<const char *> path_ptrs(n);
  std::vector<std::string> index(n);
  std::vector<gpa_plugin_device> devs(n);
  for (int i = 0; i < n; i++) {
    paths[i] = "/dev/nvidia" + std::to_string(i);
    path_ptrs[i] = paths[i].c_str();
    index[i] = std::to_string(i);
    devs[i] = {uuids[i].c_str(), index[i].c_str(), &path_ptrs[i], 1, 183359ull, i / 4, i};
  }
  gpa_plugin_config cfg = {"nvidia.com/sharedgpu", "NVIDIA_VISIBLE_DEVICES", "/", (uint32_t)replicas, 0, GPA_LIST_STRATEGY_ENVVAR,
                           GPA_ID_STRATEGY_UUID, 0, 1, parity ? 0 : GPA_FIX_HEALTH_PROPAGATION};
  gpa_plugin *plugin = nullptr;
  if (gpa_plugin_create(nullptr, &cfg, devs.data(), n, &plugin) != GPA_OK) return 5;
  std::vector<std::string> replica_ids;
  for (auto &u : uuids)
    for (int r = 0; r < replicas; r++) replica_ids.push_back(u + "-replica-" + std::to_string(r));

  std::atomic<bool> stop{false};
  std::atomic<long> errors{0};
  std::atomic<unsigned long long> n_calls{0};
  const int n_clients = 8;
  std::vector<std::vector<float>> lat(n_clients);
  std::vector<std::thread> clients;
  for (int c = 0; c < n_clients; c++)
    clients.emplace_back([&, c] {
      std::mt19937 rng(100 + c);
      char json[4096], err[256];
      lat[c].reserve(1 << 22);
      unsigned long calls_here = 0;
      while (!stop.load(std::memory_order_relaxed)) {
        const char *req[2] = {replica_ids[rng() % replica_ids.size()].c_str(), replica_ids[rng() % replica_ids.size()].c_str()};
        auto t0 = std::chrono::steady_clock::now();
        int rc = gpa_plugin_allocate(plugin, req, 2, json, sizeof json, err, sizeof err);
        auto t1 = std::chrono::steady_clock::now();
        if (rc != GPA_OK) errors++;
        n_calls.fetch_add(1, std::memory_order_relaxed);
        if ((calls_here++ & 3) == 0) lat[c].push_back(std::chrono::duration<float, std::micro>(t1 - t0).count());  // every 4th call
      }
    });
  long unhealthy_events = 0, list_entries_changed = 0, list_resends = 0;
  std::thread watcher([&] {
    char *json = (char *)malloc(1 << 16);
    while (!stop.load()) {
      int dev = -1;
      unsigned long long xid = 0;
      if (gpa_health_poll(health, 50, &dev, &xid) == 1) {
        unhealthy_events++;
        list_entries_changed += gpa_plugin_mark_unhealthy(plugin, dev);
        gpa_plugin_list_devices(plugin, json, 1 << 16);  // the ListAndWatch re-send (server.go:262)
        list_resends++;
      }
    }
    free(json);
  });
  std::mt19937 rng(4);
  const unsigned long long xids[] = {13, 31, 43, 45, 68, 48, 63, 79, 94};
  long injected = 0, expected_notifications = 0;
  unsigned gone = 0;  // GPUs that a critical XID has reached so far
  auto t_end = std::chrono::steady_clock::now() + std::chrono::duration<double>(seconds);
  while (std::chrono::steady_clock::now() < t_end) {
    int dev = (rng() % 20 == 0) ? -1 : (int)(rng() % n);
    const unsigned long long xid = xids[rng() % 9];
    inject(dev, 0x8ull, xid);
    injected++;
    if (xid != 13 && xid != 31 && xid != 43 && xid != 45 && xid != 68) {  // the application-error XIDs are skipped (nvidia.go:196-215)
      expected_notifications += dev < 0 ? n : 1;                          // an event without a UUID marks every device (:243-250)
      gone |= dev < 0 ? (1u << n) - 1u : 1u << dev;
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(20));  // 50 Hz
  }
  std::this_thread::sleep_for(std::chrono::milliseconds(200));
  stop.store(true);
  for (auto &t : clients) t.join();
  watcher.join();
  unsigned long long seen = 0, skipped = 0;
  int disabled = 0;
  gpa_health_stats(health, &seen, &skipped, &disabled);
  std::vector<float> all;
  for (auto &v : lat) all.insert(all.end(), v.begin(), v.end());
  std::sort(all.begin(), all.end());
  auto pct = [&](double p) { return all.empty() ? 0.f : all[(size_t)(p * (all.size() - 1))]; };
  const long expected_changed = parity ? 0 : (long)__builtin_popcount(gone) * replicas;
  const bool as_specified = unhealthy_events == expected_notifications && list_resends == expected_notifications &&
                            list_entries_changed == expected_changed;
  printf("{\"config\": \"4: XID churn at 50 Hz under %d Allocate clients, %d virtual devices, scripted NVML\", \"mode\": \"%s\", "
         "\"seconds\": %.1f, \"allocate_calls\": %llu, \"allocate_per_s\": %.0f, \"allocate_us_p50\": %.2f, \"allocate_us_p99\": %.2f, \"allocate_us_p999\": %.2f, "
         "\"allocate_errors\": %ld, \"xids_injected\": %ld, \"events_seen\": %llu, \"events_skipped\": %llu, \"unhealthy_notifications\": %ld, "
         "\"list_resends\": %ld, \"list_entries_changed\": %ld, \"expected_notifications\": %ld, \"expected_entries_changed\": %ld, "
         "\"as_specified\": %s}\n",
         n_clients, n * replicas, parity ? "parity (D1 kept: the re-sent list never changes)" : "fixed (GPA_FIX_HEALTH_PROPAGATION)", seconds,
         n_calls.load(), n_calls.load() / seconds, pct(0.5), pct(0.99), pct(0.999), errors.load(), injected, seen,
         skipped, unhealthy_events, list_resends, list_entries_changed, expected_notifications, expected_changed,
         as_specified ? "true" : "false");
  gpa_health_close(health);
  gpa_plugin_destroy(plugin);
  gpa_nvml_close(nv);
  return errors.load() == 0 && as_specified ? 0 : 1;
}
