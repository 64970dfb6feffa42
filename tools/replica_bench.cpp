// Native timing of the replica-mode decision (prioritizeDevices, cmd/nvidia-device-plugin/replica.go:96-198):
// the product's host C++ (gpa_prioritize_devices) next to the oracle's C restatement (ref_prioritize_devices) on the
// same seeded streams, at 32 / 128 / 256 / 1464 virtual devices.  No GPU involved: this branch has no parallel work.
//   g++ -O2 -std=c++17 tools/replica_bench.cpp -o /tmp/replica_bench -L k8s-gpu-sharing-plugin_b200 -lgpualloc -L oracle -loracle
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

extern "C" {
int gpa_prioritize_devices(const char *const *avail, int n_avail, const char *const *must, int n_must, int size, char *outbuf,
                           int outbuf_len, int *n_out, int *non_unique, char *errbuf, int errbuf_len);
int ref_prioritize_devices(const char *const *availableDeviceIDs, int n_avail, const char *const *mustIncludeDeviceIDs, int n_must,
                           int allocationSize, char *outbuf, int outbuf_len, int *n_out, char *errbuf, int errbuf_len);
}

int main(int argc, char **argv) {
  const int config3_masks = argc > 1 ? atoi(argv[1]) : 10000;  // masks per (ids, p) of the config-3 sweep
  printf("{\"what\": \"prioritizeDevices decisions/s, one host thread, fragmented availability p=0.5, sizes {1,2,4,8}\", \"cases\": [");
  bool first = true;
  for (int replicas : {4, 16, 32, 183}) {
    std::mt19937 rng(3);
    std::vector<std::string> ids;
    for (int g = 0; g < 8; g++)
      for (int r = 0; r < replicas; r++) {
        char b[96];
        snprintf(b, sizeof b, "GPU-%08x-4820-abfc-e83e-94318197576e-replica-%d", 0x10000u * (unsigned)(g * 7919 % 65521), r);
        ids.push_back(b);
      }
    const int n_req = 2000;
    std::vector<std::vector<const char *>> avail(n_req);
    std::vector<int> sizes(n_req);
    for (int q = 0; q < n_req; q++) {
      for (auto &s : ids)
        if (rng() & 1) avail[q].push_back(s.c_str());
      std::shuffle(avail[q].begin(), avail[q].end(), rng);
      sizes[q] = 1 << (rng() % 4);
    }
    std::vector<char> out1(1 << 20), out2(1 << 20);
    char err[256];
    int mism = 0;
    double t_ours = 0, t_ref = 0;
    for (int rep = 0; rep < 3; rep++) {
      auto t0 = std::chrono::steady_clock::now();
      for (int q = 0; q < n_req; q++) {
        int n_out = 0, nu = 0;
        gpa_prioritize_devices(avail[q].data(), (int)avail[q].size(), nullptr, 0, sizes[q], out1.data(), (int)out1.size(), &n_out, &nu, err,
                               sizeof err);
      }
      auto t1 = std::chrono::steady_clock::now();
      for (int q = 0; q < n_req; q++) {
        int n_out = 0;
        ref_prioritize_devices(avail[q].data(), (int)avail[q].size(), nullptr, 0, sizes[q], out2.data(), (int)out2.size(), &n_out, err,
                               sizeof err);
      }
      auto t2 = std::chrono::steady_clock::now();
      t_ours = std::chrono::duration<double>(t1 - t0).count();
      t_ref = std::chrono::duration<double>(t2 - t1).count();
    }
    for (int q = 0; q < n_req; q++) {
      int n1 = 0, n2 = 0, nu = 0;
      int r1 = gpa_prioritize_devices(avail[q].data(), (int)avail[q].size(), nullptr, 0, sizes[q], out1.data(), (int)out1.size(), &n1, &nu,
                                      err, sizeof err);
      int r2 = ref_prioritize_devices(avail[q].data(), (int)avail[q].size(), nullptr, 0, sizes[q], out2.data(), (int)out2.size(), &n2, err,
                                      sizeof err);
      bool same = n1 == n2 && ((r1 == 0) == (r2 == 0 || r2 == 1));
      if (same && n1 > 0) {
        size_t len = 0;
        for (int i = 0; i < n1; i++) len += strlen(out1.data() + len) + 1;
        same = memcmp(out1.data(), out2.data(), len) == 0;
      }
      mism += !same;
    }
    printf("%s{\"virtual_devices\": %d, \"ours_per_s\": %.0f, \"oracle_per_s\": %.0f, \"mismatches\": %d}", first ? "" : ", ", 8 * replicas,
           n_req / t_ours, n_req / t_ref, mism);
    first = false;
  }
  // SURVEY.md 8d config 3 as specified: 8 x 16 = 128 and 8 x 32 = 256 ids, every replica free independently with
  // p in {0.1, 0.5, 0.9}, seed 3, 10 000 masks x size in {1, 2, 3, 4, 8, 16}; every decision compared with the oracle
  printf("], \"config3\": [");
  first = true;
  int total_mism = 0;
  for (int replicas : {16, 32})
    for (int pi = 0; pi < 3; pi++) {
      const double p_free = pi == 0 ? 0.1 : (pi == 1 ? 0.5 : 0.9);
      std::mt19937 rng(3);
      std::uniform_real_distribution<double> uni(0.0, 1.0);
      std::vector<std::string> ids;
      for (int g = 0; g < 8; g++)
        for (int r = 0; r < replicas; r++) {
          char b[96];
          snprintf(b, sizeof b, "GPU-%08x-4820-abfc-e83e-94318197576e-replica-%d", 0x10000u * (unsigned)(g * 7919 % 65521), r);
          ids.push_back(b);
        }
      const int n_masks = config3_masks;
      const int sizes[6] = {1, 2, 3, 4, 8, 16};
      std::vector<char> out1(1 << 16), out2(1 << 16);
      char err[256];
      long decisions = 0, mism = 0, refused = 0;
      double t_ours = 0, t_ref = 0;
      std::vector<const char *> avail;
      for (int q = 0; q < n_masks; q++) {
        avail.clear();
        for (auto &s : ids)
          if (uni(rng) < p_free) avail.push_back(s.c_str());
        std::shuffle(avail.begin(), avail.end(), rng);
        for (int size : sizes) {
          int n1 = 0, n2 = 0, nu = 0;
          auto t0 = std::chrono::steady_clock::now();
          const int r1 = gpa_prioritize_devices(avail.data(), (int)avail.size(), nullptr, 0, size, out1.data(), (int)out1.size(), &n1, &nu, err,
                                                sizeof err);
          auto t1 = std::chrono::steady_clock::now();
          const int r2 = ref_prioritize_devices(avail.data(), (int)avail.size(), nullptr, 0, size, out2.data(), (int)out2.size(), &n2, err,
                                                sizeof err);
          auto t2 = std::chrono::steady_clock::now();
          t_ours += std::chrono::duration<double>(t1 - t0).count();
          t_ref += std::chrono::duration<double>(t2 - t1).count();
          bool same = ((r1 == 0) == (r2 == 0 || r2 == 1)) && (r1 != 0 || n1 == n2) && (r1 != 0 || (nu != 0) == (r2 == 1));
          if (same && r1 == 0 && n1 > 0) {
            size_t len = 0;
            for (int i = 0; i < n1; i++) len += strlen(out1.data() + len) + 1;
            same = memcmp(out1.data(), out2.data(), len) == 0;
          }
          mism += !same;
          refused += r1 != 0;
          decisions++;
        }
      }
      total_mism += (int)mism;
      printf("%s{\"virtual_devices\": %d, \"p_free\": %.1f, \"decisions\": %ld, \"no_devices_left\": %ld, \"ours_per_s\": %.0f, "
             "\"oracle_per_s\": %.0f, \"mismatches\": %ld}",
             first ? "" : ", ", 8 * replicas, p_free, decisions, refused, decisions / t_ours, decisions / t_ref, mism);
      first = false;
    }
  printf("]}\n");
  return total_mism == 0 ? 0 : 1;
}
