// What the reference pays in NVML calls for ONE GetPreferredAllocation container request, timed on the real driver:
//   cmd/nvidia-device-plugin/server.go:274,279 call gpuallocator.NewDevicesFrom twice, and each one re-enumerates the
//   whole node (vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/device.go:33-72):
//     nvml.GetDeviceCount, then per device nvml.NewDevice (nvml.go:359-447: handle by index + 12 attribute queries and
//     one sysfs read), then per ORDERED pair nvml.GetP2PLink (nvml.go:592-615: one common-ancestor query) and
//     nvml.GetNVLink (nvml.go:617-658 -> bindings.go:358-386: the state of links 0..11 and the remote PCI info of
//     every active one).
// This tool issues exactly that call sequence against libnvidia-ml.so.1 (dlopen, like nvml_dl.go:31) `reps` times and
// times it, next to what this library does instead: gpa_nvml_read_topology() once at start-up.
//   usage: nvml_reenum [reps=30] [libnvidia-ml path]
#include <dlfcn.h>
#include <nvml.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "gpualloc.h"

namespace {

struct Nvml {
  void *so = nullptr;
  unsigned long long calls = 0;
#define SYM(name) decltype(&::name) p_##name = nullptr
  SYM(nvmlInit_v2);
  SYM(nvmlShutdown);
  SYM(nvmlDeviceGetCount_v2);
  SYM(nvmlDeviceGetHandleByIndex_v2);
  SYM(nvmlDeviceGetName);
  SYM(nvmlDeviceGetUUID);
  SYM(nvmlDeviceGetMinorNumber);
  SYM(nvmlDeviceGetPowerManagementLimit);
  SYM(nvmlDeviceGetMemoryInfo);
  SYM(nvmlDeviceGetPciInfo_v3);
  SYM(nvmlDeviceGetBAR1MemoryInfo);
  SYM(nvmlDeviceGetMaxPcieLinkGeneration);
  SYM(nvmlDeviceGetMaxPcieLinkWidth);
  SYM(nvmlDeviceGetMaxClockInfo);
  SYM(nvmlDeviceGetCudaComputeCapability);
  SYM(nvmlDeviceGetTopologyCommonAncestor);
  SYM(nvmlDeviceGetNvLinkState);
  SYM(nvmlDeviceGetNvLinkRemotePciInfo_v2);
#undef SYM
  bool load(const char *path) {
    so = dlopen(path, RTLD_NOW);
    if (!so) return false;
#define GET(name)                                         \
  p_##name = (decltype(p_##name))dlsym(so, #name);       \
  if (!p_##name) {                                        \
    fprintf(stderr, "missing symbol %s\n", #name);        \
    return false;                                         \
  }
    GET(nvmlInit_v2);
    GET(nvmlShutdown);
    GET(nvmlDeviceGetCount_v2);
    GET(nvmlDeviceGetHandleByIndex_v2);
    GET(nvmlDeviceGetName);
    GET(nvmlDeviceGetUUID);
    GET(nvmlDeviceGetMinorNumber);
    GET(nvmlDeviceGetPowerManagementLimit);
    GET(nvmlDeviceGetMemoryInfo);
    GET(nvmlDeviceGetPciInfo_v3);
    GET(nvmlDeviceGetBAR1MemoryInfo);
    GET(nvmlDeviceGetMaxPcieLinkGeneration);
    GET(nvmlDeviceGetMaxPcieLinkWidth);
    GET(nvmlDeviceGetMaxClockInfo);
    GET(nvmlDeviceGetCudaComputeCapability);
    GET(nvmlDeviceGetTopologyCommonAncestor);
    GET(nvmlDeviceGetNvLinkState);
    GET(nvmlDeviceGetNvLinkRemotePciInfo_v2);
#undef GET
    return true;
  }
};

struct Dev {
  nvmlDevice_t h;
  std::string bus;
};

// gpuallocator.NewDevices(): returns the number of (pair, link) entries it would have recorded
int new_devices(Nvml &nv, int *n_out) {
  unsigned int count = 0;
  nv.calls++;
  if (nv.p_nvmlDeviceGetCount_v2(&count) != NVML_SUCCESS) return -1;
  std::vector<Dev> devs(count);
  for (unsigned int i = 0; i < count; i++) {
    // nvml.NewDevice(i): nvml.go:359-447
    nvmlDevice_t h;
    nv.calls++;
    if (nv.p_nvmlDeviceGetHandleByIndex_v2(i, &h) != NVML_SUCCESS) return -1;
    char name[NVML_DEVICE_NAME_BUFFER_SIZE], uuid[NVML_DEVICE_UUID_BUFFER_SIZE];
    unsigned int u = 0, v = 0;
    int maj = 0, min = 0;
    nvmlMemory_t mem;
    nvmlPciInfo_t pci;
    nvmlBAR1Memory_t bar1;
    nv.p_nvmlDeviceGetName(h, name, sizeof name);
    nv.p_nvmlDeviceGetUUID(h, uuid, sizeof uuid);
    nv.p_nvmlDeviceGetMinorNumber(h, &u);
    nv.p_nvmlDeviceGetPowerManagementLimit(h, &u);
    nv.p_nvmlDeviceGetMemoryInfo(h, &mem);
    nv.p_nvmlDeviceGetPciInfo_v3(h, &pci);
    nv.p_nvmlDeviceGetBAR1MemoryInfo(h, &bar1);
    nv.p_nvmlDeviceGetMaxPcieLinkGeneration(h, &u);
    nv.p_nvmlDeviceGetMaxPcieLinkWidth(h, &v);
    nv.p_nvmlDeviceGetMaxClockInfo(h, NVML_CLOCK_SM, &u);
    nv.p_nvmlDeviceGetMaxClockInfo(h, NVML_CLOCK_MEM, &v);
    nv.p_nvmlDeviceGetCudaComputeCapability(h, &maj, &min);
    nv.calls += 12;
    // numaNode(): nvml.go:314-332
    std::string path = std::string("/sys/bus/pci/devices/") + (pci.busId + 4) + "/numa_node";
    for (auto &c : path) c = (char)tolower(c);
    if (FILE *f = fopen(path.c_str(), "r")) {
      char b[32];
      if (fgets(b, sizeof b, f)) (void)atoi(b);
      fclose(f);
    }
    devs[i] = {h, pci.busId};
  }
  int entries = 0;
  for (unsigned int i = 0; i < count; i++)
    for (unsigned int j = 0; j < count; j++) {
      if (i == j) continue;
      // nvml.GetP2PLink: nvml.go:592-615
      nvmlGpuTopologyLevel_t level;
      nv.calls++;
      if (nv.p_nvmlDeviceGetTopologyCommonAncestor(devs[i].h, devs[j].h, &level) == NVML_SUCCESS) entries++;
      // nvml.GetNVLink -> deviceGetAllNvLinkRemotePciInfo: links 0..11 (NVML_NVLINK_MAX_LINKS of the vendored header)
      int links = 0;
      for (unsigned int l = 0; l < 12; l++) {
        nvmlEnableState_t st;
        nv.calls++;
        const nvmlReturn_t r = nv.p_nvmlDeviceGetNvLinkState(devs[i].h, l, &st);
        if (r != NVML_SUCCESS) continue;  // NOT_SUPPORTED / INVALID_ARGUMENT are "no such link" (bindings.go:340-343)
        if (st == NVML_FEATURE_ENABLED) {
          nvmlPciInfo_t rp;
          nv.calls++;
          if (nv.p_nvmlDeviceGetNvLinkRemotePciInfo_v2(devs[i].h, l, &rp) == NVML_SUCCESS && devs[j].bus == rp.busId) links++;
        }
      }
      if (links) entries++;
    }
  *n_out = (int)count;
  return entries;
}

double pct(std::vector<double> v, double p) {
  std::sort(v.begin(), v.end());
  return v[(size_t)((v.size() - 1) * p)];
}

}  // namespace

int main(int argc, char **argv) {
  const int reps = argc > 1 ? atoi(argv[1]) : 30;
  const char *path = argc > 2 ? argv[2] : "libnvidia-ml.so.1";
  Nvml nv;
  if (!nv.load(path)) {
    printf("{\"error\": \"cannot load %s\"}\n", path);
    return 2;
  }
  if (nv.p_nvmlInit_v2() != NVML_SUCCESS) {
    printf("{\"error\": \"nvmlInit_v2 failed\"}\n");
    return 2;
  }
  int n = 0, entries = 0;
  new_devices(nv, &n);  // warm
  std::vector<double> per_request_ms;
  unsigned long long calls_per_request = 0;
  for (int r = 0; r < reps; r++) {
    const unsigned long long c0 = nv.calls;
    auto t0 = std::chrono::steady_clock::now();
    entries = new_devices(nv, &n);  // NewDevicesFrom(available), server.go:274
    new_devices(nv, &n);            // NewDevicesFrom(required), server.go:279
    auto t1 = std::chrono::steady_clock::now();
    per_request_ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
    calls_per_request = nv.calls - c0;
  }
  // what this library does: one read at start-up (and again when a device goes unhealthy)
  gpa_nvml *g = nullptr;
  std::vector<double> ours_ms;
  unsigned long long our_calls = 0;
  if (gpa_nvml_open(argc > 2 ? path : nullptr, &g) == GPA_OK) {
    std::vector<int32_t> links((size_t)n * n * 2 + 2);
    for (int r = 0; r < reps; r++) {
      int nn = 0;
      auto t0 = std::chrono::steady_clock::now();
      gpa_nvml_read_topology(g, GPA_TOPO_PARITY, &nn, links.data(), (int)links.size(), &our_calls);
      auto t1 = std::chrono::steady_clock::now();
      ours_ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
    }
    gpa_nvml_close(g);
  }
  printf("{\"what\": \"NVML cost of one GetPreferredAllocation container request in the reference (2 x gpuallocator.NewDevices, "
         "device.go:33-72, the exact call sequence) vs one gpa_nvml_read_topology() at start-up here; real libnvidia-ml\", "
         "\"gpus\": %d, \"reps\": %d, \"reference_nvml_calls_per_request\": %llu, \"reference_link_entries\": %d, "
         "\"reference_ms_per_request\": {\"p50\": %.3f, \"p99\": %.3f, \"min\": %.3f}, \"ours_nvml_calls_once\": %llu, "
         "\"ours_ms_once\": {\"p50\": %.3f, \"p99\": %.3f, \"min\": %.3f}}\n",
         n, reps, calls_per_request, entries, pct(per_request_ms, 0.5), pct(per_request_ms, 0.99), pct(per_request_ms, 0.0),
         our_calls, ours_ms.empty() ? 0.0 : pct(ours_ms, 0.5), ours_ms.empty() ? 0.0 : pct(ours_ms, 0.99),
         ours_ms.empty() ? 0.0 : pct(ours_ms, 0.0));
  nv.p_nvmlShutdown();
  return 0;
}
