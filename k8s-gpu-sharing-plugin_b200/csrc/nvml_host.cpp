// NVML side of the path, host C++: the link matrix the allocator scores against, and the XID
// health watcher that feeds ListAndWatch.
//
// ref (topology): gpuallocator.NewDevices, vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/device.go:33-72
//                 nvml.GetP2PLink / nvml.GetNVLink, vendor/github.com/NVIDIA/gpu-monitoring-tools/bindings/go/nvml/
//                 nvml.go:592-658; deviceGetAllNvLinkRemotePciInfo, .../nvml/bindings.go:358-386
// ref (health):   checkHealth, cmd/nvidia-device-plugin/nvidia.go:181-269; nvml.WaitForEvent,
//                 .../nvml/bindings.go:174-193; RegisterEventForDevice, .../nvml/bindings.go:137-168
//
// The reference rebuilds the whole device list, ~1500 NVML calls, TWICE inside every
// GetPreferredAllocation request (cmd/nvidia-device-plugin/server.go:274,279) although the
// links never change while the daemon runs.  Here the matrix is read once and handed to
// gpa_set_topology_links(); gpa_nvml_read_topology reports how many NVML calls one build
// costs so that the saving can be quoted.
//
// NVML is reached through dlopen, like the reference does (nvml/nvml_dl.go:30-37), so the
// library loads on machines without a driver and tests can point it at a stand-in.
#include <dlfcn.h>
#include <nvml.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_set>
#include <vector>

#include "../../include/gpualloc.h"

extern "C" int gpa_get_additional_xids(const char *input, uint64_t *out, int out_cap);

namespace {

// nvml.P2PLinkType (nvml.go:132-154)
enum LinkType {
  kUnknown = 0, kCrossCPU = 1, kSameCPU = 2, kHostBridge = 3, kMultiSwitch = 4, kSingleSwitch = 5, kSameBoard = 6,
  kOneNVLink = 7  // ... kOneNVLink + 11 = TwelveNVLINKLinks
};

constexpr unsigned kRefMaxLinks = 12;  // NVML_NVLINK_MAX_LINKS of the reference's nvml.h (:273); B200 has 18

template <typename F>
bool load_sym(void *h, const char *name, F &out) {
  out = reinterpret_cast<F>(dlsym(h, name));
  return out != nullptr;
}

}  // namespace

struct gpa_nvml {
  void *handle = nullptr;
  std::string last_error;
  std::mutex mu;
  unsigned long long calls = 0;  // NVML entry points invoked (to quantify what caching saves)

  nvmlReturn_t (*Init)() = nullptr;
  nvmlReturn_t (*Shutdown)() = nullptr;
  const char *(*ErrorString)(nvmlReturn_t) = nullptr;
  nvmlReturn_t (*DeviceGetCount)(unsigned int *) = nullptr;
  nvmlReturn_t (*DeviceGetHandleByIndex)(unsigned int, nvmlDevice_t *) = nullptr;
  nvmlReturn_t (*DeviceGetUUID)(nvmlDevice_t, char *, unsigned int) = nullptr;
  nvmlReturn_t (*DeviceGetPciInfo)(nvmlDevice_t, nvmlPciInfo_t *) = nullptr;
  nvmlReturn_t (*DeviceGetMemoryInfo)(nvmlDevice_t, nvmlMemory_t *) = nullptr;
  nvmlReturn_t (*DeviceGetMinorNumber)(nvmlDevice_t, unsigned int *) = nullptr;
  nvmlReturn_t (*DeviceGetTopologyCommonAncestor)(nvmlDevice_t, nvmlDevice_t, nvmlGpuTopologyLevel_t *) = nullptr;
  nvmlReturn_t (*DeviceGetNvLinkState)(nvmlDevice_t, unsigned int, nvmlEnableState_t *) = nullptr;
  nvmlReturn_t (*DeviceGetNvLinkRemotePciInfo)(nvmlDevice_t, unsigned int, nvmlPciInfo_t *) = nullptr;
  nvmlReturn_t (*DeviceGetNvLinkRemoteDeviceType)(nvmlDevice_t, unsigned int, nvmlIntNvLinkDeviceType_t *) = nullptr;
  nvmlReturn_t (*EventSetCreate)(nvmlEventSet_t *) = nullptr;
  nvmlReturn_t (*DeviceRegisterEvents)(nvmlDevice_t, unsigned long long, nvmlEventSet_t) = nullptr;
  nvmlReturn_t (*EventSetWait)(nvmlEventSet_t, nvmlEventData_t *, unsigned int) = nullptr;
  nvmlReturn_t (*EventSetFree)(nvmlEventSet_t) = nullptr;

  std::string err(const char *what, nvmlReturn_t r) {
    return std::string(what) + ": nvml: " + (ErrorString ? ErrorString(r) : "error") + " (" + std::to_string((int)r) + ")";
  }
};

struct gpa_health {
  gpa_nvml *nv = nullptr;
  bool disabled = false;  // DP_DISABLE_HEALTHCHECKS says so: the reference's goroutine returns at once
  nvmlEventSet_t set = nullptr;
  std::vector<std::string> ids;        // watched device ids, caller order
  std::unordered_set<uint64_t> skipped;
  std::thread worker;
  std::atomic<bool> stop{false};
  std::mutex mu;
  std::condition_variable cv;
  struct Item {
    int device;  // index into ids
    unsigned long long xid;
  };
  std::deque<Item> queue;
  unsigned long long events_seen = 0, events_skipped = 0;
  int wait_timeout_ms = 5000;  // nvidia.go:235
};

extern "C" {

int gpa_nvml_open(const char *libpath, gpa_nvml **out) {
  if (!out) return GPA_ERR_INVALID;
  *out = nullptr;
  gpa_nvml *nv = new gpa_nvml();
  const char *path = (libpath && *libpath) ? libpath : "libnvidia-ml.so.1";  // nvml_dl.go:31
  nv->handle = dlopen(path, RTLD_LAZY | RTLD_GLOBAL);
  if (!nv->handle) {
    fprintf(stderr, "gpa_nvml_open: could not load NVML library %s: %s\n", path, dlerror());
    delete nv;
    return GPA_ERR_NVML;
  }
  bool ok = true;
  ok &= load_sym(nv->handle, "nvmlInit_v2", nv->Init);
  ok &= load_sym(nv->handle, "nvmlShutdown", nv->Shutdown);
  ok &= load_sym(nv->handle, "nvmlErrorString", nv->ErrorString);
  ok &= load_sym(nv->handle, "nvmlDeviceGetCount_v2", nv->DeviceGetCount);
  ok &= load_sym(nv->handle, "nvmlDeviceGetHandleByIndex_v2", nv->DeviceGetHandleByIndex);
  ok &= load_sym(nv->handle, "nvmlDeviceGetUUID", nv->DeviceGetUUID);
  ok &= load_sym(nv->handle, "nvmlDeviceGetPciInfo_v3", nv->DeviceGetPciInfo);
  ok &= load_sym(nv->handle, "nvmlDeviceGetNvLinkState", nv->DeviceGetNvLinkState);
  ok &= load_sym(nv->handle, "nvmlDeviceGetNvLinkRemotePciInfo_v2", nv->DeviceGetNvLinkRemotePciInfo);
  ok &= load_sym(nv->handle, "nvmlEventSetCreate", nv->EventSetCreate);
  ok &= load_sym(nv->handle, "nvmlDeviceRegisterEvents", nv->DeviceRegisterEvents);
  ok &= load_sym(nv->handle, "nvmlEventSetFree", nv->EventSetFree);
  // optional ones: the reference treats a missing symbol as "unknown" (bindings.go:249-252) or falls back (:100-103)
  load_sym(nv->handle, "nvmlDeviceGetTopologyCommonAncestor", nv->DeviceGetTopologyCommonAncestor);
  load_sym(nv->handle, "nvmlDeviceGetNvLinkRemoteDeviceType", nv->DeviceGetNvLinkRemoteDeviceType);
  load_sym(nv->handle, "nvmlDeviceGetMemoryInfo", nv->DeviceGetMemoryInfo);
  load_sym(nv->handle, "nvmlDeviceGetMinorNumber", nv->DeviceGetMinorNumber);
  if (!load_sym(nv->handle, "nvmlEventSetWait_v2", nv->EventSetWait)) ok &= load_sym(nv->handle, "nvmlEventSetWait", nv->EventSetWait);
  if (!ok) {
    fprintf(stderr, "gpa_nvml_open: %s lacks a required NVML symbol\n", path);
    dlclose(nv->handle);
    delete nv;
    return GPA_ERR_NVML;
  }
  nvmlReturn_t r = nv->Init();
  if (r != NVML_SUCCESS) {
    fprintf(stderr, "gpa_nvml_open: %s\n", nv->err("nvmlInit_v2", r).c_str());
    dlclose(nv->handle);
    delete nv;
    return GPA_ERR_NVML;
  }
  *out = nv;
  return GPA_OK;
}

void gpa_nvml_close(gpa_nvml *nv) {
  if (!nv) return;
  if (nv->Shutdown) nv->Shutdown();
  if (nv->handle) dlclose(nv->handle);
  delete nv;
}

const char *gpa_nvml_last_error(gpa_nvml *nv) { return nv ? nv->last_error.c_str() : ""; }

int gpa_nvml_device_count(gpa_nvml *nv, int *n) {
  if (!nv || !n) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(nv->mu);
  unsigned int c = 0;
  nvmlReturn_t r = nv->DeviceGetCount(&c);
  nv->calls++;
  if (r != NVML_SUCCESS) {
    nv->last_error = nv->err("error calling nvml.GetDeviceCount", r);
    return GPA_ERR_NVML;
  }
  *n = (int)c;
  return GPA_OK;
}

int gpa_nvml_device_uuid(gpa_nvml *nv, int index, char *buf, int cap) {
  if (!nv || !buf || cap < 2 || index < 0) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(nv->mu);
  nvmlDevice_t d;
  nvmlReturn_t r = nv->DeviceGetHandleByIndex((unsigned)index, &d);
  nv->calls++;
  if (r == NVML_SUCCESS) {
    r = nv->DeviceGetUUID(d, buf, (unsigned)cap);
    nv->calls++;
  }
  if (r != NVML_SUCCESS) {
    nv->last_error = nv->err("nvmlDeviceGetUUID", r);
    return GPA_ERR_NVML;
  }
  return GPA_OK;
}

int gpa_nvml_device_info(gpa_nvml *nv, int index, unsigned long long *total_memory_mib, int *minor, int *numa_node, char *bus_id,
                         int bus_id_cap) {
  if (!nv || index < 0) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(nv->mu);
  nvmlDevice_t d;
  nvmlReturn_t r = nv->DeviceGetHandleByIndex((unsigned)index, &d);
  nv->calls++;
  if (r != NVML_SUCCESS) {
    nv->last_error = nv->err("nvmlDeviceGetHandleByIndex", r);
    return GPA_ERR_NVML;
  }
  if (total_memory_mib) {
    // totalMemory := uint(*status.Memory.Global.Free + *status.Memory.Global.Used), both already in MiB
    // (cmd/nvidia-device-plugin/nvidia.go:90-92; nvml/bindings.go:491-520)
    *total_memory_mib = 0;
    nvmlMemory_t mem;
    if (nv->DeviceGetMemoryInfo) {
      r = nv->DeviceGetMemoryInfo(d, &mem);
      nv->calls++;
      if (r == NVML_SUCCESS) {
        *total_memory_mib = mem.free / (1024ull * 1024ull) + mem.used / (1024ull * 1024ull);
      } else if (r != NVML_ERROR_NOT_SUPPORTED) {
        nv->last_error = nv->err("nvmlDeviceGetMemoryInfo", r);
        return GPA_ERR_NVML;
      }
    }
  }
  if (minor) {
    *minor = -1;
    unsigned int m = 0;
    if (nv->DeviceGetMinorNumber) {
      r = nv->DeviceGetMinorNumber(d, &m);
      nv->calls++;
      if (r == NVML_SUCCESS) {
        *minor = (int)m;
      } else if (r != NVML_ERROR_NOT_SUPPORTED) {  // bindings.go:316-324: NOT_SUPPORTED is "no minor number"
        nv->last_error = nv->err("nvmlDeviceGetMinorNumber", r);
        return GPA_ERR_NVML;
      }
    }
  }
  if (numa_node || bus_id) {
    nvmlPciInfo_t pci;
    r = nv->DeviceGetPciInfo(d, &pci);
    nv->calls++;
    if (r != NVML_SUCCESS) {
      nv->last_error = nv->err("nvmlDeviceGetPciInfo", r);
      return GPA_ERR_NVML;
    }
    if (bus_id && bus_id_cap > 0) snprintf(bus_id, (size_t)bus_id_cap, "%s", pci.busId);
    if (numa_node) {
      // numaNode(), nvml.go:314-332: /sys/bus/pci/devices/<bus id without its first four characters, lower case>/numa_node;
      // unreadable or negative = no NUMA information
      *numa_node = -1;
      const char *root = getenv("GPA_SYSFS_PCI_DEVICES");  // tests point this at a scratch directory
      std::string path = std::string(root && *root ? root : "/sys/bus/pci/devices") + "/";
      const char *b = pci.busId;
      if (strlen(b) > 4) b += 4;
      for (; *b; b++) path += (char)tolower((unsigned char)*b);
      path += "/numa_node";
      if (FILE *f = fopen(path.c_str(), "r")) {
        char buf[32] = {0};
        if (fgets(buf, sizeof buf, f)) {
          char *end = nullptr;
          const long v = strtol(buf, &end, 10);
          if (end != buf && v >= 0) *numa_node = (int)v;
        }
        fclose(f);
      }
    }
  }
  return GPA_OK;
}

int gpa_nvml_read_topology(gpa_nvml *nv, int flags, int *n_out, int32_t *links2, int cap, unsigned long long *nvml_calls) {
  if (!nv || !n_out) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(nv->mu);
  const unsigned long long calls0 = nv->calls;
  unsigned int count = 0;
  nvmlReturn_t r = nv->DeviceGetCount(&count);
  nv->calls++;
  if (r != NVML_SUCCESS) {
    nv->last_error = nv->err("error calling nvml.GetDeviceCount", r);  // device.go:34-37
    return GPA_ERR_NVML;
  }
  const int n = (int)count;
  *n_out = n;
  if (n > GPA_MAX_TOPOLOGY) return GPA_ERR_TOO_LARGE;
  if (!links2) return GPA_OK;
  if (cap < n * n * 2) return GPA_ERR_BUFFER;
  memset(links2, 0, sizeof(int32_t) * (size_t)n * n * 2);

  std::vector<nvmlDevice_t> dev(n);
  std::vector<std::string> bus(n);
  for (int i = 0; i < n; i++) {
    r = nv->DeviceGetHandleByIndex((unsigned)i, &dev[i]);
    nv->calls++;
    nvmlPciInfo_t pci;
    memset(&pci, 0, sizeof pci);
    if (r == NVML_SUCCESS) {
      r = nv->DeviceGetPciInfo(dev[i], &pci);
      nv->calls++;
    }
    if (r != NVML_SUCCESS) {
      nv->last_error = "error creating nvml.Device " + std::to_string(i) + ": " + nv->err("", r);  // device.go:41-44
      return GPA_ERR_NVML;
    }
    bus[i] = pci.busId;
  }

  const bool aware = (flags & GPA_TOPO_NVSWITCH_AWARE) != 0;
  const unsigned max_links = aware ? NVML_NVLINK_MAX_LINKS : kRefMaxLinks;
  // per device: bus ids at the far end of its active links (bindings.go:358-386), and how many end in a switch
  std::vector<std::vector<std::string>> remote(n);
  std::vector<int> switch_links(n, 0);
  for (int i = 0; i < n; i++) {
    for (unsigned l = 0; l < max_links; l++) {
      nvmlEnableState_t st;
      r = nv->DeviceGetNvLinkState(dev[i], l, &st);
      nv->calls++;
      if (r == NVML_ERROR_NOT_SUPPORTED || r == NVML_ERROR_INVALID_ARGUMENT) continue;  // bindings.go:340-342
      if (r != NVML_SUCCESS) {
        nv->last_error = "error getting NVLink for devices: " + nv->err("nvmlDeviceGetNvLinkState", r);
        return GPA_ERR_NVML;
      }
      if (st != NVML_FEATURE_ENABLED) continue;
      nvmlPciInfo_t pci;
      memset(&pci, 0, sizeof pci);
      r = nv->DeviceGetNvLinkRemotePciInfo(dev[i], l, &pci);
      nv->calls++;
      if (r == NVML_ERROR_NOT_SUPPORTED || r == NVML_ERROR_INVALID_ARGUMENT) continue;  // bindings.go:351-353
      if (r != NVML_SUCCESS) {
        nv->last_error = "error getting NVLink for devices: " + nv->err("nvmlDeviceGetNvLinkRemotePciInfo", r);
        return GPA_ERR_NVML;
      }
      remote[i].push_back(pci.busId);
      if (aware && nv->DeviceGetNvLinkRemoteDeviceType) {
        nvmlIntNvLinkDeviceType_t ty;
        nvmlReturn_t r2 = nv->DeviceGetNvLinkRemoteDeviceType(dev[i], l, &ty);
        nv->calls++;
        if (r2 == NVML_SUCCESS && ty == NVML_NVLINK_DEVICE_TYPE_SWITCH) switch_links[i]++;
      }
    }
  }

  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (i == j) continue;  // device.go:51
      int32_t *cell = links2 + ((size_t)i * n + j) * 2;
      // nvml.GetP2PLink (nvml.go:592-615)
      int p2p = kUnknown;
      if (nv->DeviceGetTopologyCommonAncestor) {
        nvmlGpuTopologyLevel_t level;
        r = nv->DeviceGetTopologyCommonAncestor(dev[i], dev[j], &level);
        nv->calls++;
        if (r == NVML_ERROR_NOT_SUPPORTED) {
          p2p = kUnknown;  // bindings.go:256-258
        } else if (r != NVML_SUCCESS) {
          nv->last_error = "error getting P2PLink for devices (" + std::to_string(i) + ", " + std::to_string(j) + "): " +
                           nv->err("", r);  // device.go:52-55
          return GPA_ERR_NVML;
        } else {
          switch (level) {
            case NVML_TOPOLOGY_INTERNAL: p2p = kSameBoard; break;
            case NVML_TOPOLOGY_SINGLE: p2p = kSingleSwitch; break;
            case NVML_TOPOLOGY_MULTIPLE: p2p = kMultiSwitch; break;
            case NVML_TOPOLOGY_HOSTBRIDGE: p2p = kHostBridge; break;
            case NVML_TOPOLOGY_NODE: p2p = kSameCPU; break;  // NVML_TOPOLOGY_CPU
            case NVML_TOPOLOGY_SYSTEM: p2p = kCrossCPU; break;
            default:
              nv->last_error = "error getting P2PLink for devices (" + std::to_string(i) + ", " + std::to_string(j) +
                               "): unsupported P2P link type";  // ErrUnsupportedP2PLink, nvml.go:611-612
              return GPA_ERR_NVML;
          }
        }
      }
      cell[0] = p2p;  // recorded only when != Unknown (device.go:56-58): 0 means absent
      // nvml.GetNVLink (nvml.go:617-658): links whose far end IS the peer GPU
      int direct = 0;
      for (const std::string &b : remote[i])
        if (b == bus[j]) direct++;
      if (direct > 12) direct = 12;  // the switch at nvml.go:626-651 saturates at TwelveNVLINKLinks
      int links = direct;
      if (aware && direct == 0) {
        // extension (no reference counterpart; the reference stops at "TODO(klueska): Handle NVSwitch
        // semantics", nvml.go:655): two GPUs that both hang off the switch fabric reach each other over
        // min(their switch links) links
        int s = switch_links[i] < switch_links[j] ? switch_links[i] : switch_links[j];
        links = s > 12 ? 12 : s;
      }
      cell[1] = links > 0 ? kOneNVLink + links - 1 : kUnknown;
    }
  if (nvml_calls) *nvml_calls = nv->calls - calls0;
  return GPA_OK;
}

int gpa_set_topology_from_nvml(gpa_ctx *ctx, gpa_nvml *nv, int flags) {
  if (!ctx || !nv) return GPA_ERR_INVALID;
  int n = 0;
  int rc = gpa_nvml_read_topology(nv, flags, &n, nullptr, 0, nullptr);
  if (rc != GPA_OK) return rc;
  if (n < 1) return GPA_ERR_NVML;
  std::vector<int32_t> links((size_t)n * n * 2);
  rc = gpa_nvml_read_topology(nv, flags, &n, links.data(), (int)links.size(), nullptr);
  if (rc != GPA_OK) return rc;
  rc = gpa_set_topology_links(ctx, n, links.data());
  if (rc != GPA_OK) return rc;
  // Requests list devices in byte-wise sorted-UUID order (stripReplicas, cmd/nvidia-device-plugin/replica.go:43):
  // declare that order, so that batches of such requests take the fast kernel.
  std::vector<std::string> uuids(n);
  for (int i = 0; i < n; i++) {
    char buf[96];
    rc = gpa_nvml_device_uuid(nv, i, buf, sizeof buf);
    if (rc != GPA_OK) return rc;
    uuids[i] = buf;
  }
  std::vector<int32_t> order(n);
  for (int i = 0; i < n; i++) order[i] = i;
  std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return uuids[a] < uuids[b]; });
  return gpa_set_device_order(ctx, order.data(), n);
}

// ---------------------------------------------------------------------------------------
// Health watcher

static void health_loop(gpa_health *h) {
  gpa_nvml *nv = h->nv;
  while (!h->stop.load()) {  // nvidia.go:228-233
    nvmlEventData_t data;
    memset(&data, 0, sizeof data);
    data.gpuInstanceId = 0xFFFFFFFFu;  // bindings.go:176-177
    data.computeInstanceId = 0xFFFFFFFFu;
    nvmlReturn_t r = nv->EventSetWait(h->set, &data, (unsigned)h->wait_timeout_ms);
    if (r != NVML_SUCCESS) continue;  // nvidia.go:236-238 — on error Event{} has Etype 0 != XidCriticalError
    // (err == nil: the reference goes on whatever the event type; only Xid events are registered)
    std::unique_lock<std::mutex> lk(h->mu);
    h->events_seen++;
    if (h->skipped.count(data.eventData)) {  // nvidia.go:240-242
      h->events_skipped++;
      continue;
    }
    char uuid[96];
    uuid[0] = 0;
    if (data.device) nv->DeviceGetUUID(data.device, uuid, sizeof uuid);  // bindings.go:184
    if (uuid[0] == 0) {
      // nvidia.go:244-251 — no device attached to the event: every device goes unhealthy
      for (size_t i = 0; i < h->ids.size(); i++) h->queue.push_back({(int)i, data.eventData});
    } else {
      // nvidia.go:253-267 — full GPUs carry gi = ci = 0xFFFFFFFF (MIG devices are outside this path)
      for (size_t i = 0; i < h->ids.size(); i++)
        if (h->ids[i] == uuid && data.gpuInstanceId == 0xFFFFFFFFu && data.computeInstanceId == 0xFFFFFFFFu)
          h->queue.push_back({(int)i, data.eventData});
    }
    lk.unlock();
    h->cv.notify_all();
  }
}

int gpa_health_open(gpa_nvml *nv, const char *const *device_ids, int n, const char *disable_env, int wait_timeout_ms,
                    gpa_health **out) {
  if (!nv || !out || n < 0 || (n > 0 && !device_ids)) return GPA_ERR_INVALID;
  *out = nullptr;
  gpa_health *h = new gpa_health();
  h->nv = nv;
  if (wait_timeout_ms > 0) h->wait_timeout_ms = wait_timeout_ms;
  for (int i = 0; i < n; i++) h->ids.push_back(device_ids[i]);

  // nvidia.go:182-188
  std::string dis = disable_env ? disable_env : (getenv("DP_DISABLE_HEALTHCHECKS") ? getenv("DP_DISABLE_HEALTHCHECKS") : "");
  for (auto &c : dis) c = (char)tolower((unsigned char)c);
  if (dis == "all") dis = "xids";
  if (dis.find("xids") != std::string::npos) {
    h->disabled = true;
    *out = h;
    return GPA_OK;
  }
  // nvidia.go:193-208
  for (uint64_t x : {13ull, 31ull, 43ull, 45ull, 68ull}) h->skipped.insert(x);
  uint64_t extra[256];
  int ne = gpa_get_additional_xids(dis.c_str(), extra, 256);
  for (int i = 0; i < ne; i++) h->skipped.insert(extra[i]);

  std::lock_guard<std::mutex> lk(nv->mu);
  nvmlReturn_t r = nv->EventSetCreate(&h->set);  // nvidia.go:210
  if (r != NVML_SUCCESS) {
    nv->last_error = nv->err("nvmlEventSetCreate", r);
    delete h;
    return GPA_ERR_NVML;
  }
  // nvidia.go:213-226 with RegisterEventForDevice (bindings.go:137-168)
  unsigned int count = 0;
  nv->DeviceGetCount(&count);
  for (size_t di = 0; di < h->ids.size(); di++) {
    bool found = false;
    for (unsigned i = 0; i < count && !found; i++) {
      nvmlDevice_t d;
      char uuid[96];
      if (nv->DeviceGetHandleByIndex(i, &d) != NVML_SUCCESS) continue;
      if (nv->DeviceGetUUID(d, uuid, sizeof uuid) != NVML_SUCCESS) continue;
      if (h->ids[di] != uuid) continue;
      found = true;
      r = nv->DeviceRegisterEvents(d, nvmlEventTypeXidCriticalError, h->set);
      if (r == NVML_ERROR_NOT_SUPPORTED) {
        // "too old to support healthchecking ... Marking it unhealthy" (nvidia.go:220-224)
        h->queue.push_back({(int)di, 0});
      } else if (r != NVML_SUCCESS) {
        nv->last_error = nv->err("nvmlDeviceRegisterEvents", r);  // check(err) -> log.Panicln (nvidia.go:225)
        nv->EventSetFree(h->set);
        delete h;
        return GPA_ERR_NVML;
      }
    }
    if (!found) {
      nv->last_error = "nvml: device not found";  // bindings.go:167 -> check(err) panics
      nv->EventSetFree(h->set);
      delete h;
      return GPA_ERR_UNKNOWN_DEVICE;
    }
  }
  h->worker = std::thread(health_loop, h);
  *out = h;
  return GPA_OK;
}

int gpa_health_poll(gpa_health *h, int timeout_ms, int *device_index, unsigned long long *xid) {
  if (!h || !device_index) return GPA_ERR_INVALID;
  std::unique_lock<std::mutex> lk(h->mu);
  if (h->queue.empty() && timeout_ms > 0 && !h->disabled)
    h->cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), [&] { return !h->queue.empty(); });
  if (h->queue.empty()) return 0;
  *device_index = h->queue.front().device;
  if (xid) *xid = h->queue.front().xid;
  h->queue.pop_front();
  return 1;
}

int gpa_health_stats(gpa_health *h, unsigned long long *events_seen, unsigned long long *events_skipped, int *disabled) {
  if (!h) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(h->mu);
  if (events_seen) *events_seen = h->events_seen;
  if (events_skipped) *events_skipped = h->events_skipped;
  if (disabled) *disabled = h->disabled ? 1 : 0;
  return GPA_OK;
}

void gpa_health_close(gpa_health *h) {
  if (!h) return;
  h->stop.store(true);
  if (h->worker.joinable()) h->worker.join();  // returns within one wait timeout, like the goroutine at nvidia.go:229-233
  if (h->set) h->nv->EventSetFree(h->set);     // defer nvml.DeleteEventSet (nvidia.go:211)
  delete h;
}

}  // extern "C"
