// Host-side mirror of the reference's device-plugin server for the path: the four RPC
// bodies that surround the allocation decision, without the gRPC transport.
//
// ref: cmd/nvidia-device-plugin/server.go — NvidiaDevicePlugin :56-71, initialize :95-116,
// GetDevicePluginOptions :243-248, ListAndWatch :251-265, GetPreferredAllocation :268-313,
// Allocate :316-353, deviceExists / deviceReplicaExists :377-394, deviceIDsFromUUIDs :397-413,
// apiDevices / apiEnvs / apiMounts / apiDeviceSpecs :415-480.
//
// Same names, same argument meaning, same error strings.  What changes is the data layout:
// replica ids live in a hash set (the reference scans a slice per id, :387-394), the uuid ->
// NVML index map is built once (the reference re-enumerates NVML per request, :274,279) and
// the decision itself goes to the GPU through gpa_besteffort().  The reference's defects on
// this path are kept by default (bit-exact drop-in) and can be switched off one by one with
// the GPA_FIX_* bits.
#include <sys/stat.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../../include/gpualloc.h"

namespace {

struct Device {  // cmd/nvidia-device-plugin/nvidia.go:41-46 around pluginapi.Device
  std::string id;
  std::string index;
  std::vector<std::string> paths;
  unsigned long long total_memory = 0;
  bool healthy = true;
  int numa = -1;        // Topology.Nodes[0].ID, -1 = no topology (nvidia.go:169-177)
  int nvml_index = -1;  // gpuallocator.Device.Index
};

const char kListPath[] = "/var/run/nvidia-container-devices";  // deviceListAsVolumeMountsContainerPathRoot, server.go:52
const char kNullPath[] = "/dev/null";                          // deviceListAsVolumeMountsHostPath, server.go:51

std::string json_escape(const std::string &s) {
  std::string o;
  for (unsigned char c : s) {
    if (c == '"' || c == '\\') {
      o += '\\';
      o += (char)c;
    } else if (c < 0x20) {
      char b[8];
      snprintf(b, sizeof b, "\\u%04x", c);
      o += b;
    } else {
      o += (char)c;
    }
  }
  return o;
}

// filepath.Join(a, b) for the two shapes used here: cleans duplicate slashes at the seam
std::string path_join(const std::string &a, const std::string &b) {
  if (a.empty()) return b;
  if (b.empty()) return a;
  std::string r = a;
  while (r.size() > 1 && r.back() == '/') r.pop_back();
  size_t s = 0;
  while (s < b.size() && b[s] == '/') s++;
  if (r == "/") return "/" + b.substr(s);
  return r + "/" + b.substr(s);
}

int copy_out(const std::string &s, char *buf, int cap) {
  if (!buf || cap < (int)s.size() + 1) return GPA_ERR_BUFFER;
  memcpy(buf, s.c_str(), s.size() + 1);
  return GPA_OK;
}

void set_err(char *errbuf, int cap, const std::string &m) {
  if (errbuf && cap > 0) snprintf(errbuf, (size_t)cap, "%s", m.c_str());
}

std::vector<std::string> unpack(const char *buf, int n) {
  std::vector<std::string> v;
  const char *p = buf;
  for (int i = 0; i < n; i++) {
    v.emplace_back(p);
    p += v.back().size() + 1;
  }
  return v;
}

}  // namespace

struct gpa_plugin {
  gpa_ctx *ctx = nullptr;
  std::string resource_name, envvar, driver_root;
  unsigned replicas = 0;
  bool auto_replicas = false;
  int list_strategy = GPA_LIST_STRATEGY_ENVVAR;
  int id_strategy = GPA_ID_STRATEGY_UUID;
  bool pass_device_specs = false;
  bool has_policy = false;
  int fixes = 0;
  std::vector<Device> cached;    // cachedDevices
  std::vector<Device> replicas_; // deviceReplicas: struct COPIES of the cached devices (server.go:107)
  std::unordered_set<std::string> replica_ids;
  std::unordered_map<std::string, int> cached_by_id;
  std::mutex mu;
  // Which control nodes exist on the host (apiDeviceSpecs, server.go:446-463), one bit per path.  The reference stats the
  // four paths in every Allocate; here the answer is kept and refreshed at most once a second, so a node that appears
  // later (nvidia-uvm is created on first use) is still picked up, and a burst of Allocate calls pays for one round of
  // stat() calls.
  int control_mask = 0;
  std::chrono::steady_clock::time_point control_checked{};
  bool control_valid = false;
};

namespace {
const char *const kControlNodes[4] = {"/dev/nvidiactl", "/dev/nvidia-uvm", "/dev/nvidia-uvm-tools", "/dev/nvidia-modeset"};

int control_nodes_of(gpa_plugin *p) {
  const auto now = std::chrono::steady_clock::now();
  std::lock_guard<std::mutex> lk(p->mu);
  if (!p->control_valid || now - p->control_checked > std::chrono::seconds(1)) {
    int m = 0;
    for (int i = 0; i < 4; i++) {
      struct stat st;
      if (stat(kControlNodes[i], &st) == 0) m |= 1 << i;  // os.Stat(p) on the HOST path (:454)
    }
    p->control_mask = m;
    p->control_checked = now;
    p->control_valid = true;
  }
  return p->control_mask;
}
}  // namespace

extern "C" {

int gpa_plugin_create(gpa_ctx *ctx, const gpa_plugin_config *cfg, const gpa_plugin_device *devs, int n_devs, gpa_plugin **out) {
  if (!cfg || !out || n_devs < 0 || (n_devs > 0 && !devs)) return GPA_ERR_INVALID;
  *out = nullptr;
  gpa_plugin *p = new gpa_plugin();
  p->ctx = ctx;
  p->resource_name = cfg->resource_name ? cfg->resource_name : "";
  p->envvar = cfg->device_list_envvar ? cfg->device_list_envvar : "NVIDIA_VISIBLE_DEVICES";
  p->driver_root = cfg->driver_root ? cfg->driver_root : "/";
  p->replicas = cfg->replicas;
  p->auto_replicas = cfg->auto_replicas != 0;
  p->list_strategy = cfg->device_list_strategy;
  p->id_strategy = cfg->device_id_strategy;
  p->pass_device_specs = cfg->pass_device_specs != 0;
  p->has_policy = cfg->has_policy != 0;
  p->fixes = cfg->fixes;
  // initialize(), server.go:95-111
  for (int i = 0; i < n_devs; i++) {
    Device d;
    d.id = devs[i].id ? devs[i].id : "";
    d.index = devs[i].index ? devs[i].index : "";
    for (int k = 0; k < devs[i].n_paths; k++) d.paths.emplace_back(devs[i].paths[k]);
    d.total_memory = devs[i].total_memory_mib;
    d.numa = devs[i].numa_node;
    d.nvml_index = devs[i].nvml_index;
    p->cached_by_id.emplace(d.id, (int)p->cached.size());
    p->cached.push_back(d);
  }
  for (const Device &dev : p->cached) {
    unsigned reps = p->replicas;
    if (p->auto_replicas) reps = (unsigned)(dev.total_memory / 1000);  // :100-103
    for (unsigned i = 0; i < reps; i++) {
      Device r = dev;  // "This is replicating the Device struct" (:107)
      r.id = dev.id + "-replica-" + std::to_string(i);
      p->replica_ids.insert(r.id);
      p->replicas_.push_back(std::move(r));
    }
  }
  // The plugin always hands the allocator its devices in sorted-UUID order (stripReplicas, replica.go:43): tell the
  // context, when its topology covers exactly these devices, so that batched decisions take the fast kernel.
  if (ctx && p->has_policy) {
    int nt = 0;
    if (gpa_get_topology(ctx, &nt, nullptr, nullptr, 0) == GPA_OK && nt == (int)p->cached.size()) {
      std::vector<int32_t> order;
      std::vector<const Device *> by_id;
      for (const Device &d : p->cached) by_id.push_back(&d);
      std::sort(by_id.begin(), by_id.end(), [](const Device *a, const Device *b) { return a->id < b->id; });
      bool ok = true;
      for (const Device *d : by_id) {
        ok = ok && d->nvml_index >= 0 && d->nvml_index < nt;
        order.push_back(d->nvml_index);
      }
      if (ok) gpa_set_device_order(ctx, order.data(), nt);  // a non-permutation is refused there and changes nothing
    }
  }
  if (p->pass_device_specs) control_nodes_of(p);
  *out = p;
  return GPA_OK;
}

void gpa_plugin_destroy(gpa_plugin *p) { delete p; }

int gpa_plugin_get_options(gpa_plugin *p, int *pre_start_required, int *get_preferred_allocation_available) {
  if (!p) return GPA_ERR_INVALID;
  if (pre_start_required) *pre_start_required = 0;
  // server.go:244-246
  if (get_preferred_allocation_available)
    *get_preferred_allocation_available = (p->has_policy || p->replicas > 1 || p->auto_replicas) ? 1 : 0;
  return GPA_OK;
}

int gpa_plugin_device_count(gpa_plugin *p, int *n_cached, int *n_replicas) {
  if (!p) return GPA_ERR_INVALID;
  if (n_cached) *n_cached = (int)p->cached.size();
  if (n_replicas) *n_replicas = (int)p->replicas_.size();
  return GPA_OK;
}

// apiDevices(), server.go:415-421: what ListAndWatch sends (the replicas, not the cached devices)
int gpa_plugin_list_devices(gpa_plugin *p, char *json, int cap) {
  if (!p) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(p->mu);
  std::string o = "[";
  for (size_t i = 0; i < p->replicas_.size(); i++) {
    const Device &d = p->replicas_[i];
    if (i) o += ",";
    o += "{\"ID\":\"" + json_escape(d.id) + "\",\"health\":\"" + (d.healthy ? "Healthy" : "Unhealthy") + "\"";
    if (d.numa >= 0) o += ",\"topology\":{\"nodes\":[{\"ID\":" + std::to_string(d.numa) + "}]}";
    o += "}";
  }
  o += "]";
  return copy_out(o, json, cap);
}

// The body of `case d := <-m.health` (server.go:258-262).  Returns how many entries of the list
// that ListAndWatch re-sends changed: 0 in parity mode (the channel carries a pointer into
// cachedDevices, the list is built from the COPIES in deviceReplicas — SURVEY.md D1), all the
// replicas of that GPU with GPA_FIX_HEALTH_PROPAGATION.
int gpa_plugin_mark_unhealthy(gpa_plugin *p, int cached_index) {
  if (!p || cached_index < 0 || cached_index >= (int)p->cached.size()) return GPA_ERR_INVALID;
  std::lock_guard<std::mutex> lk(p->mu);
  p->cached[cached_index].healthy = false;  // d.Health = pluginapi.Unhealthy (:260)
  int changed = 0;
  if (p->fixes & GPA_FIX_HEALTH_PROPAGATION) {
    const std::string prefix = p->cached[cached_index].id + "-replica-";
    for (Device &r : p->replicas_)
      if (r.healthy && r.id.compare(0, prefix.size(), prefix) == 0) {
        r.healthy = false;
        changed++;
      }
  }
  return changed;
}

// GPA_OPT_REPLICA_TOPOLOGY (no reference counterpart): the replica branch of GetPreferredAllocation answered by the
// replica-set search instead of prioritizeDevices.  The ids are sorted byte-wise first (what prioritizeDevices' callers
// see in the end, replica.go:191), so the replicas of a GPU are contiguous and the answer does not depend on the order
// kubelet listed them in.  Errors keep the reference's kinds and texts: replica.go:121 (panic), :136, :183.
static int preferred_by_topology(gpa_plugin *p, const char *const *avail, int n_avail, const char *const *must, int n_must, int size,
                                 char *outbuf, int outbuf_len, int *n_out, char *errbuf, int errbuf_len) {
  if (n_must > size) {
    set_err(errbuf, errbuf_len, "runtime error: makeslice: cap out of range");
    return GPA_ERR_PANIC;
  }
  std::vector<std::string> ids;
  ids.reserve((size_t)n_avail);
  for (int i = 0; i < n_avail; i++) ids.emplace_back(avail[i]);
  std::sort(ids.begin(), ids.end());
  ids.erase(std::unique(ids.begin(), ids.end()), ids.end());
  std::vector<int32_t> phys(ids.size()), req;
  for (size_t i = 0; i < ids.size(); i++) {
    const size_t cut = ids[i].find("-replica-");  // stripReplica, replica.go:28-30
    auto it = p->cached_by_id.find(cut == std::string::npos ? ids[i] : ids[i].substr(0, cut));
    if (it == p->cached_by_id.end()) return GPA_ERR_UNKNOWN_DEVICE;  // cannot happen: resolve() has seen every id
    phys[i] = p->cached[it->second].nvml_index;
  }
  for (int i = 0; i < n_must; i++) {
    auto it = std::lower_bound(ids.begin(), ids.end(), std::string(must[i]));
    if (it == ids.end() || *it != must[i]) {
      set_err(errbuf, errbuf_len, std::string("device '") + must[i] + "' in mustIncludeDeviceIDs is missing from availableDeviceIDs");
      return GPA_ERR_MUST_MISSING;
    }
    const int32_t pos = (int32_t)(it - ids.begin());
    if (std::find(req.begin(), req.end(), pos) == req.end()) req.push_back(pos);
  }
  if (size <= 0) return GPA_OK;
  if ((int)ids.size() < size) {
    set_err(errbuf, errbuf_len, "no devices left to allocate");
    return GPA_ERR_NO_DEVICES_LEFT;
  }
  std::vector<int32_t> out((size_t)size + 1);
  int cnt = 0;
  int rc = gpa_subset_besteffort(p->ctx, phys.data(), (int)phys.size(), req.data(), (int)req.size(), size, 0, out.data(), &cnt);
  if (rc != GPA_OK) {
    set_err(errbuf, errbuf_len, std::string(gpa_strerror(rc)) + ": " + gpa_last_error(p->ctx));
    return rc;
  }
  std::string packed;
  for (int i = 0; i < cnt; i++) {
    packed += ids[(size_t)out[i]];
    packed.push_back('\0');
  }
  if ((int)packed.size() > outbuf_len) return GPA_ERR_BUFFER;
  if (!packed.empty()) memcpy(outbuf, packed.data(), packed.size());
  *n_out = cnt;
  return GPA_OK;
}

// One iteration of the loop over r.ContainerRequests in GetPreferredAllocation (server.go:273-311).
int gpa_plugin_get_preferred_allocation(gpa_plugin *p, const char *const *avail, int n_avail, const char *const *must,
                                        int n_must, int size, char *outbuf, int outbuf_len, int *n_out, char *errbuf,
                                        int errbuf_len) {
  if (!p || !n_out || n_avail < 0 || n_must < 0) return GPA_ERR_INVALID;
  *n_out = 0;
  set_err(errbuf, errbuf_len, "");

  // :274 — NewDevicesFrom(stripReplicas(AvailableDeviceIDs)): uuid -> device, unknown uuid is an error
  auto resolve = [&](const char *const *ids, int n, const char *what, std::vector<int> &idx) -> int {
    int cnt = 0;
    size_t need = 16;
    for (int i = 0; i < n; i++) need += strlen(ids[i]) + 1;
    std::vector<char> buf(need);
    int rc = gpa_strip_replicas(ids, n, buf.data(), (int)buf.size(), &cnt);
    if (rc != GPA_OK) return rc;
    for (const std::string &u : unpack(buf.data(), cnt)) {
      auto it = p->cached_by_id.find(u);
      if (it == p->cached_by_id.end()) {
        set_err(errbuf, errbuf_len, std::string("unable to retrieve list of ") + what + " devices: no device with uuid: " + u);
        return GPA_ERR_UNKNOWN_DEVICE;  // :276 / :281 with gpuallocator/device.go:89-91
      }
      idx.push_back(it->second);
    }
    return GPA_OK;
  };
  std::vector<int> available, required;
  int rc = GPA_OK;
  // On the replica branch the two NewDevicesFrom calls (:274, :279) only matter for their error: the devices they return
  // are not used (:285-296).  So the ids are first checked in place - a node has a handful of GPUs, a scan of their uuids
  // needs no sort, no copy and no hash - and the sorted walk, which names the FIRST unknown uuid in sort.Strings order as
  // the reference does, only runs when something is unknown.
  const bool replica_branch = p->replicas > 1 || p->auto_replicas;
  auto all_known = [&](const char *const *ids, int n) {
    for (int i = 0; i < n; i++) {
      const char *id = ids[i];
      const char *cut = strstr(id, "-replica-");
      const size_t len = cut ? (size_t)(cut - id) : strlen(id);
      bool found = false;
      for (const Device &d : p->cached)
        if (d.id.size() == len && memcmp(d.id.data(), id, len) == 0) {
          found = true;
          break;
        }
      if (!found) return false;
    }
    return true;
  };
  if (!(replica_branch && p->cached.size() <= 32 && all_known(avail, n_avail) && all_known(must, n_must))) {
    rc = resolve(avail, n_avail, "available", available);
    if (rc != GPA_OK) return rc;
    rc = resolve(must, n_must, "required", required);  // :279
    if (rc != GPA_OK) return rc;
  }

  if ((p->replicas > 1 || p->auto_replicas) && (p->fixes & GPA_OPT_REPLICA_TOPOLOGY) && p->ctx && p->has_policy)
    return preferred_by_topology(p, avail, n_avail, must, n_must, size, outbuf, outbuf_len, n_out, errbuf, errbuf_len);
  if (p->replicas > 1 || p->auto_replicas) {
    // :285-296
    int non_unique = 0;
    rc = gpa_prioritize_devices(avail, n_avail, must, n_must, size, outbuf, outbuf_len, n_out, &non_unique, errbuf, errbuf_len);
    if (rc == GPA_ERR_MUST_MISSING) {
      std::string id = errbuf ? errbuf : "";
      set_err(errbuf, errbuf_len, "device '" + id + "' in mustIncludeDeviceIDs is missing from availableDeviceIDs");
    } else if (rc == GPA_ERR_NO_DEVICES_LEFT) {
      set_err(errbuf, errbuf_len, "no devices left to allocate");
    }
    return rc;  // NonUniqueError is logged and ignored (:288-291): the ids are returned
  }
  if (!p->has_policy) {
    set_err(errbuf, errbuf_len, "GetPreferredAllocation() not implemented in this case");  // :302-304
    return GPA_ERR_NOT_IMPLEMENTED;
  }
  if (!p->ctx) return GPA_ERR_INVALID;

  // :297-301 — allocatePolicy.Allocate(available, required, size)
  std::vector<int32_t> av, rq;
  for (int i : available) av.push_back(p->cached[i].nvml_index);
  for (int i : required) {
    // `required` comes from a second NewDevices() call: fresh pointers, equal to none of the available ones
    // (SURVEY.md D4).  With GPA_FIX_REQUIRED_IDENTITY the uuid decides instead.
    rq.push_back((p->fixes & GPA_FIX_REQUIRED_IDENTITY) ? p->cached[i].nvml_index : GPA_FOREIGN);
  }
  std::vector<int32_t> out((size_t)(size > 0 ? size : 0) + 1);
  int cnt = 0;
  rc = gpa_besteffort(p->ctx, av.data(), (int)av.size(), rq.data(), (int)rq.size(), size, GPA_MODE_GO_FAITHFUL, out.data(), &cnt);
  if (rc != GPA_OK) {
    set_err(errbuf, errbuf_len, std::string(gpa_strerror(rc)) + ": " + gpa_last_error(p->ctx));
    return rc;
  }
  std::string packed;
  int written = 0;
  for (int i = 0; i < cnt; i++) {
    if (out[i] == GPA_NIL) {
      // `device.UUID` on a nil *Device (:299-301): the Go handler dies here
      set_err(errbuf, errbuf_len, "runtime error: invalid memory address or nil pointer dereference");
      return GPA_ERR_PANIC;
    }
    for (const Device &d : p->cached)
      if (d.nvml_index == out[i]) {
        // bare UUID, no "-replica-0" suffix (:300) unless GPA_FIX_REPLICA_SUFFIX
        packed += d.id;
        if (p->fixes & GPA_FIX_REPLICA_SUFFIX) packed += "-replica-0";
        packed.push_back('\0');
        written++;
        break;
      }
  }
  if ((int)packed.size() > outbuf_len) return GPA_ERR_BUFFER;
  if (!packed.empty()) memcpy(outbuf, packed.data(), packed.size());
  *n_out = written;
  return GPA_OK;
}

// One iteration of the loop over reqs.ContainerRequests in Allocate (server.go:318-350).
int gpa_plugin_allocate(gpa_plugin *p, const char *const *ids, int n, char *json, int cap, char *errbuf, int errbuf_len) {
  if (!p || n < 0 || (n > 0 && !ids)) return GPA_ERR_INVALID;
  set_err(errbuf, errbuf_len, "");
  // :319-323
  for (int i = 0; i < n; i++)
    if (!p->replica_ids.count(ids[i])) {
      set_err(errbuf, errbuf_len, "invalid allocation request for '" + p->resource_name + "': unknown device: " + ids[i]);
      return GPA_ERR_UNKNOWN_DEVICE;
    }
  // :325
  int cnt = 0;
  size_t need = 16;
  for (int i = 0; i < n; i++) need += strlen(ids[i]) + 1;
  std::vector<char> buf(need);
  int rc = gpa_strip_replicas(ids, n, buf.data(), (int)buf.size(), &cnt);
  if (rc != GPA_OK) return rc;
  std::vector<std::string> uuids = unpack(buf.data(), cnt);
  // :328-332
  for (const std::string &u : uuids)
    if (!p->cached_by_id.count(u)) {
      set_err(errbuf, errbuf_len, "invalid allocation request for '" + p->resource_name + "': unknown device: " + u);
      return GPA_ERR_UNKNOWN_DEVICE;
    }
  // deviceIDsFromUUIDs, :397-413
  std::vector<std::string> device_ids;
  if (p->id_strategy == GPA_ID_STRATEGY_UUID) {
    device_ids = uuids;
  } else if (p->id_strategy == GPA_ID_STRATEGY_INDEX) {
    for (const Device &d : p->cached)  // cachedDevices order, not request order (:404-410)
      for (const std::string &u : uuids)
        if (d.id == u) device_ids.push_back(d.index);
  }
  auto join = [](const std::vector<std::string> &v) {
    std::string s;
    for (size_t i = 0; i < v.size(); i++) s += (i ? "," : "") + v[i];
    return s;
  };
  std::string o = "{";
  // :338-344
  o += "\"envs\":{";
  if (p->list_strategy == GPA_LIST_STRATEGY_ENVVAR)
    o += "\"" + json_escape(p->envvar) + "\":\"" + json_escape(join(device_ids)) + "\"";
  else if (p->list_strategy == GPA_LIST_STRATEGY_VOLUME_MOUNTS)
    o += "\"" + json_escape(p->envvar) + "\":\"" + kListPath + "\"";
  o += "},\"mounts\":[";
  if (p->list_strategy == GPA_LIST_STRATEGY_VOLUME_MOUNTS)
    for (size_t i = 0; i < device_ids.size(); i++)  // apiMounts, :429-441
      o += std::string(i ? "," : "") + "{\"container_path\":\"" + json_escape(path_join(kListPath, device_ids[i])) +
           "\",\"host_path\":\"" + kNullPath + "\"}";
  o += "],\"devices\":[";
  if (p->pass_device_specs) {  // apiDeviceSpecs, :443-480
    bool first = true;
    auto spec = [&](const std::string &path) {
      o += std::string(first ? "" : ",") + "{\"container_path\":\"" + json_escape(path) + "\",\"host_path\":\"" +
           json_escape(path_join(p->driver_root, path)) + "\",\"permissions\":\"rw\"}";
      first = false;
    };
    const int ctl_mask = control_nodes_of(p);
    for (int i = 0; i < 4; i++)
      if (ctl_mask >> i & 1) spec(kControlNodes[i]);
    for (const Device &d : p->cached)
      for (const std::string &u : uuids)
        if (d.id == u)
          for (const std::string &path : d.paths) spec(path);
  }
  o += "]}";
  return copy_out(o, json, cap);
}

}  // extern "C"
