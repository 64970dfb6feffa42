// Replica-id helpers and the replica-aware prioritiser, host C++ behind the C ABI.
//
// ref: cmd/nvidia-device-plugin/replica.go (stripReplica :28, stripReplicas :32,
// devCount :62-84, prioritizeDevices :96-198) and getAdditionalXids,
// cmd/nvidia-device-plugin/nvidia.go:274-294.  These are string-keyed O(size x #GPUs)
// procedures with no parallel work, so they stay on the host; what
// changes against the Go code is the data layout: ids are string_views into the caller's
// buffers, physical GPUs live in one sorted vector, each group is an index window.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>

#include "../../include/gpualloc.h"

namespace {

constexpr std::string_view kJoin = "-replica-";  // replica.go:26

// everything before the first "-replica-" (strings.Split(id, joinStr)[0], replica.go:28-30)
std::string_view physical_of(std::string_view id) {
  size_t p = id.find(kJoin);
  return p == std::string_view::npos ? id : id.substr(0, p);
}

int pack_out(const std::vector<std::string_view> &v, char *outbuf, int outbuf_len, int *n_out) {
  size_t need = 0;
  for (auto s : v) need += s.size() + 1;
  if ((long long)need > (long long)outbuf_len) return GPA_ERR_BUFFER;
  char *w = outbuf;
  for (auto s : v) {
    memcpy(w, s.data(), s.size());
    w[s.size()] = 0;
    w += s.size() + 1;
  }
  *n_out = (int)v.size();
  return GPA_OK;
}

// One free replica id: the whole id plus its tail after the physical id.  Ids of one group share the physical id
// and the separator, so ordering them by the tail is ordering them byte-wise as whole strings (sort.Strings, :116)
// at a fraction of the comparison cost.
struct Replica {
  std::string_view id;
  std::string_view tail;
  uint64_t key;  // the tail's bytes after "-replica-", big-endian and zero-padded: u64 order == byte-wise string order
};

// false when the tail does not have the usual shape ("-replica-" + at most 8 bytes): the group then sorts by string
bool tail_key(std::string_view tail, uint64_t *key) {
  if (tail.size() < kJoin.size() || tail.size() > kJoin.size() + 8 || memcmp(tail.data(), kJoin.data(), kJoin.size()) != 0) return false;
  uint64_t k = 0;
  for (size_t i = kJoin.size(); i < kJoin.size() + 8; i++) k = (k << 8) | (i < tail.size() ? (unsigned char)tail[i] : 0u);
  *key = k;
  return true;
}

struct Group {
  bool keyed = true;  // every replica of the group has a packed key
  std::string_view phys;
  std::vector<Replica> names;  // free replicas; [head, names.size()) is the live window
  size_t head = 0;
  bool allocated = false;
  size_t free_count() const { return names.size() - head; }
};

}  // namespace

namespace {

// strings.TrimSpace trims every rune with unicode.IsSpace: '\t' '\n' '\v' '\f' '\r' ' ', U+0085, U+00A0 and the Unicode
// Z category / White_Space property: U+1680, U+2000..U+200A, U+2028, U+2029, U+202F, U+205F, U+3000 (as UTF-8 here).
// length in bytes of the white-space rune that STARTS at p (0 = none)
size_t space_rune_at(const unsigned char *p, size_t n) {
  if (n == 0) return 0;
  const unsigned char c = p[0];
  if (c == ' ' || (c >= '\t' && c <= '\r')) return 1;
  if (c == 0xC2 && n >= 2 && (p[1] == 0x85 || p[1] == 0xA0)) return 2;
  if (n >= 3) {
    if (c == 0xE1 && p[1] == 0x9A && p[2] == 0x80) return 3;                               // U+1680
    if (c == 0xE2 && p[1] == 0x80 && ((p[2] >= 0x80 && p[2] <= 0x8A) || p[2] == 0xA8 || p[2] == 0xA9 || p[2] == 0xAF)) return 3;
    if (c == 0xE2 && p[1] == 0x81 && p[2] == 0x9F) return 3;                               // U+205F
    if (c == 0xE3 && p[1] == 0x80 && p[2] == 0x80) return 3;                               // U+3000
  }
  return 0;
}

std::string_view go_trim_space(std::string_view v) {
  for (;;) {
    const size_t k = space_rune_at(reinterpret_cast<const unsigned char *>(v.data()), v.size());
    if (k == 0) break;
    v.remove_prefix(k);
  }
  for (;;) {
    size_t k = 0;
    for (size_t len = 1; len <= 3 && len <= v.size() && k == 0; len++) {
      const unsigned char *p = reinterpret_cast<const unsigned char *>(v.data()) + v.size() - len;
      if (space_rune_at(p, len) == len) k = len;
    }
    if (k == 0) break;
    v.remove_suffix(k);
  }
  return v;
}

}  // namespace


extern "C" {

int gpa_strip_replicas(const char *const *ids, int n, char *outbuf, int outbuf_len, int *n_out) {
  if (n < 0 || (n > 0 && !ids) || !n_out || (!outbuf && outbuf_len > 0)) return GPA_ERR_INVALID;
  std::vector<std::string_view> uniq;
  uniq.reserve((size_t)n);
  for (int i = 0; i < n; i++) uniq.push_back(physical_of(ids[i]));
  // dedupe + sort.Strings (replica.go:38-43): the first-seen order is erased by the sort
  std::sort(uniq.begin(), uniq.end());
  uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
  return pack_out(uniq, outbuf, outbuf_len, n_out);
}

int gpa_prioritize_devices(const char *const *avail, int n_avail, const char *const *must, int n_must, int size,
                           char *outbuf, int outbuf_len, int *n_out, int *non_unique, char *errbuf, int errbuf_len) {
  if (n_avail < 0 || n_must < 0 || (n_avail > 0 && !avail) || (n_must > 0 && !must) || !n_out) return GPA_ERR_INVALID;
  *n_out = 0;
  if (non_unique) *non_unique = 0;
  if (errbuf && errbuf_len > 0) errbuf[0] = 0;

  // replica.go:98-117 — group the replica ids by physical GPU, each group sorted byte-wise.  A node has a handful of
  // physical GPUs, so the group of an id is found by scanning them (length first, then bytes); the hash map only takes
  // over on unusually wide nodes.
  std::vector<Group> groups;
  groups.reserve(16);
  std::unordered_map<std::string_view, size_t> index;
  auto group_of = [&](std::string_view phys) -> Group * {
    if (groups.size() <= 32) {
      for (Group &g : groups)
        if (g.phys.size() == phys.size() && memcmp(g.phys.data(), phys.data(), phys.size()) == 0) return &g;
      return nullptr;
    }
    if (index.empty())
      for (size_t i = 0; i < groups.size(); i++) index.emplace(groups[i].phys, i);
    auto it = index.find(phys);
    return it == index.end() ? nullptr : &groups[it->second];
  };
  for (int i = 0; i < n_avail; i++) {
    std::string_view id = avail[i];
    std::string_view phys = physical_of(id);
    Group *g = group_of(phys);
    if (!g) {
      groups.emplace_back();
      g = &groups.back();
      g->phys = phys;
      g->names.reserve((size_t)(n_avail < 64 ? n_avail : 64));  // a small request must not pay for a big node
      if (!index.empty()) index.emplace(phys, groups.size() - 1);
    }
    Replica rep{id, id.substr(phys.size()), 0};
    g->keyed = tail_key(rep.tail, &rep.key) && g->keyed;
    g->names.push_back(rep);
  }
  for (auto &g : groups) {
    if (g.keyed)
      std::sort(g.names.begin(), g.names.end(), [](const Replica &a, const Replica &b) { return a.key < b.key; });
    else
      std::sort(g.names.begin(), g.names.end(), [](const Replica &a, const Replica &b) { return a.tail < b.tail; });
  }

  // replica.go:121 — make([]string, len(must), allocationSize)
  if (size < 0 || n_must > size) return GPA_ERR_PANIC;

  std::vector<std::string_view> allocated;
  allocated.reserve((size_t)size);
  bool unique = true;

  // replica.go:126-139 — must-includes; removal swaps the last free replica into the hole (:56-60)
  for (int i = 0; i < n_must; i++) {
    std::string_view id = must[i];
    Group *g = group_of(physical_of(id));
    size_t at = 0;
    bool found = false;
    if (g) {
      for (at = g->head; at < g->names.size(); at++)
        if (g->names[at].id == id) {
          found = true;
          break;
        }
    }
    if (g && g->allocated) unique = false;  // :131-134 is evaluated before the replica lookup fails (:135)
    if (!found) {
      if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%.*s", (int)id.size(), id.data());
      return GPA_ERR_MUST_MISSING;
    }
    g->names[at] = g->names.back();
    g->names.pop_back();
    g->allocated = true;
    allocated.push_back(id);
  }

  // replica.go:143-147 — physical ids in byte-wise order decide ties
  std::vector<Group *> order;
  order.reserve(groups.size());
  for (auto &g : groups) order.push_back(&g);
  std::sort(order.begin(), order.end(), [](const Group *a, const Group *b) { return a->phys < b->phys; });

  // replica.go:149-190 — prefer an untouched GPU with the most free replicas, else a touched one
  for (int i = (int)allocated.size(); i < size; i++) {
    Group *fresh = nullptr, *used = nullptr;
    size_t fresh_best = 0, used_best = 0;
    for (Group *g : order) {
      size_t c = g->free_count();
      if (g->allocated) {
        if (c > used_best) {
          used = g;
          used_best = c;
        }
      } else if (c > fresh_best) {
        fresh = g;
        fresh_best = c;
      }
    }
    Group *pick = fresh ? fresh : used;
    if (!pick) return GPA_ERR_NO_DEVICES_LEFT;
    if (pick->allocated) unique = false;
    allocated.push_back(pick->names[pick->head++].id);  // allocateAny(): the group's first free replica (:79-84)
    pick->allocated = true;
  }
  std::sort(allocated.begin(), allocated.end());  // :191
  int rc = pack_out(allocated, outbuf, outbuf_len, n_out);
  if (rc != GPA_OK) return rc;
  if (non_unique) *non_unique = unique ? 0 : 1;  // :193-197 &NonUniqueError{}
  return GPA_OK;
}

int gpa_get_additional_xids(const char *input, uint64_t *out, int out_cap) {
  if (!input || (!out && out_cap > 0)) return GPA_ERR_INVALID;
  std::string_view s = input;
  if (s.empty()) return 0;  // nvidia.go:275-277
  int n = 0;
  size_t pos = 0;
  for (;;) {
    size_t comma = s.find(',', pos);
    std::string_view tok = s.substr(pos, comma == std::string_view::npos ? std::string_view::npos : comma - pos);
    tok = go_trim_space(tok);
    if (!tok.empty()) {
      // strconv.ParseUint(tok, 10, 64): decimal digits only, no sign, no underscore, <= 2^64-1
      bool ok = true;
      uint64_t v = 0;
      for (char ch : tok) {
        if (ch < '0' || ch > '9') {
          ok = false;
          break;
        }
        uint64_t d = (uint64_t)(ch - '0');
        if (v > (UINT64_MAX - d) / 10) {
          ok = false;
          break;
        }
        v = v * 10 + d;
      }
      if (ok) {
        if (n >= out_cap) return GPA_ERR_BUFFER;
        out[n++] = v;
      }
    }
    if (comma == std::string_view::npos) break;
    pos = comma + 1;
  }
  return n;
}

// parseResourceConfig, cmd/nvidia-device-plugin/main.go:171-203
int gpa_parse_resource_config(const char *config, gpa_resource_variant *out, int cap, int *n_out, char *errbuf, int errbuf_len) {
  if (!config || !n_out || cap < 0 || (cap > 0 && !out)) return GPA_ERR_INVALID;
  *n_out = 0;
  auto fail = [&](const char *msg) {
    if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%s", msg);
    return GPA_ERR_INVALID;
  };
  if (errbuf && errbuf_len > 0) errbuf[0] = 0;
  std::string_view s = config;
  int n = 0;
  size_t pos = 0;
  for (;;) {
    const size_t comma = s.find(',', pos);
    std::string_view entry = go_trim_space(s.substr(pos, comma == std::string_view::npos ? std::string_view::npos : comma - pos));  // :175
    if (!entry.empty()) {                                                                                                          // :176-178
      std::string_view parts[3];
      int np = 0;
      size_t q = 0;
      for (;;) {  // strings.Split(entry, ":")
        const size_t colon = entry.find(':', q);
        if (np < 3) parts[np] = entry.substr(q, colon == std::string_view::npos ? std::string_view::npos : colon - q);
        np++;
        if (colon == std::string_view::npos) break;
        q = colon + 1;
      }
      if (np != 3) return fail("an entry must have three parts separated by a semicolon");  // :180-182
      // strconv.ParseInt(parts[2], 10, 32): optional sign, decimal digits only, range of int32   (:185-188)
      std::string_view num = parts[2];
      bool neg = false;
      if (!num.empty() && (num[0] == '+' || num[0] == '-')) {
        neg = num[0] == '-';
        num.remove_prefix(1);
      }
      if (num.empty()) return fail("replica must be an integer");
      long long v = 0;
      for (char ch : num) {
        if (ch < '0' || ch > '9') return fail("replica must be an integer");
        v = v * 10 + (ch - '0');
        if (v > 2147483648ll) return fail("replica must be an integer");  // out of range is an error too
      }
      if (neg) v = -v;
      if (v > 2147483647ll) return fail("replica must be an integer");
      gpa_resource_variant var;
      memset(&var, 0, sizeof var);
      if (parts[0].size() >= sizeof var.orig || parts[1].size() >= sizeof var.name) return GPA_ERR_BUFFER;
      memcpy(var.orig, parts[0].data(), parts[0].size());
      memcpy(var.name, parts[1].data(), parts[1].size());
      var.auto_replicas = v == -1;            // :189-193
      if (v == -1) v = 1;
      var.replicas = (unsigned long long)v;   // uint(replicas): a negative count wraps, as in Go
      int slot = -1;
      for (int i = 0; i < n; i++)
        if (strcmp(out[i].orig, var.orig) == 0) slot = i;  // the map keeps one variant per origName (:195)
      if (slot < 0) {
        if (n >= cap) return GPA_ERR_BUFFER;
        slot = n++;
      }
      out[slot] = var;
    }
    if (comma == std::string_view::npos) break;
    pos = comma + 1;
  }
  *n_out = n;
  return GPA_OK;
}

// resourceConfiguration.Get, mig-strategy.go:66-76
void gpa_resource_config_get(const gpa_resource_variant *variants, int n, const char *orig, gpa_resource_variant *out) {
  if (!out) return;
  memset(out, 0, sizeof *out);
  if (!orig) return;
  for (int i = 0; i < n && variants; i++)
    if (strcmp(variants[i].orig, orig) == 0) {
      *out = variants[i];
      return;
    }
  snprintf(out->orig, sizeof out->orig, "%s", orig);
  snprintf(out->name, sizeof out->name, "%s", orig);
}

}  // extern "C"
