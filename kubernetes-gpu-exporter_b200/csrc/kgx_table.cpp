// kgx_table.cpp — cold path: turn the caller's join table (podPIDMap,
// main.go:84-114) into the device plan.  See kgx_table.hpp.
#include "kgx_table.hpp"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <string_view>
#include <unordered_map>

namespace kgx {

namespace {

// main.go:22-35, sorted by name as Registry.Gather emits them (O1)
const std::vector<FamilyDef> kRFamilies = {
    {"docker_gpu_memory_perc_usage", "GPU memory in percentage used by pod", V_R_PERC},
    {"pod_gpu_memory_usage", "GPU memory used by Kubernetes Pod", V_R_USED},
};
// X mode (north-star extension): the reference's two names over the wider
// label set plus four more; sorted by name.
const std::vector<FamilyDef> kXFamilies = {
    {"docker_gpu_memory_perc_usage", "GPU memory in percentage used by pod", V_X_PERC},
    {"pod_gpu_last_sample_timestamp_seconds", "Unix time of the newest sample of the pod", V_X_TS},
    {"pod_gpu_mem_utilization", "Mean memory utilization of the pod's processes in percent", V_X_MEM},
    {"pod_gpu_memory_usage", "GPU memory used by Kubernetes Pod", V_X_SUM},
    {"pod_gpu_memory_usage_max", "Largest GPU memory used by one process of the pod", V_X_MAX},
    {"pod_gpu_sm_utilization", "Mean SM utilization of the pod's processes in percent", V_X_SM},
};
const char* const kRLabels[] = {"pid", "pod"};                          // main.go:27,34 (O2: sorted)
const char* const kXLabels[] = {"container", "gpu", "namespace", "pod"};  // O2

// utf8.ValidString: client_golang panics on invalid label values
bool valid_utf8(std::string_view s) {
  const unsigned char* p = reinterpret_cast<const unsigned char*>(s.data());
  size_t n = s.size(), i = 0;
  while (i < n) {
    unsigned char c = p[i];
    if (c < 0x80) { ++i; continue; }
    int len; uint32_t cp, min;
    if ((c & 0xE0) == 0xC0) { len = 2; cp = c & 0x1F; min = 0x80; }
    else if ((c & 0xF0) == 0xE0) { len = 3; cp = c & 0x0F; min = 0x800; }
    else if ((c & 0xF8) == 0xF0) { len = 4; cp = c & 0x07; min = 0x10000; }
    else return false;
    if (i + len > n) return false;
    for (int k = 1; k < len; ++k) {
      if ((p[i + k] & 0xC0) != 0x80) return false;
      cp = (cp << 6) | (p[i + k] & 0x3F);
    }
    if (cp < min || cp > 0x10FFFF || (cp >= 0xD800 && cp <= 0xDFFF)) return false;
    i += len;
  }
  return true;
}

// expfmt writeEscapedString for label values (G2)
void append_escaped(std::string& out, std::string_view s) {
  for (char c : s) {
    if (c == '\\') out += "\\\\";
    else if (c == '\n') out += "\\n";
    else if (c == '"') out += "\\\"";
    else out += c;
  }
}

// Orders u32 values as their decimal strings compare byte-wise ("10" < "9").
inline uint64_t decimal_string_key(uint32_t v) {
  static const uint64_t p10[11] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull,
                                   10000000ull, 100000000ull, 1000000000ull, 10000000000ull};
  int nd = 1;
  while (nd < 10 && v >= p10[nd]) ++nd;
  return (uint64_t(v) * p10[10 - nd]) << 4 | uint64_t(nd);
}

std::string_view sv(const char* blob, const uint64_t* off, uint64_t i) {
  return std::string_view(blob + off[i], size_t(off[i + 1] - off[i]));
}

uint32_t table_capacity(size_t entries) {
  size_t cap = entries * 2;
  if (cap < 8) cap = 8;
  return uint32_t((cap + 7) & ~size_t(7));
}

// Insert (key -> val) cells of one node; `cells` sorted by key so equal keys
// are adjacent and the "more" bit can be chained.
void fill_node_table(std::vector<Entry>& tab, uint32_t off, uint32_t cap,
                     const std::vector<std::pair<uint32_t, uint32_t>>& cells) {
  for (uint32_t i = 0; i < cap; ++i) tab[off + i] = Entry{0, kEmpty};
  uint32_t prev_pos = 0;
  for (size_t i = 0; i < cells.size(); ++i) {
    uint32_t key = cells[i].first;
    uint32_t pos = uint32_t((uint64_t(hash_pid(key)) * cap) >> 32);
    while (tab[off + pos].val != kEmpty) pos = pos + 1 == cap ? 0 : pos + 1;
    tab[off + pos] = Entry{key, cells[i].second};
    if (i > 0 && cells[i - 1].first == key) tab[off + prev_pos].val |= kMore;
    prev_pos = pos;
  }
}

}  // namespace

const std::vector<FamilyDef>& families(int mode) { return mode == KGX_MODE_X ? kXFamilies : kRFamilies; }
const char* const* label_names(int mode, int* n) {
  if (mode == KGX_MODE_X) { *n = 4; return kXLabels; }
  *n = 2; return kRLabels;
}

int build_plan(int mode, const kgx_table* t, const std::vector<Carried>& carried, Plan* out, std::string* err) {
  auto fail = [&](int code, const char* msg) { if (err) *err = msg; return code; };
  if (!t || t->n_nodes == 0 || t->gpus_per_node == 0) return fail(KGX_E_INVAL, "table: n_nodes and gpus_per_node must be >= 1");
  if (t->n_podkeys && (!t->podkey_node || !t->pod_blob || !t->pod_off)) return fail(KGX_E_INVAL, "table: podkey arrays missing");
  if (t->n_entries && (!t->entry_podkey || !t->entry_pid)) return fail(KGX_E_INVAL, "table: entry arrays missing");
  const bool x = mode == KGX_MODE_X;
  if (x && t->n_podkeys && (!t->namespace_blob || !t->namespace_off || !t->container_blob || !t->container_off))
    return fail(KGX_E_INVAL, "table: X mode needs namespace and container strings");
  if (x && (!t->gpu_uuid_blob || !t->gpu_uuid_off)) return fail(KGX_E_INVAL, "table: X mode needs gpu uuid strings");
  if (t->n_podkeys >= 0x7fffffffull || t->n_entries >= 0x7fffffffull) return fail(KGX_E_LIMIT, "table: more than 2^31-1 podkeys or entries");

  Plan& p = *out;
  p = Plan();
  p.mode = mode; p.n_nodes = t->n_nodes; p.G = t->gpus_per_node;
  const uint32_t NN = p.n_nodes, G = p.G;
  const size_t npk = size_t(t->n_podkeys);

  for (size_t i = 0; i < npk; ++i) {
    if (t->podkey_node[i] >= NN) return fail(KGX_E_INVAL, "table: podkey_node out of range");
    if (t->pod_off[i + 1] < t->pod_off[i]) return fail(KGX_E_INVAL, "table: pod_off not ascending");
    if (!valid_utf8(sv(t->pod_blob, t->pod_off, i))) return fail(KGX_E_INVAL, "table: pod name is not valid UTF-8");
    if (x) {
      if (!valid_utf8(sv(t->namespace_blob, t->namespace_off, i)) || !valid_utf8(sv(t->container_blob, t->container_off, i)))
        return fail(KGX_E_INVAL, "table: namespace/container is not valid UTF-8");
    }
  }
  for (uint64_t e = 0; e < t->n_entries; ++e)
    if (t->entry_podkey[e] >= npk) return fail(KGX_E_INVAL, "table: entry_podkey out of range");

  // group podkeys and entries by node
  std::vector<std::vector<uint32_t>> node_pks(NN);
  std::vector<uint32_t> pk_local(npk);  // position of a podkey among the podkeys of its node
  for (size_t i = 0; i < npk; ++i) { pk_local[i] = uint32_t(node_pks[t->podkey_node[i]].size()); node_pks[t->podkey_node[i]].push_back(uint32_t(i)); }
  std::vector<uint32_t> pk_count(npk, 0);  // len(pids) of each podkey (R literal)
  std::vector<uint64_t> node_ent_off(NN + 1, 0);
  for (uint64_t e = 0; e < t->n_entries; ++e) { ++pk_count[t->entry_podkey[e]]; ++node_ent_off[t->podkey_node[t->entry_podkey[e]] + 1]; }
  for (uint32_t n = 0; n < NN; ++n) node_ent_off[n + 1] += node_ent_off[n];
  std::vector<uint32_t> ent_sorted(size_t(t->n_entries));  // entry indices grouped by node
  {
    std::vector<uint64_t> cur(node_ent_off.begin(), node_ent_off.end() - 1);
    for (uint64_t e = 0; e < t->n_entries; ++e) ent_sorted[cur[t->podkey_node[t->entry_podkey[e]]]++] = uint32_t(e);
  }
  std::vector<std::vector<uint32_t>> node_carried(NN);
  if (!x)
    for (size_t i = 0; i < carried.size(); ++i) {
      if (carried[i].node >= NN) return fail(KGX_E_INVAL, "carried series on unknown node");
      node_carried[carried[i].node].push_back(uint32_t(i));
    }

  p.node_slot_base.assign(NN + 1, 0);
  p.node_tab.resize(NN);
  p.A_off.assign(1, 0);
  p.B_off.assign(1, 0);
  p.G_off.assign(1, 0);
  p.carried_slot.assign(carried.size(), 0);
  uint64_t tab_total = 0, slot_total = 0;
  std::vector<std::pair<uint32_t, uint32_t>> cells;

  if (x) {
    p.grank.assign(size_t(NN) * G, 0);
    p.node_a_base.assign(size_t(NN) + 1, 0);
    for (size_t g = 0; g < size_t(NN) * G; ++g) {
      std::string_view u = sv(t->gpu_uuid_blob, t->gpu_uuid_off, g);
      if (!valid_utf8(u)) return fail(KGX_E_INVAL, "table: gpu uuid is not valid UTF-8");
      append_escaped(p.G_blob, u);
      p.G_off.push_back(p.G_blob.size());
    }
  } else {
    p.G_off.assign(size_t(NN) * G + 1, 0);
  }

  // pass 1: per node ordering, slots, cells
  std::vector<std::vector<std::pair<uint32_t, uint32_t>>> node_cells(NN);
  for (uint32_t n = 0; n < NN; ++n) {
    const std::vector<uint32_t>& pks = node_pks[n];
    p.node_slot_base[n] = uint32_t(slot_total);
    cells.clear();
    if (x) {
      // O3 over (container, gpu, namespace, pod): podkeys ordered by
      // (container, namespace, pod); a slot run per (container, gpu)
      std::vector<uint32_t> ord(pks);
      auto key_cmp = [&](uint32_t a, uint32_t b) {
        int c = sv(t->container_blob, t->container_off, a).compare(sv(t->container_blob, t->container_off, b));
        if (c) return c < 0;
        c = sv(t->namespace_blob, t->namespace_off, a).compare(sv(t->namespace_blob, t->namespace_off, b));
        if (c) return c < 0;
        return sv(t->pod_blob, t->pod_off, a).compare(sv(t->pod_blob, t->pod_off, b)) < 0;
      };
      std::sort(ord.begin(), ord.end(), key_cmp);
      for (size_t i = 1; i < ord.size(); ++i)
        if (!key_cmp(ord[i - 1], ord[i])) return fail(KGX_E_INVAL, "table: duplicate (container, namespace, pod) on one node");
      // gpu label order within the node
      std::vector<uint32_t> gord(G);
      std::iota(gord.begin(), gord.end(), 0u);
      auto gsv = [&](uint32_t g) { return sv(t->gpu_uuid_blob, t->gpu_uuid_off, size_t(n) * G + g); };
      std::sort(gord.begin(), gord.end(), [&](uint32_t a, uint32_t b) { return gsv(a) < gsv(b); });
      for (uint32_t r = 1; r < G; ++r)
        if (gsv(gord[r - 1]) == gsv(gord[r])) return fail(KGX_E_INVAL, "table: duplicate gpu uuid on one node");
      for (uint32_t r = 0; r < G; ++r) p.grank[size_t(n) * G + gord[r]] = r;
      const uint32_t a_base = uint32_t(p.n_a);
      p.node_a_base[n] = a_base;
      p.max_node_pks = std::max<uint32_t>(p.max_node_pks, uint32_t(pks.size()));
      std::unordered_map<uint32_t, uint32_t> a_of_pk;
      a_of_pk.reserve(pks.size() * 2);
      size_t i = 0;
      while (i < ord.size()) {
        size_t j = i + 1;
        std::string_view ct = sv(t->container_blob, t->container_off, ord[i]);
        while (j < ord.size() && sv(t->container_blob, t->container_off, ord[j]) == ct) ++j;
        const uint32_t len = uint32_t(j - i);
        for (uint32_t r = 0; r < G; ++r)
          p.runs.push_back(Run{uint32_t(slot_total) + r * len, len, a_base + uint32_t(i), uint32_t(size_t(n) * G + gord[r]), n});
        for (size_t q = i; q < j; ++q) {
          p.pk_slot0.push_back(uint32_t(slot_total) + uint32_t(q - i));
          p.pk_stride.push_back(len);
          a_of_pk[ord[q]] = a_base + uint32_t(q);
          // label pieces (O2 order: container, gpu, namespace, pod)
          p.A_blob += "{container=\""; append_escaped(p.A_blob, ct); p.A_blob += "\",gpu=\"";
          p.A_off.push_back(p.A_blob.size());
          p.B_blob += "\",namespace=\""; append_escaped(p.B_blob, sv(t->namespace_blob, t->namespace_off, ord[q]));
          p.B_blob += "\",pod=\""; append_escaped(p.B_blob, sv(t->pod_blob, t->pod_off, ord[q])); p.B_blob += "\"}";
          p.B_off.push_back(p.B_blob.size());
        }
        slot_total += uint64_t(len) * G;
        if (slot_total >= 0xfffffff0ull) return fail(KGX_E_LIMIT, "table: more than 2^32 potential series");
        i = j;
      }
      p.n_a += ord.size();
      // cells: distinct (pid, podkey)
      for (uint64_t k = node_ent_off[n]; k < node_ent_off[n + 1]; ++k) {
        uint32_t e = ent_sorted[k];
        cells.emplace_back(t->entry_pid[e], a_of_pk[t->entry_podkey[e]]);
      }
      std::sort(cells.begin(), cells.end());
      cells.erase(std::unique(cells.begin(), cells.end()), cells.end());
    } else {
      // pods of the node: the table's, plus pods that only carried series name
      std::vector<std::string_view> names;
      std::unordered_map<std::string_view, uint32_t> by_name;
      for (uint32_t pk : pks) {
        std::string_view nm = sv(t->pod_blob, t->pod_off, pk);
        if (!by_name.emplace(nm, uint32_t(names.size())).second)
          return fail(KGX_E_INVAL, "table: duplicate pod name on one node (podPIDMap keys are unique)");
        names.push_back(nm);
      }
      for (uint32_t ci : node_carried[n]) {
        std::string_view nm = carried[ci].pod;
        if (by_name.emplace(nm, uint32_t(names.size())).second) names.push_back(nm);
      }
      std::vector<uint32_t> nord(names.size()), nrank(names.size());
      std::iota(nord.begin(), nord.end(), 0u);
      std::sort(nord.begin(), nord.end(), [&](uint32_t a, uint32_t b) { return names[a] < names[b]; });
      for (uint32_t r = 0; r < nord.size(); ++r) nrank[nord[r]] = r;
      const uint32_t pod_base = uint32_t(p.pod_names.size());
      for (auto nm : names) { p.pod_names.emplace_back(nm); p.pod_node.push_back(n); }
      // pairs (pid label, pod): one packed 64-bit order key each = O3's (pid string, pod string) order.
      // src: -1 = from the table (joinable), otherwise the index of the carried series it stands for.
      struct Pair { uint64_t key; uint32_t pid, pod; int32_t src; };
      if (names.size() >= (1u << 24)) return fail(KGX_E_LIMIT, "table: more than 2^24 pods on one node");
      auto order_key = [&](uint32_t pid, uint32_t li) { return (decimal_string_key(pid) << 24) | nrank[li]; };
      std::vector<Pair> pairs;
      if (mode == KGX_MODE_R_LITERAL) {
        // main.go:144: `for pid := range pids` -> indices 0..len(pids)-1
        for (size_t li = 0; li < pks.size(); ++li)
          for (uint32_t i = 0; i < pk_count[pks[li]]; ++i) pairs.push_back(Pair{order_key(i, uint32_t(li)), i, uint32_t(li), -1});
      } else {
        pairs.reserve(size_t(node_ent_off[n + 1] - node_ent_off[n]) + node_carried[n].size());
        for (uint64_t k = node_ent_off[n]; k < node_ent_off[n + 1]; ++k) {
          const uint32_t e = ent_sorted[k];
          const uint32_t li = pk_local[t->entry_podkey[e]];  // podkeys of a node are `names` in table order
          pairs.push_back(Pair{order_key(t->entry_pid[e], li), t->entry_pid[e], li, -1});
        }
      }
      for (uint32_t ci : node_carried[n]) {
        const uint32_t li = by_name[std::string_view(carried[ci].pod)];
        pairs.push_back(Pair{order_key(carried[ci].pid, li), carried[ci].pid, li, int32_t(ci)});
      }
      std::sort(pairs.begin(), pairs.end(), [](const Pair& a, const Pair& b) { return a.key != b.key ? a.key < b.key : a.src < b.src; });
      // one slot per distinct key; a table copy (src -1) sorts first and makes the slot joinable
      size_t n_new = 0;
      p.A_blob.reserve(p.A_blob.size() + pairs.size() * 40);
      for (size_t i = 0; i < pairs.size();) {
        const uint32_t slot = uint32_t(slot_total + n_new);
        if (pairs[i].src < 0) cells.emplace_back(pairs[i].pid, slot);
        size_t j = i;
        for (; j < pairs.size() && pairs[j].key == pairs[i].key; ++j)
          if (pairs[j].src >= 0) p.carried_slot[size_t(pairs[j].src)] = slot;
        p.slot_pid.push_back(pairs[i].pid);
        p.slot_pod.push_back(pod_base + pairs[i].pod);
        char num[12];  // fmt.Sprintf("%d", pid), main.go:147
        int nd = 0;
        for (uint32_t v = pairs[i].pid; nd == 0 || v; v /= 10) num[11 - nd++] = char('0' + v % 10);
        p.A_blob += "{pid=\""; p.A_blob.append(num + 12 - nd, size_t(nd)); p.A_blob += "\",pod=\"";
        append_escaped(p.A_blob, names[pairs[i].pod]); p.A_blob += "\"}";
        p.A_off.push_back(p.A_blob.size());
        ++n_new;
        i = j;
      }
      if (slot_total + n_new >= 0x7ffffff0ull) return fail(KGX_E_LIMIT, "table: more than 2^31 potential series");
      p.runs.push_back(Run{uint32_t(slot_total), uint32_t(n_new), uint32_t(slot_total), 0xFFFFFFFFu, n});
      slot_total += n_new;
      p.n_a += n_new;
      std::sort(cells.begin(), cells.end());
    }
    p.node_tab[n] = NodeTab{uint32_t(tab_total), table_capacity(cells.size())};
    p.max_node_cells = std::max(p.max_node_cells, p.node_tab[n].cap);
    tab_total += p.node_tab[n].cap;
    if (tab_total >= 0xfffffff0ull) return fail(KGX_E_LIMIT, "table: hash table beyond 2^32 cells");
    p.n_table_entries += cells.size();
    node_cells[n] = cells;
  }
  p.node_slot_base[NN] = uint32_t(slot_total);
  if (x) p.node_a_base[NN] = uint32_t(p.n_a);
  p.n_slots = slot_total;
  if (!x) p.B_off.assign(size_t(p.n_a) + 1, 0);
  // pass 2: the open-addressing table
  p.tab.resize(size_t(tab_total));
  for (uint32_t n = 0; n < NN; ++n) fill_node_table(p.tab, p.node_tab[n].off, p.node_tab[n].cap, node_cells[n]);
  // longest label text any slot can have
  uint64_t max_g = 0;
  for (size_t g = 0; g + 1 < p.G_off.size(); ++g) max_g = std::max<uint64_t>(max_g, p.G_off[g + 1] - p.G_off[g]);
  uint64_t max_ab = 0;
  for (size_t a = 0; a < size_t(p.n_a); ++a)
    max_ab = std::max<uint64_t>(max_ab, (p.A_off[a + 1] - p.A_off[a]) + (p.B_off[a + 1] - p.B_off[a]));
  if (max_ab + max_g > 0x3fffffffull) return fail(KGX_E_LIMIT, "table: label text too long");
  p.max_label_len = uint32_t(max_ab + max_g);
  return KGX_OK;
}

}  // namespace kgx
