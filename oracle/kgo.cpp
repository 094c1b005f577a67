// kgo.cpp — ORACLE (test infrastructure): CPU restatement of the reference's
// join -> value calc -> upsert -> gather -> text-encode path.
//
// What is restated and from where (citations into /root/reference):
//   * schema: two gauge families, help text, labels pid,pod      main.go:21-36
//   * join loop and predicate (literal: slice INDEX == Pid;
//     intended: PID value equality)                              main.go:141-145
//   * values float64(used), float64(used)/float64(total)*100     main.go:147,149
//   * upsert: GaugeVec.WithLabelValues(...).Set(v): get-or-create
//     a child per label-value tuple, last write wins, children
//     are never deleted                                          main.go:147,150
//   * the pod->PID table is rebuilt each cycle, children persist main.go:84,113
//   * gather + text exposition served by promhttp                main.go:69-71
// Third-party rules (client_golang v1.20.2 registry.go / internal/metric.go,
// common v0.55.0 expfmt — go.mod:7,36; not vendored): O1 families sorted by
// name, empty families omitted; O2 label pairs sorted by name; O3 series sorted
// by label values; G1 grammar; G2 escaping.  See kgo.h for the parity status.
//
// X mode (key pod,namespace,container,gpu; sum/max/last/count) is the
// north-star extension: its values are a builder decision, only
// its encoding follows the rules above.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>
#include "kgo.h"

namespace {

struct Child {            // one GaugeVec child pair (both vecs are Set together)
  std::string pid_label;  // fmt.Sprintf("%d", pid), main.go:147,150
  uint32_t podkey;        // the pod whose name is the pod label
  std::string pod;        // copy of the name: children outlive the table
  double used;            // podGpuMemoryUsed value
  double perc;            // podGpuMemoryPercUsed value
};

struct Acc {  // X mode per (podkey, gpu)
  uint64_t sum_used = 0, max_used = 0, last_ts = 0, sum_sm = 0, sum_mem = 0, count = 0;
};

struct Series {
  std::string_view lv[4];
  int nl;
  double v;
};

struct FamilyDef {
  const char* name;
  const char* help;
};
// main.go:22-35
const FamilyDef kRFamilies[2] = {
    {"pod_gpu_memory_usage", "GPU memory used by Kubernetes Pod"},
    {"docker_gpu_memory_perc_usage", "GPU memory in percentage used by pod"},
};
// X mode: the two reference names with the wider label set plus four new ones
const FamilyDef kXFamilies[6] = {
    {"pod_gpu_memory_usage", "GPU memory used by Kubernetes Pod"},
    {"docker_gpu_memory_perc_usage", "GPU memory in percentage used by pod"},
    {"pod_gpu_memory_usage_max", "Largest GPU memory used by one process of the pod"},
    {"pod_gpu_sm_utilization", "Mean SM utilization of the pod's processes in percent"},
    {"pod_gpu_mem_utilization", "Mean memory utilization of the pod's processes in percent"},
    {"pod_gpu_last_sample_timestamp_seconds", "Unix time of the newest sample of the pod"},
};

}  // namespace

struct kgo_ctx {
  int mode, join, threads;
  bool have_table = false;
  uint32_t n_nodes = 0, G = 0;
  std::vector<uint32_t> pk_node;
  std::vector<std::string> pod, ns, ct, uuid;
  std::vector<std::vector<uint32_t>> pk_pids;   // podPIDMap values, entry order
  std::vector<std::vector<uint32_t>> node_pks;  // podkeys of each node
  // per node: pid -> podkeys listing it (KGO_JOIN_HASH)
  std::vector<std::unordered_map<uint32_t, std::vector<uint32_t>>> pid_index;
  // R modes: per node GaugeVec children keyed by label values joined with 0xff
  std::vector<std::unordered_map<std::string, Child>> children;
  // X mode
  std::vector<Acc> acc;  // [n_podkeys * G]
  std::vector<uint64_t> gpu_total;  // memoryInfo.Total of the latest ingest (X render)
  std::string body;
  std::vector<uint64_t> body_off;
};

namespace {

void append_escaped(std::string& out, std::string_view s, bool is_label) {
  for (char c : s) {
    if (c == '\\') out += "\\\\";
    else if (c == '\n') out += "\\n";
    else if (c == '"' && is_label) out += "\\\"";
    else out += c;
  }
}

// expfmt MetricFamilyToText for a gauge family (G1)
void encode_family(std::string& out, const char* name, const char* help, const char* const* label_names,
                   std::vector<Series>& series) {
  if (series.empty()) return;  // a family without metrics is not gathered at all
  // O3: internal.MetricSorter — label values compared pairwise as strings
  std::sort(series.begin(), series.end(), [](const Series& a, const Series& b) {
    for (int i = 0; i < a.nl; ++i) {
      int c = a.lv[i].compare(b.lv[i]);
      if (c != 0) return c < 0;
    }
    return false;
  });
  out += "# HELP "; out += name; out += ' '; append_escaped(out, help, false); out += '\n';
  out += "# TYPE "; out += name; out += " gauge\n";
  char num[32];
  for (const Series& s : series) {
    out += name;
    out += '{';
    for (int i = 0; i < s.nl; ++i) {
      if (i) out += ',';
      out += label_names[i]; out += "=\""; append_escaped(out, s.lv[i], true); out += '"';
    }
    out += "} ";
    out.append(num, size_t(kgo_format_value(s.v, num)));
    out += '\n';
  }
}

double family_value_x(int fam, const Acc& a, uint64_t total) {
  switch (fam) {
    case 0: return double(a.sum_used);
    case 1: return (double(a.sum_used) / double(total)) * 100;
    case 2: return double(a.max_used);
    case 3: return double(a.sum_sm) / double(a.count);
    case 4: return double(a.sum_mem) / double(a.count);
    default: return double(a.last_ts) / 1e6;
  }
}

void render_node(kgo_ctx* c, uint32_t node, const std::vector<uint64_t>& gpu_total, std::string& out) {
  if (c->mode != KGX_MODE_X) {
    static const char* const names[2] = {"pid", "pod"};  // O2: already sorted
    // O1: families sorted by name -> docker_gpu_memory_perc_usage first
    int order[2] = {0, 1};
    std::sort(order, order + 2, [](int a, int b) { return std::strcmp(kRFamilies[a].name, kRFamilies[b].name) < 0; });
    for (int f : order) {
      std::vector<Series> series;
      series.reserve(c->children[node].size());
      for (auto& kv : c->children[node]) {
        const Child& ch = kv.second;
        Series s; s.nl = 2; s.lv[0] = ch.pid_label; s.lv[1] = ch.pod;
        s.v = f == 0 ? ch.used : ch.perc;
        series.push_back(s);
      }
      encode_family(out, kRFamilies[f].name, kRFamilies[f].help, names, series);
    }
  } else {
    static const char* const names[4] = {"container", "gpu", "namespace", "pod"};  // O2
    int order[6] = {0, 1, 2, 3, 4, 5};
    std::sort(order, order + 6, [](int a, int b) { return std::strcmp(kXFamilies[a].name, kXFamilies[b].name) < 0; });
    for (int f : order) {
      std::vector<Series> series;
      for (uint32_t pk : c->node_pks[node])
        for (uint32_t g = 0; g < c->G; ++g) {
          const Acc& a = c->acc[size_t(pk) * c->G + g];
          if (a.count == 0) continue;
          Series s; s.nl = 4;
          s.lv[0] = c->ct[pk]; s.lv[1] = c->uuid[size_t(node) * c->G + g]; s.lv[2] = c->ns[pk]; s.lv[3] = c->pod[pk];
          s.v = family_value_x(f, a, gpu_total[size_t(node) * c->G + g]);
          series.push_back(s);
        }
      encode_family(out, kXFamilies[f].name, kXFamilies[f].help, names, series);
    }
  }
}

// GaugeVec.WithLabelValues(pid, pod).Set(v) for both vecs (main.go:147-150)
void upsert(kgo_ctx* c, uint32_t node, uint32_t pid_label_num, uint32_t pk, uint64_t used, uint64_t total) {
  std::string pid_label = std::to_string(pid_label_num);  // fmt.Sprintf("%d", pid)
  std::string key = pid_label; key += char(0xff); key += c->pod[pk];
  Child& ch = c->children[node][key];
  if (ch.pid_label.empty()) { ch.pid_label = pid_label; ch.podkey = pk; ch.pod = c->pod[pk]; }
  ch.used = double(used);                               // main.go:147
  ch.perc = (double(used) / double(total)) * 100;       // main.go:149
}

void ingest_node(kgo_ctx* c, const kgx_cols* s, uint32_t node, const uint64_t* idx, size_t n) {
  const uint32_t G = c->G;
  for (size_t ii = 0; ii < n; ++ii) {
    size_t i = idx ? size_t(idx[ii]) : ii;
    uint32_t pid = s->pid[i], gpu = s->gpu[i];
    if (gpu >= G) continue;  // not a device of this node: no row of main.go:123 produces it
    uint64_t total = s->gpu_total[size_t(node) * G + gpu];
    if (c->mode == KGX_MODE_X) {
      auto fold = [&](uint32_t pk) {
        Acc& a = c->acc[size_t(pk) * G + gpu];
        a.sum_used += s->used[i];
        a.max_used = std::max(a.max_used, s->used[i]);
        a.last_ts = std::max(a.last_ts, s->ts_us[i]);
        a.sum_sm += s->sm_util[i];
        a.sum_mem += s->mem_util[i];
        a.count += 1;
      };
      if (c->join == KGO_JOIN_LOOPS) {
        for (uint32_t pk : c->node_pks[node])
          for (uint32_t p : c->pk_pids[pk])
            if (p == pid) { fold(pk); break; }  // a process belongs to a container once
      } else {
        auto it = c->pid_index[node].find(pid);
        if (it != c->pid_index[node].end()) for (uint32_t pk : it->second) fold(pk);
      }
      continue;
    }
    if (c->join == KGO_JOIN_LOOPS) {
      // main.go:143-151, literally (pod order is Go map order: irrelevant,
      // every pod writes a distinct (pid,pod) child)
      for (uint32_t pk : c->node_pks[node]) {
        const std::vector<uint32_t>& pids = c->pk_pids[pk];
        for (size_t j = 0; j < pids.size(); ++j) {
          if (c->mode == KGX_MODE_R_LITERAL) {
            if (j == size_t(pid)) upsert(c, node, uint32_t(j), pk, s->used[i], total);  // main.go:144-145
          } else {
            if (pids[j] == pid) upsert(c, node, pid, pk, s->used[i], total);
          }
        }
      }
    } else if (c->mode == KGX_MODE_R_LITERAL) {
      for (uint32_t pk : c->node_pks[node])
        if (size_t(pid) < c->pk_pids[pk].size()) upsert(c, node, pid, pk, s->used[i], total);
    } else {
      auto it = c->pid_index[node].find(pid);
      if (it != c->pid_index[node].end()) for (uint32_t pk : it->second) upsert(c, node, pid, pk, s->used[i], total);
    }
  }
}

template <class F>
void parallel_nodes(int threads, uint32_t n_nodes, F f) {
  if (threads <= 1 || n_nodes <= 1) { for (uint32_t n = 0; n < n_nodes; ++n) f(n); return; }
  std::vector<std::thread> pool;
  uint32_t t = std::min<uint32_t>(uint32_t(threads), n_nodes);
  for (uint32_t w = 0; w < t; ++w)
    pool.emplace_back([=] { for (uint32_t n = w; n < n_nodes; n += t) f(n); });
  for (auto& th : pool) th.join();
}

}  // namespace

extern "C" {

kgo_ctx* kgo_create(int mode, int join, int threads) {
  if (mode < 0 || mode > 2) return nullptr;
  kgo_ctx* c = new kgo_ctx();
  c->mode = mode; c->join = join; c->threads = threads < 1 ? 1 : threads;
  return c;
}
void kgo_destroy(kgo_ctx* c) { delete c; }

int kgo_set_table(kgo_ctx* c, const kgx_table* t) {
  if (!c || !t || t->n_nodes == 0 || t->gpus_per_node == 0) return KGX_E_INVAL;
  const bool x = c->mode == KGX_MODE_X;
  if (x && (!t->namespace_off || !t->container_off || !t->gpu_uuid_off)) return KGX_E_INVAL;
  const bool first = !c->have_table;
  if (!first && (t->n_nodes != c->n_nodes || t->gpus_per_node != c->G)) return KGX_E_INVAL;
  c->n_nodes = t->n_nodes; c->G = t->gpus_per_node;
  size_t npk = size_t(t->n_podkeys);
  c->pk_node.assign(t->podkey_node, t->podkey_node + npk);
  c->pod.resize(npk); c->ns.assign(npk, ""); c->ct.assign(npk, "");
  for (size_t i = 0; i < npk; ++i) {
    if (c->pk_node[i] >= c->n_nodes) return KGX_E_INVAL;
    c->pod[i].assign(t->pod_blob + t->pod_off[i], t->pod_blob + t->pod_off[i + 1]);
    if (x) {
      c->ns[i].assign(t->namespace_blob + t->namespace_off[i], t->namespace_blob + t->namespace_off[i + 1]);
      c->ct[i].assign(t->container_blob + t->container_off[i], t->container_blob + t->container_off[i + 1]);
    }
  }
  c->pk_pids.assign(npk, {});
  for (uint64_t e = 0; e < t->n_entries; ++e) {
    if (t->entry_podkey[e] >= npk) return KGX_E_INVAL;
    c->pk_pids[t->entry_podkey[e]].push_back(t->entry_pid[e]);
  }
  c->node_pks.assign(c->n_nodes, {});
  for (size_t i = 0; i < npk; ++i) c->node_pks[c->pk_node[i]].push_back(uint32_t(i));
  c->pid_index.assign(c->n_nodes, {});
  for (size_t i = 0; i < npk; ++i) {
    auto& index = c->pid_index[c->pk_node[i]];
    for (uint32_t p : c->pk_pids[i]) {
      auto& v = index[p];
      if (v.empty() || v.back() != uint32_t(i)) {
        if (std::find(v.begin(), v.end(), uint32_t(i)) == v.end()) v.push_back(uint32_t(i));
      }
    }
  }
  c->uuid.assign(size_t(c->n_nodes) * c->G, "");
  if (x)
    for (size_t i = 0; i < c->uuid.size(); ++i)
      c->uuid[i].assign(t->gpu_uuid_blob + t->gpu_uuid_off[i], t->gpu_uuid_blob + t->gpu_uuid_off[i + 1]);
  if (first) c->children.assign(c->n_nodes, {});  // R: children persist across tables
  if (x) c->acc.assign(npk * c->G, Acc());
  c->have_table = true;
  return KGX_OK;
}

int kgo_ingest(kgo_ctx* c, const kgx_cols* s) {
  if (!c || !s) return KGX_E_INVAL;
  if (!c->have_table) return KGX_E_STATE;
  if (!s->gpu_total) return KGX_E_INVAL;
  if (s->n && (!s->pid || !s->gpu || !s->used)) return KGX_E_INVAL;
  if (c->mode == KGX_MODE_X && s->n && (!s->sm_util || !s->mem_util || !s->ts_us)) return KGX_E_INVAL;
  // bucket sample indices by node, keeping sample order inside a node
  std::vector<std::vector<uint64_t>> by_node(c->n_nodes);
  for (uint64_t i = 0; i < s->n; ++i) {
    uint32_t node = s->node ? s->node[i] : 0;
    if (node >= c->n_nodes) continue;
    by_node[node].push_back(i);
  }
  parallel_nodes(c->threads, c->n_nodes,
                 [&](uint32_t n) { ingest_node(c, s, n, by_node[n].data(), by_node[n].size()); });
  c->gpu_total.assign(s->gpu_total, s->gpu_total + size_t(c->n_nodes) * c->G);
  return KGX_OK;
}

int kgo_reset_window(kgo_ctx* c) {
  if (!c) return KGX_E_INVAL;
  if (c->mode != KGX_MODE_X) return KGX_E_STATE;
  std::fill(c->acc.begin(), c->acc.end(), Acc());
  return KGX_OK;
}

int kgo_expire(kgo_ctx* c, uint64_t older_than_ts_us, uint64_t* n_expired) {
  if (!c) return KGX_E_INVAL;
  if (n_expired) *n_expired = 0;
  if (c->mode != KGX_MODE_X) return KGX_E_STATE;
  uint64_t n = 0;
  for (Acc& a : c->acc)
    if (a.count && a.last_ts < older_than_ts_us) { a = Acc(); ++n; }
  if (n_expired) *n_expired = n;
  return KGX_OK;
}

int kgo_render(kgo_ctx* c, const uint8_t** out, size_t* len) {
  if (!c || !out || !len) return KGX_E_INVAL;
  if (!c->have_table) return KGX_E_STATE;
  if (c->gpu_total.empty()) c->gpu_total.assign(size_t(c->n_nodes) * c->G, 0);
  std::vector<std::string> bodies(c->n_nodes);
  parallel_nodes(c->threads, c->n_nodes, [&](uint32_t n) { render_node(c, n, c->gpu_total, bodies[n]); });
  c->body.clear();
  c->body_off.assign(size_t(c->n_nodes) + 1, 0);
  for (uint32_t n = 0; n < c->n_nodes; ++n) {
    c->body_off[n] = c->body.size();
    c->body += bodies[n];
  }
  c->body_off[c->n_nodes] = c->body.size();
  *out = reinterpret_cast<const uint8_t*>(c->body.data());
  *len = c->body.size();
  return KGX_OK;
}

int kgo_body_offsets(kgo_ctx* c, uint64_t* off, size_t n) {
  if (!c || !off || n != c->body_off.size()) return KGX_E_INVAL;
  std::copy(c->body_off.begin(), c->body_off.end(), off);
  return KGX_OK;
}

uint64_t kgo_series_count(kgo_ctx* c) {
  uint64_t n = 0;
  if (c->mode != KGX_MODE_X) { for (auto& m : c->children) n += m.size(); }
  else { for (const Acc& a : c->acc) n += a.count != 0; }
  return n;
}

// Values of one family (index in sorted family order) in exposition order.
int kgo_read_values(kgo_ctx* c, int family, double* values, size_t cap, size_t* n_out) {
  if (!c || !n_out) return KGX_E_INVAL;
  const bool x = c->mode == KGX_MODE_X;
  const int nf = x ? 6 : 2;
  if (family < 0 || family >= nf) return KGX_E_INVAL;
  std::vector<int> order(nf);
  for (int i = 0; i < nf; ++i) order[i] = i;
  const FamilyDef* defs = x ? kXFamilies : kRFamilies;
  std::sort(order.begin(), order.end(), [&](int a, int b) { return std::strcmp(defs[a].name, defs[b].name) < 0; });
  const int f = order[family];
  size_t n = 0;
  for (uint32_t node = 0; node < c->n_nodes; ++node) {
    std::vector<Series> series;
    if (!x) {
      for (auto& kv : c->children[node]) {
        Series s; s.nl = 2; s.lv[0] = kv.second.pid_label; s.lv[1] = kv.second.pod;
        s.v = f == 0 ? kv.second.used : kv.second.perc;
        series.push_back(s);
      }
    } else {
      for (uint32_t pk : c->node_pks[node])
        for (uint32_t g = 0; g < c->G; ++g) {
          const Acc& a = c->acc[size_t(pk) * c->G + g];
          if (!a.count) continue;
          Series s; s.nl = 4;
          s.lv[0] = c->ct[pk]; s.lv[1] = c->uuid[size_t(node) * c->G + g]; s.lv[2] = c->ns[pk]; s.lv[3] = c->pod[pk];
          s.v = family_value_x(f, a, c->gpu_total.empty() ? 0 : c->gpu_total[size_t(node) * c->G + g]);
          series.push_back(s);
        }
    }
    std::sort(series.begin(), series.end(), [](const Series& a, const Series& b) {
      for (int i = 0; i < a.nl; ++i) { int d = a.lv[i].compare(b.lv[i]); if (d) return d < 0; }
      return false;
    });
    for (const Series& s : series) { if (n < cap && values) values[n] = s.v; ++n; }
  }
  *n_out = n;
  return n <= cap ? KGX_OK : KGX_E_NOSPACE;
}

}  // extern "C"
