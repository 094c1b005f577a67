// kgx_build.cuh — cold path on the device: turn the caller's join table
// (podPIDMap, main.go:84-114) into everything the hot path needs — the order
// Registry.Gather would emit every possible series in (rules O2/O3 of SURVEY.md
// 8c), their slots, the pre-escaped label pieces and the per-node
// open-addressing table — without the host string sorts of kgx_table.cpp.
// Same results as build_plan() there (the table's cell layout may differ: cells
// are inserted concurrently; what a probe finds does not).
//
// Sorting, scans and stream compaction are library calls (CUB, shipped with the
// CUDA toolkit); the string comparators, validation, slot arithmetic, label
// escaping and the hash insert are the kernels below.  This is table-change
// work, not the per-scrape path.
#pragma once
#include <cub/cub.cuh>

#include "kgx_kernels.cuh"

namespace kgx {
namespace build {

// ---- the caller's table, uploaded as is ------------------------------------------------
struct TableIn {
  uint32_t n_nodes, G;
  uint32_t npk, n_entries;
  const uint32_t* podkey_node;
  const uint8_t* pod_blob; const unsigned long long* pod_off;
  const uint8_t* ns_blob; const unsigned long long* ns_off;
  const uint8_t* ct_blob; const unsigned long long* ct_off;
  const uint32_t* entry_podkey; const uint32_t* entry_pid;
  const uint8_t* uuid_blob; const unsigned long long* uuid_off;
};

// first error wins; the host maps the id to (KGX_E_*, message)
enum BuildErr : uint32_t {
  BE_OK = 0, BE_NODE_RANGE, BE_POD_UTF8, BE_NSCT_UTF8, BE_ENTRY_RANGE, BE_UUID_UTF8, BE_DUP_KEY, BE_DUP_UUID, BE_DUP_POD,
  BE_PODS_PER_NODE, BE_SLOTS, BE_CELLS, BE_LABEL_LEN, BE_COUNT
};
__device__ __forceinline__ void set_err(uint32_t* err, uint32_t id) { atomicCAS(err, 0u, id); }

// running maxima the host wants back
enum BuildStat { BS_MAX_AB = 0, BS_MAX_G, BS_MAX_NODE_PKS, BS_MAX_NODE_CELLS, BS_MAX_WORDS, BS_COUNT };

struct Str { const uint8_t* p; uint32_t n; };
__device__ __forceinline__ Str str_at(const uint8_t* blob, const unsigned long long* off, uint64_t i) {
  return Str{blob + off[i], uint32_t(off[i + 1] - off[i])};
}
// byte-wise order, a proper prefix first: Go's string comparison, which is what O3 sorts by
__device__ inline int cmp_str(Str a, Str b) {
  const uint32_t n = a.n < b.n ? a.n : b.n;
  for (uint32_t i = 0; i < n; ++i) {
    const int d = int(a.p[i]) - int(b.p[i]);
    if (d) return d;
  }
  return a.n < b.n ? -1 : (a.n > b.n ? 1 : 0);
}
// utf8.ValidString: client_golang panics on label values that are not valid UTF-8
__device__ inline bool valid_utf8(Str s) {
  uint32_t i = 0;
  while (i < s.n) {
    const uint32_t c = s.p[i];
    if (c < 0x80) { ++i; continue; }
    uint32_t len, cp, lo;
    if ((c & 0xE0) == 0xC0) { len = 2; cp = c & 0x1F; lo = 0x80; }
    else if ((c & 0xF0) == 0xE0) { len = 3; cp = c & 0x0F; lo = 0x800; }
    else if ((c & 0xF8) == 0xF0) { len = 4; cp = c & 0x07; lo = 0x10000; }
    else return false;
    if (i + len > s.n) return false;
    for (uint32_t k = 1; k < len; ++k) {
      const uint32_t b = s.p[i + k];
      if ((b & 0xC0) != 0x80) return false;
      cp = (cp << 6) | (b & 0x3F);
    }
    if (cp < lo || cp > 0x10FFFF || (cp >= 0xD800 && cp <= 0xDFFF)) return false;
    i += len;
  }
  return true;
}
// expfmt escapes backslash, double quote and newline in label values (G2)
__device__ inline uint32_t esc_len(Str s) {
  uint32_t n = s.n;
  for (uint32_t i = 0; i < s.n; ++i) n += (s.p[i] == '\\' || s.p[i] == '"' || s.p[i] == '\n');
  return n;
}
__device__ inline uint8_t* esc_copy(uint8_t* d, Str s) {
  for (uint32_t i = 0; i < s.n; ++i) {
    const uint8_t ch = s.p[i];
    if (ch == '\\') { *d++ = '\\'; *d++ = '\\'; }
    else if (ch == '"') { *d++ = '\\'; *d++ = '"'; }
    else if (ch == '\n') { *d++ = '\\'; *d++ = 'n'; }
    else *d++ = ch;
  }
  return d;
}
__device__ inline uint8_t* lit_copy(uint8_t* d, const char* s) {
  while (*s) *d++ = uint8_t(*s++);
  return d;
}
__device__ __forceinline__ uint32_t table_capacity_dev(uint32_t entries) {  // load factor <= 0.5, multiple of 8 cells
  uint32_t cap = entries * 2;
  if (cap < 8) cap = 8;
  return (cap + 7u) & ~7u;
}
// first index in [0, n) with a[i] >= v (a ascending)
template <class T>
__device__ __forceinline__ uint32_t lower_bound_dev(const T* a, uint32_t n, T v) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (a[mid] < v) lo = mid + 1; else hi = mid; }
  return lo;
}

// ---- validation ------------------------------------------------------------------------
__global__ void k_validate_podkeys(TableIn t, bool x, uint32_t* err) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= t.npk) return;
  if (t.podkey_node[i] >= t.n_nodes) { set_err(err, BE_NODE_RANGE); return; }
  if (!valid_utf8(str_at(t.pod_blob, t.pod_off, i))) { set_err(err, BE_POD_UTF8); return; }
  if (x && (!valid_utf8(str_at(t.ns_blob, t.ns_off, i)) || !valid_utf8(str_at(t.ct_blob, t.ct_off, i)))) set_err(err, BE_NSCT_UTF8);
}
__global__ void k_validate_entries(TableIn t, uint32_t* err) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < t.n_entries && t.entry_podkey[e] >= t.npk) set_err(err, BE_ENTRY_RANGE);
}
__global__ void k_iota(uint32_t* a, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = i;
}

// ---- X mode: order of the podkeys -------------------------------------------------------
// O3 over (container, gpu, namespace, pod) within a node: podkeys by (container, namespace, pod)
struct XLess {
  TableIn t;
  __device__ bool operator()(uint32_t a, uint32_t b) const {
    const uint32_t na = t.podkey_node[a], nb = t.podkey_node[b];
    if (na != nb) return na < nb;
    int c = cmp_str(str_at(t.ct_blob, t.ct_off, a), str_at(t.ct_blob, t.ct_off, b));
    if (c) return c < 0;
    c = cmp_str(str_at(t.ns_blob, t.ns_off, a), str_at(t.ns_blob, t.ns_off, b));
    if (c) return c < 0;
    return cmp_str(str_at(t.pod_blob, t.pod_off, a), str_at(t.pod_blob, t.pod_off, b)) < 0;
  }
};
// a = position in sorted order.  A slot run starts where the node or the container changes.
__global__ void k_x_heads(TableIn t, const uint32_t* ord, uint32_t* head_run, uint32_t* node_sorted, uint32_t* a_of_pk, uint32_t* err) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= t.npk) return;
  const uint32_t pk = ord[a], node = t.podkey_node[pk];
  node_sorted[a] = node;
  a_of_pk[pk] = a;
  bool head = true;
  if (a > 0) {
    const uint32_t q = ord[a - 1];
    if (t.podkey_node[q] == node && cmp_str(str_at(t.ct_blob, t.ct_off, q), str_at(t.ct_blob, t.ct_off, pk)) == 0) {
      head = false;
      if (cmp_str(str_at(t.ns_blob, t.ns_off, q), str_at(t.ns_blob, t.ns_off, pk)) == 0 &&
          cmp_str(str_at(t.pod_blob, t.pod_off, q), str_at(t.pod_blob, t.pod_off, pk)) == 0)
        set_err(err, BE_DUP_KEY);
    }
  }
  head_run[a] = head ? 1u : 0u;
}
// run_id_incl = inclusive scan of head_run
__global__ void k_x_run_starts(const uint32_t* head_run, const uint32_t* run_id_incl, uint32_t n, uint32_t* run_start) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a < n && head_run[a]) run_start[run_id_incl[a] - 1] = a;
}
__global__ void k_x_run_slots(const uint32_t* run_start, uint32_t n_runs, uint32_t n_a, uint32_t G, unsigned long long* run_slots) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_runs) return;
  const uint32_t len = (r + 1 < n_runs ? run_start[r + 1] : n_a) - run_start[r];
  run_slots[r] = uint64_t(len) * G;
}
// per sorted podkey: first slot, stride, lengths of its two label pieces
__global__ void k_x_podkeys(TableIn t, const uint32_t* ord, const uint32_t* run_id_incl, const uint32_t* run_start, uint32_t n_runs,
                            const unsigned long long* run_slot, uint32_t* pk_slot0, uint32_t* pk_stride, unsigned long long* lenA,
                            unsigned long long* lenB, uint32_t* stats) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= t.npk) return;
  const uint32_t r = run_id_incl[a] - 1, s = run_start[r];
  const uint32_t len = (r + 1 < n_runs ? run_start[r + 1] : t.npk) - s;
  pk_slot0[a] = uint32_t(run_slot[r]) + (a - s);
  pk_stride[a] = len;
  const uint32_t pk = ord[a];
  // {container="<c>",gpu="   and   ",namespace="<ns>",pod="<pod>"}   (O2: label names ascending)
  const uint32_t la = 12 + esc_len(str_at(t.ct_blob, t.ct_off, pk)) + 7;
  const uint32_t lb = 13 + esc_len(str_at(t.ns_blob, t.ns_off, pk)) + 7 + esc_len(str_at(t.pod_blob, t.pod_off, pk)) + 2;
  lenA[a] = la; lenB[a] = lb;
  atomicMax(&stats[BS_MAX_AB], la + lb);
}
__global__ void k_x_pieces(TableIn t, const uint32_t* ord, const unsigned long long* A_off, const unsigned long long* B_off,
                           uint8_t* A_blob, uint8_t* B_blob) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= t.npk) return;
  const uint32_t pk = ord[a];
  uint8_t* d = A_blob + A_off[a];
  d = lit_copy(d, "{container=\""); d = esc_copy(d, str_at(t.ct_blob, t.ct_off, pk)); lit_copy(d, "\",gpu=\"");
  d = B_blob + B_off[a];
  d = lit_copy(d, "\",namespace=\""); d = esc_copy(d, str_at(t.ns_blob, t.ns_off, pk));
  d = lit_copy(d, "\",pod=\""); d = esc_copy(d, str_at(t.pod_blob, t.pod_off, pk)); lit_copy(d, "\"}");
}
// first sorted index of every node (nodes without podkeys get their successor's)
__global__ void k_node_base(const uint32_t* node_sorted, uint32_t n, uint32_t n_nodes, uint32_t* node_base, uint32_t* stats, int stat_id) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n_nodes) return;
  node_base[k] = k == n_nodes ? n : lower_bound_dev(node_sorted, n, k);
  if (stat_id >= 0 && k < n_nodes) {
    const uint32_t next = k + 1 == n_nodes ? n : lower_bound_dev(node_sorted, n, k + 1);
    atomicMax(&stats[stat_id], next - node_base[k]);
  }
}
// first slot of every node; words of the series index a node spans
__global__ void k_x_node_slot_base(const uint32_t* node_a_base, const uint32_t* pk_slot0, uint32_t n_a, uint32_t n_slots, uint32_t n_nodes,
                                   uint32_t* node_slot_base) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n_nodes) return;
  const uint32_t a = k == n_nodes ? n_a : node_a_base[k];
  node_slot_base[k] = a == n_a ? n_slots : pk_slot0[a];
}
__global__ void k_max_words(const uint32_t* node_slot_base, uint32_t n_nodes, uint32_t* stats) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_nodes) atomicMax(&stats[BS_MAX_WORDS], ((node_slot_base[k + 1] + 31) >> 5) - (node_slot_base[k] >> 5));
}
// gpu label order inside a node: one thread per node, insertion sort of G uuid strings (G is small)
__global__ void k_x_gpu_order(TableIn t, uint32_t* gord, uint32_t* grank, unsigned long long* lenG, uint32_t* stats, uint32_t* err) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= t.n_nodes) return;
  const uint32_t G = t.G;
  uint32_t* o = gord + size_t(n) * G;
  for (uint32_t g = 0; g < G; ++g) {
    const Str u = str_at(t.uuid_blob, t.uuid_off, size_t(n) * G + g);
    if (!valid_utf8(u)) { set_err(err, BE_UUID_UTF8); return; }
    const uint32_t l = esc_len(u);
    lenG[size_t(n) * G + g] = l;
    atomicMax(&stats[BS_MAX_G], l);
    uint32_t j = g;
    while (j > 0) {
      const int c = cmp_str(str_at(t.uuid_blob, t.uuid_off, size_t(n) * G + o[j - 1]), u);
      if (c == 0) { set_err(err, BE_DUP_UUID); return; }
      if (c < 0) break;
      o[j] = o[j - 1];
      --j;
    }
    o[j] = g;
  }
  for (uint32_t r = 0; r < G; ++r) grank[size_t(n) * G + o[r]] = r;
}
__global__ void k_x_uuid_pieces(TableIn t, const unsigned long long* G_off, uint8_t* G_blob) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < t.n_nodes * t.G) esc_copy(G_blob + G_off[g], str_at(t.uuid_blob, t.uuid_off, g));
}
// one Run per (container run, gpu in label order): the inverse map slot -> label pieces
__global__ void k_x_runs(const uint32_t* run_start, const unsigned long long* run_slot, const uint32_t* node_sorted, const uint32_t* gord,
                         uint32_t n_runs, uint32_t n_a, uint32_t G, Run* runs) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_runs * G) return;
  const uint32_t r = i / G, j = i % G;
  const uint32_t s = run_start[r], len = (r + 1 < n_runs ? run_start[r + 1] : n_a) - s, node = node_sorted[s];
  runs[i] = Run{uint32_t(run_slot[r]) + j * len, len, s, node * G + gord[size_t(node) * G + j], node};
}
__global__ void k_x_cell_keys(TableIn t, const uint32_t* a_of_pk, unsigned long long* keys) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < t.n_entries) keys[e] = (uint64_t(a_of_pk[t.entry_podkey[e]]) << 32) | t.entry_pid[e];
}

// ---- the open-addressing table (both modes) ----------------------------------------------
// cells: distinct keys (target << 32 | pid), ascending, so the cells of a node are contiguous.
// base[n] = first target of node n (X: sorted podkey index, R: slot).
__global__ void k_cell_bounds(const unsigned long long* cells, uint32_t n_cells, const uint32_t* base, uint32_t n_nodes, uint32_t* cell_lo) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n_nodes) return;
  cell_lo[k] = k == n_nodes ? n_cells : lower_bound_dev(cells, n_cells, (unsigned long long)base[k] << 32);
}
__global__ void k_node_caps(const uint32_t* cell_lo, uint32_t n_nodes, unsigned long long* caps, uint32_t* stats) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  const uint32_t cap = table_capacity_dev(cell_lo[n + 1] - cell_lo[n]);
  caps[n] = cap;
  atomicMax(&stats[BS_MAX_NODE_CELLS], cap);
}
__global__ void k_node_tabs(const unsigned long long* caps, const unsigned long long* cap_off, uint32_t n_nodes, NodeTab* node_tab) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < n_nodes) node_tab[n] = NodeTab{uint32_t(cap_off[n]), uint32_t(caps[n])};
}
__global__ void k_tab_clear(Entry* tab, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) tab[i] = Entry{0u, kEmpty};
}
// Concurrent linear-probing insert: claim the first empty cell at or after the hash position with
// one 64-bit CAS.  Cells never go away, so no empty cell ever appears between a key's hash position
// and the key: the invariant the probe relies on.
__global__ void k_tab_insert(const unsigned long long* cells, uint32_t n_cells, const uint32_t* cell_lo, uint32_t n_nodes,
                             const NodeTab* node_tab, Entry* tab) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_cells) return;
  uint32_t lo = 0, hi = n_nodes;  // the node whose cell range holds i: last n with cell_lo[n] <= i
  while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (cell_lo[mid] <= i) lo = mid; else hi = mid; }
  const NodeTab nt = node_tab[lo];
  const uint32_t pid = uint32_t(cells[i]), target = uint32_t(cells[i] >> 32);
  const unsigned long long empty = (unsigned long long)kEmpty << 32, mine = ((unsigned long long)target << 32) | pid;
  unsigned long long* slice = reinterpret_cast<unsigned long long*>(tab + nt.off);  // Entry{key, val} = val << 32 | key
  uint32_t pos = uint32_t((uint64_t(dev_hash_pid(pid)) * nt.cap) >> 32);
  for (;;) {
    if (atomicCAS(&slice[pos], empty, mine) == empty) return;
    pos = pos + 1 == nt.cap ? 0 : pos + 1;
  }
}
// A cell gets the "more" bit iff the same key occurs again later on its probe path (before the next empty cell).
// cap_off: [n_nodes+1] first cell of every node's slice.
__global__ void k_tab_more(const unsigned long long* cap_off, uint32_t n_nodes, uint32_t n_cells_total, Entry* tab) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_cells_total) return;
  const Entry e = tab[i];
  if (e.val == kEmpty) return;
  uint32_t lo = 0, hi = n_nodes;  // last n with cap_off[n] <= i
  while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (cap_off[mid] <= i) lo = mid; else hi = mid; }
  const uint32_t off = uint32_t(cap_off[lo]), cap = uint32_t(cap_off[lo + 1]) - off;
  uint32_t q = i - off;
  for (;;) {
    q = q + 1 == cap ? 0 : q + 1;
    const Entry f = tab[off + q];
    if (f.val == kEmpty) return;
    if (f.key == e.key) { atomicOr(&tab[i].val, kMore); return; }
  }
}

// ---- R modes ---------------------------------------------------------------------------
// Series persist forever (main.go has no Delete/Reset), so a new table is merged with the
// series that already have a value: the previous generation's slots (pid, pod) and pods
// (node, name), all still on the device.
struct OldGen {
  const uint32_t* slot_pid; const uint32_t* slot_pod;      // [old slots]
  const uint32_t* pods_node; const uint8_t* pods_blob; const unsigned long long* pods_off;  // [old pods]
  uint32_t n_pods;
};
// series with a value -> carried list (order = rank order)
__global__ void k_r_carried_flags(const unsigned long long* seq, uint32_t S, uint32_t* flag) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < S) flag[r] = seq[r] != 0;
}
__global__ void k_r_carried_gather(const uint32_t* flag, const uint32_t* pos, uint32_t S, const uint32_t* ser_slot, const unsigned long long* used,
                                   const unsigned long long* total, OldGen og, uint32_t* c_pid, uint32_t* c_oldpod, unsigned long long* c_used,
                                   unsigned long long* c_total, uint32_t* old_used) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= S || !flag[r]) return;
  const uint32_t i = pos[r], s = ser_slot[r], pod = og.slot_pod[s];
  c_pid[i] = og.slot_pid[s]; c_oldpod[i] = pod; c_used[i] = used[r]; c_total[i] = total[r];
  old_used[pod] = 1;
}
__global__ void k_compact_ids(const uint32_t* flag, const uint32_t* pos, uint32_t n, uint32_t* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && flag[i]) out[pos[i]] = i;
}
// A pod-name reference: j < npk is pod j of the new table, otherwise old pod old_list[j - npk].
struct NameRefs {
  TableIn t; OldGen og; const uint32_t* old_list;
  __device__ uint32_t node(uint32_t j) const { return j < t.npk ? t.podkey_node[j] : og.pods_node[old_list[j - t.npk]]; }
  __device__ Str str(uint32_t j) const {
    return j < t.npk ? str_at(t.pod_blob, t.pod_off, j) : str_at(og.pods_blob, og.pods_off, old_list[j - t.npk]);
  }
};
struct NameLess {  // (node, name), the table's copy before the carried one
  NameRefs r;
  __device__ bool operator()(uint32_t a, uint32_t b) const {
    const uint32_t na = r.node(a), nb = r.node(b);
    if (na != nb) return na < nb;
    const int c = cmp_str(r.str(a), r.str(b));
    if (c) return c < 0;
    return a < b;
  }
};
// head[i]: first reference of a distinct (node, name); podPIDMap keys are unique, so two table pods of one name are an error
__global__ void k_r_name_heads(NameRefs r, const uint32_t* ord, uint32_t n, uint32_t* head, uint32_t* err) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  bool h = true;
  if (i > 0 && r.node(ord[i - 1]) == r.node(ord[i]) && cmp_str(r.str(ord[i - 1]), r.str(ord[i])) == 0) {
    h = false;
    if (ord[i - 1] < r.t.npk && ord[i] < r.t.npk) set_err(err, BE_DUP_POD);
  }
  head[i] = h;
}
// new pod ids = distinct names in (node, name) order
__global__ void k_r_names(NameRefs r, const uint32_t* ord, const uint32_t* head, const uint32_t* id_incl, uint32_t n, uint32_t* pod_of_pk,
                          uint32_t* pod_of_old, uint32_t* pods_node, unsigned long long* pods_len, uint32_t* pods_ref,
                          uint32_t* node_name_base) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t j = ord[i], id = id_incl[i] - 1, node = r.node(j);
  if (j < r.t.npk) pod_of_pk[j] = id; else pod_of_old[r.old_list[j - r.t.npk]] = id;
  if (head[i]) {
    pods_node[id] = node; pods_len[id] = r.str(j).n; pods_ref[id] = j;
    if (i == 0 || r.node(ord[i - 1]) != node) node_name_base[node] = id;
  }
}
__global__ void k_r_pods_copy(NameRefs r, const uint32_t* pods_ref, const unsigned long long* pods_off, uint32_t n_pods, uint8_t* blob) {
  const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= n_pods) return;
  const Str s = r.str(pods_ref[id]);
  uint8_t* d = blob + pods_off[id];
  for (uint32_t k = 0; k < s.n; ++k) d[k] = s.p[k];
}
// Orders u32 values as their decimal strings compare byte-wise ("10" < "9"): digits left-aligned, length as tie-break
__device__ __forceinline__ unsigned long long decimal_string_key_dev(uint32_t v) {
  const unsigned long long p10[11] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull,
                                      1000000000ull, 10000000000ull};
  int nd = 1;
  while (nd < 10 && v >= p10[nd]) ++nd;
  return ((unsigned long long)v * p10[10 - nd]) << 4 | (unsigned long long)nd;
}
// A candidate series: (node, key) with key = (decimal order of pid, rank of the pod name in its node, carried?).
struct Pairs { uint32_t* node; unsigned long long* key; uint32_t* pid; uint32_t* pod; int32_t* src; };
__device__ __forceinline__ void put_pair(const Pairs& p, uint32_t j, uint32_t node, uint32_t pid, uint32_t pod, uint32_t rank, int32_t src,
                                         uint32_t* err) {
  if (rank >= (1u << 24)) { set_err(err, BE_PODS_PER_NODE); rank &= 0xffffffu; }
  p.node[j] = node; p.pid[j] = pid; p.pod[j] = pod; p.src[j] = src;
  p.key[j] = (decimal_string_key_dev(pid) << 25) | ((unsigned long long)rank << 1) | (src >= 0 ? 1ull : 0ull);
}
// intended join: one pair per table entry (main.go:144-145 read as PID equality)
__global__ void k_r_pairs_entries(TableIn t, const uint32_t* pod_of_pk, const uint32_t* node_name_base, Pairs p, uint32_t* err) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= t.n_entries) return;
  const uint32_t pk = t.entry_podkey[e], node = t.podkey_node[pk], pod = pod_of_pk[pk];
  put_pair(p, e, node, t.entry_pid[e], pod, pod - node_name_base[node], -1, err);
}
// literal join: `for pid := range pids` yields the indices 0..len(pids)-1 (main.go:144)
__global__ void k_r_count_entries(TableIn t, uint32_t* cnt) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < t.n_entries) atomicAdd(&cnt[t.entry_podkey[e]], 1u);
}
__global__ void k_r_pairs_literal(TableIn t, const uint32_t* pair_off, const uint32_t* pod_of_pk, const uint32_t* node_name_base, Pairs p,
                                  uint32_t* err) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= t.n_entries) return;
  uint32_t lo = 0, hi = t.npk;  // last podkey with pair_off <= j (podkeys without entries share an offset: take the last)
  while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (pair_off[mid] <= j) lo = mid; else hi = mid; }
  const uint32_t pk = lo, node = t.podkey_node[pk], pod = pod_of_pk[pk];
  put_pair(p, j, node, j - pair_off[pk], pod, pod - node_name_base[node], -1, err);
}
__global__ void k_r_pairs_carried(const uint32_t* c_pid, const uint32_t* c_oldpod, uint32_t n_c, const uint32_t* pod_of_old,
                                  const uint32_t* pods_node, const uint32_t* node_name_base, uint32_t first, Pairs p, uint32_t* err) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_c) return;
  const uint32_t pod = pod_of_old[c_oldpod[i]], node = pods_node[pod];
  put_pair(p, first + i, node, c_pid[i], pod, pod - node_name_base[node], int32_t(i), err);
}
struct PairLess {
  const uint32_t* node; const unsigned long long* key;
  __device__ bool operator()(uint32_t a, uint32_t b) const {
    if (node[a] != node[b]) return node[a] < node[b];
    return key[a] < key[b];
  }
};
__global__ void k_r_slot_heads(const uint32_t* idx, uint32_t n, const uint32_t* node, const unsigned long long* key, uint32_t* head) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  head[i] = i == 0 || node[idx[i - 1]] != node[idx[i]] || (key[idx[i - 1]] >> 1) != (key[idx[i]] >> 1);
}
// One slot per distinct (node, pid, pod).  The table's copy sorts first: it makes the slot joinable.
__global__ void k_r_slots(const uint32_t* idx, const uint32_t* head, const uint32_t* slot_incl, uint32_t n, Pairs p, uint32_t* slot_pid,
                          uint32_t* slot_pod, uint32_t* slot_node, unsigned long long* lenA, uint32_t* cell_flag,
                          unsigned long long* cell_key, uint32_t* carried_slot) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t j = idx[i], s = slot_incl[i] - 1;
  if (p.src[j] >= 0) carried_slot[p.src[j]] = s;
  if (!head[i]) return;
  const uint32_t pid = p.pid[j], pod = p.pod[j];
  slot_pid[s] = pid; slot_pod[s] = pod; slot_node[s] = p.node[j];
  cell_flag[s] = p.src[j] < 0;
  cell_key[s] = ((unsigned long long)s << 32) | pid;
  uint32_t nd = 1;
  for (uint32_t v = pid; v >= 10; v /= 10) ++nd;
  lenA[s] = 6 + nd + 7 + 2;  // {pid="<pid>",pod="<pod>"} without the pod name: k_r_label_len adds its escaped length
}
__global__ void k_r_label_len(const uint32_t* slot_pod, uint32_t n_slots, const uint8_t* pods_blob, const unsigned long long* pods_off,
                              unsigned long long* lenA, uint32_t* stats) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  const uint32_t l = uint32_t(lenA[s]) + esc_len(str_at(pods_blob, pods_off, slot_pod[s]));
  lenA[s] = l;
  atomicMax(&stats[BS_MAX_AB], l);
}
__global__ void k_r_pieces(const uint32_t* slot_pid, const uint32_t* slot_pod, uint32_t n_slots, const uint8_t* pods_blob,
                           const unsigned long long* pods_off, const unsigned long long* A_off, uint8_t* A_blob) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  uint8_t* d = lit_copy(A_blob + A_off[s], "{pid=\"");
  char num[10];  // fmt.Sprintf("%d", pid), main.go:147
  uint32_t nd = 0;
  for (uint32_t v = slot_pid[s]; nd == 0 || v; v /= 10) num[nd++] = char('0' + v % 10);
  while (nd) *d++ = uint8_t(num[--nd]);
  d = lit_copy(d, "\",pod=\"");
  d = esc_copy(d, str_at(pods_blob, pods_off, slot_pod[s]));
  lit_copy(d, "\"}");
}
__global__ void k_r_runs(const uint32_t* node_slot_base, uint32_t n_nodes, Run* runs) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  const uint32_t s = node_slot_base[n];
  runs[n] = Run{s, node_slot_base[n + 1] - s, s, 0xFFFFFFFFu, n};
}

}  // namespace build
}  // namespace kgx
