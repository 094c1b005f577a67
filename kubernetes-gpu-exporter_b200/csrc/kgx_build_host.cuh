// kgx_build_host.cuh — host driver of the device-side table build (kernels in
// kgx_build.cuh).  Included by kgx_api.cu after kgx_ctx and its helpers.
//
// Two phases.  Phase 1 works in scratch memory only: upload, validation, the
// sorts, every count.  All KGX_E_INVAL / KGX_E_LIMIT answers are known at its
// end, so a rejected table leaves the ctx exactly as it was (like the host
// builder).  Phase 2 writes the ctx-owned buffers.
#pragma once

namespace {

using namespace kgx::build;

// Scratch for one build: bump allocation out of chunks the ctx keeps, so that a table that
// changes every cycle (churn) does not pay cudaMalloc/cudaFree for its ~40 temporaries.
struct Arena {
  kgx_ctx* c;
  explicit Arena(kgx_ctx* c_) : c(c_) { c->arena_i = 0; c->arena_off = 0; }
  void* take(size_t bytes) {
    bytes = (std::max<size_t>(bytes, 1) + 255) & ~size_t(255);
    while (c->arena_i < c->arena.size()) {
      DBuf& b = c->arena[c->arena_i];
      if (c->arena_off + bytes <= b.cap) { void* p = static_cast<char*>(b.p) + c->arena_off; c->arena_off += bytes; return p; }
      ++c->arena_i; c->arena_off = 0;
    }
    DBuf nb;
    const size_t want = std::max<size_t>(bytes + bytes / 4, size_t(8) << 20);
    if (cudaMalloc(&nb.p, want) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    nb.cap = want;
    c->device_bytes += want;
    c->arena.push_back(nb);
    c->arena_i = c->arena.size() - 1;
    c->arena_off = bytes;
    return nb.p;
  }
};
#define KGX_TAKE(T, name, count)                                                            \
  T* name = static_cast<T*>(ar.take(size_t(count) * sizeof(T)));                            \
  if (!name) return fail(c, KGX_E_NOMEM, "table build: out of device memory (scratch)")

template <class F>
int cub_run(kgx_ctx* c, Arena& ar, F&& f) {
  size_t bytes = 0;
  KGX_CUDA(c, f(static_cast<void*>(nullptr), bytes));
  void* tmp = ar.take(bytes);
  if (!tmp) return fail(c, KGX_E_NOMEM, "table build: out of device memory (scratch)");
  KGX_CUDA(c, f(tmp, bytes));
  return KGX_OK;
}
// exclusive prefix sums of in[0..n) into out[0..n]; out[n] = total.  `in` must have n+1 elements (the last is cleared here).
template <class T>
int scan_with_total(kgx_ctx* c, Arena& ar, T* in, T* out, uint32_t n) {
  KGX_CUDA(c, cudaMemsetAsync(in + n, 0, sizeof(T), c->stream));
  return cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceScan::ExclusiveSum(t, b, in, out, int(n + 1), c->stream); });
}
template <class T>
int read_back(kgx_ctx* c, T* host, const T* dev, size_t n = 1) {
  KGX_CUDA(c, cudaMemcpyAsync(host, dev, n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGX_OK;
}
template <class T>
int to_device(kgx_ctx* c, T* dev, const void* host, size_t n) {
  if (n) KGX_CUDA(c, cudaMemcpyAsync(dev, host, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  return KGX_OK;
}

int build_error(kgx_ctx* c, uint32_t id) {
  switch (id) {
    case BE_NODE_RANGE: return fail(c, KGX_E_INVAL, "table: podkey_node out of range");
    case BE_POD_UTF8: return fail(c, KGX_E_INVAL, "table: pod name is not valid UTF-8");
    case BE_NSCT_UTF8: return fail(c, KGX_E_INVAL, "table: namespace/container is not valid UTF-8");
    case BE_ENTRY_RANGE: return fail(c, KGX_E_INVAL, "table: entry_podkey out of range");
    case BE_UUID_UTF8: return fail(c, KGX_E_INVAL, "table: gpu uuid is not valid UTF-8");
    case BE_DUP_KEY: return fail(c, KGX_E_INVAL, "table: duplicate (container, namespace, pod) on one node");
    case BE_DUP_UUID: return fail(c, KGX_E_INVAL, "table: duplicate gpu uuid on one node");
    case BE_DUP_POD: return fail(c, KGX_E_INVAL, "table: duplicate pod name on one node (podPIDMap keys are unique)");
    case BE_PODS_PER_NODE: return fail(c, KGX_E_LIMIT, "table: more than 2^24 pods on one node");
    default: return fail(c, KGX_E_CUDA, "table build: unknown device error");
  }
}

struct BuildSmall {  // a few device words: error id, running maxima, counts written by CUB
  uint32_t* err; uint32_t* stats; uint32_t* n_sel;
};

// cells (target << 32 | pid, ascending, distinct) -> node_tab + tab.  base = first target of every node.
int build_hash_table(kgx_ctx* c, Arena& ar, const BuildSmall& sm, const unsigned long long* cells, uint32_t n_cells, const uint32_t* base,
                     uint32_t NN) {
  KGX_TAKE(uint32_t, cell_lo, NN + 1);
  KGX_TAKE(unsigned long long, caps, NN + 1);
  KGX_TAKE(unsigned long long, cap_off, NN + 1);
  k_cell_bounds<<<cdiv(NN + 1, 256), 256, 0, c->stream>>>(cells, n_cells, base, NN, cell_lo);
  k_node_caps<<<cdiv(NN, 256), 256, 0, c->stream>>>(cell_lo, NN, caps, sm.stats);
  KGX_TRY(scan_with_total(c, ar, caps, cap_off, NN));
  unsigned long long tab_total = 0;
  KGX_TRY(read_back(c, &tab_total, cap_off + NN));
  if (tab_total >= 0xfffffff0ull) return fail(c, KGX_E_LIMIT, "table: hash table beyond 2^32 cells");
  KGX_TRY(dev_reserve(c, c->d_node_tab, size_t(NN) * sizeof(NodeTab)));
  KGX_TRY(dev_reserve(c, c->d_tab, size_t(tab_total) * sizeof(Entry)));
  k_node_tabs<<<cdiv(NN, 256), 256, 0, c->stream>>>(caps, cap_off, NN, ptr<NodeTab>(c->d_node_tab));
  k_tab_clear<<<cdiv(tab_total, 256), 256, 0, c->stream>>>(ptr<Entry>(c->d_tab), uint32_t(tab_total));
  if (n_cells)
    k_tab_insert<<<cdiv(n_cells, 256), 256, 0, c->stream>>>(cells, n_cells, cell_lo, NN, ptr<NodeTab>(c->d_node_tab), ptr<Entry>(c->d_tab));
  k_tab_more<<<cdiv(tab_total, 256), 256, 0, c->stream>>>(cap_off, NN, uint32_t(tab_total), ptr<Entry>(c->d_tab));
  c->stats.kernel_launches += 6;
  c->plan.n_table_entries = n_cells;
  return launch_check(c, "table build: hash table");
}

// Argument checks of build_plan() plus the one thing that must be known before copying: the offsets
// ascend (their last element is the blob size).
int check_table_args(kgx_ctx* c, const kgx_table* t, bool x) {
  if (!t || t->n_nodes == 0 || t->gpus_per_node == 0) return fail(c, KGX_E_INVAL, "table: n_nodes and gpus_per_node must be >= 1");
  if (t->n_podkeys && (!t->podkey_node || !t->pod_blob || !t->pod_off)) return fail(c, KGX_E_INVAL, "table: podkey arrays missing");
  if (t->n_entries && (!t->entry_podkey || !t->entry_pid)) return fail(c, KGX_E_INVAL, "table: entry arrays missing");
  if (x && t->n_podkeys && (!t->namespace_blob || !t->namespace_off || !t->container_blob || !t->container_off))
    return fail(c, KGX_E_INVAL, "table: X mode needs namespace and container strings");
  if (x && (!t->gpu_uuid_blob || !t->gpu_uuid_off)) return fail(c, KGX_E_INVAL, "table: X mode needs gpu uuid strings");
  if (t->n_podkeys >= 0x7fffffffull || t->n_entries >= 0x7fffffffull) return fail(c, KGX_E_LIMIT, "table: more than 2^31-1 podkeys or entries");
  if (uint64_t(t->n_nodes) * t->gpus_per_node >= 0x7fffffffull) return fail(c, KGX_E_LIMIT, "table: more than 2^31-1 gpus");
  auto ascending = [](const uint64_t* off, uint64_t n) {
    bool ok = true;
    for (uint64_t i = 0; i < n; ++i) ok &= off[i + 1] >= off[i];
    return ok;
  };
  if (t->n_podkeys && !ascending(t->pod_off, t->n_podkeys)) return fail(c, KGX_E_INVAL, "table: pod_off not ascending");
  if (x && t->n_podkeys && (!ascending(t->namespace_off, t->n_podkeys) || !ascending(t->container_off, t->n_podkeys)))
    return fail(c, KGX_E_INVAL, "table: namespace_off/container_off not ascending");
  if (x && !ascending(t->gpu_uuid_off, uint64_t(t->n_nodes) * t->gpus_per_node)) return fail(c, KGX_E_INVAL, "table: gpu_uuid_off not ascending");
  return KGX_OK;
}

// strings of one label: offsets rebased to 0 on the way (the caller's blob may start anywhere)
int upload_strings(kgx_ctx* c, Arena& ar, const char* blob, const uint64_t* off, uint64_t n, const uint8_t** d_blob,
                   const unsigned long long** d_off) {
  KGX_TAKE(unsigned long long, o, n + 1);
  const uint64_t first = n ? off[0] : 0, bytes = n ? off[n] - off[0] : 0;
  KGX_TAKE(uint8_t, b, bytes + 16);
  if (n == 0) { KGX_CUDA(c, cudaMemsetAsync(o, 0, 8, c->stream)); }
  else if (first == 0) { KGX_TRY(to_device(c, o, off, n + 1)); }
  else {
    std::vector<unsigned long long> tmp(n + 1);
    for (uint64_t i = 0; i <= n; ++i) tmp[i] = off[i] - first;
    KGX_TRY(to_device(c, o, tmp.data(), n + 1));
    KGX_CUDA(c, cudaStreamSynchronize(c->stream));  // tmp dies here
  }
  KGX_TRY(to_device(c, b, blob + first, bytes));
  *d_blob = b; *d_off = o;
  return KGX_OK;
}

int upload_table(kgx_ctx* c, Arena& ar, const kgx_table* t, bool x, TableIn* in) {
  TableIn& d = *in;
  d = TableIn{};
  d.n_nodes = t->n_nodes; d.G = t->gpus_per_node; d.npk = uint32_t(t->n_podkeys); d.n_entries = uint32_t(t->n_entries);
  KGX_TAKE(uint32_t, pkn, d.npk); KGX_TRY(to_device(c, pkn, t->podkey_node, d.npk)); d.podkey_node = pkn;
  KGX_TAKE(uint32_t, epk, d.n_entries); KGX_TRY(to_device(c, epk, t->entry_podkey, d.n_entries)); d.entry_podkey = epk;
  KGX_TAKE(uint32_t, epid, d.n_entries); KGX_TRY(to_device(c, epid, t->entry_pid, d.n_entries)); d.entry_pid = epid;
  KGX_TRY(upload_strings(c, ar, t->pod_blob, t->pod_off, d.npk, &d.pod_blob, &d.pod_off));
  if (x) {
    KGX_TRY(upload_strings(c, ar, t->namespace_blob, t->namespace_off, d.npk, &d.ns_blob, &d.ns_off));
    KGX_TRY(upload_strings(c, ar, t->container_blob, t->container_off, d.npk, &d.ct_blob, &d.ct_off));
    KGX_TRY(upload_strings(c, ar, t->gpu_uuid_blob, t->gpu_uuid_off, uint64_t(d.n_nodes) * d.G, &d.uuid_blob, &d.uuid_off));
  }
  return KGX_OK;
}

// ---- X mode -------------------------------------------------------------------------------
int build_table_device_x(kgx_ctx* c, Arena& ar, const TableIn& in, const BuildSmall& sm) {
  const uint32_t NN = in.n_nodes, G = in.G, npk = in.npk, E = in.n_entries, NG = NN * G;
  const int B = 256;
  // ---------- phase 1: scratch only
  k_validate_podkeys<<<cdiv(std::max(npk, 1u), B), B, 0, c->stream>>>(in, true, sm.err);
  k_validate_entries<<<cdiv(std::max(E, 1u), B), B, 0, c->stream>>>(in, sm.err);
  KGX_TAKE(uint32_t, gord, NG); KGX_TAKE(uint32_t, grank, NG); KGX_TAKE(unsigned long long, lenG, NG + 1);
  k_x_gpu_order<<<cdiv(NN, 64), 64, 0, c->stream>>>(in, gord, grank, lenG, sm.stats, sm.err);
  uint32_t e = 0;
  KGX_TRY(read_back(c, &e, sm.err));  // strings are about to be compared: they must be valid first
  if (e) return build_error(c, e);
  KGX_TAKE(uint32_t, ord, npk);
  k_iota<<<cdiv(std::max(npk, 1u), B), B, 0, c->stream>>>(ord, npk);
  KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceMergeSort::SortKeys(t, b, ord, int(npk), XLess{in}, c->stream); }));
  KGX_TAKE(uint32_t, head_run, npk + 1); KGX_TAKE(uint32_t, node_sorted, npk + 1); KGX_TAKE(uint32_t, a_of_pk, npk + 1);
  KGX_TAKE(uint32_t, run_id_incl, npk + 1);
  k_x_heads<<<cdiv(std::max(npk, 1u), B), B, 0, c->stream>>>(in, ord, head_run, node_sorted, a_of_pk, sm.err);
  KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceScan::InclusiveSum(t, b, head_run, run_id_incl, int(npk), c->stream); }));
  uint32_t n_runs = 0;
  if (npk) KGX_TRY(read_back(c, &n_runs, run_id_incl + npk - 1));
  KGX_TRY(read_back(c, &e, sm.err));
  if (e) return build_error(c, e);
  KGX_TAKE(uint32_t, run_start, n_runs + 1); KGX_TAKE(unsigned long long, run_slots, n_runs + 1); KGX_TAKE(unsigned long long, run_slot, n_runs + 1);
  if (npk) k_x_run_starts<<<cdiv(npk, B), B, 0, c->stream>>>(head_run, run_id_incl, npk, run_start);
  if (n_runs) k_x_run_slots<<<cdiv(n_runs, B), B, 0, c->stream>>>(run_start, n_runs, npk, G, run_slots);
  KGX_TRY(scan_with_total(c, ar, run_slots, run_slot, n_runs));
  // distinct (pid, podkey) cells, grouped by node (sorted podkey indices are node-contiguous)
  KGX_TAKE(unsigned long long, keys, E + 1); KGX_TAKE(unsigned long long, keys_sorted, E + 1); KGX_TAKE(unsigned long long, cells, E + 1);
  if (E) {
    k_x_cell_keys<<<cdiv(E, B), B, 0, c->stream>>>(in, a_of_pk, keys);
    KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceRadixSort::SortKeys(t, b, keys, keys_sorted, int(E), 0, 64, c->stream); }));
    KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceSelect::Unique(t, b, keys_sorted, cells, sm.n_sel, int(E), c->stream); }));
  }
  unsigned long long n_slots = 0;
  uint32_t n_cells = 0;
  KGX_TRY(read_back(c, &n_slots, run_slot + n_runs));
  if (E) KGX_TRY(read_back(c, &n_cells, sm.n_sel));
  if (n_slots >= 0xfffffff0ull) return fail(c, KGX_E_LIMIT, "table: more than 2^32 potential series");
  if (2ull * n_cells + 8ull * NN >= 0xfffffff0ull) return fail(c, KGX_E_LIMIT, "table: hash table beyond 2^32 cells");
  KGX_TAKE(unsigned long long, lenA, npk + 1); KGX_TAKE(unsigned long long, lenB, npk + 1);
  KGX_TAKE(uint32_t, slot0_tmp, npk + 1); KGX_TAKE(uint32_t, stride_tmp, npk + 1);
  if (npk)
    k_x_podkeys<<<cdiv(npk, B), B, 0, c->stream>>>(in, ord, run_id_incl, run_start, n_runs, run_slot, slot0_tmp, stride_tmp, lenA, lenB, sm.stats);
  uint32_t st[BS_COUNT] = {};
  KGX_TRY(read_back(c, st, sm.stats, BS_COUNT));
  if (uint64_t(st[BS_MAX_AB]) + st[BS_MAX_G] > 0x3fffffffull) return fail(c, KGX_E_LIMIT, "table: label text too long");
  // ---------- phase 2: the ctx-owned buffers
  c->build_dirty = true;  // from here on the previous table's buffers are being overwritten
  KGX_TRY(dev_reserve(c, c->d_pk_slot0, size_t(npk) * 4));
  KGX_TRY(dev_reserve(c, c->d_pk_stride, size_t(npk) * 4));
  KGX_CUDA(c, cudaMemcpyAsync(c->d_pk_slot0.p, slot0_tmp, size_t(npk) * 4, cudaMemcpyDeviceToDevice, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(c->d_pk_stride.p, stride_tmp, size_t(npk) * 4, cudaMemcpyDeviceToDevice, c->stream));
  KGX_TRY(dev_reserve(c, c->d_A_off, (size_t(npk) + 1) * 8));
  KGX_TRY(dev_reserve(c, c->d_B_off, (size_t(npk) + 1) * 8));
  KGX_TRY(dev_reserve(c, c->d_G_off, (size_t(NG) + 1) * 8));
  KGX_TRY(scan_with_total(c, ar, lenA, ptr<unsigned long long>(c->d_A_off), npk));
  KGX_TRY(scan_with_total(c, ar, lenB, ptr<unsigned long long>(c->d_B_off), npk));
  KGX_TRY(scan_with_total(c, ar, lenG, ptr<unsigned long long>(c->d_G_off), NG));
  unsigned long long bytesA = 0, bytesB = 0, bytesG = 0;
  KGX_CUDA(c, cudaMemcpyAsync(&bytesA, ptr<unsigned long long>(c->d_A_off) + npk, 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_CUDA(c, cudaMemcpyAsync(&bytesB, ptr<unsigned long long>(c->d_B_off) + npk, 8, cudaMemcpyDeviceToHost, c->stream));
  KGX_TRY(read_back(c, &bytesG, ptr<unsigned long long>(c->d_G_off) + NG));
  KGX_TRY(dev_reserve(c, c->d_A_blob, size_t(bytesA)));
  KGX_TRY(dev_reserve(c, c->d_B_blob, size_t(bytesB)));
  KGX_TRY(dev_reserve(c, c->d_G_blob, size_t(bytesG)));
  if (npk)
    k_x_pieces<<<cdiv(npk, B), B, 0, c->stream>>>(in, ord, ptr<unsigned long long>(c->d_A_off), ptr<unsigned long long>(c->d_B_off),
                                                  ptr<uint8_t>(c->d_A_blob), ptr<uint8_t>(c->d_B_blob));
  k_x_uuid_pieces<<<cdiv(NG, B), B, 0, c->stream>>>(in, ptr<unsigned long long>(c->d_G_off), ptr<uint8_t>(c->d_G_blob));
  KGX_TRY(dev_reserve(c, c->d_node_a_base, (size_t(NN) + 1) * 4));
  KGX_TRY(dev_reserve(c, c->d_node_slot_base, (size_t(NN) + 1) * 4));
  k_node_base<<<cdiv(NN + 1, B), B, 0, c->stream>>>(node_sorted, npk, NN, ptr<uint32_t>(c->d_node_a_base), sm.stats, int(BS_MAX_NODE_PKS));
  k_x_node_slot_base<<<cdiv(NN + 1, B), B, 0, c->stream>>>(ptr<uint32_t>(c->d_node_a_base), ptr<uint32_t>(c->d_pk_slot0), npk,
                                                           uint32_t(n_slots), NN, ptr<uint32_t>(c->d_node_slot_base));
  k_max_words<<<cdiv(NN, B), B, 0, c->stream>>>(ptr<uint32_t>(c->d_node_slot_base), NN, sm.stats);
  KGX_TRY(dev_reserve(c, c->d_runs, size_t(n_runs) * G * sizeof(Run)));
  if (n_runs) k_x_runs<<<cdiv(uint64_t(n_runs) * G, B), B, 0, c->stream>>>(run_start, run_slot, node_sorted, gord, n_runs, npk, G, ptr<Run>(c->d_runs));
  KGX_TRY(dev_reserve(c, c->d_grank, size_t(NG) * 4));
  KGX_CUDA(c, cudaMemcpyAsync(c->d_grank.p, grank, size_t(NG) * 4, cudaMemcpyDeviceToDevice, c->stream));
  c->stats.kernel_launches += 16;
  KGX_TRY(launch_check(c, "table build (X)"));
  KGX_TRY(build_hash_table(c, ar, sm, cells, n_cells, ptr<uint32_t>(c->d_node_a_base), NN));
  c->plan.n_slots = n_slots; c->plan.n_a = npk;
  c->n_runs = n_runs * G;
  return KGX_OK;
}

// ---- R modes --------------------------------------------------------------------------------
// What survives the table change: every series that has a value, as (pid, pod of the previous
// generation, used, total) — taken out of the old state before anything is overwritten.
struct CarriedDev {
  uint32_t n = 0;
  uint32_t* pid = nullptr; uint32_t* oldpod = nullptr; unsigned long long* used = nullptr; unsigned long long* total = nullptr;
  uint32_t* slot = nullptr;       // [n] slot in the new table (filled by the build)
  uint32_t* old_list = nullptr;   // old pods that still name a series
  uint32_t n_old_used = 0;
};

int collect_carried_device(kgx_ctx* c, Arena& ar, const OldGen& og, CarriedDev* out) {
  CarriedDev& cd = *out;
  if (!c->have_table || c->n_series == 0) return KGX_OK;
  const uint32_t S = c->n_series;
  const int B = 256;
  KGX_TAKE(uint32_t, flag, S + 1); KGX_TAKE(uint32_t, pos, S + 1);
  k_r_carried_flags<<<cdiv(S, B), B, 0, c->stream>>>(ptr<unsigned long long>(c->d_acc[0]), S, flag);
  KGX_TRY(scan_with_total(c, ar, flag, pos, S));
  KGX_TRY(read_back(c, &cd.n, pos + S));
  if (cd.n == 0) return KGX_OK;
  KGX_TAKE(uint32_t, pid, cd.n); KGX_TAKE(uint32_t, oldpod, cd.n); KGX_TAKE(unsigned long long, used, cd.n);
  KGX_TAKE(unsigned long long, total, cd.n); KGX_TAKE(uint32_t, slot, cd.n);
  KGX_TAKE(uint32_t, old_used, og.n_pods + 1); KGX_TAKE(uint32_t, opos, og.n_pods + 1);
  KGX_CUDA(c, cudaMemsetAsync(old_used, 0, (size_t(og.n_pods) + 1) * 4, c->stream));
  k_r_carried_gather<<<cdiv(S, B), B, 0, c->stream>>>(flag, pos, S, ptr<uint32_t>(c->d_ser_slot), ptr<unsigned long long>(c->d_acc[1]),
                                                     ptr<unsigned long long>(c->d_acc[2]), og, pid, oldpod, used, total, old_used);
  KGX_TRY(scan_with_total(c, ar, old_used, opos, og.n_pods));
  KGX_TRY(read_back(c, &cd.n_old_used, opos + og.n_pods));
  KGX_TAKE(uint32_t, old_list, cd.n_old_used);
  k_compact_ids<<<cdiv(std::max(og.n_pods, 1u), B), B, 0, c->stream>>>(old_used, opos, og.n_pods, old_list);
  cd.pid = pid; cd.oldpod = oldpod; cd.used = used; cd.total = total; cd.slot = slot; cd.old_list = old_list;
  c->stats.kernel_launches += 3;
  return launch_check(c, "table build: carried series");
}

int build_table_device_r(kgx_ctx* c, Arena& ar, const TableIn& in, const BuildSmall& sm, const OldGen& og, CarriedDev& cd) {
  const uint32_t NN = in.n_nodes, G = in.G, npk = in.npk, E = in.n_entries, NG = NN * G;
  const int B = 256;
  const int ng = c->gen ^ 1;  // the generation being built
  // ---------- phase 1
  k_validate_podkeys<<<cdiv(std::max(npk, 1u), B), B, 0, c->stream>>>(in, false, sm.err);
  k_validate_entries<<<cdiv(std::max(E, 1u), B), B, 0, c->stream>>>(in, sm.err);
  uint32_t e = 0;
  KGX_TRY(read_back(c, &e, sm.err));
  if (e) return build_error(c, e);
  // pod names of the new generation: the table's pods and the pods that only carried series name
  const NameRefs refs{in, og, cd.old_list};
  const uint32_t R = npk + cd.n_old_used;
  KGX_TAKE(uint32_t, ord, R + 1); KGX_TAKE(uint32_t, head, R + 1); KGX_TAKE(uint32_t, id_incl, R + 1);
  uint32_t n_pods = 0;
  if (R) {
    k_iota<<<cdiv(R, B), B, 0, c->stream>>>(ord, R);
    KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceMergeSort::SortKeys(t, b, ord, int(R), NameLess{refs}, c->stream); }));
    k_r_name_heads<<<cdiv(R, B), B, 0, c->stream>>>(refs, ord, R, head, sm.err);
    KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceScan::InclusiveSum(t, b, head, id_incl, int(R), c->stream); }));
    KGX_TRY(read_back(c, &n_pods, id_incl + R - 1));
  }
  KGX_TRY(read_back(c, &e, sm.err));
  if (e) return build_error(c, e);
  KGX_TAKE(uint32_t, pod_of_pk, npk + 1); KGX_TAKE(uint32_t, pod_of_old, og.n_pods + 1);
  KGX_TAKE(unsigned long long, pods_len, n_pods + 1); KGX_TAKE(uint32_t, pods_ref, n_pods + 1); KGX_TAKE(uint32_t, node_name_base, NN);
  KGX_TRY(dev_reserve(c, c->d_pods_node[ng], size_t(n_pods) * 4));   // the other generation's buffers: the old one stays intact
  KGX_TRY(dev_reserve(c, c->d_pods_off[ng], (size_t(n_pods) + 1) * 8));
  if (R)
    k_r_names<<<cdiv(R, B), B, 0, c->stream>>>(refs, ord, head, id_incl, R, pod_of_pk, pod_of_old, ptr<uint32_t>(c->d_pods_node[ng]), pods_len,
                                              pods_ref, node_name_base);
  KGX_TRY(scan_with_total(c, ar, pods_len, ptr<unsigned long long>(c->d_pods_off[ng]), n_pods));
  unsigned long long pods_bytes = 0;
  KGX_TRY(read_back(c, &pods_bytes, ptr<unsigned long long>(c->d_pods_off[ng]) + n_pods));
  KGX_TRY(dev_reserve(c, c->d_pods_blob[ng], size_t(pods_bytes)));
  if (n_pods)
    k_r_pods_copy<<<cdiv(n_pods, B), B, 0, c->stream>>>(refs, pods_ref, ptr<unsigned long long>(c->d_pods_off[ng]), n_pods,
                                                        ptr<uint8_t>(c->d_pods_blob[ng]));
  // candidate series: the table's (pid, pod) pairs and the carried ones
  const uint32_t N = E + cd.n;
  if (uint64_t(E) + cd.n >= 0x7ffffff0ull) return fail(c, KGX_E_LIMIT, "table: more than 2^31 potential series");
  KGX_TAKE(uint32_t, p_node, N + 1); KGX_TAKE(unsigned long long, p_key, N + 1); KGX_TAKE(uint32_t, p_pid, N + 1);
  KGX_TAKE(uint32_t, p_pod, N + 1); KGX_TAKE(int32_t, p_src, N + 1);
  const Pairs P{p_node, p_key, p_pid, p_pod, p_src};
  if (E) {
    if (c->mode == KGX_MODE_R_LITERAL) {
      KGX_TAKE(uint32_t, cnt, npk + 1); KGX_TAKE(uint32_t, pair_off, npk + 1);
      KGX_CUDA(c, cudaMemsetAsync(cnt, 0, (size_t(npk) + 1) * 4, c->stream));
      k_r_count_entries<<<cdiv(E, B), B, 0, c->stream>>>(in, cnt);
      KGX_TRY(scan_with_total(c, ar, cnt, pair_off, npk));
      k_r_pairs_literal<<<cdiv(E, B), B, 0, c->stream>>>(in, pair_off, pod_of_pk, node_name_base, P, sm.err);
    } else {
      k_r_pairs_entries<<<cdiv(E, B), B, 0, c->stream>>>(in, pod_of_pk, node_name_base, P, sm.err);
    }
  }
  if (cd.n)
    k_r_pairs_carried<<<cdiv(cd.n, B), B, 0, c->stream>>>(cd.pid, cd.oldpod, cd.n, pod_of_old, ptr<uint32_t>(c->d_pods_node[ng]), node_name_base,
                                                         E, P, sm.err);
  KGX_TAKE(uint32_t, idx, N + 1); KGX_TAKE(uint32_t, shead, N + 1); KGX_TAKE(uint32_t, slot_incl, N + 1);
  uint32_t n_slots = 0;
  if (N) {
    k_iota<<<cdiv(N, B), B, 0, c->stream>>>(idx, N);
    KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceMergeSort::SortKeys(t, b, idx, int(N), PairLess{p_node, p_key}, c->stream); }));
    k_r_slot_heads<<<cdiv(N, B), B, 0, c->stream>>>(idx, N, p_node, p_key, shead);
    KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) { return cub::DeviceScan::InclusiveSum(t, b, shead, slot_incl, int(N), c->stream); }));
    KGX_TRY(read_back(c, &n_slots, slot_incl + N - 1));
  }
  KGX_TRY(read_back(c, &e, sm.err));
  if (e) return build_error(c, e);
  KGX_TRY(dev_reserve(c, c->d_slot_pid[ng], size_t(n_slots) * 4));
  KGX_TRY(dev_reserve(c, c->d_slot_pod[ng], size_t(n_slots) * 4));
  KGX_TAKE(uint32_t, slot_node, n_slots + 1); KGX_TAKE(unsigned long long, lenA, n_slots + 1);
  KGX_TAKE(uint32_t, cell_flag, n_slots + 1); KGX_TAKE(unsigned long long, cell_key, n_slots + 1); KGX_TAKE(unsigned long long, cells, n_slots + 1);
  uint32_t n_cells = 0;
  if (N) {
    k_r_slots<<<cdiv(N, B), B, 0, c->stream>>>(idx, shead, slot_incl, N, P, ptr<uint32_t>(c->d_slot_pid[ng]), ptr<uint32_t>(c->d_slot_pod[ng]),
                                              slot_node, lenA, cell_flag, cell_key, cd.slot);
    k_r_label_len<<<cdiv(n_slots, B), B, 0, c->stream>>>(ptr<uint32_t>(c->d_slot_pod[ng]), n_slots, ptr<uint8_t>(c->d_pods_blob[ng]),
                                                        ptr<unsigned long long>(c->d_pods_off[ng]), lenA, sm.stats);
    KGX_TRY(cub_run(c, ar, [&](void* t, size_t& b) {
      return cub::DeviceSelect::Flagged(t, b, cell_key, cell_flag, cells, sm.n_sel, int(n_slots), c->stream);
    }));
    KGX_TRY(read_back(c, &n_cells, sm.n_sel));
  }
  uint32_t st[BS_COUNT] = {};
  KGX_TRY(read_back(c, st, sm.stats, BS_COUNT));
  if (uint64_t(st[BS_MAX_AB]) > 0x3fffffffull) return fail(c, KGX_E_LIMIT, "table: label text too long");
  if (2ull * n_cells + 8ull * NN >= 0xfffffff0ull) return fail(c, KGX_E_LIMIT, "table: hash table beyond 2^32 cells");
  // ---------- phase 2
  c->build_dirty = true;
  KGX_TRY(dev_reserve(c, c->d_A_off, (size_t(n_slots) + 1) * 8));
  KGX_TRY(dev_reserve(c, c->d_B_off, (size_t(n_slots) + 1) * 8));
  KGX_TRY(dev_reserve(c, c->d_G_off, (size_t(NG) + 1) * 8));
  KGX_CUDA(c, cudaMemsetAsync(c->d_B_off.p, 0, (size_t(n_slots) + 1) * 8, c->stream));
  KGX_CUDA(c, cudaMemsetAsync(c->d_G_off.p, 0, (size_t(NG) + 1) * 8, c->stream));
  KGX_TRY(scan_with_total(c, ar, lenA, ptr<unsigned long long>(c->d_A_off), n_slots));
  unsigned long long bytesA = 0;
  KGX_TRY(read_back(c, &bytesA, ptr<unsigned long long>(c->d_A_off) + n_slots));
  KGX_TRY(dev_reserve(c, c->d_A_blob, size_t(bytesA)));
  KGX_TRY(dev_reserve(c, c->d_B_blob, 0)); KGX_TRY(dev_reserve(c, c->d_G_blob, 0));
  KGX_TRY(dev_reserve(c, c->d_pk_slot0, 0)); KGX_TRY(dev_reserve(c, c->d_pk_stride, 0));
  KGX_TRY(dev_reserve(c, c->d_grank, 0)); KGX_TRY(dev_reserve(c, c->d_node_a_base, 0));
  if (n_slots)
    k_r_pieces<<<cdiv(n_slots, B), B, 0, c->stream>>>(ptr<uint32_t>(c->d_slot_pid[ng]), ptr<uint32_t>(c->d_slot_pod[ng]), n_slots,
                                                     ptr<uint8_t>(c->d_pods_blob[ng]), ptr<unsigned long long>(c->d_pods_off[ng]),
                                                     ptr<unsigned long long>(c->d_A_off), ptr<uint8_t>(c->d_A_blob));
  KGX_TRY(dev_reserve(c, c->d_node_slot_base, (size_t(NN) + 1) * 4));
  k_node_base<<<cdiv(NN + 1, B), B, 0, c->stream>>>(slot_node, n_slots, NN, ptr<uint32_t>(c->d_node_slot_base), sm.stats, -1);
  k_max_words<<<cdiv(NN, B), B, 0, c->stream>>>(ptr<uint32_t>(c->d_node_slot_base), NN, sm.stats);
  KGX_TRY(dev_reserve(c, c->d_runs, size_t(NN) * sizeof(Run)));
  k_r_runs<<<cdiv(NN, B), B, 0, c->stream>>>(ptr<uint32_t>(c->d_node_slot_base), NN, ptr<Run>(c->d_runs));
  c->stats.kernel_launches += 16;
  KGX_TRY(launch_check(c, "table build (R)"));
  KGX_TRY(build_hash_table(c, ar, sm, cells, n_cells, ptr<uint32_t>(c->d_node_slot_base), NN));
  c->plan.n_slots = n_slots; c->plan.n_a = n_slots;
  c->n_runs = NN;
  c->gen = ng; c->n_pods = n_pods;
  return KGX_OK;
}

}  // namespace
