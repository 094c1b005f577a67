// kgx_table.hpp — host-side (cold path) plan of everything the device needs to
// join, reduce and render without sorting or string work in the hot loop.
//
// Built from the caller's join table — the flat form of `podPIDMap`
// (main.go:84-114) — whenever it changes.  It fixes, once, the order in which
// Registry.Gather would emit every series the table CAN produce (rules O2/O3 of
// label pairs sorted by name, series sorted by label values as
// byte strings), so that the hot path only ever scans in slot order.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/kgx.h"

namespace kgx {

struct Entry {      // one open-addressing table cell (8 bytes)
  uint32_t key;     // PID (R literal: slice index)
  uint32_t val;     // bit31: another cell with this key follows in the probe
                    // chain; bits 0..30: R slot id / X sorted podkey index;
                    // 0xFFFFFFFF = empty
};
constexpr uint32_t kEmpty = 0xFFFFFFFFu;
constexpr uint32_t kMore = 0x80000000u;

struct NodeTab { uint32_t off, cap; };  // a node's slice of the table

struct Run {             // a maximal range of slots sharing node (and, in X
  uint32_t slot_start;   // mode, container and gpu): inverse map slot -> labels
  uint32_t len;
  uint32_t a_first;      // label piece index of the run's first slot
  uint32_t gidx;         // node*G+gpu, or 0xFFFFFFFF when gpu is not a label
  uint32_t node;
};

struct FamilyDef { const char* name; const char* help; int value_kind; };
// value kinds
enum { V_R_USED = 0, V_R_PERC, V_X_SUM, V_X_PERC, V_X_MAX, V_X_SM, V_X_MEM, V_X_TS };

struct Plan {
  int mode = 0;
  uint32_t n_nodes = 0, G = 0;
  uint64_t n_slots = 0, n_a = 0;
  std::vector<uint32_t> node_slot_base;  // [n_nodes+1]
  std::vector<Run> runs;                 // ascending slot_start
  std::vector<NodeTab> node_tab;         // [n_nodes]
  std::vector<Entry> tab;
  std::vector<uint32_t> pk_slot0, pk_stride;  // X: per sorted podkey
  std::vector<uint32_t> grank;                // X: [n_nodes*G] order of the gpu label value in its node
  std::vector<uint32_t> node_a_base;          // X: [n_nodes+1] first sorted-podkey index of each node
  uint32_t max_node_cells = 0, max_node_pks = 0;  // largest per-node table slice / podkey count
  // label pieces: text of slot s = A[a] + GL[gidx] + B[a]
  std::string A_blob, B_blob, G_blob;
  std::vector<uint64_t> A_off, B_off, G_off;
  uint32_t max_label_len = 0;
  uint64_t n_table_entries = 0;
  // R modes: identity of each slot (for carry-over across tables)
  std::vector<uint32_t> slot_pid;        // label value of pid
  std::vector<uint32_t> slot_pod;        // index into pod_names
  std::vector<std::string> pod_names;    // per (node-scoped) pod
  std::vector<uint32_t> pod_node;
  std::vector<uint32_t> carried_slot;    // slot of each carried series (same order as the input)
};

// A series that already has a value and must survive a table change (R modes:
// GaugeVec children are never deleted, main.go has no Delete/Reset).
struct Carried { uint32_t node, pid; std::string pod; uint64_t used, total; };

// Returns KGX_OK or KGX_E_*; err receives a message.
int build_plan(int mode, const kgx_table* t, const std::vector<Carried>& carried, Plan* out, std::string* err);

// families of a mode in exposition order (O1: sorted by name)
const std::vector<FamilyDef>& families(int mode);
const char* const* label_names(int mode, int* n);

// 32-bit mixer used by both the host builder and the device probe
inline uint32_t hash_pid(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

}  // namespace kgx
