/*
 * kgx.h — C ABI of libkgx: the B200-native join → reduce → render path of a
 * per-pod GPU metrics exporter.
 *
 * This is the drop-in boundary for the hot path of the reference exporter
 * (reference: /root/reference/main.go).  The reference has no FFI of its own:
 * its hot path is the inline loop main.go:141-154 (join + value calc +
 * GaugeVec upsert) and the promhttp render behind main.go:69-71.  Each entry
 * point below names the reference lines it replaces.  A Go host keeps the
 * NVML polling (main.go:117-138), the pod/PID discovery (main.go:77-114) and
 * the prometheus.Collector / http.Handler surface, and calls these functions
 * through cgo.
 *
 * Conventions
 *   - plain C, no C++ types, no exceptions across the boundary;
 *   - every function returns 0 (KGX_OK) or a negative KGX_E_* code and never
 *     aborts the process (the reference calls log.Fatalf / panic instead,
 *     main.go:47,59,79,119);
 *   - the caller owns every buffer it passes in; the library copies what it
 *     needs before returning and owns all device memory;
 *   - one kgx_ctx per GPU, no global state; calls on one ctx may come from
 *     any number of threads (the reference polls in one goroutine and serves
 *     scrapes in others, main.go:68-74).  Calls that touch the series state
 *     are serialised by the library: one writer, any number of readers, and
 *     a reader sees the last COMPLETED ingest.  kgx_ingest copies its batch
 *     to the GPU before it takes that lock, so a scrape never waits for a
 *     batch in transit, only for a join that is actually running;
 *   - kgx_last_error is per calling thread;
 *   - there is NO CPU fallback: without a CUDA device kgx_create fails with
 *     KGX_E_CUDA.
 */
#ifndef KGX_H
#define KGX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 2: kgx_ingest_device returns with the join in flight; kgx_last_error is per calling thread; kgx_body_node and
 *    kgx_read_labels added.  Struct layouts are those of version 1. */
#define KGX_ABI_VERSION 2

/* ---- error codes -------------------------------------------------------- */
#define KGX_OK 0
#define KGX_E_INVAL (-1)   /* bad argument, bad UTF-8 in a label, bad ids     */
#define KGX_E_NOMEM (-2)   /* host or device allocation failed                */
#define KGX_E_CUDA (-3)    /* CUDA runtime error; text via kgx_last_error     */
#define KGX_E_NOSPACE (-4) /* caller's output buffer too small; *len = needed */
#define KGX_E_STATE (-5)   /* call order (e.g. ingest before set_table)       */
#define KGX_E_LIMIT (-6)   /* size beyond a documented maximum                */

/* ---- result modes (SURVEY.md section 8 "Modes") ---------------------------
 * R modes reproduce main.go exactly: families and help strings of
 * main.go:22-35, labels pid,pod, values of main.go:147,149, last write wins
 * in sample order, series persist forever (no Delete/Reset in main.go).
 *   KGX_MODE_R_INTENDED  join predicate "PID value equality" (what
 *                        main.go:144-145 means to do)
 *   KGX_MODE_R_LITERAL   join predicate exactly as written at main.go:144-145:
 *                        `for pid := range pids` yields slice INDICES, so a
 *                        process matches pod P iff Pid < len(pids of P) and
 *                        the pid label is that index.
 * X mode is the north-star extension (not in the reference): key
 * (pod, namespace, container, gpu); integer sum / max / last-timestamp /
 * count reductions over a window of ingests.
 */
#define KGX_MODE_R_INTENDED 0
#define KGX_MODE_R_LITERAL 1
#define KGX_MODE_X 2

typedef struct kgx_ctx kgx_ctx;

typedef struct kgx_opts {
  int32_t abi_version;   /* must be KGX_ABI_VERSION                           */
  int32_t mode;          /* KGX_MODE_*                                        */
  void* stream;          /* cudaStream_t to run on, or NULL: library-owned    */
  int32_t profile;       /* 1: record per-kernel CUDA-event times             */
  int32_t reserved;
} kgx_opts;

/* The join table: the device-side form of `podPIDMap map[string][]string`
 * (main.go:84,109,113), widened by node id and, in X mode, namespace and
 * container.  A "podkey" is one map key: the pod name in R modes
 * (main.go:113 keys by name only), (pod, namespace, container) in X mode.
 * Entries are the PID lists flattened: entry i says "pid entry_pid[i] is in
 * the list of podkey entry_podkey[i]"; duplicates are allowed, as the
 * reference appends one `ps` output per container (main.go:109).  Strings are
 * raw (unescaped) UTF-8, given as one blob plus n+1 byte offsets.
 */
typedef struct kgx_table {
  uint32_t n_nodes;        /* >= 1; node ids are 0..n_nodes-1                  */
  uint32_t gpus_per_node;  /* G >= 1; gpu ids are 0..G-1 within a node         */
  uint64_t n_podkeys;
  const uint32_t* podkey_node;   /* [n_podkeys]                                */
  const char* pod_blob;
  const uint64_t* pod_off;       /* [n_podkeys+1]                              */
  const char* namespace_blob;    /* X mode only, else NULL                     */
  const uint64_t* namespace_off; /* [n_podkeys+1]                              */
  const char* container_blob;    /* X mode only, else NULL                     */
  const uint64_t* container_off; /* [n_podkeys+1]                              */
  uint64_t n_entries;
  const uint32_t* entry_podkey;  /* [n_entries]                                */
  const uint32_t* entry_pid;     /* [n_entries]                                */
  const char* gpu_uuid_blob;     /* X mode only: value of the gpu label        */
  const uint64_t* gpu_uuid_off;  /* [n_nodes*G+1], index node*G+gpu            */
} kgx_table;

/* One batch of per-process samples in column form: what main.go:123-138 reads
 * from NVML each cycle (nvml.h nvmlProcessInfo_t: pid, usedGpuMemory;
 * nvmlMemory_t: total), in the loop order of main.go:123,141 (device index
 * ascending, then process order) — in R modes the LAST sample in this order
 * wins for a series, as repeated GaugeVec.Set does.  X mode adds the
 * utilisation and timestamp columns (nvml.h nvmlProcessUtilizationSample_t).
 * All pointers are host memory (pageable or from kgx_host_alloc).
 */
typedef struct kgx_cols {
  uint64_t n;
  const uint32_t* pid;       /* [n]                                            */
  const uint32_t* gpu;       /* [n] device index within the node               */
  const uint64_t* used;      /* [n] usedGpuMemory bytes                        */
  const uint32_t* node;      /* [n] or NULL (= all node 0)                     */
  const uint32_t* sm_util;   /* [n] X mode                                     */
  const uint32_t* mem_util;  /* [n] X mode                                     */
  const uint64_t* ts_us;     /* [n] X mode                                     */
  const uint64_t* node_off;  /* optional [n_nodes+1]: samples of node k are
                                rows node_off[k]..node_off[k+1]; lets the join
                                stage each node's table slice in shared memory */
  const uint64_t* gpu_total; /* [n_nodes*G] memoryInfo.Total, main.go:129,149  */
} kgx_cols;

typedef struct kgx_stats {
  uint64_t n_series;        /* series with a value after the last ingest       */
  uint64_t n_slots;         /* potential series the table can produce          */
  uint64_t n_matched;       /* samples of the last ingest that joined          */
  uint64_t n_dropped;       /* samples of the last ingest with no table entry  */
  uint64_t rendered_bytes;  /* size of the last rendered body                  */
  uint64_t kernel_launches; /* kernels launched by this ctx so far             */
  uint64_t rebuilds;        /* series-index rebuilds (churn) so far            */
  uint64_t device_bytes;    /* device memory currently held                    */
} kgx_stats;

/* Per-kernel device times (opts.profile = 1), accumulated since the last
 * kgx_profile_reset.  Index with KGX_K_*. */
#define KGX_K_INGEST 0   /* join + reduce                                      */
#define KGX_K_FINALIZE 1 /* R modes: fetch values of last-writer samples       */
#define KGX_K_INDEX 2    /* series-index rebuild kernels (churn only)          */
#define KGX_K_RENDER 3   /* exposition writer                                  */
#define KGX_K_OTHER 4    /* clears, tile bookkeeping                           */
#define KGX_K_COUNT 5
typedef struct kgx_profile {
  double ms[KGX_K_COUNT];
  uint64_t launches[KGX_K_COUNT];
} kgx_profile;

/* ---- lifecycle ---------------------------------------------------------- */

/* Create a context on CUDA device `device`.
 * Replaces: the registry + two GaugeVecs of main.go:21-36,40-42 (the schema
 * is fixed by `mode`). */
int kgx_create(int device, const kgx_opts* opts, kgx_ctx** out);
void kgx_destroy(kgx_ctx* ctx);

/* ---- cold path ---------------------------------------------------------- */

/* Install the join table (podPIDMap of main.go:84-114).  Copies the caller's
 * arrays to the GPU and builds there the output order of every series the
 * table can produce (Registry.Gather's sort, see DESIGN.md "Order"), the
 * pre-escaped label bytes and the open-addressing hash table.  (With
 * KGX_HOST_TABLE_BUILD=1 in the environment at kgx_create the same plan is
 * built by a host routine and uploaded; results are identical.)
 * A table that is rejected (KGX_E_INVAL: ids out of range, label values that
 * are not valid UTF-8, duplicate keys; KGX_E_LIMIT on its sizes) leaves the ctx
 * as it was.  One limit is only known once the old table has been released —
 * label text of a single series too long for a render tile (about 200 KB):
 * that KGX_E_LIMIT leaves the ctx without a table (kgx_ingest -> KGX_E_STATE).
 * R modes: series already exported keep their last value across calls, as
 * GaugeVec children do (main.go has no Delete/Reset).  X mode: clears the
 * window. */
int kgx_set_table(kgx_ctx* ctx, const kgx_table* table);

/* ---- hot path ----------------------------------------------------------- */

/* Join one batch of samples to the table and fold it into the per-series
 * state.  Replaces the loop at main.go:141-154 (join :141-145, values
 * :147,149, upsert :147,150). */
int kgx_ingest(kgx_ctx* ctx, const kgx_cols* cols);

/* Same, with every column pointer being DEVICE memory already resident on
 * ctx's GPU (no host<->device copy).  Used to time the device pipeline.
 * Contract (checked by kgx_ingest on the host, NOT here, where the offsets are
 * device memory): node_off, when given, has n_nodes+1 entries, starts at 0,
 * never decreases and ends at n.  Offsets beyond n make the join read past the
 * columns; rows outside [node_off[0], node_off[n_nodes]) are ignored.  When
 * both `node` and `node_off` are given, node_off decides the node of a row.
 * Unlike kgx_ingest, this call RETURNS WITH THE JOIN IN FLIGHT on ctx's stream:
 * the caller's device columns must stay valid, and unmodified, until the next
 * call on ctx that delivers results (kgx_render*, kgx_read_*, kgx_get_stats,
 * kgx_sync, the next ingest ...).  That call also settles the batch: it reads
 * the join's statistics and, if the batch brought series the index did not
 * hold yet, rebuilds the index and folds what the first pass skipped (which
 * allocates; call kgx_sync first if the next call must not, as
 * kgx_render_noalloc).  A render queued behind an unsettled batch costs no
 * extra host round trip: steady state is join kernel, writer kernel, one
 * synchronisation. */
int kgx_ingest_device(kgx_ctx* ctx, const kgx_cols* cols);

/* R modes: ingest straight from what NVML hands back, without building columns
 * on the host.  `infos` is every device's nvmlProcessInfo_t array
 * (DeviceGetComputeRunningProcesses, main.go:135; struct at nvml.h:294-303)
 * concatenated in the loop order of main.go:123 — node 0 device 0, node 0
 * device 1, ... — and dev_off[node*G+gpu] .. dev_off[node*G+gpu+1] is the row
 * range of one device (n_nodes*G+1 entries); gpu_total as in kgx_cols.  The
 * library unpacks on the GPU and then runs the same join as kgx_ingest. */
typedef struct kgx_nvml_process_info {  /* = nvmlProcessInfo_t, 24 bytes */
  uint32_t pid;
  uint32_t pad_;
  uint64_t used_gpu_memory;
  uint32_t gpu_instance_id;
  uint32_t compute_instance_id;
} kgx_nvml_process_info;
int kgx_ingest_nvml(kgx_ctx* ctx, const kgx_nvml_process_info* infos, const uint64_t* dev_off,
                    const uint64_t* gpu_total);

/* X mode: the same for the wider sample.  NVML reports a process twice: memory
 * in nvmlProcessInfo_t (as above) and utilisation in
 * nvmlProcessUtilizationSample_t (nvmlDeviceGetProcessUtilization; struct at
 * nvml.h:1277-1285).  `utils` is every device's sample array concatenated in
 * the same device order, util_off[node*G+gpu] .. util_off[node*G+gpu+1] the
 * samples of one device.  A sample row is one process-info row joined, on the
 * GPU, with the utilisation sample of the same pid on the same device (the last
 * one in array order if a pid occurs twice; sm = mem = timestamp = 0 if none);
 * utilisation samples without a process-info row are ignored.  The reference
 * never calls this NVML function: X mode is the extension. */
typedef struct kgx_nvml_process_utilization_sample {  /* = nvmlProcessUtilizationSample_t, 32 bytes */
  uint32_t pid;
  uint32_t pad_;
  uint64_t time_stamp;  /* CPU timestamp in microseconds */
  uint32_t sm_util;
  uint32_t mem_util;
  uint32_t enc_util;
  uint32_t dec_util;
} kgx_nvml_process_utilization_sample;
int kgx_ingest_nvml_x(kgx_ctx* ctx, const kgx_nvml_process_info* infos, const uint64_t* dev_off,
                      const kgx_nvml_process_utilization_sample* utils, const uint64_t* util_off, const uint64_t* gpu_total);

/* X mode: start a new aggregation window (forget sums/max/last/count).
 * No-op error KGX_E_STATE in R modes, whose series never reset. */
int kgx_reset_window(kgx_ctx* ctx);

/* X mode: explicit staleness policy.  The reference
 * keeps a series for ever once it has been Set (main.go has no Delete/Reset,
 * rule a10) and the R modes stay faithful to that: there this call returns
 * KGX_E_STATE.  In X mode it drops every series whose newest sample is older
 * than `older_than_ts_us` (same clock as kgx_cols.ts_us): the series leaves the
 * exposition, its window state is forgotten, and its place in the series
 * index, label text and render tiles is reclaimed; a later sample for the same
 * key starts a new series from zero.  *n_expired (may be NULL) = exported
 * series dropped.  Nothing happens unless the embedder calls it. */
int kgx_expire(kgx_ctx* ctx, uint64_t older_than_ts_us, uint64_t* n_expired);

/* Render the Prometheus text exposition (format 0.0.4) of the current state
 * into ctx-owned device memory.  Replaces what promhttp.HandlerFor serves at
 * main.go:69-71 (Registry.Gather + expfmt text encoder).  One body per node,
 * concatenated in node order.  *len receives the total size. */
int kgx_render_device(kgx_ctx* ctx, size_t* len);

/* Render (as above) and copy the bytes to host memory `out` of capacity
 * `cap`.  If cap is too small returns KGX_E_NOSPACE with *len = needed. */
int kgx_render(kgx_ctx* ctx, uint8_t* out, size_t cap, size_t* len);

/* Render without allocating anything.  (No counterpart in the reference: Go's
 * promhttp handler allocates freely.  It exists for embedders that cannot call
 * cudaMalloc when a scrape arrives: the exporter may share its GPU with a job
 * that has reserved every byte of device memory, and a latency-bound handler
 * cannot afford the device-wide stall that cudaMalloc/cudaFree bring.)
 * Same bytes as kgx_render_device, but written into `d_out`, DEVICE memory the
 * caller owns (16-byte aligned, `cap` bytes), and with the writer's scratch in
 * a fixed array inside the library image instead of ctx-owned buffers.  That
 * array takes KGX_NOALLOC_SCRATCH_BYTES on every GPU that loads the library,
 * for the life of the process.  It makes no cudaMalloc/cudaFree (nor any other
 * allocating call) and leaves the ctx-owned body of kgx_render untouched.
 *   - Limits: at most KGX_NOALLOC_MAX_TILES render tiles (a tile is up to 128
 *     series of one family of one node) and KGX_NOALLOC_MAX_NODES nodes;
 *     beyond that it returns KGX_E_LIMIT and kgx_render must be used.
 *   - `cap` must cover the worst-case body size; if it does not (or d_out is
 *     NULL) it returns KGX_E_NOSPACE with that size in *len.
 *   - The array's device address is looked up once per GPU
 *     (cudaGetSymbolAddress, with that GPU current): when the first ctx is
 *     created on it, or, should that have failed, by the first call here.
 *     Create the ctx — and to be safe make the first call — outside any
 *     stream capture.
 *   - All calls, from every ctx on that GPU, share the one array: it is cleared
 *     at the start of each call, and two calls that overlap on the GPU (two
 *     ctxs on different streams) race.  Serialise them. */
#define KGX_NOALLOC_MAX_TILES 65536
#define KGX_NOALLOC_MAX_NODES 4096
#define KGX_NOALLOC_SCRATCH_BYTES ((8 + KGX_NOALLOC_MAX_TILES + KGX_NOALLOC_MAX_NODES + 1) * 8) /* 557 128 */
int kgx_render_noalloc(kgx_ctx* ctx, uint8_t* d_out, size_t cap, size_t* len);

/* Split form of kgx_render for a pipelined exporter: kgx_render_begin renders,
 * returns the size, and STARTS the device->host copy of the body into `out`
 * (pinned memory from kgx_host_alloc, or the copy cannot overlap) on a second
 * stream; kgx_render_end waits until the bytes are in `out`.  Between the two
 * the caller may already kgx_reset_window / kgx_ingest the next scrape: its
 * host->device copies and kernels overlap the outgoing body on the full-duplex
 * link.  kgx_render_begin may also be called while the previous body is still
 * in flight: the library then renders into a second device buffer (allocated
 * on first need, same size as the first) and queues the new copy behind the
 * old one, so the link never idles; kgx_render_end waits for the OLDEST copy in
 * flight.  At most two may be outstanding (a third begin waits for the first).
 * kgx_render, kgx_set_table and an index rebuild wait for every copy in flight.
 * The reference serves a scrape and polls concurrently too (main.go:68-72). */
int kgx_render_begin(kgx_ctx* ctx, uint8_t* out, size_t cap, size_t* len);
int kgx_render_end(kgx_ctx* ctx);

/* Device pointer and size of the last rendered text (valid until the next
 * render on this ctx); for peer copies / NCCL gathers of bodies. */
int kgx_body_device(kgx_ctx* ctx, const uint8_t** dptr, size_t* len);

/* Byte offset of every node's body inside the last rendered text:
 * off[k]..off[k+1] is node k's /metrics body; off has n_nodes+1 entries. */
int kgx_body_offsets(kgx_ctx* ctx, uint64_t* off, size_t n);

/* One node's /metrics body out of the last rendered text, copied to host
 * memory `out`: what the per-node exporter of main.go:69-71 serves for one
 * scrape.  KGX_E_STATE before any render; KGX_E_NOSPACE with *len = needed if
 * cap is too small.  The body stays valid until the next render on this ctx. */
int kgx_body_node(kgx_ctx* ctx, uint32_t node, uint8_t* out, size_t cap, size_t* len);

/* A byte range of the last rendered text, copied to host memory `out`: up to
 * `cap` bytes from `offset`; *n receives how many were copied (0 at the end of
 * the body).  For a handler that streams the body with chunked transfer
 * encoding instead of holding all of it in host memory: kgx_render_device,
 * then kgx_body_read chunk after chunk into the http.ResponseWriter (what
 * promhttp does with its pooled buffer behind main.go:69-71).
 * The first chunk is out one render and one chunk copy after the scrape
 * arrived, whatever the body's size.  KGX_E_STATE before any render,
 * KGX_E_INVAL if offset lies beyond the body.  The body stays valid until the
 * next render on this ctx and no lock is held between calls: the handler
 * serialises its scrapes (promhttp's MaxRequestsInFlight = 1 does the same). */
int kgx_body_read(kgx_ctx* ctx, uint64_t offset, uint8_t* out, size_t cap, size_t* n);

/* ---- Collector read-back (prometheus.Collector.Collect) ------------------ */

/* Values of family `family` (index in sorted family order) for every series
 * that has a value, in exposition order; *n receives the count.  With
 * slots != NULL also returns each series' slot id, resolvable to label values
 * with kgx_slot_labels.  Replaces GaugeVec.Collect reached from main.go:41-42. */
int kgx_read_values(kgx_ctx* ctx, int family, double* values, uint32_t* slots,
                    size_t cap, size_t* n);
int kgx_family_count(kgx_ctx* ctx);
/* name/help/label names of a family; pointers valid for ctx's lifetime. */
int kgx_family_info(kgx_ctx* ctx, int family, const char** name,
                    const char** help, const char* const** label_names,
                    int* n_labels);
/* Label text of EVERY series that has a value, in one call: `blob` receives the
 * pre-escaped texts `{a="..",b=".."}` back to back in exposition order (the
 * order of kgx_read_values: entry i belongs to value i of any family), off[i]
 * .. off[i+1] the bytes of series i (n+1 offsets).  *n and *blob_len receive
 * the counts; with blob == off == NULL only those are returned, and if a
 * buffer is too small the call returns KGX_E_NOSPACE with both set.  Two
 * device->host copies in the common case; replaces one kgx_slot_labels round
 * trip per series in Collector.Collect (GaugeVec.Collect, main.go:41-42). */
int kgx_read_labels(kgx_ctx* ctx, char* blob, size_t blob_cap, uint64_t* off,
                    size_t off_cap, size_t* n, size_t* blob_len);
/* The pre-escaped label text `{a="..",b=".."}` of a slot; copies at most cap
 * bytes, *len = full length. */
int kgx_slot_labels(kgx_ctx* ctx, uint32_t slot, char* out, size_t cap,
                    size_t* len);

/* ---- utilities ---------------------------------------------------------- */

/* Format n doubles on the GPU exactly as expfmt writes sample values
 * (strconv.AppendFloat(f,'g',-1,64) plus the 0/1/-1/NaN/Inf special cases),
 * one value per 32-byte record: bytes 0..23 text, byte 31 length.  Test and
 * diagnostic entry point for the device float printer. */
int kgx_format_values(kgx_ctx* ctx, const double* values, size_t n,
                      uint8_t* out32);

/* Pinned host memory for sample columns and output (faster PCIe copies). */
void* kgx_host_alloc(size_t bytes);
void kgx_host_free(void* p);

int kgx_get_stats(kgx_ctx* ctx, kgx_stats* out);
int kgx_get_profile(kgx_ctx* ctx, kgx_profile* out);
int kgx_profile_reset(kgx_ctx* ctx);
/* Turn per-kernel event timing on (1) or off (0) after creation. */
int kgx_set_profile(kgx_ctx* ctx, int on);
/* Block until all work queued by this ctx has finished. */
int kgx_sync(kgx_ctx* ctx);

const char* kgx_strerror(int code);
/* Detail of the last error of the CALLING THREAD (valid until its next call into the library). */
const char* kgx_last_error(kgx_ctx* ctx);
int kgx_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* KGX_H */
