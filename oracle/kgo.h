/*
 * kgo.h — CPU ORACLE for the join → reduce → render path.  TEST INFRASTRUCTURE.
 *
 * This is a CPU restatement of what /root/reference/main.go does on the hot
 * path (main.go:141-154 join/values/upsert, main.go:69-71 render) together
 * with the third-party encoder rules it depends on (client_golang v1.20.2,
 * prometheus/common v0.55.0 expfmt, Go strconv; go.mod:3,7,36) — none of which
 * is vendored under /root/reference, so they are restated from their
 * published behaviour (rules O1-O3, G1-G2, F1-F5, V1-V2).
 *
 * PARITY STATUS: "parity unpinned" against Go itself — the reference has no
 * tests, no golden /metrics files, and no Go toolchain exists on the build
 * box.  The oracle is pinned instead against (a) every known-answer vector of
 * SURVEY.md section 8c, (b) the hand-derived whole-body goldens in tests/golden,
 * and (c) for the shortest-float digits, CPython's repr() on millions of
 * doubles (an independent shortest-round-trip implementation).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product (libkgx.so) never
 * links, loads or calls it.
 *
 * It shares the plain-C input structs of include/kgx.h so the same synthetic
 * inputs can be fed to both sides.
 */
#ifndef KGO_H
#define KGO_H
#include "../include/kgx.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct kgo_ctx kgo_ctx;

/* join strategy */
#define KGO_JOIN_LOOPS 0 /* the nested loops of main.go:141-145, literally     */
#define KGO_JOIN_HASH 1  /* same results via a per-node hash map of PIDs       */

kgo_ctx* kgo_create(int mode, int join, int threads);
void kgo_destroy(kgo_ctx*);
int kgo_set_table(kgo_ctx*, const kgx_table*);
int kgo_ingest(kgo_ctx*, const kgx_cols*);
int kgo_reset_window(kgo_ctx*);
/* X mode, explicit expiry (mirrors kgx_expire): series whose newest sample is older than the
 * cutoff stop being exported and lose their window state.  R modes: KGX_E_STATE (main.go never deletes). */
int kgo_expire(kgo_ctx*, uint64_t older_than_ts_us, uint64_t* n_expired);
/* Render; *out stays valid until the next call on the ctx. */
int kgo_render(kgo_ctx*, const uint8_t** out, size_t* len);
int kgo_body_offsets(kgo_ctx*, uint64_t* off, size_t n);
int kgo_read_values(kgo_ctx*, int family, double* values, size_t cap, size_t* n);
uint64_t kgo_series_count(kgo_ctx*);

/* expfmt writeFloat: returns the length, writes at most 32 bytes. */
int kgo_format_value(double v, char* out);
/* batch variant for tests: one 32-byte record per value (text, len in [31]). */
void kgo_format_values(const double* v, size_t n, uint8_t* out32);
/* the 128-bit power-of-ten table the oracle derives for itself (k in
 * [-292,324]); lets tests cross-check it with an independent big-int model. */
void kgo_pow10_entry(int k, uint64_t* hi, uint64_t* lo);
/* escape a label value / help text as expfmt does; returns the length. */
size_t kgo_escape(const char* in, size_t n, int is_label, char* out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
