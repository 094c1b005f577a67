// Collector and /metrics handler over libkgx: the drop-in for the two GaugeVecs
// registered at main.go:40-42 and for promhttp.HandlerFor at main.go:69-70.
//
// NOT COMPILED ON THE BUILD BOX (no Go toolchain); see INTEGRATION.md.
package kgx

import (
	"fmt"
	"net/http"
	"strings"
	"sync"

	"github.com/prometheus/client_golang/prometheus"
	"github.com/prometheus/common/expfmt"
)

// Collector satisfies prometheus.Collector with the same fully-qualified
// names, help strings and label names as main.go:22-35.
type Collector struct {
	ctx   *Ctx
	descs []*prometheus.Desc
	fams  []FamilyInfo
}

func NewCollector(ctx *Ctx) *Collector {
	c := &Collector{ctx: ctx, fams: ctx.Families()}
	for _, f := range c.fams {
		c.descs = append(c.descs, prometheus.NewDesc(f.Name, f.Help, f.LabelNames, nil))
	}
	return c
}

func (c *Collector) Describe(ch chan<- *prometheus.Desc) {
	for _, d := range c.descs {
		ch <- d
	}
}

// Collect reads the reduced values back through the ABI (kgx_read_values) and
// emits const metrics; used for protobuf/OpenMetrics scrapes.  Text scrapes
// take the GPU-rendered body instead (Handler below).
func (c *Collector) Collect(ch chan<- prometheus.Metric) {
	// label text of all series in one call; entry i belongs to value i of every family
	texts, err := c.ctx.ReadLabels()
	if err != nil {
		for _, d := range c.descs {
			ch <- prometheus.NewInvalidMetric(d, err)
		}
		return
	}
	labels := make([][]string, len(texts))
	for i, t := range texts {
		labels[i] = labelValues(t)
	}
	for f, d := range c.descs {
		vals, _, err := c.ctx.ReadValues(f)
		if err != nil || len(vals) != len(labels) {
			ch <- prometheus.NewInvalidMetric(d, fmt.Errorf("kgx read-back: %v (%d values, %d label sets)", err, len(vals), len(labels)))
			continue
		}
		for i, v := range vals {
			ch <- prometheus.MustNewConstMetric(d, prometheus.GaugeValue, v, labels[i]...)
		}
	}
}

// labelValues parses `{a="x",b="y"}` (pre-escaped by the library) back to values.
func labelValues(s string) []string {
	var out []string
	s = strings.TrimSuffix(strings.TrimPrefix(s, "{"), "}")
	for len(s) > 0 {
		i := strings.Index(s, `="`)
		s = s[i+2:]
		var b strings.Builder
		j := 0
		for ; j < len(s) && s[j] != '"'; j++ {
			if s[j] == '\\' && j+1 < len(s) {
				j++
				if s[j] == 'n' {
					b.WriteByte('\n')
				} else {
					b.WriteByte(s[j])
				}
			} else {
				b.WriteByte(s[j])
			}
		}
		out = append(out, b.String())
		s = strings.TrimPrefix(s[j+1:], ",")
	}
	return out
}

// Handler serves the GPU-rendered text exposition; any other negotiated
// format falls back to the registry path (Collector.Collect).
//
// The Content-Type is whatever promhttp itself would send: promhttp.HandlerFor
// (main.go:69) calls expfmt.Negotiate on the request header and writes the
// result verbatim, so the same call here gives the same header, including the
// `escaping=` parameter of common v0.55.0, whatever its default is.
//
// net/http serves scrapes concurrently (an HA Prometheus pair, a curl during a
// scrape), so every request renders into a buffer of its own, taken from a pool
// and returned only after the body has been written.
func Handler(ctx *Ctx, fallback http.Handler) http.Handler {
	pool := sync.Pool{New: func() any { b := make([]byte, 0, 1<<16); return &b }}
	return http.HandlerFunc(func(w http.ResponseWriter, r *http.Request) {
		ct := expfmt.Negotiate(r.Header)
		if ct.FormatType() != expfmt.TypeTextPlain {
			fallback.ServeHTTP(w, r)
			return
		}
		bp := pool.Get().(*[]byte)
		body, err := ctx.Render((*bp)[:0])
		if err != nil { // promhttp.HandlerOpts{} default: HTTP 500
			pool.Put(bp)
			http.Error(w, err.Error(), http.StatusInternalServerError)
			return
		}
		w.Header().Set("Content-Type", string(ct))
		w.Write(body)
		*bp = body[:0] // keep the (possibly grown) backing array for the next scrape
		pool.Put(bp)
	})
}

// StreamingHandler serves the same bytes as Handler with chunked transfer
// encoding (net/http switches to it by itself when no Content-Length is set and
// the body outgrows its buffer): the body stays in device memory and crosses
// `chunkBytes` at a time, so a fleet-sized body (gigabytes) needs one chunk of
// host memory and its first bytes are on the wire after one render and one
// chunk copy.  One scrape at a time (the body on the device belongs to the
// scrape being served), like promhttp's MaxRequestsInFlight = 1; a second
// scrape waits.
func StreamingHandler(ctx *Ctx, fallback http.Handler, chunkBytes int) http.Handler {
	var mu sync.Mutex
	if chunkBytes <= 0 {
		chunkBytes = 8 << 20
	}
	chunk := make([]byte, chunkBytes)
	return http.HandlerFunc(func(w http.ResponseWriter, r *http.Request) {
		ct := expfmt.Negotiate(r.Header)
		if ct.FormatType() != expfmt.TypeTextPlain {
			fallback.ServeHTTP(w, r)
			return
		}
		mu.Lock()
		defer mu.Unlock()
		w.Header().Set("Content-Type", string(ct))
		if n, err := ctx.WriteBodyTo(w, chunk); err != nil && n == 0 {
			// nothing written yet: the header is still ours (promhttp.HandlerOpts{} default: HTTP 500)
			w.Header().Del("Content-Type")
			http.Error(w, err.Error(), http.StatusInternalServerError)
		}
	})
}
