// Collector and /metrics handler over libkgx: the drop-in for the two GaugeVecs
// registered at main.go:40-42 and for promhttp.HandlerFor at main.go:69-70.
//
// NOT COMPILED ON THE BUILD BOX (no Go toolchain); see INTEGRATION.md.
package kgx

import (
	"net/http"
	"strings"

	"github.com/prometheus/client_golang/prometheus"
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
	for f, d := range c.descs {
		vals, slots, err := c.ctx.ReadValues(f)
		if err != nil {
			ch <- prometheus.NewInvalidMetric(d, err)
			continue
		}
		for i, v := range vals {
			ch <- prometheus.MustNewConstMetric(d, prometheus.GaugeValue, v, labelValues(c.ctx.SlotLabels(slots[i]))...)
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
func Handler(ctx *Ctx, fallback http.Handler) http.Handler {
	var buf []byte
	return http.HandlerFunc(func(w http.ResponseWriter, r *http.Request) {
		if a := r.Header.Get("Accept"); strings.Contains(a, "protobuf") || strings.Contains(a, "openmetrics") {
			fallback.ServeHTTP(w, r)
			return
		}
		body, err := ctx.Render(buf)
		if err != nil { // promhttp.HandlerOpts{} default: HTTP 500
			http.Error(w, err.Error(), http.StatusInternalServerError)
			return
		}
		buf = body[:0]
		w.Header().Set("Content-Type", "text/plain; version=0.0.4; charset=utf-8; escaping=values")
		w.Write(body)
	})
}
