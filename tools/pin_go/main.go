// pin_go — produce Go-made golden files for the parity tests.
//
// Runs the reference's own encoder stack on fixed inputs: two GaugeVecs with the
// names, help strings and label names of the reference (main.go:22-35), a
// registry (main.go:39-42) and promhttp.HandlerFor(reg, promhttp.HandlerOpts{})
// (main.go:69) scraped through net/http/httptest, at the module versions of the
// reference's go.mod (client_golang v1.20.2, common v0.55.0).  For every case of
// tests/golden/go_inputs.json it writes tests/golden/go/<name>.txt (the body)
// and <name>.hdr (the Content-Type).  See tools/pin_go/README.md.
package main

import (
	"encoding/json"
	"math"
	"net/http/httptest"
	"os"
	"path/filepath"
	"strconv"

	"github.com/prometheus/client_golang/prometheus"
	"github.com/prometheus/client_golang/prometheus/promhttp"
)

type sample struct{ Pid, Pod, Used, Perc string } // Used/Perc: float64 bit patterns, hex
type gcase struct {
	Name    string
	Samples []sample
}

func f64(hex string) float64 {
	u, err := strconv.ParseUint(hex, 16, 64)
	if err != nil {
		panic(err)
	}
	return math.Float64frombits(u)
}

func main() {
	root := os.Args[1] // repository root
	raw, err := os.ReadFile(filepath.Join(root, "tests", "golden", "go_inputs.json"))
	if err != nil {
		panic(err)
	}
	var cases []gcase
	if err := json.Unmarshal(raw, &cases); err != nil {
		panic(err)
	}
	out := filepath.Join(root, "tests", "golden", "go")
	os.MkdirAll(out, 0o755)
	for _, c := range cases {
		used := prometheus.NewGaugeVec(prometheus.GaugeOpts{Name: "pod_gpu_memory_usage", Help: "GPU memory used by Kubernetes Pod"}, []string{"pid", "pod"})
		perc := prometheus.NewGaugeVec(prometheus.GaugeOpts{Name: "docker_gpu_memory_perc_usage", Help: "GPU memory in percentage used by pod"}, []string{"pid", "pod"})
		reg := prometheus.NewRegistry()
		reg.MustRegister(used)
		reg.MustRegister(perc)
		for _, s := range c.Samples {
			used.WithLabelValues(s.Pid, s.Pod).Set(f64(s.Used))
			perc.WithLabelValues(s.Pid, s.Pod).Set(f64(s.Perc))
		}
		rec := httptest.NewRecorder()
		promhttp.HandlerFor(reg, promhttp.HandlerOpts{}).ServeHTTP(rec, httptest.NewRequest("GET", "/metrics", nil))
		os.WriteFile(filepath.Join(out, c.Name+".txt"), rec.Body.Bytes(), 0o644)
		os.WriteFile(filepath.Join(out, c.Name+".hdr"), []byte(rec.Header().Get("Content-Type")+"\n"), 0o644)
	}
}
