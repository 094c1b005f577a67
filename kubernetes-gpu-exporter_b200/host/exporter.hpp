// exporter.hpp — C++ host side above the C ABI (include/kgx.h), mirroring the
// reference's interface for the hot path so it can be dropped behind the same
// call sites (the Go shim in go/kgx does the same in Go; Go is not installed on
// the build box, C++ is).  Names follow /root/reference/main.go:
//
//   PodPIDMap            podPIDMap := make(map[string][]string)      main.go:84
//   ProcessInfo, Memory  the nvml fields the loop reads               main.go:129,135,145-149
//   Exporter::Cycle      the loop body after discovery and sampling   main.go:141-154
//   Describe / Collect   prometheus.Collector, as the two GaugeVecs
//                        registered at main.go:41-42 satisfy it
//   ServeHTTP            promhttp.HandlerFor(reg, HandlerOpts{})      main.go:69-70
//
// Error behaviour: the reference log.Fatalf's / panics (main.go:47,59,79,119);
// here every failure is a std::runtime_error carrying kgx_last_error.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/kgx.h"

namespace kgx_host {

using PodPIDMap = std::map<std::string, std::vector<std::string>>;  // pod name -> PIDs as `ps` printed them

struct ProcessInfo { uint32_t Pid; uint64_t UsedGpuMemory; };  // nvml.ProcessInfo (main.go:145,147)
struct Memory { uint64_t Total; };                            // nvml.Memory (main.go:149)
struct Device { Memory memoryInfo; std::vector<ProcessInfo> processInfos; };  // one iteration of main.go:123

struct Desc { std::string fqName, help; std::vector<std::string> variableLabels; };
struct Metric { const Desc* desc; double value; std::vector<std::string> labelValues; };
struct Response { int status; std::string contentType, body; };

class Exporter {
 public:
  // literal=true keeps main.go:144-145 exactly as written (`for pid := range pids` yields indices)
  explicit Exporter(int device = 0, bool literal = false) : literal_(literal) {
    kgx_opts o{};
    o.abi_version = KGX_ABI_VERSION;
    o.mode = literal ? KGX_MODE_R_LITERAL : KGX_MODE_R_INTENDED;
    int rc = kgx_create(device, &o, &ctx_);
    if (rc != KGX_OK) throw std::runtime_error(std::string("kgx_create: ") + kgx_strerror(rc) + " (no CPU fallback)");
    for (int f = 0; f < kgx_family_count(ctx_); ++f) {
      const char *name, *help;
      const char* const* names;
      int nl;
      kgx_family_info(ctx_, f, &name, &help, &names, &nl);
      Desc d{name, help, {}};
      for (int i = 0; i < nl; ++i) d.variableLabels.push_back(names[i]);
      descs_.push_back(d);
    }
  }
  ~Exporter() { kgx_destroy(ctx_); }
  Exporter(const Exporter&) = delete;
  Exporter& operator=(const Exporter&) = delete;

  // One poll cycle on one node: hand over the map (main.go:113) and the samples in the
  // reference's loop order — device index ascending, then process order (main.go:123,141).
  void Cycle(const PodPIDMap& podPIDMap, const std::vector<Device>& devices) {
    std::string pod_blob;
    std::vector<uint64_t> pod_off{0};
    std::vector<uint32_t> podkey_node, entry_podkey, entry_pid;
    uint32_t pk = 0;
    for (const auto& kv : podPIDMap) {
      pod_blob += kv.first;
      pod_off.push_back(pod_blob.size());
      podkey_node.push_back(0);
      for (const std::string& s : kv.second) {
        char* end = nullptr;
        unsigned long v = std::strtoul(s.c_str(), &end, 10);
        const bool numeric = end && *end == 0 && !s.empty();
        // a field that is not a number can never equal a Pid; in literal mode only the list LENGTH matters
        if (!numeric && !literal_) continue;
        entry_pid.push_back(numeric ? uint32_t(v) : 0u);
        entry_podkey.push_back(pk);
      }
      ++pk;
    }
    pod_blob.push_back('\0');
    kgx_table t{};
    t.n_nodes = 1;
    t.gpus_per_node = uint32_t(devices.empty() ? 1 : devices.size());
    t.n_podkeys = pk;
    t.podkey_node = podkey_node.data();
    t.pod_blob = pod_blob.data();
    t.pod_off = pod_off.data();
    t.n_entries = entry_pid.size();
    t.entry_podkey = entry_podkey.data();
    t.entry_pid = entry_pid.data();
    check(kgx_set_table(ctx_, &t));
    std::vector<uint32_t> pid, gpu;
    std::vector<uint64_t> used, total(t.gpus_per_node, 0);
    for (size_t di = 0; di < devices.size(); ++di) {
      total[di] = devices[di].memoryInfo.Total;
      for (const ProcessInfo& p : devices[di].processInfos) { pid.push_back(p.Pid); gpu.push_back(uint32_t(di)); used.push_back(p.UsedGpuMemory); }
    }
    kgx_cols c{};
    c.n = pid.size();
    c.pid = pid.data(); c.gpu = gpu.data(); c.used = used.data(); c.gpu_total = total.data();
    check(kgx_ingest(ctx_, &c));
  }

  // prometheus.Collector
  void Describe(std::vector<const Desc*>& ch) const { for (const Desc& d : descs_) ch.push_back(&d); }
  void Collect(std::vector<Metric>& ch) {
    // label text of every series in one call (two device->host copies), parsed once and shared by all families:
    // entry i of kgx_read_labels belongs to value i of every family
    size_t n_lbl = 0, blob_len = 0;
    check(kgx_read_labels(ctx_, nullptr, 0, nullptr, 0, &n_lbl, &blob_len));
    std::string blob(blob_len, '\0');
    std::vector<uint64_t> off(n_lbl + 1);
    check(kgx_read_labels(ctx_, blob.empty() ? nullptr : &blob[0], blob.size(), off.data(), off.size(), &n_lbl, &blob_len));
    std::vector<std::vector<std::string>> labels(n_lbl);
    for (size_t i = 0; i < n_lbl; ++i) labels[i] = label_values(blob.data() + off[i], size_t(off[i + 1] - off[i]));
    for (size_t f = 0; f < descs_.size(); ++f) {
      size_t n = 0;
      kgx_read_values(ctx_, int(f), nullptr, nullptr, 0, &n);
      std::vector<double> v(n);
      if (n) check(kgx_read_values(ctx_, int(f), v.data(), nullptr, n, &n));
      if (n != n_lbl) throw std::runtime_error("Collect: values and labels disagree");
      for (size_t i = 0; i < n; ++i) ch.push_back(Metric{&descs_[f], v[i], labels[i]});
    }
  }

  // GET /metrics with the default text format
  Response ServeHTTP() {
    size_t len = 0;
    int rc = kgx_render(ctx_, reinterpret_cast<uint8_t*>(&buf_[0]), buf_.size(), &len);
    if (rc == KGX_E_NOSPACE) { buf_.resize(len + len / 8 + 64); rc = kgx_render(ctx_, reinterpret_cast<uint8_t*>(&buf_[0]), buf_.size(), &len); }
    if (rc != KGX_OK) return Response{500, "text/plain; charset=utf-8", kgx_last_error(ctx_)};  // HandlerOpts{}: errors -> HTTP 500
    return Response{200, "text/plain; version=0.0.4; charset=utf-8; escaping=values", buf_.substr(0, len)};
  }

  // GET /metrics with chunked transfer encoding: the body stays on the device and is handed to `write` one chunk at a
  // time (Go side: StreamingHandler in go/kgx/collector.go).  Returns the status code; 500 only before the first chunk.
  template <class Write>
  int ServeHTTPChunked(size_t chunk_bytes, Write&& write) {
    size_t len = 0;
    if (kgx_render_device(ctx_, &len) != KGX_OK) return 500;
    if (chunk_.size() < chunk_bytes) chunk_.resize(chunk_bytes);
    for (uint64_t at = 0;;) {
      size_t n = 0;
      if (kgx_body_read(ctx_, at, reinterpret_cast<uint8_t*>(&chunk_[0]), chunk_bytes, &n) != KGX_OK) return at ? 200 : 500;
      if (n == 0) return 200;
      write(chunk_.data(), n);
      at += n;
    }
  }

  kgx_ctx* ctx() { return ctx_; }

 private:
  void check(int rc) { if (rc != KGX_OK) throw std::runtime_error(std::string(kgx_strerror(rc)) + ": " + kgx_last_error(ctx_)); }

  // `{a="x",b="y"}` (escaped as expfmt writes it) back to raw label values
  static std::vector<std::string> label_values(const char* p, size_t len) {
    const std::string s(p, len);
    std::vector<std::string> out;
    size_t i = 0;
    while ((i = s.find("=\"", i)) != std::string::npos) {
      i += 2;
      std::string v;
      for (; i < s.size() && s[i] != '"'; ++i) {
        if (s[i] == '\\' && i + 1 < s.size()) { ++i; v += s[i] == 'n' ? '\n' : s[i]; }
        else v += s[i];
      }
      out.push_back(v);
    }
    return out;
  }

  kgx_ctx* ctx_ = nullptr;
  bool literal_ = false;
  std::vector<Desc> descs_;
  std::string buf_ = std::string(1 << 16, '\0');
  std::string chunk_;
};

}  // namespace kgx_host
