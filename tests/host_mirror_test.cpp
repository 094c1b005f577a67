// Drives the C++ host mirror (kubernetes-gpu-exporter_b200/host/exporter.hpp)
// the way main.go drives its GaugeVecs and handler, and checks the served body
// against the hand-derived golden of SURVEY.md section 8c.  Needs a GPU.
#include <chrono>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

#include "../kubernetes-gpu-exporter_b200/host/exporter.hpp"

using namespace kgx_host;

#define CHECK(c) do { if (!(c)) { std::fprintf(stderr, "CHECK failed: %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

// Collect over a million series (the size of BASELINE config 3): values and label text come back through the bulk
// read-back calls, not one call per series.  Prints the time of the library calls and of the whole Collect.
static int big_collect() {
  const uint64_t GiB = 1ull << 30;
  const uint32_t n_pods = 131072, per_pod = 8, n_dev = 8;
  PodPIDMap m;
  std::vector<Device> devs(n_dev);
  uint32_t pid = 300;
  for (uint32_t p = 0; p < n_pods; ++p) {
    std::vector<std::string>& v = m["pod-" + std::to_string(p)];
    for (uint32_t k = 0; k < per_pod; ++k, ++pid) {
      v.push_back(std::to_string(pid));
      devs[pid % n_dev].processInfos.push_back(ProcessInfo{pid, (uint64_t(pid) % 4096 + 1) * (2ull << 20)});
    }
  }
  for (Device& d : devs) d.memoryInfo.Total = 192 * GiB;
  Exporter ex(0, false);
  ex.Cycle(m, devs);
  std::vector<Metric> ms;
  ex.Collect(ms);  // warm-up (buffers)
  ms.clear();
  const auto t0 = std::chrono::steady_clock::now();
  size_t n = 0, nb = 0;
  kgx_read_labels(ex.ctx(), nullptr, 0, nullptr, 0, &n, &nb);
  std::string blob(nb, '\0');
  std::vector<uint64_t> off(n + 1);
  CHECK(kgx_read_labels(ex.ctx(), &blob[0], blob.size(), off.data(), off.size(), &n, &nb) == KGX_OK);
  std::vector<double> v0(n), v1(n);
  size_t n0 = 0, n1 = 0;
  CHECK(kgx_read_values(ex.ctx(), 0, v0.data(), nullptr, n, &n0) == KGX_OK && kgx_read_values(ex.ctx(), 1, v1.data(), nullptr, n, &n1) == KGX_OK);
  const auto t1 = std::chrono::steady_clock::now();
  ex.Collect(ms);
  const auto t2 = std::chrono::steady_clock::now();
  const double lib_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
  const double all_ms = std::chrono::duration<double, std::milli>(t2 - t1).count();
  std::printf("collect of %zu series: read-back through the C ABI %.1f ms, whole Collect (with label parsing) %.1f ms\n", n, lib_ms, all_ms);
  CHECK(n == size_t(n_pods) * per_pod && n0 == n && n1 == n && ms.size() == 2 * n);
  CHECK(lib_ms < 100.0);
  // series order is by (pid as a string, pod): the first one is pid "1000000"'s neighbourhood; check a few by lookup
  size_t seen = 0;
  for (size_t i = 0; i < n; i += 50021) {
    const Metric& a = ms[i];          // family 0 (percentage)
    const Metric& b = ms[n + i];      // family 1 (bytes), same series
    CHECK(a.labelValues == b.labelValues && a.labelValues.size() == 2);
    const uint32_t p = uint32_t(std::stoul(a.labelValues[0]));
    const uint64_t used = (uint64_t(p) % 4096 + 1) * (2ull << 20);
    CHECK(a.labelValues[1] == "pod-" + std::to_string((p - 300) / per_pod));
    CHECK(b.value == double(used) && a.value == double(used) / double(192 * GiB) * 100);
    CHECK(b.value == v1[i] && a.value == v0[i]);
    ++seen;
  }
  CHECK(seen > 10);
  std::puts("ok");
  return 0;
}

// One node at BASELINE config 1's size (64 pods, 256 processes) through the C ABI without an interpreter in the way:
// what one poll + scrape costs when the step is launch-bound.  Prints microseconds per call pair.
static int small_latency() {
  const uint64_t GiB = 1ull << 30;
  Exporter ex(0, false);
  kgx_ctx* c = ex.ctx();
  PodPIDMap m;
  std::vector<Device> devs(8);
  uint32_t pid = 1000;
  for (uint32_t p = 0; p < 64; ++p)
    for (uint32_t k = 0; k < 4; ++k, ++pid) {
      m["pod-" + std::to_string(p)].push_back(std::to_string(pid));
      devs[pid % 8].processInfos.push_back(ProcessInfo{pid, (uint64_t(pid) % 4096 + 1) * (2ull << 20)});
    }
  for (Device& d : devs) d.memoryInfo.Total = 192 * GiB;
  ex.Cycle(m, devs);  // installs the table, first ingest
  std::vector<uint32_t> pids, gpus;
  std::vector<uint64_t> used, total(8, 192 * GiB);
  for (size_t di = 0; di < devs.size(); ++di)
    for (const ProcessInfo& p : devs[di].processInfos) { pids.push_back(p.Pid); gpus.push_back(uint32_t(di)); used.push_back(p.UsedGpuMemory); }
  kgx_cols cols{};
  cols.n = pids.size(); cols.pid = pids.data(); cols.gpu = gpus.data(); cols.used = used.data(); cols.gpu_total = total.data();
  std::vector<uint8_t> body(1 << 20);
  size_t len = 0;
  for (int i = 0; i < 200; ++i) { CHECK(kgx_ingest(c, &cols) == KGX_OK); CHECK(kgx_render(c, body.data(), body.size(), &len) == KGX_OK); }
  const int reps = 2000;
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < reps; ++i) { kgx_ingest(c, &cols); kgx_render(c, body.data(), body.size(), &len); }
  const auto t1 = std::chrono::steady_clock::now();
  for (int i = 0; i < reps; ++i) kgx_render(c, body.data(), body.size(), &len);
  const auto t2 = std::chrono::steady_clock::now();
  const double pair_us = std::chrono::duration<double, std::micro>(t1 - t0).count() / reps;
  const double render_us = std::chrono::duration<double, std::micro>(t2 - t1).count() / reps;
  std::printf("256 samples, %zu bytes of text: kgx_ingest + kgx_render %.1f us per cycle, kgx_render alone %.1f us (host buffers, C++ caller)\n",
              len, pair_us, render_us);
  CHECK(len > 30000 && pair_us < 2000.0);
  std::puts("ok");
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  if (std::string(argv[1]) == "big") return big_collect();
  if (std::string(argv[1]) == "small") return small_latency();
  std::ifstream f(argv[1], std::ios::binary);
  std::stringstream golden;
  golden << f.rdbuf();
  const uint64_t GiB = 1ull << 30;

  Exporter ex(0, /*literal=*/false);
  std::vector<const Desc*> descs;
  ex.Describe(descs);
  CHECK(descs.size() == 2);
  CHECK(descs[0]->fqName == "docker_gpu_memory_perc_usage" && descs[1]->fqName == "pod_gpu_memory_usage");  // Gather order
  CHECK(descs[1]->help == "GPU memory used by Kubernetes Pod" && descs[1]->variableLabels == (std::vector<std::string>{"pid", "pod"}));

  CHECK(ex.ServeHTTP().body.empty());  // nothing Set yet: no family is gathered
  PodPIDMap m{{"trainer-0", {"1", "4242", "77"}}, {"idle", {"5", "6", "ps:"}}};
  std::vector<Device> devs{{{80 * GiB}, {{4242, 1 * GiB}}}, {{80 * GiB}, {{999, 5 * GiB}}}};
  ex.Cycle(m, devs);
  Response r = ex.ServeHTTP();
  CHECK(r.status == 200 && r.contentType.rfind("text/plain; version=0.0.4", 0) == 0);
  CHECK(r.body == golden.str());

  std::vector<Metric> ms;
  ex.Collect(ms);
  CHECK(ms.size() == 2);
  CHECK(ms[0].desc->fqName == "docker_gpu_memory_perc_usage" && ms[0].value == 1.25);
  CHECK(ms[1].value == 1073741824.0 && ms[1].labelValues == (std::vector<std::string>{"4242", "trainer-0"}));

  // next cycle: the pod is gone from the map, its child keeps its last value (no Delete/Reset in main.go);
  // the same PID on two devices: the later device wins (main.go:123 loop order)
  PodPIDMap m2{{"web \"a\"", {"7"}}};
  std::vector<Device> devs2{{{80 * GiB}, {{7, 100}}}, {{40 * GiB}, {{7, 200}}}};
  ex.Cycle(m2, devs2);
  r = ex.ServeHTTP();
  CHECK(r.body.find("pod_gpu_memory_usage{pid=\"4242\",pod=\"trainer-0\"} 1.073741824e+09\n") != std::string::npos);
  CHECK(r.body.find("pod_gpu_memory_usage{pid=\"7\",pod=\"web \\\"a\\\"\"} 200\n") != std::string::npos);
  {  // the same scrape served with chunked transfer encoding, 7 bytes at a time: the same bytes
    std::string chunked;
    int n_chunks = 0;
    CHECK(ex.ServeHTTPChunked(7, [&](const char* p, size_t n) { chunked.append(p, n); ++n_chunks; }) == 200);
    CHECK(chunked == r.body && n_chunks == int((r.body.size() + 6) / 7));
  }

  // the predicate exactly as written at main.go:144-145
  Exporter lit(0, /*literal=*/true);
  lit.Cycle(PodPIDMap{{"p", {"1000", "2000", "3000"}}}, {{{80 * GiB}, {{0, 10}, {2, 30}, {5, 60}}}});
  r = lit.ServeHTTP();
  CHECK(r.body.find("pod_gpu_memory_usage{pid=\"0\",pod=\"p\"} 10\n") != std::string::npos);
  CHECK(r.body.find("pod_gpu_memory_usage{pid=\"2\",pod=\"p\"} 30\n") != std::string::npos);
  CHECK(r.body.find("pid=\"5\"") == std::string::npos);
  std::puts("ok");
  return 0;
}
