// Drives the C++ host mirror (kubernetes-gpu-exporter_b200/host/exporter.hpp)
// the way main.go drives its GaugeVecs and handler, and checks the served body
// against the hand-derived golden of SURVEY.md section 8c.  Needs a GPU.
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

#include "../kubernetes-gpu-exporter_b200/host/exporter.hpp"

using namespace kgx_host;

#define CHECK(c) do { if (!(c)) { std::fprintf(stderr, "CHECK failed: %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

int main(int argc, char** argv) {
  if (argc < 2) return 2;
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
