// benchmark_main.cpp -- C++20 host program over the drop-in API: the reference's
// benchmarks/benchmark_main.cpp call pattern (ParticleSystem::initialize, a second calculator from the
// bare factory, a timed loop of computeForces / update), but with the device synchronised around the
// timed region (the reference times asynchronous launches).  Prints one JSON object
// in the reference's BenchmarkRunRecord shape (src/utils/performance_observability.cpp:102-144).
//
//   benchmark_main [--particles N] [--iterations K] [--benchmark force.direct_n2|force.barnes_hut|
//                   force.spatial_hash|integration.direct_n2|all]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "nbody/particle_system.hpp"

using namespace nbody;
using Clock = std::chrono::steady_clock;

struct Record {
  std::string name, method;
  size_t n;
  int iters;
  double ms;
  double interactions_per_s;
};

static double elapsed_ms(Clock::time_point a, Clock::time_point b) {
  return std::chrono::duration<double, std::milli>(b - a).count();
}

static Record run_force(const char* name, ForceMethod m, const char* mname, size_t n, int iters) {
  ParticleSystem sys;
  SimulationConfig cfg;
  cfg.particle_count = n;
  cfg.init_distribution = InitDistribution::SPHERICAL;
  cfg.force_method = m;
  sys.initialize(cfg);
  auto calc = createForceCalculator(m, cfg);  // bare factory, as benchmarks/benchmark_main.cpp:120
  calc->computeForces(sys.getDeviceData());   // warm-up
  sys.synchronize();
  auto t0 = Clock::now();
  for (int i = 0; i < iters; i++) calc->computeForces(sys.getDeviceData());
  sys.synchronize();
  double ms = elapsed_ms(t0, Clock::now()) / iters;
  double ips = (m == ForceMethod::DIRECT_N2) ? double(n) * double(n) / (ms * 1e-3) : 0.0;
  return {name, mname, n, iters, ms, ips};
}

static Record run_integration(size_t n, int iters) {
  ParticleSystem sys;
  SimulationConfig cfg;
  cfg.particle_count = n;
  cfg.init_distribution = InitDistribution::SPHERICAL;
  sys.initialize(cfg);
  sys.update(cfg.dt);
  sys.synchronize();
  auto t0 = Clock::now();
  for (int i = 0; i < iters; i++) sys.update(cfg.dt);
  sys.synchronize();
  double ms = elapsed_ms(t0, Clock::now()) / iters;
  return {"integration.direct_n2", "direct_n2", n, iters, ms, double(n) * double(n) / (ms * 1e-3)};
}

int main(int argc, char** argv) {
  size_t n = 4096;  // benchmarks/benchmark_main.cpp:18-23 defaults
  int iters = 5;
  std::string which = "all";
  if (const char* e = getenv("NBODY_BENCHMARK_PARTICLES")) n = strtoull(e, nullptr, 10);
  if (const char* e = getenv("NBODY_BENCHMARK_ITERATIONS")) iters = atoi(e);
  for (int i = 1; i < argc; i++) {
    if (!strcmp(argv[i], "--particles") && i + 1 < argc) n = strtoull(argv[++i], nullptr, 10);
    else if (!strcmp(argv[i], "--iterations") && i + 1 < argc) iters = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--benchmark") && i + 1 < argc) which = argv[++i];
    else if (!strcmp(argv[i], "--list")) {
      puts("force.direct_n2\nforce.barnes_hut\nforce.spatial_hash\nintegration.direct_n2");
      return 0;
    }
  }
  try {
    std::vector<Record> recs;
    auto want = [&](const char* b) { return which == "all" || which == b; };
    if (want("force.direct_n2")) recs.push_back(run_force("force.direct_n2", ForceMethod::DIRECT_N2, "direct_n2", n, iters));
    if (want("force.barnes_hut")) recs.push_back(run_force("force.barnes_hut", ForceMethod::BARNES_HUT, "barnes_hut", n, iters));
    if (want("force.spatial_hash"))
      recs.push_back(run_force("force.spatial_hash", ForceMethod::SPATIAL_HASH, "spatial_hash", n, iters));
    if (want("integration.direct_n2")) recs.push_back(run_integration(n, iters));
    printf("{\"benchmarks\":[");
    for (size_t i = 0; i < recs.size(); i++) {
      const Record& r = recs[i];
      printf("%s{\"benchmark_name\":\"%s\",\"force_method\":\"%s\",\"particle_count\":%zu,\"iterations\":%d,"
             "\"metrics\":{\"wall_time_ms\":%.6f,\"interactions_per_second\":%.6e},\"parameters\":{},\"phase_timings\":[]}",
             i ? "," : "", r.name.c_str(), r.method.c_str(), r.n, r.iters, r.ms, r.interactions_per_s);
    }
    printf("]}\n");
  } catch (const std::exception& e) {
    fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
  return 0;
}
