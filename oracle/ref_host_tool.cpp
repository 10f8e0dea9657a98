// ref_host_tool.cpp -- TEST INFRASTRUCTURE.  A tiny driver linked against the reference's UNMODIFIED
// host sources (src/utils/{serialization,simulation_state,error_handling,performance_observability}.cpp,
// compiled where they lie under /root/reference by oracle/build_ref_host.sh) so the tests can
//   write  <file>  : have the reference's Serializer write a fixed checkpoint
//   read   <file>  : have the reference's Serializer load a checkpoint and print it as JSON
//   record         : print the reference's serializeBenchmarkRunRecords() of a fixed record
// The fixed inputs are restated in tests/test_checkpoint_format.py.
#include <cstdio>
#include <cstring>
#include <iostream>

#include "nbody/performance_observability.hpp"
#include "nbody/serialization.hpp"

using namespace nbody;

static SimulationState fixed_state() {
  SimulationState s;
  s.particle_count = 5;
  s.simulation_time = 1.25f;
  s.dt = 0.001f;
  s.G = 1.0f;
  s.softening = 0.01f;
  s.force_method = ForceMethod::BARNES_HUT;
  for (int i = 0; i < 5; i++) {
    s.pos_x.push_back(i + 0.5f);
    s.pos_y.push_back(-2.0f * i);
    s.pos_z.push_back(0.25f * i * i);
    s.vel_x.push_back(0.1f * i);
    s.vel_y.push_back(-0.2f * i);
    s.vel_z.push_back(0.3f);
    s.mass.push_back(1.0f + i);
  }
  return s;
}

static void print_array(const char* name, const std::vector<float>& v, bool last = false) {
  printf("\"%s\":[", name);
  for (size_t i = 0; i < v.size(); i++) printf("%s%.9g", i ? "," : "", v[i]);
  printf("]%s", last ? "" : ",");
}

int main(int argc, char** argv) {
  try {
    if (argc >= 3 && !strcmp(argv[1], "write")) {
      Serializer::save(argv[2], fixed_state());
      return 0;
    }
    if (argc >= 3 && !strcmp(argv[1], "read")) {
      if (!Serializer::validateFile(argv[2])) {
        fprintf(stderr, "validateFile: false\n");
        return 2;
      }
      SimulationState s = Serializer::load(argv[2]);
      printf("{\"particle_count\":%zu,\"simulation_time\":%.9g,\"dt\":%.9g,\"G\":%.9g,\"softening\":%.9g,\"force_method\":%d,",
             s.particle_count, s.simulation_time, s.dt, s.G, s.softening, (int)s.force_method);
      print_array("pos_x", s.pos_x); print_array("pos_y", s.pos_y); print_array("pos_z", s.pos_z);
      print_array("vel_x", s.vel_x); print_array("vel_y", s.vel_y); print_array("vel_z", s.vel_z);
      print_array("mass", s.mass, true);
      printf("}\n");
      return 0;
    }
    if (argc >= 2 && !strcmp(argv[1], "record")) {
      BenchmarkRunRecord r;
      r.benchmark_name = "force.direct_n2 \"quoted\"\\path\tTab";
      r.force_method = ForceMethod::SPATIAL_HASH;
      r.particle_count = 4096;
      r.iterations = 7;
      r.metrics["wall_time_ms"] = 12.3456789;
      r.metrics["interactions_per_second"] = 2.427e12;
      r.parameters["theta"] = 0.5;
      r.parameters["dt"] = 0.001;
      PhaseTiming p;
      p.name = "serialization.save";
      p.total_duration = Milliseconds(0.07);
      p.samples = 3;
      r.phase_timings.push_back(p);
      BenchmarkRunRecord r2;
      r2.benchmark_name = "integration.direct_n2";
      std::cout << serializeBenchmarkRunRecords({r, r2}) << "\n";
      return 0;
    }
  } catch (const std::exception& e) {
    fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
  fprintf(stderr, "usage: ref_host_tool write|read <file> | record\n");
  return 64;
}
