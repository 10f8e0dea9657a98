// test_host_api.cpp -- the reference's own test cases re-expressed against the C++ drop-in headers
// (tests/test_force_calculation.cpp, test_integrator.cpp, test_barnes_hut.cpp, test_spatial_hash.cpp,
// test_validation.cpp).  Plain asserts (GoogleTest is not in this image).  Needs a B200; run by
// tests/test_gpu_cpp_host.py.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "nbody/force_calculator.hpp"
#include "nbody/integrator.hpp"
#include "nbody/particle_data.hpp"
#include "nbody/particle_system.hpp"
#include "nbody/renderer.hpp"
#include "nbody/serialization.hpp"

using namespace nbody;

#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) {                                                         \
      fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);    \
      exit(1);                                                             \
    }                                                                      \
  } while (0)

template <class F>
static bool throws_validation(F&& f) {
  try { f(); } catch (const ValidationException&) { return true; } catch (...) { return false; }
  return false;
}

static void two_body_unit_force() {  // tests/test_force_calculation.cpp:13-30 + :62-96 on the device
  Vec3 f = computeGravitationalForceCPU(Vec3(0, 0, 0), Vec3(1, 0, 0), 1, 1, 1, 0);
  CHECK(std::fabs(f.length() - 1.0f) < 1e-6f);
  ParticleData h, d;
  ParticleDataManager::allocateHost(h, 2);
  ParticleDataManager::allocateDevice(d, 2);
  h.pos_x[1] = 1.0f;
  h.mass[0] = h.mass[1] = 1.0f;
  ParticleDataManager::copyToDevice(d, h);
  DirectForceCalculator calc;
  calc.setSofteningParameter(0.0f);
  calc.computeForces(&d);
  ParticleDataManager::copyToHost(h, d);
  CHECK(std::fabs(h.acc_x[0] - 1.0f) < 1e-6f && std::fabs(h.acc_x[1] + 1.0f) < 1e-6f);
  ParticleDataManager::freeDevice(d);
  ParticleDataManager::freeHost(h);
}

static void integrator_cases() {  // tests/test_integrator.cpp:15-84, 90-162
  ParticleData h, d;
  ParticleDataManager::allocateHost(h, 2);
  ParticleDataManager::allocateDevice(d, 2);
  h.mass[0] = 1; h.mass[1] = 2;
  h.vel_x[0] = 1; h.vel_x[1] = 2;
  ParticleDataManager::copyToDevice(d, h);
  Integrator integ;
  CHECK(std::fabs(integ.computeKineticEnergy(&d) - 4.5f) < 1e-5f);
  integ.updatePositions(&d, 0.1f);
  ParticleDataManager::copyToHost(h, d);
  CHECK(std::fabs(h.pos_x[0] - 0.1f) < 1e-6f);
  // circular two-body orbit: drift < 1 % over 100 steps
  h.pos_x[0] = -2.5f; h.pos_x[1] = 2.5f; h.mass[1] = 1;
  const float v = std::sqrt(1.0f / 10.0f);
  h.vel_x[0] = h.vel_x[1] = 0; h.vel_y[0] = -v; h.vel_y[1] = v;
  ParticleDataManager::copyToDevice(d, h);
  DirectForceCalculator calc;
  calc.setSofteningParameter(0.01f);
  calc.computeForces(&d);
  const double e0 = integ.computeTotalEnergyF64(&d, 1.0f, 0.01f);
  for (int i = 0; i < 100; i++) integ.integrate(&d, &calc, 0.001f);
  const double e1 = integ.computeTotalEnergyF64(&d, 1.0f, 0.01f);
  CHECK(std::fabs((e1 - e0) / e0) < 0.01);
  // conserved-quantity diagnostics: the equal-mass binary keeps P = 0, L = m (x cross v) summed, M = 2
  const nb_diagnostics_t dg = integ.computeDiagnostics(&d, 1.0f, 0.01f);
  CHECK(std::fabs(dg.total_mass - 2.0) < 1e-12);
  CHECK(std::fabs(dg.momentum[0]) < 1e-6 && std::fabs(dg.momentum[1]) < 1e-6 && std::fabs(dg.momentum[2]) < 1e-12);
  CHECK(std::fabs(dg.angular_momentum[2] - 5.0 * v) < 1e-4);    // 2 bodies x m r v with r = 2.5
  CHECK(std::fabs(dg.total_energy - e1) < 1e-9 && dg.has_potential == 1);
  CHECK(std::fabs(dg.center_of_mass[0]) < 1e-6 && std::fabs(dg.center_of_mass[1]) < 1e-6);
  ParticleDataManager::freeDevice(d);
  ParticleDataManager::freeHost(h);
}

static void particle_system_cases() {
  ParticleSystem sys;
  SimulationConfig cfg;
  cfg.particle_count = 3000;
  sys.initialize(cfg);
  const double e0 = sys.computeTotalEnergyF64();
  for (int i = 0; i < 20; i++) sys.update(cfg.dt);
  CHECK(std::fabs((sys.computeTotalEnergyF64() - e0) / e0) < 1e-3);
  CHECK(std::fabs(sys.getSimulationTime() - 20 * cfg.dt) < 1e-6f);
  // pause / resume (tests/test_serialization.cpp:171-217)
  SimulationState a = sys.getState();
  sys.pause();
  sys.update(cfg.dt);
  SimulationState b = sys.getState();
  CHECK(a.pos_x == b.pos_x && a.vel_z == b.vel_z && a.simulation_time == b.simulation_time);
  sys.resume();
  // state round trip
  ParticleSystem other;
  other.setState(a);
  SimulationState c = other.getState();
  CHECK(c.pos_x == a.pos_x && c.mass == a.mass && c.particle_count == a.particle_count);
  // Barnes-Hut and spatial hash through the same facade
  sys.setForceMethod(ForceMethod::BARNES_HUT);
  sys.setBarnesHutTheta(0.5f);
  sys.update(cfg.dt);
  auto* bh = dynamic_cast<BarnesHutCalculator*>(sys.getForceCalculator());
  CHECK(bh && bh->getTree() && bh->getTree()->getNodeCount() == 2 * 3000 - 1);
  CHECK(std::fabs(bh->getTree()->getRootMass(3000) - 3000.0f) < 3.0f);  // tests/test_barnes_hut.cpp:38-62
  sys.setForceMethod(ForceMethod::SPATIAL_HASH);
  sys.update(cfg.dt);
  auto* sh = dynamic_cast<SpatialHashCalculator*>(sys.getForceCalculator());
  std::vector<int> cs, ce, pc, si;
  sh->getGrid()->copyCellDataToHost(3000, cs, ce, pc, si);
  long total = 0;
  for (size_t k = 0; k < cs.size(); k++) total += ce[k] - cs[k];
  CHECK(total == 3000);  // tests/test_spatial_hash.cpp:89-130
  SimulationState s = sys.getState();
  for (float x : s.pos_x) CHECK(std::isfinite(x));
}

static void validation_cases() {  // tests/test_validation.cpp
  SimulationConfig c;
  c.dt = 0.0f;            CHECK(throws_validation([&] { validateSimulationConfig(c); })); c.dt = 0.001f;
  c.particle_count = 0;   CHECK(throws_validation([&] { validateSimulationConfig(c); })); c.particle_count = 10;
  c.softening = -1.0f;    CHECK(throws_validation([&] { validateSimulationConfig(c); })); c.softening = 0.0f;
  c.cuda_block_size = 2048; CHECK(throws_validation([&] { validateSimulationConfig(c); })); c.cuda_block_size = 1024;
  validateSimulationConfig(c);
  ParticleSystem sys;
  CHECK(throws_validation([&] { sys.setGravitationalConstant(-1.0f); }));
  CHECK(throws_validation([&] { sys.setBarnesHutTheta(3.0f); }));
  CHECK(throws_validation([&] { sys.setTimeStep(2.0f); }));
  bool threw = false;
  try {
    SimulationConfig big;
    big.particle_count = 200;
    big.force_method = ForceMethod::SPATIAL_HASH;
    big.spatial_hash_cell_size = 0.001f;  // > 1e8 cells -> std::runtime_error (force_spatial_hash.cu:252-254)
    sys.initialize(big);
  } catch (const std::runtime_error&) { threw = true; }
  CHECK(threw);
}

static void graph_replay_matches_eager() {
  for (ForceMethod m : {ForceMethod::DIRECT_N2, ForceMethod::BARNES_HUT}) {
    SimulationConfig cfg;
    cfg.particle_count = 6000;
    cfg.force_method = m;
    cfg.init_distribution = InitDistribution::SPHERICAL;
    ParticleSystem eager, graphed;
    eager.initialize(cfg);
    graphed.initialize(cfg);
    graphed.setUseCudaGraph(true);
    for (int i = 0; i < 5; i++) {
      eager.update(cfg.dt);
      graphed.update(cfg.dt);
    }
    CHECK(graphed.graphNodeCount() >= 3);
    SimulationState a = eager.getState(), b = graphed.getState();
    for (std::size_t i = 0; i < a.particle_count; i++)
      CHECK(a.pos_x[i] == b.pos_x[i] && a.vel_y[i] == b.vel_y[i] && a.pos_z[i] == b.pos_z[i]);
  }
}

static void checkpoint_round_trip_through_the_system() {  // tests/test_serialization.cpp of the reference
  SimulationConfig cfg;
  cfg.particle_count = 777;
  ParticleSystem a;
  a.initialize(cfg);
  a.update(cfg.dt);
  const char* path = "/tmp/nbody_b200_test_checkpoint.nbody";
  a.saveState(path);
  CHECK(Serializer::validateFile(path));
  ParticleSystem b;
  b.loadState(path);
  CHECK(b.getState() == a.getState());
  CHECK(b.getParticleCount() == 777);
  std::remove(path);
}

static void standalone_tree_and_grid() {  // the call sites of tests/test_barnes_hut.cpp:14-62 and test_spatial_hash.cpp:15-130
  const std::size_t n = 500;
  ParticleData h, d;
  ParticleDataManager::allocateHost(h, n);
  ParticleDataManager::allocateDevice(d, n);
  SphericalDistParams sp; sp.radius = 10.0f;
  ParticleInitializer::initSpherical(h, sp);
  ParticleDataManager::copyToDevice(d, h);
  {  // positions-only copies (particle_data.hpp:29-35)
    std::vector<float> x(n), y(n), z(n);
    ParticleDataManager::copyPositionsToHost(x.data(), y.data(), z.data(), d);
    CHECK(x[7] == h.pos_x[7] && y[n - 1] == h.pos_y[n - 1] && z[0] == h.pos_z[0]);
    x[7] += 1.0f;
    ParticleDataManager::copyPositionsToDevice(d, x.data(), y.data(), z.data());
    std::vector<float> x2(n), y2(n), z2(n);
    ParticleDataManager::copyPositionsToHost(x2.data(), y2.data(), z2.data(), d);
    CHECK(x2[7] == h.pos_x[7] + 1.0f);
    ParticleDataManager::copyPositionsToDevice(d, h.pos_x, h.pos_y, h.pos_z);
  }
  BarnesHutTree tree(n);
  tree.build(&d);
  CHECK(tree.getNodeCount() > 0);
  tree.copyNodesToHost();
  CHECK(tree.verifyMassConservation(&h));
  CHECK(tree.verifyTreeStructure());
  CHECK(tree.getMaxDepth() > 0 && tree.getMaxDepth() <= 11);
  const OctreeNode* nodes = tree.getNodes();
  CHECK(!nodes[0].is_leaf && nodes[0].particle_count == static_cast<int>(n) && nodes[0].children[2] == -1);
  CHECK(std::fabs(nodes[0].total_mass - static_cast<float>(n)) < 0.5f);
  tree.computeForces(&d, 0.5f, 1.0f, 0.01f);
  CHECK(tree.getInteractionCount() > 0);

  CHECK(SpatialHashGrid::getCellIndex(0.5f, 0.5f, 0.5f, 1.0f).x == 0);   // tests/test_spatial_hash.cpp:38-51
  const int3 c = SpatialHashGrid::getCellIndex(2.5f, 3.5f, 4.5f, 1.0f);
  CHECK(c.x == 2 && c.y == 3 && c.z == 4);
  CHECK(SpatialHashGrid::hashCell(make_int3(-1, 0, 0), make_int3(4, 4, 4)) == 3);
  SpatialHashGrid grid(n, 2.0f);
  grid.build(&d);
  CHECK(grid.getTotalCells() > 0 && grid.getCellSize() == 2.0f);
  const int3 dims = grid.getGridDims();
  CHECK(dims.x * dims.y * dims.z == grid.getTotalCells());
  CHECK(grid.verifyCellAssignment(&h));
  std::vector<int> cs, ce, pc, si;
  grid.copyCellDataToHost(cs, ce, pc, si);   // :89-130: every particle in exactly one cell
  std::vector<int> seen(n, 0);
  for (int cell = 0; cell < grid.getTotalCells(); cell++)
    for (int k = cs[cell]; k < ce[cell]; k++) { CHECK(pc[si[k]] == cell); seen[si[k]]++; }
  bool once = true;
  for (std::size_t i = 0; i < n; i++) once = once && seen[i] == 1;
  CHECK(once);
  grid.computeForces(&d, 4.0f, 1.0f, 0.01f);
  ParticleDataManager::freeDevice(d);
  ParticleDataManager::freeHost(h);
}

static void headless_interop() {  // src/render/cuda_gl_interop.cpp + particle_system.cpp:320-345, without GL
  ParticleSystem sys;
  sys.initializeWithDistribution(1000, InitDistribution::DISK);
  CHECK(sys.getParticleCount() == 1000 && sys.getInterop() == nullptr);
  sys.initializeInterop();
  CudaGLInterop* io = sys.getInterop();
  CHECK(io != nullptr && io->getParticleCount() == 1000 && io->getPositionVBO() == 0);
  SimulationState st = sys.getState();
  std::vector<float> xyz(3000);
  for (int i = 0; i < 1000; i++) { xyz[3 * i] = st.pos_x[i]; xyz[3 * i + 1] = st.pos_y[i]; xyz[3 * i + 2] = st.pos_z[i]; }
  CHECK(io->verifyDataIntegrity(xyz.data(), xyz.size()));
  sys.update(0.001f);  // update() refreshes the interop buffers (:123-126)
  CHECK(!io->verifyDataIntegrity(xyz.data(), xyz.size()));
  st = sys.getState();
  for (int i = 0; i < 1000; i++) { xyz[3 * i] = st.pos_x[i]; xyz[3 * i + 1] = st.pos_y[i]; xyz[3 * i + 2] = st.pos_z[i]; }
  CHECK(io->verifyDataIntegrity(xyz.data(), xyz.size()));
  CHECK(io->mapPositionBuffer() != nullptr && io->isMapped());
  io->unmapAllBuffers();
  CHECK(!io->isMapped());
  // the interop follows a re-initialisation to another size (:96-99)
  sys.initializeWithDistribution(500, InitDistribution::UNIFORM);
  CHECK(sys.getInterop() != nullptr && sys.getInterop()->getParticleCount() == 500);
  // free-standing copy kernels into a caller's buffer
  void* buf = nullptr;
  CHECK(nb_device_alloc(Engine::get(), 500 * 3 * sizeof(float), &buf) == 0);
  launchCopyVelocitiesToVBOKernel(static_cast<float*>(buf), sys.getDeviceData(), 256);
  std::vector<float> v(1500, 1.0f);
  CHECK(nb_device_download(Engine::get(), v.data(), buf, v.size() * sizeof(float), nullptr) == 0);
  st = sys.getState();
  bool same = true;
  for (int i = 0; i < 500; i++) same = same && v[3 * i] == st.vel_x[i] && v[3 * i + 1] == st.vel_y[i] && v[3 * i + 2] == st.vel_z[i];
  CHECK(same);
  nb_device_free(Engine::get(), buf);
}

static void headless_renderer() {
  const int w = 64, h = 36;
  Renderer r;
  r.initialize(w, h);
  const Mat4 pr = r.getCamera().getProjectionMatrix();
  CHECK(std::fabs(pr.at(1, 1) - 2.4142135f) < 1e-5f && pr.at(2, 3) == -1.0f);
  // one body on the centre of pixel (w/2, h/2) at depth 50: t = 0.5 -> grey, alpha 1
  ParticleData hp, dp;
  ParticleDataManager::allocateHost(hp, 1);
  ParticleDataManager::allocateDevice(dp, 1);
  hp.pos_x[0] = 0.5f * 50.0f / (w / 2 * pr.at(0, 0));
  hp.pos_y[0] = 0.5f * 50.0f / (h / 2 * pr.at(1, 1));
  hp.mass[0] = 1.0f;
  ParticleDataManager::copyToDevice(dp, hp);
  r.render(&dp);
  const auto px = r.readPixels();
  const std::uint8_t* c = &px[((h / 2) * w + w / 2) * 4];
  CHECK(c[0] == 128 && c[1] == 128 && c[2] == 128 && c[3] == 255);
  CHECK(px[0] == 5 && px[1] == 5 && px[2] == 13 && px[3] == 255);  // clear colour
  const Vec3 col = ColorMapper::velocityToColor(5.0f, 10.0f);
  CHECK(col.x == 0.5f && col.y == 0.0f && col.z == 0.5f);
  ParticleDataManager::freeDevice(dp);
  ParticleDataManager::freeHost(hp);
}

int main() {
  two_body_unit_force();
  integrator_cases();
  particle_system_cases();
  validation_cases();
  graph_replay_matches_eager();
  standalone_tree_and_grid();
  headless_interop();
  headless_renderer();
  checkpoint_round_trip_through_the_system();
  puts("cpp host api: all checks passed");
  return 0;
}
