// integration/b200_backend.cpp -- the reference-side binding of INTEGRATION.md section A, as a real
// translation unit: compiled against the REFERENCE's unmodified headers, it replaces
// src/cuda/{force_direct,force_barnes_hut,force_spatial_hash,integrator}.cu.
// integration/build_dropin.sh links it with the reference's own particle_system.cpp and
// benchmarks/benchmark_main.cpp, so the reference's benchmark program runs on these kernels.
#include "nbody/force_calculator.hpp"   // reference headers, unchanged
#include "nbody/integrator.hpp"
#include "nbody/error_handling.hpp"
#include "nbody/barnes_hut_tree.hpp"
#include "nbody/spatial_hash_grid.hpp"
#include "nbody_b200.h"                  // this repo
#include <stdexcept>

namespace nbody {
static nb_context* ctx() {               // one context per process; the reference is single-device
  static nb_context* c = [] { nb_context* p = nullptr;
    if (nb_context_create(0, 0, &p)) throw CudaException(nb_last_error(), __FILE__, __LINE__);
    return p; }();
  return c;
}
static void check(int rc, const char* f, int l) {           // return code -> reference exception
  if (rc == NB_OK) return;
  if (rc == NB_ERR_INVALID) throw ValidationException(nb_last_error());
  if (rc == NB_ERR_RESOURCE) throw ResourceException(nb_last_error(), 0, 0);
  if (rc == NB_ERR_GRID_TOO_LARGE) throw std::runtime_error(nb_last_error());
  throw CudaException(nb_last_error(), f, l);
}
#define B200(x) check((x), __FILE__, __LINE__)
// ParticleData (types.hpp:234-276) and nb_particles have the same layout: pass by cast.
static nb_particles* P(ParticleData* d) { return reinterpret_cast<nb_particles*>(d); }
static const nb_particles* P(const ParticleData* d) { return reinterpret_cast<const nb_particles*>(d); }

// force_direct.cu:100-106
DirectForceCalculator::DirectForceCalculator(int block_size) : block_size_(block_size) {}
void DirectForceCalculator::computeForces(ParticleData* d) {
  B200(nb_direct_forces(ctx(), P(d), G_, softening_eps2_, nullptr));
}
// force_barnes_hut.cu:521-532 (tree_ stays null: the tree lives in the context; see getTree() note)
BarnesHutCalculator::BarnesHutCalculator(float theta) : theta_(theta) {}
BarnesHutCalculator::~BarnesHutCalculator() = default;
void BarnesHutCalculator::computeForces(ParticleData* d) {
  B200(nb_bh_build(ctx(), P(d), nullptr));
  B200(nb_bh_forces(ctx(), P(d), theta_, G_, softening_eps_, nullptr));
}
// force_spatial_hash.cu:363-377
SpatialHashCalculator::SpatialHashCalculator(float c, float r) : cell_size_(c), cutoff_radius_(r) {}
SpatialHashCalculator::~SpatialHashCalculator() = default;
void SpatialHashCalculator::computeForces(ParticleData* d) {
  B200(nb_sh_build(ctx(), P(d), cell_size_, nullptr));
  B200(nb_sh_forces(ctx(), P(d), cutoff_radius_, G_, softening_eps_, nullptr));
}
// force_spatial_hash.cu:380-401: createForceCalculator is pure host code; restated because its
// translation unit is one of the four removed.
std::unique_ptr<ForceCalculator> createForceCalculator(ForceMethod method, const SimulationConfig& cfg) {
  if (method == ForceMethod::BARNES_HUT) {
    auto c = std::make_unique<BarnesHutCalculator>(cfg.barnes_hut_theta);
    c->setGravitationalConstant(cfg.G);
    c->setSofteningParameter(cfg.softening);
    return c;
  }
  if (method == ForceMethod::SPATIAL_HASH) {
    auto c = std::make_unique<SpatialHashCalculator>(cfg.spatial_hash_cell_size, cfg.spatial_hash_cutoff);
    c->setGravitationalConstant(cfg.G);
    c->setSofteningParameter(cfg.softening);
    return c;
  }
  return std::make_unique<DirectForceCalculator>(cfg.cuda_block_size);
}
// The host mirror classes are never instantiated by this backend (tree_/grid_ stay null); their
// special members only have to exist for std::unique_ptr.
BarnesHutTree::BarnesHutTree(size_t n) : d_nodes_(nullptr), d_sorted_indices_(nullptr), d_morton_codes_(nullptr),
    max_particles_(n), max_nodes_(0), node_count_(0), max_depth_(0) {}
BarnesHutTree::~BarnesHutTree() {}
SpatialHashGrid::SpatialHashGrid(size_t n, float c) : d_cell_start_(nullptr), d_cell_end_(nullptr),
    d_particle_cell_(nullptr), d_sorted_indices_(nullptr), d_cell_counts_(nullptr), max_particles_(n), cell_size_(c),
    total_cells_(0) { grid_dims_ = make_int3(0, 0, 0); }
SpatialHashGrid::~SpatialHashGrid() {}

// integrator.cu:205-293
Integrator::Integrator(int block_size) : block_size_(block_size) {}
Integrator::~Integrator() = default;
void Integrator::ensureScratchBuffer(size_t) {}
void Integrator::storeOldAccelerations(ParticleData* d) { B200(nb_store_accelerations(ctx(), P(d), nullptr)); }
void Integrator::updatePositions(ParticleData* d, float dt) { B200(nb_update_positions(ctx(), P(d), dt, nullptr)); }
void Integrator::updateVelocities(ParticleData* d, float dt) { B200(nb_update_velocities(ctx(), P(d), dt, nullptr)); }
void Integrator::integrate(ParticleData* d, ForceCalculator* fc, float dt) {
  if (fc->getMethod() == ForceMethod::DIRECT_N2 && dynamic_cast<DirectForceCalculator*>(fc)) {
    // fused store+drift+pack / force / reduce+kick; same post-state as the 4 calls below
    const float e = fc->getSofteningParameter();
    B200(nb_step_direct(ctx(), P(d), dt, fc->getGravitationalConstant(), e * e, nullptr));
    return;
  }
  storeOldAccelerations(d); updatePositions(d, dt); fc->computeForces(d); updateVelocities(d, dt);
}
float Integrator::computeKineticEnergy(const ParticleData* d) {
  double v; B200(nb_kinetic_energy(ctx(), P(d), &v, nullptr)); return float(v);
}
float Integrator::computePotentialEnergy(const ParticleData* d, float G, float eps) {
  double v; B200(nb_potential_energy(ctx(), P(d), G, eps, &v, nullptr)); return float(v);
}
float Integrator::computeTotalEnergy(const ParticleData* d, float G, float eps) {
  return computeKineticEnergy(d) + computePotentialEnergy(d, G, eps);
}
}  // namespace nbody
