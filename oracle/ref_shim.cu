// oracle/ref_shim.cu -- extern "C" doorway into the UNMODIFIED reference CUDA sources
// (/root/reference/src/cuda/*.cu), compiled for sm_100 by oracle/build_ref.sh into
// oracle/_ref/libnbody_ref.so.  TEST INFRASTRUCTURE ONLY: it lets tests/ and bench.py run the
// reference's own kernels on the B200 box on the same inputs, as the parity oracle and as the
// "reference on the same box" timing.  No reference source is copied into this repo: this file
// only #includes the reference's public headers and calls its classes.
//
// `#define private public` exposes BarnesHutTree's / SpatialHashGrid's intermediate buffers
// (Morton keys, sort permutation, bbox) without touching the reference sources; class layout is
// unaffected.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define private public
#include "nbody/barnes_hut_tree.hpp"
#include "nbody/spatial_hash_grid.hpp"
#undef private
#include "nbody/error_handling.hpp"
#include "nbody/force_calculator.hpp"
#include "nbody/integrator.hpp"
#include "nbody/particle_data.hpp"

using namespace nbody;

namespace {
thread_local char g_err[512] = "";

struct DeviceState {
  ParticleData d{}, h{};
  ~DeviceState() {
    if (d.pos_x) ParticleDataManager::freeDevice(d);
  }
};

// host view over caller arrays in ParticleData order (0-2 pos, 3-5 vel, 6-8 acc, 9-11 acc_old, 12 mass)
ParticleData host_view(float* const a[13], size_t n) {
  ParticleData h;
  h.pos_x = a[0]; h.pos_y = a[1]; h.pos_z = a[2];
  h.vel_x = a[3]; h.vel_y = a[4]; h.vel_z = a[5];
  h.acc_x = a[6]; h.acc_y = a[7]; h.acc_z = a[8];
  h.acc_old_x = a[9]; h.acc_old_y = a[10]; h.acc_old_z = a[11];
  h.mass = a[12];
  h.count = n;
  return h;
}

std::unique_ptr<ForceCalculator> make_calc(int method, float G, float eps, float theta, float cell, float cutoff) {
  SimulationConfig cfg;
  cfg.G = G;
  cfg.softening = eps;
  cfg.barnes_hut_theta = theta;
  cfg.spatial_hash_cell_size = cell;
  cfg.spatial_hash_cutoff = cutoff;
  auto c = createForceCalculator(static_cast<ForceMethod>(method), cfg);
  c->setGravitationalConstant(G);  // as ParticleSystem::createForceCalculator does
  c->setSofteningParameter(eps);
  return c;
}
}  // namespace

#define REF_TRY try {
#define REF_CATCH                                  \
  }                                                \
  catch (const std::exception& e) {                \
    strncpy(g_err, e.what(), sizeof g_err - 1);    \
    return -1;                                     \
  }                                                \
  return 0;

extern "C" {

const char* ref_last_error() { return g_err; }

// forces with any of the three calculators; iters timed launches (device-synchronised), ms = mean.
int ref_forces(float* const arrays[13], size_t n, int method, float G, float eps, float theta, float cell,
               float cutoff, int iters, float* ms_out) {
  REF_TRY
  DeviceState s;
  ParticleData h = host_view(arrays, n);
  ParticleDataManager::allocateDevice(s.d, n);
  ParticleDataManager::copyToDevice(s.d, h);
  auto calc = make_calc(method, G, eps, theta, cell, cutoff);
  calc->computeForces(&s.d);
  CUDA_CHECK(cudaDeviceSynchronize());
  if (iters > 0) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int k = 0; k < iters; k++) calc->computeForces(&s.d);
    cudaEventRecord(e1);
    CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms_out) *ms_out = ms / iters;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
  ParticleDataManager::copyToHost(h, s.d);
  REF_CATCH
}

// `steps` Velocity-Verlet steps through Integrator::integrate; state in/out on the host.
// energies[0..1] = KE, PE before; [2..3] after (the reference's fp32 diagnostics), if non-null.
int ref_run(float* const arrays[13], size_t n, int method, float dt, float G, float eps, float theta, float cell,
            float cutoff, int steps, int compute_initial_forces, float* energies, float* ms_out) {
  REF_TRY
  DeviceState s;
  ParticleData h = host_view(arrays, n);
  ParticleDataManager::allocateDevice(s.d, n);
  ParticleDataManager::copyToDevice(s.d, h);
  auto calc = make_calc(method, G, eps, theta, cell, cutoff);
  Integrator integ(256);
  if (compute_initial_forces) calc->computeForces(&s.d);
  if (energies) {
    energies[0] = integ.computeKineticEnergy(&s.d);
    energies[1] = integ.computePotentialEnergy(&s.d, G, eps);
  }
  CUDA_CHECK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  for (int k = 0; k < steps; k++) integ.integrate(&s.d, calc.get(), dt);
  cudaEventRecord(e1);
  CUDA_CHECK(cudaEventSynchronize(e1));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  if (ms_out) *ms_out = steps > 0 ? ms / steps : 0.f;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (energies) {
    energies[2] = integ.computeKineticEnergy(&s.d);
    energies[3] = integ.computePotentialEnergy(&s.d, G, eps);
  }
  ParticleDataManager::copyToHost(h, s.d);
  CUDA_CHECK(cudaMemcpy(h.acc_old_x, s.d.acc_old_x, n * sizeof(float), cudaMemcpyDeviceToHost));
  CUDA_CHECK(cudaMemcpy(h.acc_old_y, s.d.acc_old_y, n * sizeof(float), cudaMemcpyDeviceToHost));
  CUDA_CHECK(cudaMemcpy(h.acc_old_z, s.d.acc_old_z, n * sizeof(float), cudaMemcpyDeviceToHost));
  REF_CATCH
}

// Barnes-Hut front half + host octree of the reference: bbox, Morton keys (before and after the
// Thrust sort), sort permutation, node count and nodes (76-byte OctreeNode, max_nodes entries).
int ref_bh_tree(float* const arrays[13], size_t n, float bbox[6], uint32_t* keys_unsorted, uint32_t* keys_sorted,
                int32_t* perm, int* node_count, void* nodes_out, int max_nodes_out) {
  REF_TRY
  DeviceState s;
  ParticleData h = host_view(arrays, n);
  ParticleDataManager::allocateDevice(s.d, n);
  ParticleDataManager::copyToDevice(s.d, h);
  BarnesHutTree tree(n);
  tree.computeBoundingBox(&s.d);
  tree.computeMortonCodes(&s.d);
  if (keys_unsorted)
    CUDA_CHECK(cudaMemcpy(keys_unsorted, tree.d_morton_codes_, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  tree.sortParticlesByMorton();
  if (keys_sorted)
    CUDA_CHECK(cudaMemcpy(keys_sorted, tree.d_morton_codes_, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  if (perm) CUDA_CHECK(cudaMemcpy(perm, tree.d_sorted_indices_, n * sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (bbox) {
    bbox[0] = tree.bbox_min_.x; bbox[1] = tree.bbox_min_.y; bbox[2] = tree.bbox_min_.z;
    bbox[3] = tree.bbox_max_.x; bbox[4] = tree.bbox_max_.y; bbox[5] = tree.bbox_max_.z;
  }
  tree.buildTreeGPU(&s.d);
  tree.computeCentersOfMass(&s.d);
  if (node_count) *node_count = tree.getNodeCount();
  if (nodes_out) {
    int c = tree.getNodeCount() < max_nodes_out ? tree.getNodeCount() : max_nodes_out;
    memcpy(nodes_out, tree.getNodes(), (size_t)c * sizeof(OctreeNode));
  }
  REF_CATCH
}

// spatial hash grid of the reference: padded bbox, dims, per-particle cell, cell_start/end, sorted
// indices (cell arrays must hold dims product entries: call once with them NULL to get dims).
int ref_sh_grid(float* const arrays[13], size_t n, float cell_size, float bbox_padded[6], int32_t dims[3],
                int32_t* particle_cell, int32_t* cell_start, int32_t* cell_end, int32_t* sorted_indices) {
  REF_TRY
  DeviceState s;
  ParticleData h = host_view(arrays, n);
  ParticleDataManager::allocateDevice(s.d, n);
  ParticleDataManager::copyToDevice(s.d, h);
  SpatialHashGrid grid(n, cell_size);
  grid.build(&s.d);
  CUDA_CHECK(cudaDeviceSynchronize());
  int3 gd = grid.getGridDims();
  dims[0] = gd.x; dims[1] = gd.y; dims[2] = gd.z;
  if (bbox_padded) {
    bbox_padded[0] = grid.bbox_min_.x; bbox_padded[1] = grid.bbox_min_.y; bbox_padded[2] = grid.bbox_min_.z;
    bbox_padded[3] = grid.bbox_max_.x; bbox_padded[4] = grid.bbox_max_.y; bbox_padded[5] = grid.bbox_max_.z;
  }
  if (particle_cell && cell_start && cell_end && sorted_indices) {
    std::vector<int> cs, ce, pc, si;
    grid.copyCellDataToHost(cs, ce, pc, si);
    memcpy(cell_start, cs.data(), cs.size() * sizeof(int));
    memcpy(cell_end, ce.data(), ce.size() * sizeof(int));
    memcpy(particle_cell, pc.data(), n * sizeof(int));
    memcpy(sorted_indices, si.data(), n * sizeof(int));
  }
  REF_CATCH
}

// the reference's only CPU force code: computeGravitationalForceCPU (src/cuda/force_direct.cu:109-117)
void ref_pair_force_cpu(const float p1[3], const float p2[3], float m1, float m2, float G, float eps, float out[3]) {
  Vec3 f = computeGravitationalForceCPU(Vec3(p1[0], p1[1], p1[2]), Vec3(p2[0], p2[1], p2[2]), m1, m2, G, eps);
  out[0] = f.x; out[1] = f.y; out[2] = f.z;
}

// device Morton helper of the reference, evaluated on the host (it is __host__ __device__)
uint32_t ref_morton_code(float x, float y, float z) { return computeMortonCode(x, y, z); }

void ref_cell_index(float x, float y, float z, float cell_size, int32_t out[3]) {
  int3 c = SpatialHashGrid::getCellIndex(x, y, z, cell_size);
  out[0] = c.x; out[1] = c.y; out[2] = c.z;
}

}  // extern "C"
