// nbody.hpp -- C++20 host layer of the B200-native engine: the reference's class names
// (namespace nbody: ParticleData, SimulationConfig, ForceCalculator family, Integrator,
// ParticleDataManager, ParticleInitializer, ParticleSystem, exceptions) as thin inline wrappers over
// the C ABI in include/nbody_b200.h.  Header-only; link with -lnbody_b200.  No CUDA headers are
// needed by host code: device memory goes through nb_particles_alloc_device.
//
// Source compatibility: code written against the reference's include/nbody/{types,
// force_calculator,integrator,particle_data,particle_system,error_handling,simulation_state,
// serialization,performance_observability}.hpp compiles against
// the forwarding headers next to this file.  Reference file:line for each piece is cited inline.
#pragma once

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <initializer_list>
#include <istream>
#include <map>
#include <memory>
#include <mutex>
#include <ostream>
#include <random>
#include <sstream>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

#include "../nbody_b200.h"

#ifndef __CUDACC__  // include/nbody/types.hpp:20-35: host builds get the reference's stand-in, nvcc brings its own
struct int3 {
  int x, y, z;
};
inline constexpr int3 make_int3(int x, int y, int z) { return {x, y, z}; }
#endif

namespace nbody {

// ---- enums / PODs (include/nbody/types.hpp:66-86, 123-198, 234-313, 343-382) -----------------------
enum class ForceMethod { DIRECT_N2, BARNES_HUT, SPATIAL_HASH };
enum class InitDistribution { UNIFORM, SPHERICAL, DISK };

struct Vec3 {
  float x = 0, y = 0, z = 0;
  constexpr Vec3() = default;
  constexpr Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
  constexpr Vec3 operator+(const Vec3& v) const { return {x + v.x, y + v.y, z + v.z}; }
  constexpr Vec3 operator-(const Vec3& v) const { return {x - v.x, y - v.y, z - v.z}; }
  constexpr Vec3 operator*(float s) const { return {x * s, y * s, z * s}; }
  constexpr Vec3 operator/(float s) const { return {x / s, y / s, z / s}; }
  Vec3& operator+=(const Vec3& v) { x += v.x; y += v.y; z += v.z; return *this; }
  Vec3& operator-=(const Vec3& v) { x -= v.x; y -= v.y; z -= v.z; return *this; }
  constexpr float dot(const Vec3& v) const { return x * v.x + y * v.y + z * v.z; }
  constexpr float length2() const { return dot(*this); }
  float length() const { return std::sqrt(length2()); }
  Vec3 normalized() const { const float l = length(); return l > 0 ? *this / l : Vec3(); }
};
inline constexpr Vec3 operator*(float s, const Vec3& v) { return v * s; }

// Layout-identical to nb_particles (13 float* + count, 112 bytes): passed to the C ABI by cast.
struct ParticleData {
  float *pos_x = nullptr, *pos_y = nullptr, *pos_z = nullptr;
  float *vel_x = nullptr, *vel_y = nullptr, *vel_z = nullptr;
  float *acc_x = nullptr, *acc_y = nullptr, *acc_z = nullptr;
  float *acc_old_x = nullptr, *acc_old_y = nullptr, *acc_old_z = nullptr;
  float* mass = nullptr;
  std::size_t count = 0;
};
static_assert(sizeof(ParticleData) == sizeof(nb_particles) && sizeof(ParticleData) == 112);
inline nb_particles* as_nb(ParticleData* p) { return reinterpret_cast<nb_particles*>(p); }
inline const nb_particles* as_nb(const ParticleData* p) { return reinterpret_cast<const nb_particles*>(p); }

struct SimulationConfig {
  std::size_t particle_count = 10000;
  InitDistribution init_distribution = InitDistribution::SPHERICAL;
  ForceMethod force_method = ForceMethod::DIRECT_N2;
  float dt = 0.001f;
  float G = 1.0f;
  float softening = 0.1f;
  float barnes_hut_theta = 0.5f;
  float spatial_hash_cell_size = 1.0f;
  float spatial_hash_cutoff = 2.0f;
  int cuda_block_size = 256;  // accepted and validated; a hint only for the sm_100a kernels
};
static_assert(sizeof(SimulationConfig) == 48);

struct UniformDistParams { Vec3 min_bounds, max_bounds; float min_mass = 1.0f, max_mass = 1.0f; };
struct SphericalDistParams { Vec3 center; float radius = 10.0f, min_mass = 1.0f, max_mass = 1.0f; };
struct DiskDistParams {
  Vec3 center;
  float radius = 10.0f, thickness = 1.0f, min_mass = 1.0f, max_mass = 1.0f, rotation_speed = 1.0f;
};

// ---- exceptions (include/nbody/error_handling.hpp:29-102) --------------------------------------------
class CudaException : public std::runtime_error {
 public:
  CudaException(const char* msg, const char* file, int line)
      : std::runtime_error(format(msg, file, line)), msg_(msg ? msg : ""), file_(file ? file : ""), line_(line) {}
  const char* getErrorMsg() const { return msg_.c_str(); }
  const char* getFile() const { return file_.c_str(); }
  int getLine() const { return line_; }

 private:
  static std::string format(const char* m, const char* f, int l) {
    std::ostringstream o;
    o << "CUDA Error: " << (m ? m : "") << " at " << (f ? f : "?") << ":" << l;
    return o.str();
  }
  std::string msg_, file_;
  int line_;
};

class ResourceException : public std::runtime_error {
 public:
  ResourceException(const char* msg, std::size_t required, std::size_t available)
      : std::runtime_error(std::string(msg) + " - Required: " + std::to_string(required) +
                           " bytes, Available: " + std::to_string(available) + " bytes"),
        required_(required), available_(available) {}
  std::size_t getRequired() const { return required_; }
  std::size_t getAvailable() const { return available_; }

 private:
  std::size_t required_, available_;
};

class ValidationException : public std::runtime_error {
 public:
  explicit ValidationException(const std::string& m) : std::runtime_error("Validation Error: " + m) {}
};

// C-ABI return code -> the exception type the reference would have thrown
inline void nb_throw(int rc, const char* file, int line) {
  if (rc == NB_OK) return;
  const char* m = nb_last_error();
  switch (rc) {
    case NB_ERR_INVALID: throw ValidationException(m);
    case NB_ERR_RESOURCE: throw ResourceException(m, 0, 0);
    case NB_ERR_GRID_TOO_LARGE: throw std::runtime_error(m);  // force_spatial_hash.cu:252-254
    default:
      if (rc > 0) throw CudaException(m, file, line);
      throw std::runtime_error(m);
  }
}
#define NB_CALL(expr) ::nbody::nb_throw((expr), __FILE__, __LINE__)

// ---- validation (src/utils/error_handling.cpp:46-123) -------------------------------------------------
inline bool nb_finite(float v) { return !(std::isnan(v) || std::isinf(v)); }
inline void validateParticleCountRange(std::size_t n) {
  if (n == 0) throw ValidationException("Particle count must be greater than 0");
  if (n > 100000000) throw ValidationException("Particle count exceeds maximum supported (100M)");
}
inline void validateTimeStep(float dt) {
  if (dt <= 0) throw ValidationException("Time step must be positive");
  if (!nb_finite(dt)) throw ValidationException("Time step must be a finite number");
  if (dt > 1.0f) throw ValidationException("Time step is too large (max 1.0)");
}
inline void validateSoftening(float e) {
  if (e < 0) throw ValidationException("Softening parameter must be non-negative");
  if (!nb_finite(e)) throw ValidationException("Softening parameter must be a finite number");
}
inline void validateTheta(float t) {
  if (t < 0 || t > 2.0f) throw ValidationException("Barnes-Hut theta must be between 0 and 2");
  if (!nb_finite(t)) throw ValidationException("Barnes-Hut theta must be a finite number");
}
inline void validateSimulationConfig(const SimulationConfig& c) {
  validateParticleCountRange(c.particle_count);
  validateTimeStep(c.dt);
  validateSoftening(c.softening);
  if (c.force_method == ForceMethod::BARNES_HUT) validateTheta(c.barnes_hut_theta);
  if (c.G <= 0 || !nb_finite(c.G)) throw ValidationException("Gravitational constant must be positive and finite");
  if (c.force_method == ForceMethod::SPATIAL_HASH) {
    if (c.spatial_hash_cell_size <= 0 || !nb_finite(c.spatial_hash_cell_size))
      throw ValidationException("Spatial hash cell size must be positive and finite");
    if (c.spatial_hash_cutoff <= 0 || !nb_finite(c.spatial_hash_cutoff))
      throw ValidationException("Spatial hash cutoff must be positive and finite");
  }
  if (c.cuda_block_size <= 0 || c.cuda_block_size > 1024)
    throw ValidationException("CUDA block size must be between 1 and 1024");
}
inline void validateResourceRequirements(std::size_t n) {  // error_handling.cpp:25-44
  std::size_t total = 0;
  NB_CALL(nb_device_info(0, nullptr, nullptr, nullptr, &total, nullptr));
  const std::size_t need = n * sizeof(float) * 13 * 2;
  if (need > total * 0.8) throw ResourceException("Insufficient GPU memory", need, static_cast<std::size_t>(total * 0.8));
}
inline void validateParticleCount(std::size_t n) {
  validateParticleCountRange(n);
  validateResourceRequirements(n);
}

// ---- one engine context per process/device (the reference is single-device, device 0) -----------------
class Engine {
 public:
  static nb_context* get() {
    static Engine e;
    return e.ctx_;
  }
  // stream the compute wrappers launch on: the legacy default stream (what the reference uses
  // everywhere) unless a graph capture has redirected it
  static nb_stream& stream() {
    static thread_local nb_stream s = nullptr;
    return s;
  }

 private:
  Engine() { NB_CALL(nb_context_create(0, 0, &ctx_)); }
  ~Engine() { nb_context_destroy(ctx_); }
  nb_context* ctx_ = nullptr;
};

// ---- ParticleDataManager / ParticleInitializer (src/cuda/particle_init.cu:143-357) ----------------------
class ParticleDataManager {
 public:
  static void allocateDevice(ParticleData& d, std::size_t n) { NB_CALL(nb_particles_alloc_device(as_nb(&d), n)); }
  static void freeDevice(ParticleData& d) { nb_particles_free_device(as_nb(&d)); d = ParticleData(); }
  static void allocateHost(ParticleData& h, std::size_t n) {
    h = ParticleData();
    float** f[13] = {&h.pos_x, &h.pos_y, &h.pos_z, &h.vel_x, &h.vel_y, &h.vel_z, &h.acc_x,
                     &h.acc_y, &h.acc_z, &h.acc_old_x, &h.acc_old_y, &h.acc_old_z, &h.mass};
    try {
      for (auto p : f) *p = new float[n ? n : 1]();
    } catch (const std::bad_alloc&) {
      for (auto p : f) delete[] *p;
      h = ParticleData();
      throw ResourceException("Failed to allocate host memory for particles", n * 13 * sizeof(float), 0);
    }
    h.count = n;
  }
  static void freeHost(ParticleData& h) {
    for (float* p : {h.pos_x, h.pos_y, h.pos_z, h.vel_x, h.vel_y, h.vel_z, h.acc_x, h.acc_y, h.acc_z, h.acc_old_x,
                     h.acc_old_y, h.acc_old_z, h.mass})
      delete[] p;
    h = ParticleData();
  }
  static void copyToDevice(ParticleData& d, const ParticleData& h) {
    NB_CALL(nb_particles_upload(Engine::get(), as_nb(&d), as_nb(&h), nullptr));
    NB_CALL(nb_synchronize(Engine::get(), nullptr));
  }
  static void copyToHost(ParticleData& h, const ParticleData& d) {
    NB_CALL(nb_particles_download(Engine::get(), as_nb(&h), as_nb(&d), nullptr));
  }
  // positions only (src/cuda/particle_init.cu:285-299)
  static void copyPositionsToHost(float* h_pos_x, float* h_pos_y, float* h_pos_z, const ParticleData& d) {
    const std::size_t bytes = d.count * sizeof(float);
    NB_CALL(nb_device_download(Engine::get(), h_pos_x, d.pos_x, bytes, nullptr));
    NB_CALL(nb_device_download(Engine::get(), h_pos_y, d.pos_y, bytes, nullptr));
    NB_CALL(nb_device_download(Engine::get(), h_pos_z, d.pos_z, bytes, nullptr));
  }
  static void copyPositionsToDevice(ParticleData& d, const float* h_pos_x, const float* h_pos_y, const float* h_pos_z) {
    const std::size_t bytes = d.count * sizeof(float);
    NB_CALL(nb_device_upload(Engine::get(), d.pos_x, h_pos_x, bytes, nullptr));
    NB_CALL(nb_device_upload(Engine::get(), d.pos_y, h_pos_y, bytes, nullptr));
    NB_CALL(nb_device_upload(Engine::get(), d.pos_z, h_pos_z, bytes, nullptr));
  }
};

class ParticleInitializer {
 public:
  static void initUniform(ParticleData& h, const UniformDistParams& p, unsigned seed = 42) {
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> dx(p.min_bounds.x, p.max_bounds.x), dy(p.min_bounds.y, p.max_bounds.y),
        dz(p.min_bounds.z, p.max_bounds.z), dm(p.min_mass, p.max_mass);
    for (std::size_t i = 0; i < h.count; i++) {
      h.pos_x[i] = dx(rng); h.pos_y[i] = dy(rng); h.pos_z[i] = dz(rng);
      h.vel_x[i] = h.vel_y[i] = h.vel_z[i] = 0.0f;
      h.mass[i] = dm(rng);
    }
    zeroAccelerations(h);
  }
  static void initSpherical(ParticleData& h, const SphericalDistParams& p, unsigned seed = 42) {
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> u(0.0f, 1.0f), dm(p.min_mass, p.max_mass);
    for (std::size_t i = 0; i < h.count; i++) {
      const float r = std::cbrt(u(rng)) * p.radius, th = u(rng) * 2.0f * 3.14159265f, ph = std::acos(2.0f * u(rng) - 1.0f);
      h.pos_x[i] = p.center.x + r * std::sin(ph) * std::cos(th);
      h.pos_y[i] = p.center.y + r * std::sin(ph) * std::sin(th);
      h.pos_z[i] = p.center.z + r * std::cos(ph);
      h.vel_x[i] = h.vel_y[i] = h.vel_z[i] = 0.0f;
      h.mass[i] = dm(rng);
    }
    zeroAccelerations(h);
  }
  static void initDisk(ParticleData& h, const DiskDistParams& p, unsigned seed = 42) {
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> u(0.0f, 1.0f), dm(p.min_mass, p.max_mass);
    for (std::size_t i = 0; i < h.count; i++) {
      const float r = std::sqrt(u(rng)) * p.radius, th = u(rng) * 2.0f * 3.14159265f, z = (u(rng) - 0.5f) * p.thickness;
      h.pos_x[i] = p.center.x + r * std::cos(th);
      h.pos_y[i] = p.center.y + r * std::sin(th);
      h.pos_z[i] = p.center.z + z;
      const float v = p.rotation_speed * std::sqrt(r);
      h.vel_x[i] = -v * std::sin(th); h.vel_y[i] = v * std::cos(th); h.vel_z[i] = 0.0f;
      h.mass[i] = dm(rng);
    }
    zeroAccelerations(h);
  }
  static void zeroVelocities(ParticleData& h) {
    std::fill_n(h.vel_x, h.count, 0.0f); std::fill_n(h.vel_y, h.count, 0.0f); std::fill_n(h.vel_z, h.count, 0.0f);
  }
  static void zeroAccelerations(ParticleData& h) {
    for (float* a : {h.acc_x, h.acc_y, h.acc_z, h.acc_old_x, h.acc_old_y, h.acc_old_z}) std::fill_n(a, h.count, 0.0f);
  }
};

// ---- force calculators (include/nbody/force_calculator.hpp) ----------------------------------------------
class ForceCalculator {
 public:
  virtual ~ForceCalculator() = default;
  virtual void computeForces(ParticleData* d_particles) = 0;
  virtual ForceMethod getMethod() const = 0;
  void setSofteningParameter(float eps) { softening_eps_ = eps; softening_eps2_ = eps * eps; }
  void setGravitationalConstant(float G) { G_ = G; }
  float getSofteningParameter() const noexcept { return softening_eps_; }
  float getSofteningSquared() const noexcept { return softening_eps2_; }
  float getGravitationalConstant() const noexcept { return G_; }

 protected:
  float softening_eps_ = 0.01f, softening_eps2_ = 0.0001f, G_ = 1.0f;
};

class DirectForceCalculator : public ForceCalculator {
 public:
  explicit DirectForceCalculator(int block_size = 256) : block_size_(block_size) {}
  void computeForces(ParticleData* d) override {  // src/cuda/force_direct.cu:103-106
    NB_CALL(nb_direct_forces(Engine::get(), as_nb(d), G_, softening_eps2_, Engine::stream()));
  }
  ForceMethod getMethod() const noexcept override { return ForceMethod::DIRECT_N2; }
  void setBlockSize(int s) { block_size_ = s; }
  int getBlockSize() const noexcept { return block_size_; }

 private:
  int block_size_;
};

// include/nbody/barnes_hut_tree.hpp:9-30.  The reference's node record, here filled from the device LBVH on request
// (copyNodesToHost): a binary tree, so children[0] / children[1] are the left / right child and the other six are -1;
// center / half_size are those of the node's Morton cell; node 0 is the root; leaves follow the n-1 internal nodes.
struct OctreeNode {
  Vec3 center;
  float half_size = 0;
  Vec3 center_of_mass;
  float total_mass = 0;
  int children[8] = {-1, -1, -1, -1, -1, -1, -1, -1};
  int particle_index = -1;
  bool is_leaf = true;
  int particle_count = 0;
};

// include/nbody/barnes_hut_tree.hpp:36-78 -- same public surface; the tree itself lives on the device (nb_bh_build),
// one per engine context, and is read back on demand.
class BarnesHutTree {
 public:
  explicit BarnesHutTree(std::size_t max_particles = 0) : max_particles_(max_particles) {}
  void build(const ParticleData* d) {  // src/cuda/force_barnes_hut.cu:282-289
    NB_CALL(nb_bh_build(Engine::get(), as_nb(d), Engine::stream()));
    n_ = d->count;
    h_nodes_.clear();
  }
  void computeForces(ParticleData* d, float theta, float G, float eps) {  // :488-498
    NB_CALL(nb_bh_forces(Engine::get(), as_nb(d), theta, G, eps, Engine::stream()));
  }
  int getNodeCount() const { int c = 0; NB_CALL(nb_bh_stats(Engine::get(), nullptr, &c)); return c; }
  int getMaxDepth() const { return max_depth_; }  // octree levels spanned by the Morton prefixes; after copyNodesToHost
  const OctreeNode* getNodes() const { return h_nodes_.data(); }
  std::uint64_t getInteractionCount() const { std::uint64_t v = 0; NB_CALL(nb_bh_stats(Engine::get(), &v, nullptr)); return v; }
  void copyNodesToHost() {  // :500-503
    const std::size_t n = bodies();
    if (n == 0) { h_nodes_.clear(); return; }
    std::vector<std::int32_t> perm(n), left(n > 1 ? n - 1 : 1), right(left.size()), prefix(left.size());
    std::vector<float> com(4 * (2 * n - 1));
    float bb[6];
    NB_CALL(nb_bh_export(Engine::get(), nullptr, perm.data(), left.data(), right.data(), nullptr, prefix.data(), com.data(), bb));
    h_nodes_.assign(2 * n - 1, OctreeNode());
    const float side = std::max(std::max(bb[3] - bb[0], bb[4] - bb[1]), bb[5] - bb[2]) + 0.002f;  // :339-344
    const Vec3 root_center((bb[0] + bb[3]) * 0.5f, (bb[1] + bb[4]) * 0.5f, (bb[2] + bb[5]) * 0.5f);
    max_depth_ = 0;
    for (std::size_t i = 0; i < 2 * n - 1; i++) {
      OctreeNode& o = h_nodes_[i];
      o.center_of_mass = Vec3(com[4 * i], com[4 * i + 1], com[4 * i + 2]);
      o.total_mass = com[4 * i + 3];
      o.center = root_center;
      o.half_size = side * 0.5f;
      if (n > 1 && i < n - 1) {
        o.is_leaf = false;
        o.children[0] = left[i];
        o.children[1] = right[i];
        const int level = (prefix[i] + 2) / 3;  // octree level of the node's common Morton prefix
        o.half_size = side * 0.5f / static_cast<float>(1 << std::min(level, 30));
        max_depth_ = std::max(max_depth_, level);
      } else {
        o.particle_index = perm[n > 1 ? i - (n - 1) : 0];
        o.particle_count = 1;
      }
    }
    if (n > 1) countBodies(0);  // particle_count of the internal nodes, bottom-up
  }
  bool verifyTreeStructure() const {  // :505-509, plus what it says it would check: every body exactly once
    if (h_nodes_.empty()) return getNodeCount() > 0;
    const std::size_t n = bodies();
    std::vector<char> seen(n, 0);
    for (const OctreeNode& o : h_nodes_)
      if (o.is_leaf) {
        if (o.particle_index < 0 || static_cast<std::size_t>(o.particle_index) >= n || seen[o.particle_index]) return false;
        seen[o.particle_index] = 1;
      }
    return h_nodes_[0].particle_count == static_cast<int>(n);
  }
  bool verifyMassConservation(const ParticleData* h) const {  // :511-519
    float total_mass = 0.0f;
    for (std::size_t i = 0; i < h->count; i++) total_mass += h->mass[i];
    const float root_mass = h_nodes_.empty() ? getRootMass(h->count) : h_nodes_[0].total_mass;
    return std::fabs(total_mass - root_mass) < 0.001f * total_mass;
  }
  // sorted Morton keys + permutation, LBVH topology, node centres of mass (x,y,z,m)
  void exportTree(std::size_t n, std::vector<std::uint32_t>& keys, std::vector<std::int32_t>& perm,
                  std::vector<std::int32_t>& left, std::vector<std::int32_t>& right, std::vector<float>& node_com) const {
    keys.resize(n); perm.resize(n); left.resize(n > 1 ? n - 1 : 1); right.resize(left.size()); node_com.resize(4 * (2 * n - 1));
    NB_CALL(nb_bh_export(Engine::get(), keys.data(), perm.data(), left.data(), right.data(), nullptr, nullptr,
                         node_com.data(), nullptr));
  }
  float getRootMass(std::size_t n) const {
    std::vector<float> com(4 * (2 * n - 1));
    NB_CALL(nb_bh_export(Engine::get(), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, com.data(), nullptr));
    return com[3];
  }

 private:
  std::size_t bodies() const { return n_ ? n_ : static_cast<std::size_t>((getNodeCount() + 1) / 2); }
  int countBodies(int i) {  // explicit stack: the LBVH of sorted duplicates can be deep
    std::vector<std::pair<int, int>> st;  // (node, state)
    st.emplace_back(i, 0);
    while (!st.empty()) {
      auto [node, state] = st.back();
      OctreeNode& o = h_nodes_[node];
      if (o.is_leaf) { st.pop_back(); continue; }
      if (state == 0) {
        st.back().second = 1;
        st.emplace_back(o.children[0], 0);
        st.emplace_back(o.children[1], 0);
      } else {
        o.particle_count = h_nodes_[o.children[0]].particle_count + h_nodes_[o.children[1]].particle_count;
        st.pop_back();
      }
    }
    return h_nodes_[i].particle_count;
  }
  std::size_t max_particles_ = 0, n_ = 0;
  int max_depth_ = 0;
  std::vector<OctreeNode> h_nodes_;
};

class BarnesHutCalculator : public ForceCalculator {
 public:
  explicit BarnesHutCalculator(float theta = 0.5f) : theta_(theta) {}
  void computeForces(ParticleData* d) override {  // src/cuda/force_barnes_hut.cu:526-532
    if (!tree_) tree_ = std::make_unique<BarnesHutTree>(d->count);
    tree_->build(d);
    tree_->computeForces(d, theta_, G_, softening_eps_);
  }
  ForceMethod getMethod() const noexcept override { return ForceMethod::BARNES_HUT; }
  void setTheta(float t) { theta_ = t; }
  float getTheta() const noexcept { return theta_; }
  BarnesHutTree* getTree() noexcept { return tree_.get(); }

 private:
  std::unique_ptr<BarnesHutTree> tree_;
  float theta_;
};

// include/nbody/spatial_hash_grid.hpp:13-57 -- same public surface; the grid lives on the device (nb_sh_build), one
// per engine context.
class SpatialHashGrid {
 public:
  explicit SpatialHashGrid(std::size_t max_particles = 0, float cell_size = 1.0f)
      : max_particles_(max_particles), cell_size_(cell_size) {}
  void build(const ParticleData* d) {  // src/cuda/force_spatial_hash.cu:235-303
    NB_CALL(nb_sh_build(Engine::get(), as_nb(d), cell_size_, Engine::stream()));
    n_ = d->count;
    std::int32_t dims[3];
    NB_CALL(nb_sh_export(Engine::get(), dims, bbox_, nullptr, nullptr, nullptr, nullptr));
    grid_dims_ = make_int3(dims[0], dims[1], dims[2]);
  }
  void computeForces(ParticleData* d, float cutoff, float G, float eps) {  // :305-316
    NB_CALL(nb_sh_forces(Engine::get(), as_nb(d), cutoff, G, eps, Engine::stream()));
  }
  void setCellSize(float s) { cell_size_ = s; }
  int3 getGridDims() const { return grid_dims_; }
  float getCellSize() const { return cell_size_; }
  int getTotalCells() const { int t = 0; NB_CALL(nb_sh_total_cells(Engine::get(), &t)); return t; }
  void copyCellDataToHost(std::vector<int>& cell_start, std::vector<int>& cell_end, std::vector<int>& particle_cells,
                          std::vector<int>& sorted_indices) const {  // :318-334
    copyCellDataToHost(n_, cell_start, cell_end, particle_cells, sorted_indices);
  }
  void copyCellDataToHost(std::size_t n, std::vector<int>& cell_start, std::vector<int>& cell_end,
                          std::vector<int>& particle_cells, std::vector<int>& sorted_indices) const {
    const int t = getTotalCells();
    cell_start.resize(t); cell_end.resize(t); particle_cells.resize(n); sorted_indices.resize(n);
    NB_CALL(nb_sh_export(Engine::get(), nullptr, nullptr, particle_cells.data(), cell_start.data(), cell_end.data(),
                         sorted_indices.data()));
  }
  bool verifyCellAssignment(const ParticleData* h) const {  // :336-362
    for (std::size_t i = 0; i < h->count; i++) {
      int3 c = getCellIndex(h->pos_x[i] - bbox_[0], h->pos_y[i] - bbox_[1], h->pos_z[i] - bbox_[2], cell_size_);
      c.x = std::max(0, std::min(c.x, grid_dims_.x - 1));
      c.y = std::max(0, std::min(c.y, grid_dims_.y - 1));
      c.z = std::max(0, std::min(c.z, grid_dims_.z - 1));
      const float lo[3] = {bbox_[0] + c.x * cell_size_, bbox_[1] + c.y * cell_size_, bbox_[2] + c.z * cell_size_};
      const float p[3] = {h->pos_x[i], h->pos_y[i], h->pos_z[i]};
      for (int a = 0; a < 3; a++)
        if (p[a] < lo[a] || p[a] > lo[a] + cell_size_) return false;
    }
    return true;
  }
  static int3 getCellIndex(float x, float y, float z, float cell_size) {  // :14-17
    return make_int3(static_cast<int>(std::floor(x / cell_size)), static_cast<int>(std::floor(y / cell_size)),
                     static_cast<int>(std::floor(z / cell_size)));
  }
  static int hashCell(int3 cell, int3 grid_dims) {  // :19-25, wraps negative indices
    const int cx = ((cell.x % grid_dims.x) + grid_dims.x) % grid_dims.x;
    const int cy = ((cell.y % grid_dims.y) + grid_dims.y) % grid_dims.y;
    const int cz = ((cell.z % grid_dims.z) + grid_dims.z) % grid_dims.z;
    return cx + cy * grid_dims.x + cz * grid_dims.x * grid_dims.y;
  }

 private:
  std::size_t max_particles_ = 0, n_ = 0;
  float cell_size_ = 1.0f;
  int3 grid_dims_ = {0, 0, 0};
  float bbox_[6] = {0, 0, 0, 0, 0, 0};
};

class SpatialHashCalculator : public ForceCalculator {
 public:
  explicit SpatialHashCalculator(float cell_size = 1.0f, float cutoff_radius = 2.0f)
      : cell_size_(cell_size), cutoff_radius_(cutoff_radius) {}
  void computeForces(ParticleData* d) override {  // src/cuda/force_spatial_hash.cu:370-377
    if (!grid_) grid_ = std::make_unique<SpatialHashGrid>(d->count, cell_size_);
    grid_->setCellSize(cell_size_);
    grid_->build(d);
    grid_->computeForces(d, cutoff_radius_, G_, softening_eps_);
  }
  ForceMethod getMethod() const noexcept override { return ForceMethod::SPATIAL_HASH; }
  void setCellSize(float s) { cell_size_ = s; }
  void setCutoffRadius(float r) { cutoff_radius_ = r; }
  float getCellSize() const noexcept { return cell_size_; }
  float getCutoffRadius() const noexcept { return cutoff_radius_; }
  SpatialHashGrid* getGrid() noexcept { return grid_.get(); }

 private:
  std::unique_ptr<SpatialHashGrid> grid_;
  float cell_size_, cutoff_radius_;
};

// src/cuda/force_spatial_hash.cu:380-401: the bare factory leaves Direct at G = 1, eps = 0.01.
inline std::unique_ptr<ForceCalculator> createForceCalculator(ForceMethod method, const SimulationConfig& cfg) {
  switch (method) {
    case ForceMethod::BARNES_HUT: {
      auto c = std::make_unique<BarnesHutCalculator>(cfg.barnes_hut_theta);
      c->setGravitationalConstant(cfg.G);
      c->setSofteningParameter(cfg.softening);
      return c;
    }
    case ForceMethod::SPATIAL_HASH: {
      auto c = std::make_unique<SpatialHashCalculator>(cfg.spatial_hash_cell_size, cfg.spatial_hash_cutoff);
      c->setGravitationalConstant(cfg.G);
      c->setSofteningParameter(cfg.softening);
      return c;
    }
    default: return std::make_unique<DirectForceCalculator>(cfg.cuda_block_size);
  }
}

// ---- integrator (include/nbody/integrator.hpp, src/cuda/integrator.cu:205-293) ------------------------------
class Integrator {
 public:
  explicit Integrator(int block_size = 256) : block_size_(block_size) {}
  // Velocity Verlet.  With our DirectForceCalculator the step runs fused (store+drift+pack, force,
  // reduce+kick); with any other ForceCalculator it is the reference's 4-call sequence (:224-238).
  void integrate(ParticleData* d, ForceCalculator* fc, float dt) {
    if (auto* direct = dynamic_cast<DirectForceCalculator*>(fc)) {
      NB_CALL(nb_step_direct(Engine::get(), as_nb(d), dt, direct->getGravitationalConstant(),
                             direct->getSofteningSquared(), Engine::stream()));
      return;
    }
    storeOldAccelerations(d);
    updatePositions(d, dt);
    fc->computeForces(d);
    updateVelocities(d, dt);
  }
  void updatePositions(ParticleData* d, float dt) { NB_CALL(nb_update_positions(Engine::get(), as_nb(d), dt, Engine::stream())); }
  void updateVelocities(ParticleData* d, float dt) { NB_CALL(nb_update_velocities(Engine::get(), as_nb(d), dt, Engine::stream())); }
  void storeOldAccelerations(ParticleData* d) { NB_CALL(nb_store_accelerations(Engine::get(), as_nb(d), Engine::stream())); }
  float computeKineticEnergy(const ParticleData* d) {
    double ke = 0;
    NB_CALL(nb_kinetic_energy(Engine::get(), as_nb(d), &ke, nullptr));
    return static_cast<float>(ke);
  }
  float computePotentialEnergy(const ParticleData* d, float G, float eps) {
    double pe = 0;
    NB_CALL(nb_potential_energy(Engine::get(), as_nb(d), G, eps, &pe, nullptr));
    return static_cast<float>(pe);
  }
  // fp64 variants (the drift criterion is evaluated in double)
  double computeTotalEnergyF64(const ParticleData* d, float G, float eps) {
    double ke = 0, pe = 0;
    NB_CALL(nb_kinetic_energy(Engine::get(), as_nb(d), &ke, nullptr));
    NB_CALL(nb_potential_energy(Engine::get(), as_nb(d), G, eps, &pe, nullptr));
    return ke + pe;
  }
  float computeTotalEnergy(const ParticleData* d, float G, float eps) {
    return computeKineticEnergy(d) + computePotentialEnergy(d, G, eps);
  }
  // total mass, centre of mass, momenta, KE and (optionally) PE, total energy, virial ratio -- fp64, one call
  nb_diagnostics_t computeDiagnostics(const ParticleData* d, float G, float eps, bool with_potential = true) {
    nb_diagnostics_t out{};
    NB_CALL(nb_diagnostics(Engine::get(), as_nb(d), G, eps, with_potential ? 1 : 0, &out, nullptr));
    return out;
  }
  void ensureScratchBuffer(std::size_t) {}  // scratch lives in the engine context
  void setBlockSize(int s) { block_size_ = s; }
  int getBlockSize() const noexcept { return block_size_; }

 private:
  int block_size_;
};

// ---- ParticleSystem (src/core/particle_system.cpp) ---------------------------------------------------------
struct SimulationState {  // include/nbody/simulation_state.hpp
  std::size_t particle_count = 0;
  float simulation_time = 0, dt = 0.001f, G = 1.0f, softening = 0.01f;
  ForceMethod force_method = ForceMethod::DIRECT_N2;
  std::vector<float> pos_x, pos_y, pos_z, vel_x, vel_y, vel_z, mass;

  void serialize(std::ostream& out) const;             // src/utils/simulation_state.cpp:42-44
  static SimulationState deserialize(std::istream& in);  // :46-48
  // equality within 1e-6 absolute on every scalar and component (:7-40)
  bool operator==(const SimulationState& o) const {
    auto near = [](float a, float b) { return !(std::abs(a - b) > 1e-6f); };
    if (particle_count != o.particle_count || force_method != o.force_method) return false;
    if (!near(simulation_time, o.simulation_time) || !near(dt, o.dt) || !near(G, o.G) || !near(softening, o.softening))
      return false;
    const std::vector<float>* mine[7] = {&pos_x, &pos_y, &pos_z, &vel_x, &vel_y, &vel_z, &mass};
    const std::vector<float>* theirs[7] = {&o.pos_x, &o.pos_y, &o.pos_z, &o.vel_x, &o.vel_y, &o.vel_z, &o.mass};
    for (int a = 0; a < 7; a++)
      for (std::size_t i = 0; i < particle_count; i++)
        if (!near((*mine[a])[i], (*theirs[a])[i])) return false;
    return true;
  }
};

// ---- .nbody checkpoint format (include/nbody/serialization.hpp, src/utils/serialization.cpp) ----------------
// 56-byte header {magic "NBOD", version 1, u64 count, time, dt, G, softening, u32 method, 4 reserved
// words} followed by seven fp32 arrays of `count` values: pos x/y/z, vel x/y/z, mass.  Files written
// by either implementation load in the other (tests/golden/ref_checkpoint.nbody).
constexpr std::uint32_t NBODY_MAGIC = 0x4E424F44;
constexpr std::uint32_t NBODY_VERSION = 1;
constexpr std::uint64_t MAX_PARTICLE_COUNT = 100'000'000;

struct FileHeader {
  std::uint32_t magic, version;
  std::uint64_t particle_count;
  float simulation_time, dt, G, softening;
  std::uint32_t force_method;
  std::uint32_t reserved[4];
};
static_assert(sizeof(FileHeader) == 56, "FileHeader must match the reference's 56-byte header");

class Serializer {
 public:
  static void save(const std::string& filename, const SimulationState& state) {
    std::ofstream f(filename, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file for writing: " + filename);
    save(f, state);
  }
  static SimulationState load(const std::string& filename) {
    std::ifstream f(filename, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file for reading: " + filename);
    return load(f);
  }
  static void save(std::ostream& out, const SimulationState& st) {
    FileHeader h{};
    h.magic = NBODY_MAGIC; h.version = NBODY_VERSION; h.particle_count = st.particle_count;
    h.simulation_time = st.simulation_time; h.dt = st.dt; h.G = st.G; h.softening = st.softening;
    h.force_method = static_cast<std::uint32_t>(st.force_method);
    out.write(reinterpret_cast<const char*>(&h), sizeof h);
    for (const std::vector<float>* a : {&st.pos_x, &st.pos_y, &st.pos_z, &st.vel_x, &st.vel_y, &st.vel_z, &st.mass})
      out.write(reinterpret_cast<const char*>(a->data()), static_cast<std::streamsize>(a->size() * sizeof(float)));
  }
  static SimulationState load(std::istream& in) {
    const FileHeader h = readHeader(in);
    if (h.particle_count > MAX_PARTICLE_COUNT)
      throw ValidationException("Particle count (" + std::to_string(h.particle_count) + ") exceeds maximum allowed (" +
                                std::to_string(MAX_PARTICLE_COUNT) + ")");
    SimulationState st;
    st.particle_count = h.particle_count; st.simulation_time = h.simulation_time; st.dt = h.dt; st.G = h.G;
    st.softening = h.softening; st.force_method = static_cast<ForceMethod>(h.force_method);
    for (std::vector<float>* a : {&st.pos_x, &st.pos_y, &st.pos_z, &st.vel_x, &st.vel_y, &st.vel_z, &st.mass}) {
      a->resize(st.particle_count);
      const auto bytes = static_cast<std::streamsize>(st.particle_count * sizeof(float));
      in.read(reinterpret_cast<char*>(a->data()), bytes);
      if (!in || in.gcount() != bytes)
        throw std::runtime_error("Failed to read particle data: file may be truncated or corrupted");
    }
    return st;
  }
  static bool validateFile(const std::string& filename) {
    std::ifstream f(filename, std::ios::binary);
    return f ? validateStream(f) : false;
  }
  static bool validateStream(std::istream& in) {
    try {
      const FileHeader h = readHeader(in);
      return h.magic == NBODY_MAGIC && h.version == NBODY_VERSION;
    } catch (...) {
      return false;
    }
  }

 private:
  static FileHeader readHeader(std::istream& in) {
    FileHeader h;
    in.read(reinterpret_cast<char*>(&h), sizeof h);
    if (!in || in.gcount() != static_cast<std::streamsize>(sizeof h))
      throw std::runtime_error("Failed to read file header: file may be truncated or corrupted");
    if (h.magic != NBODY_MAGIC) throw std::runtime_error("Invalid file format: wrong magic number");
    if (h.version != NBODY_VERSION) throw std::runtime_error("Unsupported file version");
    return h;
  }
};
inline void SimulationState::serialize(std::ostream& out) const { Serializer::save(out, *this); }
inline SimulationState SimulationState::deserialize(std::istream& in) { return Serializer::load(in); }

// ---- benchmark records (include/nbody/performance_observability.hpp, src/utils/performance_observability.cpp) ----
using Milliseconds = std::chrono::duration<double, std::milli>;
struct PhaseTiming {
  std::string name;
  Milliseconds total_duration{0.0};
  std::size_t samples = 0;
};
class PhaseProfiler {  // accumulates per-name totals in first-seen order (:52-77)
 public:
  void record(std::string_view name, Milliseconds d) {
    std::lock_guard<std::mutex> lock(mutex_);
    for (auto& p : phases_)
      if (p.name == name) {
        p.total_duration += d;
        p.samples++;
        return;
      }
    phases_.push_back(PhaseTiming{std::string(name), d, 1});
  }
  std::vector<PhaseTiming> snapshot() const {
    std::lock_guard<std::mutex> lock(mutex_);
    return phases_;
  }
  void reset() {
    std::lock_guard<std::mutex> lock(mutex_);
    phases_.clear();
  }

 private:
  mutable std::mutex mutex_;
  std::vector<PhaseTiming> phases_;
};
class ScopedPhaseProfile {
 public:
  ScopedPhaseProfile(PhaseProfiler& p, std::string name)
      : profiler_(p), name_(std::move(name)), t0_(std::chrono::steady_clock::now()) {}
  ~ScopedPhaseProfile() {
    profiler_.record(name_, std::chrono::duration_cast<Milliseconds>(std::chrono::steady_clock::now() - t0_));
  }
  ScopedPhaseProfile(const ScopedPhaseProfile&) = delete;
  ScopedPhaseProfile& operator=(const ScopedPhaseProfile&) = delete;

 private:
  PhaseProfiler& profiler_;
  std::string name_;
  std::chrono::steady_clock::time_point t0_;
};
inline PhaseProfiler& globalPhaseProfiler() {
  static PhaseProfiler p;
  return p;
}
inline std::vector<PhaseTiming> consumeGlobalPhaseSnapshot() {
  auto v = globalPhaseProfiler().snapshot();
  globalPhaseProfiler().reset();
  return v;
}
struct BenchmarkRunRecord {
  std::string benchmark_name;
  ForceMethod force_method = ForceMethod::DIRECT_N2;
  std::size_t particle_count = 0, iterations = 0;
  std::map<std::string, double> metrics, parameters;
  std::vector<PhaseTiming> phase_timings;
};
inline std::string forceMethodToString(ForceMethod m) {
  switch (m) {
    case ForceMethod::DIRECT_N2: return "direct_n2";
    case ForceMethod::BARNES_HUT: return "barnes_hut";
    case ForceMethod::SPATIAL_HASH: return "spatial_hash";
  }
  return "unknown";
}
namespace detail {
inline void jsonString(std::ostream& o, std::string_view v) {  // the five escapes of :10-35
  o << '"';
  for (char ch : v) {
    if (ch == '\\') o << "\\\\";
    else if (ch == '"') o << "\\\"";
    else if (ch == '\n') o << "\\n";
    else if (ch == '\r') o << "\\r";
    else if (ch == '\t') o << "\\t";
    else o << ch;
  }
  o << '"';
}
inline void jsonNumberMap(std::ostream& o, const std::map<std::string, double>& m) {
  o << '{';
  const char* sep = "";
  for (const auto& [k, v] : m) {
    o << sep;
    jsonString(o, k);
    o << ':' << v;  // default ostream formatting (6 significant digits), as the reference
    sep = ",";
  }
  o << '}';
}
}  // namespace detail
// byte-identical to serializeBenchmarkRunRecord (:102-133) for the same record
inline std::string serializeBenchmarkRunRecord(const BenchmarkRunRecord& r) {
  std::ostringstream o;
  o << "{\"benchmark_name\":";
  detail::jsonString(o, r.benchmark_name);
  o << ",\"force_method\":\"" << forceMethodToString(r.force_method) << "\",\"particle_count\":" << r.particle_count
    << ",\"iterations\":" << r.iterations << ",\"metrics\":";
  detail::jsonNumberMap(o, r.metrics);
  o << ",\"parameters\":";
  detail::jsonNumberMap(o, r.parameters);
  o << ",\"phase_timings\":[";
  const char* sep = "";
  for (const auto& ph : r.phase_timings) {
    o << sep << "{\"name\":";
    detail::jsonString(o, ph.name);
    o << ",\"total_duration_ms\":" << ph.total_duration.count() << ",\"samples\":" << ph.samples << '}';
    sep = ",";
  }
  o << "]}";
  return o.str();
}
inline std::string serializeBenchmarkRunRecords(const std::vector<BenchmarkRunRecord>& rs) {
  std::string out = "{\"benchmarks\":[";
  for (std::size_t i = 0; i < rs.size(); i++) out += (i ? "," : "") + serializeBenchmarkRunRecord(rs[i]);
  return out + "]}";
}
inline void writeBenchmarkRunRecords(const std::string& path, const std::vector<BenchmarkRunRecord>& rs) {
  std::ofstream f(path);
  if (!f) throw std::runtime_error("Failed to open benchmark output file: " + path);
  f << serializeBenchmarkRunRecords(rs) << "\n";
}
#define NBODY_PROFILE_SCOPE(name) ::nbody::ScopedPhaseProfile nbody_scoped_phase_profile_##__LINE__(::nbody::globalPhaseProfiler(), (name))

// ---- headless render classes (include/nbody/camera.hpp, renderer.hpp; src/render/camera.cpp, renderer.cpp) ----
// Same names and members as the reference's, with Vec3 / Mat4 in place of glm types and a device
// RGBA8 framebuffer (PPM file on request) in place of the GL context.  render() runs the CUDA splat
// kernels (nb_render_splats) on the device ParticleData directly; the reference copies it into GL
// buffers first (CudaGLInterop, K16 = nb_interop_pack).
enum class ColorMode { DEPTH, VELOCITY, DENSITY };
struct Mat4 {
  float m[16] = {0};  // column-major, m[col * 4 + row], like glm::mat4
  float& at(int col, int row) { return m[col * 4 + row]; }
  float at(int col, int row) const { return m[col * 4 + row]; }
};
inline Mat4 operator*(const Mat4& a, const Mat4& b) {  // column-major product, like glm
  Mat4 r;
  for (int c = 0; c < 4; c++)
    for (int row = 0; row < 4; row++) {
      float v = 0.0f;
      for (int k = 0; k < 4; k++) v += a.at(k, row) * b.at(c, k);
      r.at(c, row) = v;
    }
  return r;
}
inline Vec3 cross(const Vec3& a, const Vec3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline Mat4 lookAt(const Vec3& eye, const Vec3& center, const Vec3& up) {  // glm::lookAt, right-handed
  const Vec3 f = (center - eye).normalized(), s = cross(f, up).normalized(), u = cross(s, f);
  Mat4 r;
  r.at(0, 0) = s.x; r.at(1, 0) = s.y; r.at(2, 0) = s.z;
  r.at(0, 1) = u.x; r.at(1, 1) = u.y; r.at(2, 1) = u.z;
  r.at(0, 2) = -f.x; r.at(1, 2) = -f.y; r.at(2, 2) = -f.z;
  r.at(3, 0) = -s.dot(eye); r.at(3, 1) = -u.dot(eye); r.at(3, 2) = f.dot(eye); r.at(3, 3) = 1.0f;
  return r;
}
inline Mat4 perspective(float fovy_rad, float aspect, float z_near, float z_far) {  // glm::perspective, RH, -1..1
  const float t = std::tan(fovy_rad / 2.0f);
  Mat4 r;
  r.at(0, 0) = 1.0f / (aspect * t);
  r.at(1, 1) = 1.0f / t;
  r.at(2, 2) = -(z_far + z_near) / (z_far - z_near);
  r.at(2, 3) = -1.0f;
  r.at(3, 2) = -(2.0f * z_far * z_near) / (z_far - z_near);
  return r;
}
class Camera {  // camera.hpp:9-88
 public:
  Camera(float fov = 45.0f, float aspect = 16.0f / 9.0f, float z_near = 0.1f, float z_far = 1000.0f)
      : fov_(fov), aspect_(aspect), near_(z_near), far_(z_far) {}
  Mat4 getViewMatrix() const { return lookAt(position_, target_, up_); }
  Mat4 getProjectionMatrix() const { return perspective(fov_ * 3.14159265358979323846f / 180.0f, aspect_, near_, far_); }
  Mat4 getViewProjectionMatrix() const { return getProjectionMatrix() * getViewMatrix(); }  // camera.cpp:32-34
  void setPosition(const Vec3& p) { position_ = p; }
  void setTarget(const Vec3& t) { target_ = t; }
  void setUp(const Vec3& u) { up_ = u; }
  Vec3 getPosition() const { return position_; }
  Vec3 getTarget() const { return target_; }
  Vec3 getUp() const { return up_; }
  Vec3 getForward() const { return (target_ - position_).normalized(); }
  Vec3 getRight() const { return cross(getForward(), up_).normalized(); }
  void setFOV(float f) { fov_ = f; }
  void setAspectRatio(float a) { aspect_ = a; }
  void setNearFar(float n, float f) { near_ = n; far_ = f; }
  float getFOV() const { return fov_; }
  float getAspectRatio() const { return aspect_; }
  float getNear() const { return near_; }
  float getFar() const { return far_; }
  void rotate(float yaw, float pitch) {  // camera.cpp:46-65
    const Vec3 d = position_ - target_;
    const float dist = d.length();
    const float theta = std::atan2(d.x, d.z) + yaw;
    const float phi = std::clamp(std::acos(d.y / dist) + pitch, 0.1f, 3.04f);
    position_ = target_ + Vec3(dist * std::sin(phi) * std::sin(theta), dist * std::cos(phi), dist * std::sin(phi) * std::cos(theta));
  }
  void orbit(float yaw, float pitch) { rotate(yaw, pitch); }
  void pan(float dx, float dy) {  // :67-75
    const Vec3 right = getRight(), up = cross(right, getForward()).normalized(), off = right * dx + up * dy;
    position_ += off;
    target_ += off;
  }
  void zoom(float delta) {  // :77-85
    const Vec3 d = target_ - position_;
    position_ = target_ - d.normalized() * std::clamp(d.length() - delta, 1.0f, 1000.0f);
  }
  void reset() { position_ = Vec3(0, 0, 50); target_ = Vec3(0, 0, 0); up_ = Vec3(0, 1, 0); }
  void setOrbitDistance(float dist) { position_ = target_ + (position_ - target_).normalized() * dist; }
  float getOrbitDistance() const { return (position_ - target_).length(); }

 private:
  Vec3 position_{0, 0, 50}, target_{0, 0, 0}, up_{0, 1, 0};
  float fov_, aspect_, near_, far_;
};
struct ColorMapper {  // camera.cpp:104-128
  static Vec3 gradientMap(float t, const Vec3& lo, const Vec3& hi) { return lo * (1.0f - t) + hi * t; }
  static Vec3 velocityToColor(float v, float vmax) { return gradientMap(std::clamp(v / vmax, 0.0f, 1.0f), {0, 0, 1}, {1, 0, 0}); }
  static Vec3 depthToColor(float d, float dmax) { return gradientMap(std::clamp(d / dmax, 0.0f, 1.0f), {1, 0.5f, 0}, {0, 0.5f, 1}); }
  static Vec3 densityToColor(float d, float dmax) { return gradientMap(std::clamp(d / dmax, 0.0f, 1.0f), {0.2f, 0.2f, 0.8f}, {1, 1, 0.2f}); }
};
class Renderer {  // renderer.hpp:20-80
 public:
  Renderer() = default;
  ~Renderer() { cleanup(); }
  Renderer(const Renderer&) = delete;
  Renderer& operator=(const Renderer&) = delete;
  void initialize(int width, int height) {  // renderer.cpp:93-121 without GLEW / shader compilation / VAO
    cleanup();
    width_ = width; height_ = height;
    NB_CALL(nb_device_alloc(Engine::get(), static_cast<std::size_t>(width) * height * 4, &framebuffer_));
    camera_.setPosition(Vec3(0, 0, 50));
    camera_.setTarget(Vec3(0, 0, 0));
    camera_.setAspectRatio(static_cast<float>(width) / static_cast<float>(height));
    is_initialized_ = true;
  }
  void cleanup() {
    if (framebuffer_) nb_device_free(Engine::get(), framebuffer_);
    framebuffer_ = nullptr;
    is_initialized_ = false;
  }
  void onResize(int width, int height) { initialize(width, height); }
  // one frame into the device framebuffer; asynchronous apart from one 8-byte readback
  void render(const ParticleData* d_particles) {
    if (!is_initialized_) return;  // :185-186
    nb_render_params p{};
    const Mat4 v = camera_.getViewMatrix(), pr = camera_.getProjectionMatrix();
    std::copy(v.m, v.m + 16, p.view);
    std::copy(pr.m, pr.m + 16, p.proj);
    p.point_size = point_size_; p.max_depth = max_depth_; p.max_velocity = max_velocity_;
    p.color_mode = static_cast<int>(color_mode_); p.width = width_; p.height = height_;
    p.point_size_min = 1.0f; p.point_size_max = 63.0f;
    p.clear_color[0] = 0.02f; p.clear_color[1] = 0.02f; p.clear_color[2] = 0.05f; p.clear_color[3] = 1.0f;  // :9-11
    NB_CALL(nb_render_splats(Engine::get(), as_nb(d_particles), &p, static_cast<std::uint8_t*>(framebuffer_), Engine::stream()));
  }
  std::vector<std::uint8_t> readPixels() const {  // height x width x RGBA, row 0 = bottom row (glReadPixels order)
    std::vector<std::uint8_t> px(static_cast<std::size_t>(width_) * height_ * 4);
    if (framebuffer_) NB_CALL(nb_device_download(Engine::get(), px.data(), framebuffer_, px.size(), Engine::stream()));
    return px;
  }
  void savePPM(const std::string& filename) const {
    const auto px = readPixels();
    std::ofstream f(filename, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to open file for writing: " + filename);
    f << "P6\n" << width_ << ' ' << height_ << "\n255\n";
    for (int y = height_ - 1; y >= 0; y--)
      for (int x = 0; x < width_; x++) f.write(reinterpret_cast<const char*>(&px[(static_cast<std::size_t>(y) * width_ + x) * 4]), 3);
  }
  void setColorMode(ColorMode m) { color_mode_ = m; }
  ColorMode getColorMode() const { return color_mode_; }
  void setPointSize(float s) { point_size_ = s; }
  float getPointSize() const { return point_size_; }
  void setCamera(const Camera& c) { camera_ = c; }
  unsigned int getShaderProgram() const { return 0; }  // no GL program object: the shader math is the CUDA kernels
  void setMaxDepth(float d) { max_depth_ = d; }
  void setMaxVelocity(float v) { max_velocity_ = v; }
  Camera& getCamera() { return camera_; }
  const Camera& getCamera() const { return camera_; }
  bool isInitialized() const { return is_initialized_; }
  int getWidth() const { return width_; }
  int getHeight() const { return height_; }

 private:
  ColorMode color_mode_ = ColorMode::DEPTH;  // renderer.cpp:79-88
  float point_size_ = 2.0f, max_depth_ = 100.0f, max_velocity_ = 10.0f;
  int width_ = 1280, height_ = 720;
  bool is_initialized_ = false;
  void* framebuffer_ = nullptr;
  Camera camera_;
};

// ---- CudaGLInterop (include/nbody/cuda_gl_interop.hpp, src/render/cuda_gl_interop.cpp) -------------------------
// Headless: the two "VBOs" are plain device buffers of interleaved xyz floats (no GL names: getPositionVBO() is 0,
// like the reference's own no-graphics build), "mapping" hands out the device pointer, and updatePositions /
// updateVelocities are the reference's copy kernels (:7-28) as one launch of nb_interop_pack.
using GLuint = unsigned int;
class CudaGLInterop {
 public:
  CudaGLInterop() = default;
  ~CudaGLInterop() { cleanup(); }
  CudaGLInterop(const CudaGLInterop&) = delete;
  CudaGLInterop& operator=(const CudaGLInterop&) = delete;
  void initialize(std::size_t particle_count) {  // :64-96
    cleanup();
    particle_count_ = particle_count;
    if (particle_count == 0) return;
    void *p = nullptr, *v = nullptr;
    NB_CALL(nb_device_alloc(Engine::get(), particle_count * 3 * sizeof(float), &p));
    pos_ = static_cast<float*>(p);
    NB_CALL(nb_device_alloc(Engine::get(), particle_count * 3 * sizeof(float), &v));
    vel_ = static_cast<float*>(v);
  }
  void cleanup() {
    if (pos_) nb_device_free(Engine::get(), pos_);
    if (vel_) nb_device_free(Engine::get(), vel_);
    pos_ = vel_ = nullptr;
    mapped_ = vel_mapped_ = false;
    particle_count_ = 0;
  }
  float* mapPositionBuffer() { mapped_ = pos_ != nullptr; return pos_; }
  float* mapVelocityBuffer() { vel_mapped_ = vel_ != nullptr; return vel_; }
  void unmapPositionBuffer() { mapped_ = false; }
  void unmapVelocityBuffer() { vel_mapped_ = false; }
  void unmapAllBuffers() { mapped_ = vel_mapped_ = false; }
  bool isMapped() const { return mapped_; }
  bool isVelocityMapped() const { return vel_mapped_; }
  GLuint getPositionVBO() const { return 0; }
  GLuint getVelocityVBO() const { return 0; }
  std::size_t getParticleCount() const { return particle_count_; }
  void updatePositions(const ParticleData* d) {  // map, copy kernel, unmap
    checkCount(d);
    NB_CALL(nb_interop_pack(Engine::get(), as_nb(d), pos_, nullptr, Engine::stream()));
  }
  void updateVelocities(const ParticleData* d) {
    checkCount(d);
    NB_CALL(nb_interop_pack(Engine::get(), as_nb(d), nullptr, vel_, Engine::stream()));
  }
  // compares the position buffer with `count` floats of expected interleaved data (bit-exact copy)
  bool verifyDataIntegrity(const float* expected, std::size_t count) {
    if (!pos_ || !expected || count > particle_count_ * 3) return false;
    std::vector<float> got(count);
    NB_CALL(nb_device_download(Engine::get(), got.data(), pos_, count * sizeof(float), Engine::stream()));
    return std::memcmp(got.data(), expected, count * sizeof(float)) == 0;
  }

 private:
  void checkCount(const ParticleData* d) const {
    if (!d || d->count != particle_count_ || !pos_) throw ValidationException("interop buffers are not sized for these particles");
  }
  float *pos_ = nullptr, *vel_ = nullptr;
  bool mapped_ = false, vel_mapped_ = false;
  std::size_t particle_count_ = 0;
};
inline void launchCopyPositionsToVBOKernel(float* d_vbo, const ParticleData* d, int /*block_size*/) {
  NB_CALL(nb_interop_pack(Engine::get(), as_nb(d), d_vbo, nullptr, Engine::stream()));
}
inline void launchCopyVelocitiesToVBOKernel(float* d_vbo, const ParticleData* d, int /*block_size*/) {
  NB_CALL(nb_interop_pack(Engine::get(), as_nb(d), nullptr, d_vbo, Engine::stream()));
}

class ParticleSystem {
 public:
  ParticleSystem() = default;
  ~ParticleSystem() { release(); }
  ParticleSystem(const ParticleSystem&) = delete;
  ParticleSystem& operator=(const ParticleSystem&) = delete;

  void initialize(const SimulationConfig& cfg) {  // :40-100
    validateSimulationConfig(cfg);
    validateResourceRequirements(cfg.particle_count);
    config_ = cfg;
    dt_ = cfg.dt; G_ = cfg.G; softening_ = cfg.softening; force_method_ = cfg.force_method; paused_ = false;
    release();
    allocate(cfg.particle_count);
    switch (cfg.init_distribution) {
      case InitDistribution::UNIFORM: {
        UniformDistParams p; p.min_bounds = Vec3(-10, -10, -10); p.max_bounds = Vec3(10, 10, 10);
        ParticleInitializer::initUniform(h_, p);
        break;
      }
      case InitDistribution::SPHERICAL: {
        SphericalDistParams p; p.radius = 10.0f;
        ParticleInitializer::initSpherical(h_, p);
        break;
      }
      case InitDistribution::DISK: {
        DiskDistParams p; p.radius = 10.0f; p.thickness = 1.0f; p.rotation_speed = 0.5f;
        ParticleInitializer::initDisk(h_, p);
        break;
      }
    }
    start();
    time_ = 0.0f;
    resizeInterop();  // :96-99
  }
  void initializeWithDistribution(std::size_t particle_count, InitDistribution dist) {  // :102-107
    SimulationConfig config;
    config.particle_count = particle_count;
    config.init_distribution = dist;
    initialize(config);
  }
  void update(float dt) {  // :115-127
    if (!initialized_ || paused_) return;
    if (use_graph_ && force_method_ != ForceMethod::SPATIAL_HASH) updateGraphed(dt);
    else integrator_->integrate(&d_, calc_.get(), dt);
    time_ += dt;
    if (interop_) updateInteropBuffer();
  }
  // rendering interop (particle_system.cpp:320-345), headless: device buffers instead of GL buffer objects
  CudaGLInterop* getInterop() noexcept { return interop_.get(); }
  const CudaGLInterop* getInterop() const noexcept { return interop_.get(); }
  void initializeInterop() {
    if (!initialized_) throw ValidationException("ParticleSystem is not initialized");
    interop_ = std::make_unique<CudaGLInterop>();
    interop_->initialize(d_.count);
    updateInteropBuffer();
  }
  void updateInteropBuffer() {
    if (!interop_) return;
    interop_->updatePositions(&d_);
    interop_->updateVelocities(&d_);
  }
  // CUDA-graph replay of the steady-state step: the first step with a given parameter set runs
  // eagerly (sizes all scratch), the second is captured on a side stream, later ones replay.
  void setUseCudaGraph(bool on) {
    use_graph_ = on;
    if (!on) dropGraph();
  }
  std::size_t graphNodeCount() const {
    std::size_t n = 0;
    if (graph_) NB_CALL(nb_graph_node_count(graph_, &n));
    return n;
  }
  void reset() { if (initialized_) initialize(config_); }
  void pause() { paused_ = true; }
  void resume() { paused_ = false; }
  bool isPaused() const { return paused_; }
  bool isInitialized() const { return initialized_; }

  void setForceMethod(ForceMethod m) {  // :137-143
    if (m != force_method_) { force_method_ = m; config_.force_method = m; makeCalculator(); }
  }
  void setGravitationalConstant(float G) {
    if (G <= 0 || !nb_finite(G)) throw ValidationException("Gravitational constant must be positive and finite");
    G_ = G; config_.G = G;
    if (calc_) calc_->setGravitationalConstant(G);
  }
  void setSofteningParameter(float e) {
    validateSoftening(e);
    softening_ = e; config_.softening = e;
    if (calc_) calc_->setSofteningParameter(e);
  }
  void setTimeStep(float dt) { validateTimeStep(dt); dt_ = dt; config_.dt = dt; }
  void setBarnesHutTheta(float t) {
    validateTheta(t);
    config_.barnes_hut_theta = t;
    if (auto* bh = dynamic_cast<BarnesHutCalculator*>(calc_.get())) bh->setTheta(t);
  }
  void setSpatialHashCellSize(float s) {
    if (s <= 0 || !nb_finite(s)) throw ValidationException("Spatial hash cell size must be positive and finite");
    config_.spatial_hash_cell_size = s;
    if (auto* sh = dynamic_cast<SpatialHashCalculator*>(calc_.get())) sh->setCellSize(s);
  }
  void setSpatialHashCutoff(float c) {
    if (c <= 0 || !nb_finite(c)) throw ValidationException("Spatial hash cutoff must be positive and finite");
    config_.spatial_hash_cutoff = c;
    if (auto* sh = dynamic_cast<SpatialHashCalculator*>(calc_.get())) sh->setCutoffRadius(c);
  }

  std::size_t getParticleCount() const { return count_; }
  float getSimulationTime() const { return time_; }
  float getTimeStep() const { return dt_; }
  float getGravitationalConstant() const { return G_; }
  float getSofteningParameter() const { return softening_; }
  ForceMethod getForceMethod() const { return force_method_; }
  ParticleData* getDeviceData() { return &d_; }
  const ParticleData* getDeviceData() const { return &d_; }
  ForceCalculator* getForceCalculator() { return calc_.get(); }
  void copyToHost(ParticleData& h) const { ParticleDataManager::copyToHost(h, d_); }

  SimulationState getState() const {  // :213-238
    SimulationState s;
    s.particle_count = count_; s.simulation_time = time_; s.dt = dt_; s.G = G_; s.softening = softening_;
    s.force_method = force_method_;
    ParticleData h;
    ParticleDataManager::allocateHost(h, count_);
    ParticleDataManager::copyToHost(h, d_);
    s.pos_x.assign(h.pos_x, h.pos_x + count_); s.pos_y.assign(h.pos_y, h.pos_y + count_);
    s.pos_z.assign(h.pos_z, h.pos_z + count_); s.vel_x.assign(h.vel_x, h.vel_x + count_);
    s.vel_y.assign(h.vel_y, h.vel_y + count_); s.vel_z.assign(h.vel_z, h.vel_z + count_);
    s.mass.assign(h.mass, h.mass + count_);
    ParticleDataManager::freeHost(h);
    return s;
  }
  void setState(const SimulationState& s) {  // :240-292
    SimulationConfig c = config_;
    c.particle_count = s.particle_count; c.dt = s.dt; c.G = s.G; c.softening = s.softening; c.force_method = s.force_method;
    validateSimulationConfig(c);
    validateResourceRequirements(s.particle_count);
    release();
    allocate(s.particle_count);
    std::copy(s.pos_x.begin(), s.pos_x.end(), h_.pos_x); std::copy(s.pos_y.begin(), s.pos_y.end(), h_.pos_y);
    std::copy(s.pos_z.begin(), s.pos_z.end(), h_.pos_z); std::copy(s.vel_x.begin(), s.vel_x.end(), h_.vel_x);
    std::copy(s.vel_y.begin(), s.vel_y.end(), h_.vel_y); std::copy(s.vel_z.begin(), s.vel_z.end(), h_.vel_z);
    std::copy(s.mass.begin(), s.mass.end(), h_.mass);
    ParticleInitializer::zeroAccelerations(h_);
    config_ = c;
    time_ = s.simulation_time; dt_ = s.dt; G_ = s.G; softening_ = s.softening; force_method_ = s.force_method;
    start();
    paused_ = false;
    resizeInterop();  // :288-291
  }
  // src/core/particle_system.cpp:294-302: .nbody checkpoint through the Serializer
  void saveState(const std::string& filename) const { Serializer::save(filename, getState()); }
  void loadState(const std::string& filename) { setState(Serializer::load(filename)); }
  float computeKineticEnergy() const { return integrator_ ? integrator_->computeKineticEnergy(&d_) : 0.0f; }
  float computePotentialEnergy() const { return integrator_ ? integrator_->computePotentialEnergy(&d_, G_, softening_) : 0.0f; }
  float computeTotalEnergy() const { return computeKineticEnergy() + computePotentialEnergy(); }
  double computeTotalEnergyF64() const { return integrator_ ? integrator_->computeTotalEnergyF64(&d_, G_, softening_) : 0.0; }
  nb_diagnostics_t computeDiagnostics(bool with_potential = true) const {
    if (!integrator_) throw ValidationException("ParticleSystem is not initialized");
    return integrator_->computeDiagnostics(&d_, G_, softening_, with_potential);
  }
  void synchronize() const { NB_CALL(nb_synchronize(Engine::get(), nullptr)); }

 private:
  void allocate(std::size_t n) {
    count_ = n;
    ParticleDataManager::allocateDevice(d_, n);
    ParticleDataManager::allocateHost(h_, n);
    allocated_ = true;
  }
  void resizeInterop() {  // an interop that exists follows the system through re-initialisation
    if (!interop_) return;
    interop_->initialize(d_.count);
    updateInteropBuffer();
  }
  void release() {
    dropGraph();
    if (allocated_) { ParticleDataManager::freeDevice(d_); ParticleDataManager::freeHost(h_); }
    allocated_ = initialized_ = false;
  }
  struct GraphKey {
    float dt = 0, G = 0, eps = 0, theta = 0;
    int method = -1;
    std::size_t count = 0;
    const void *pos = nullptr, *calc = nullptr;
    // spelled out (not "= default") so that the headers also compile as C++17, the reference's own standard
    bool operator==(const GraphKey& o) const {
      return dt == o.dt && G == o.G && eps == o.eps && theta == o.theta && method == o.method && count == o.count &&
             pos == o.pos && calc == o.calc;
    }
  };
  GraphKey graphKey(float dt) const {
    GraphKey k;
    k.dt = dt; k.G = calc_->getGravitationalConstant(); k.eps = calc_->getSofteningParameter();
    k.theta = config_.barnes_hut_theta; k.method = static_cast<int>(force_method_); k.count = count_;
    k.pos = d_.pos_x; k.calc = calc_.get();
    return k;
  }
  void dropGraph() {
    if (graph_) nb_graph_destroy(graph_);
    graph_ = nullptr;
    graph_valid_ = warm_valid_ = false;
  }
  void updateGraphed(float dt) {
    const GraphKey key = graphKey(dt);
    // a stale graph (context scratch reallocated since capture, e.g. by a renderer frame with more fragments or a
    // larger second system) must not be replayed: drop it, run this step eagerly, capture again
    if (graph_ && nb_graph_is_stale(graph_)) dropGraph();
    if (graph_ && graph_valid_ && key == graph_key_) {
      NB_CALL(nb_graph_launch(graph_, Engine::stream()));
      return;
    }
    if (!(warm_valid_ && key == warm_key_)) {
      dropGraph();
      integrator_->integrate(&d_, calc_.get(), dt);
      warm_key_ = key;
      warm_valid_ = true;
      return;
    }
    if (!graph_stream_) NB_CALL(nb_stream_create(Engine::get(), &graph_stream_));
    nb_stream outer = Engine::stream();
    Engine::stream() = graph_stream_;
    int rc = nb_graph_begin_capture(Engine::get(), graph_stream_);
    if (rc == 0) {
      try {
        integrator_->integrate(&d_, calc_.get(), dt);
      } catch (...) {
        nb_graph* dead = nullptr;
        nb_graph_end_capture(Engine::get(), graph_stream_, &dead);
        nb_graph_destroy(dead);
        Engine::stream() = outer;
        throw;
      }
      rc = nb_graph_end_capture(Engine::get(), graph_stream_, &graph_);
    }
    Engine::stream() = outer;
    NB_CALL(rc);
    graph_key_ = key;
    graph_valid_ = true;
    NB_CALL(nb_graph_launch(graph_, Engine::stream()));
  }
  void makeCalculator() {  // :109-113
    calc_ = createForceCalculator(force_method_, config_);
    calc_->setGravitationalConstant(G_);
    calc_->setSofteningParameter(softening_);
  }
  void start() {
    ParticleDataManager::copyToDevice(d_, h_);
    makeCalculator();
    integrator_ = std::make_unique<Integrator>(config_.cuda_block_size);
    calc_->computeForces(&d_);  // a(0), :91 / :283
    initialized_ = true;
  }

  std::size_t count_ = 0;
  float dt_ = 0.001f, G_ = 1.0f, softening_ = 0.01f, time_ = 0.0f;
  ForceMethod force_method_ = ForceMethod::DIRECT_N2;
  bool paused_ = false, initialized_ = false, allocated_ = false;
  std::unique_ptr<CudaGLInterop> interop_;
  bool use_graph_ = false, graph_valid_ = false, warm_valid_ = false;
  nb_graph* graph_ = nullptr;
  nb_stream graph_stream_ = nullptr;
  GraphKey graph_key_, warm_key_;
  SimulationConfig config_;
  ParticleData d_, h_;
  std::unique_ptr<ForceCalculator> calc_;
  mutable std::unique_ptr<Integrator> integrator_;
};

// src/cuda/force_direct.cu:109-117 (the reference's only CPU force code; part of its public header)
inline Vec3 computeGravitationalForceCPU(const Vec3& p1, const Vec3& p2, float /*m1*/, float m2, float G, float eps) {
  const Vec3 r = p2 - p1;
  const float d2 = r.length2() + eps * eps;
  const float inv = 1.0f / std::sqrt(d2);
  return r * (G * m2 * inv * inv * inv);
}

}  // namespace nbody
