// common.cuh -- shared device helpers (packed f32x2 math, mbarrier/TMA bulk copy) and the
// context definition.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <utility>
#include <vector>

#include "../../include/nbody_b200.h"

namespace nb {

// ---- error plumbing -----------------------------------------------------------------------
void set_error(const char* fmt, ...);
uint64_t scratch_generation();  // counts reallocations of live scratch buffers (stale-graph detection)
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define NB_CUDA(call)                                                       \
  do {                                                                      \
    cudaError_t e__ = (call);                                               \
    if (e__ != cudaSuccess) return nb::cuda_fail(e__, #call, __FILE__, __LINE__); \
  } while (0)
// launch check: cudaGetLastError only (async), like CUDA_CHECK_KERNEL in Release builds
// (include/nbody/error_handling.hpp:116-123).
#define NB_LAUNCH_CHECK(ctx)                                                \
  do {                                                                      \
    (ctx)->launches++;                                                      \
    cudaError_t e__ = cudaGetLastError();                                   \
    if (e__ != cudaSuccess) return nb::cuda_fail(e__, "kernel launch", __FILE__, __LINE__); \
  } while (0)

// ---- packed fp32x2 (Blackwell FFMA2 / FADD2 / FMUL2) -----------------------------------------
typedef unsigned long long f2;  // two fp32 in one 64-bit register pair {lo, hi}

__device__ __forceinline__ f2 pk(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk(f2 v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
  f2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
  f2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f2 sub2(f2 a, f2 b) {
  f2 d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
  f2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// single MUFU.RSQ, no denormal fix-up code (the reference's rsqrtf adds an FSETP/FMUL guard,
// SURVEY.md A.4).  Max relative error 2^-22.
__device__ __forceinline__ float rsq(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// Softening below which the force kernels use their guarded instantiation (EPS0): the un-guarded pair body relies
// on w * d == w * 0 == 0 for the self pair, which needs w = G m (eps^2)^-3/2 finite -- eps^2 = 1e-12 keeps it below
// 1e18 * G m; under that, a pair whose r^2 + eps^2 does not exceed eps^2 (the body itself, an exact coincidence)
// contributes 0 by a select, as at eps = 0 (the reference skips j == i by index, force_direct.cu:60).
constexpr float NB_EPS2_GUARD = 1e-12f;

// ---- mbarrier + 1-D TMA bulk copy (cp.async.bulk, SASS UBLKCP) --------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-B aligned.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- device buffer that only grows ---------------------------------------------------------
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes);
  void release();
  template <typename T>
  T* as() const { return static_cast<T*>(p); }
};

}  // namespace nb

namespace nb {
// RAII event pair around the dominant kernel of a path when profiling is on.
struct ProfScope {
  nb_context* ctx;
  cudaStream_t st;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  ProfScope(nb_context* c, cudaStream_t s);
  ~ProfScope();
};
// Named stage timer (nb_profile_stages_enable): CUDA events on the launching stream around one stage of a
// path (bbox, morton, sort, ...); bench.py's `paths` block reads the per-stage sums.  Off: no events, no cost.
struct StageScope {
  nb_context* ctx;
  cudaStream_t st;
  const char* name;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  StageScope(nb_context* c, cudaStream_t s, const char* n);
  ~StageScope();
};
struct StageRec {
  const char* name;
  cudaEvent_t e0, e1;
};
}  // namespace nb
#define NB_STAGE_CAT2_(a, b) a##b
#define NB_STAGE_CAT_(a, b) NB_STAGE_CAT2_(a, b)
#define NB_STAGE(ctx, st, name) nb::StageScope NB_STAGE_CAT_(nb_stage_scope_, __LINE__)(ctx, st, name)

// One pass of the direct kernel (see direct.cu): n_itiles target tiles x UJ source tiles, cut into
// `chunks` = n_itiles * C chunks that G CTAs claim dynamically; chunk c writes partial slot slot_base + c.
struct nb_pass {
  long long U;       // work units = n_itiles * UJ
  long long chunks;  // n_itiles * C
  int UJ;            // source tiles per target tile
  int C;             // chunks per target tile
  int G;             // gridDim.x of that launch
  int n_itiles;
  size_t slot_base;  // first partial slot of this pass
};

struct nb_context {
  int device = 0;
  int sm_count = 0;
  int cc_major = 0, cc_minor = 0;
  uint64_t launches = 0;
  bool prof = false;  // nb_profile_enable
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_ev;
  std::vector<cudaEvent_t> prof_pool;
  bool prof_stages = false;  // nb_profile_stages_enable
  std::vector<nb::StageRec> stage_ev;
  uint64_t capture_generation = 0;  // scratch generation (c_api.cu) when the running stream capture began
  // direct N^2
  nb::DevBuf src;       // packed float4 (x,y,z,G*m), tile padded
  nb::DevBuf partial;   // per-chunk partial accelerations
  nb::DevBuf direct_work;  // chunk counters of the two passes
  nb_pass pass[2];
  int n_pass = 0;
  int direct_occ[4] = {0, 0, 0, 0};  // resident CTAs/SM per kernel config (lazy)
  // energy
  nb::DevBuf red;       // per-CTA fp64 partials
  double* h_red = nullptr;  // pinned
  // host-buffer entry points
  nb_particles stage{};  // device mirror for nb_host_*
  size_t stage_cap = 0;
  cudaStream_t own_stream = nullptr;
  // Barnes-Hut
  struct BH* bh = nullptr;
  // spatial hash
  struct SH* sh = nullptr;
  // splat renderer
  struct RenderState* render = nullptr;
  // radix sort scratch
  nb::DevBuf sort_keys2, sort_vals2, sort_hist;
  int* h_sort_err = nullptr;  // host-mapped: a look-back wait gave up (sort.cu)
  int* d_sort_err = nullptr;
  // bbox scratch (6 ordered ints) + pinned readback
  nb::DevBuf bbox_dev;
  float* h_bbox = nullptr;
  // multi-GPU
  size_t mg_n_total = 0, mg_begin = 0, mg_count = 0, mg_padded = 0;
  int mg_rank = 0, mg_nranks = 1;
  struct PeerState* peer = nullptr;     // CUDA-IPC exchange blocks (mgpu.cu): direct N^2 sources
  struct PeerState* peer_bh = nullptr;  // ... and sharded Barnes-Hut accelerations
  size_t bhs_n = 0;
  int bhs_rank = 0, bhs_nranks = 1;
};
