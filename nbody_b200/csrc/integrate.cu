// integrate.cu -- stand-alone Velocity-Verlet kernels and the energy diagnostics.
//
// Replaces storeAccelerationsKernel / updatePositionsKernel / updateVelocitiesKernel and the
// KE / PE kernels of the reference (src/cuda/integrator.cu:11-119).  These are the *semantic
// definition* used when Integrator::integrate runs with a foreign ForceCalculator; with our own
// direct calculator the fused path in direct.cu does the same arithmetic in two fewer launches.
// All three are HBM-bound: 24 / 48 / 48 bytes per body.
#include "common.cuh"

namespace nb {

__global__ void __launch_bounds__(256) store_acc_kernel(nb_particles d, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  d.acc_old_x[i] = d.acc_x[i];
  d.acc_old_y[i] = d.acc_y[i];
  d.acc_old_z[i] = d.acc_z[i];
}

// src/cuda/integrator.cu:11-21 with the reference's rounding sequence made explicit
__global__ void __launch_bounds__(256) drift_kernel(nb_particles d, size_t n, float dt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float dt2_half = __fmul_rn(__fmul_rn(0.5f, dt), dt);
  d.pos_x[i] = __fadd_rn(d.pos_x[i], __fadd_rn(__fmul_rn(d.vel_x[i], dt), __fmul_rn(d.acc_x[i], dt2_half)));
  d.pos_y[i] = __fadd_rn(d.pos_y[i], __fadd_rn(__fmul_rn(d.vel_y[i], dt), __fmul_rn(d.acc_y[i], dt2_half)));
  d.pos_z[i] = __fadd_rn(d.pos_z[i], __fadd_rn(__fmul_rn(d.vel_z[i], dt), __fmul_rn(d.acc_z[i], dt2_half)));
}

// src/cuda/integrator.cu:24-36
__global__ void __launch_bounds__(256) kick_kernel(nb_particles d, size_t n, float dt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float dt_half = __fmul_rn(0.5f, dt);
  d.vel_x[i] = __fadd_rn(d.vel_x[i], __fmul_rn(__fadd_rn(d.acc_old_x[i], d.acc_x[i]), dt_half));
  d.vel_y[i] = __fadd_rn(d.vel_y[i], __fmul_rn(__fadd_rn(d.acc_old_y[i], d.acc_y[i]), dt_half));
  d.vel_z[i] = __fadd_rn(d.vel_z[i], __fmul_rn(__fadd_rn(d.acc_old_z[i], d.acc_z[i]), dt_half));
}

int launch_store_acc(nb_context* ctx, const nb_particles* d, cudaStream_t st) {
  const size_t n = d->count;
  if (!n) return NB_OK;
  store_acc_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(*d, n);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}
int launch_drift(nb_context* ctx, const nb_particles* d, float dt, cudaStream_t st) {
  const size_t n = d->count;
  if (!n) return NB_OK;
  drift_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(*d, n, dt);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}
int launch_kick(nb_context* ctx, const nb_particles* d, float dt, cudaStream_t st) {
  const size_t n = d->count;
  if (!n) return NB_OK;
  kick_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(*d, n, dt);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

// ---------------------------------------------------------------------------------------------
// energy: fp64 accumulation, deterministic (per-CTA partials summed on the host in index order)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double block_sum(double v, double* sh) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double t = 0.0;
  if (w == 0) {
    t = (lane < (int)(blockDim.x >> 5)) ? sh[lane] : 0.0;
    t = warp_sum(t);
  }
  return t;  // valid in thread 0
}

// src/cuda/integrator.cu:51-78: ke_i = 0.5*m*(vx^2+vy^2+vz^2)
__global__ void __launch_bounds__(256) kinetic_kernel(nb_particles d, size_t n, double* __restrict__ out) {
  __shared__ double sh[8];
  double acc = 0.0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double vx = d.vel_x[i], vy = d.vel_y[i], vz = d.vel_z[i];
    acc += 0.5 * (double)d.mass[i] * (vx * vx + vy * vy + vz * vz);
  }
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) out[blockIdx.x] = t;
}

// src/cuda/integrator.cu:81-119 re-tiled: CTA (bi, bj) handles target tile bi against source
// chunk bj; every ordered pair i != j is visited once and the result halved (= sum over j > i).
// Pair terms m_j / sqrt(r^2+eps^2) are fp32 (rsqrt), sums are fp64.
constexpr int PE_T = 256;
constexpr int PE_JT = 256;
__global__ void __launch_bounds__(PE_T) potential_kernel(nb_particles d, size_t n, float eps2, int jchunk_tiles,
                                                         double* __restrict__ out) {
  __shared__ float4 tile[PE_JT];
  __shared__ double sh[8];
  const size_t i = (size_t)blockIdx.x * PE_T + threadIdx.x;
  const bool live = i < n;
  const float xi = live ? d.pos_x[i] : 0.f, yi = live ? d.pos_y[i] : 0.f, zi = live ? d.pos_z[i] : 0.f;
  const size_t j_begin = (size_t)blockIdx.y * jchunk_tiles * PE_JT;
  size_t j_end = j_begin + (size_t)jchunk_tiles * PE_JT;
  if (j_end > n) j_end = n;
  double acc = 0.0;
  for (size_t jb = j_begin; jb < j_end; jb += PE_JT) {
    const size_t j = jb + threadIdx.x;
    tile[threadIdx.x] = (j < n) ? make_float4(d.pos_x[j], d.pos_y[j], d.pos_z[j], d.mass[j]) : make_float4(0, 0, 0, 0);
    __syncthreads();
    float s = 0.f;
#pragma unroll 8
    for (int k = 0; k < PE_JT; k++) {
      const float4 p = tile[k];
      const float dx = p.x - xi, dy = p.y - yi, dz = p.z - zi;
      const float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps2)));
      float t = p.w * rsq(r2);
      t = (jb + k == i || !(r2 > 0.f)) ? 0.f : t;  // j == i excluded (integrator.cu:97: j starts at i+1)
      s += t;
    }
    acc += (double)s;
    __syncthreads();
  }
  acc = live ? acc * (double)d.mass[i] : 0.0;
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) out[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = t;
}

static int sum_partials(nb_context* ctx, size_t count, double* result, cudaStream_t st) {
  if (!ctx->h_red) NB_CUDA(cudaMallocHost(&ctx->h_red, sizeof(double) * 65536));
  double total = 0.0;
  const double* dev = ctx->red.as<double>();
  for (size_t off = 0; off < count; off += 65536) {
    const size_t c = (count - off < 65536) ? count - off : 65536;
    NB_CUDA(cudaMemcpyAsync(ctx->h_red, dev + off, sizeof(double) * c, cudaMemcpyDeviceToHost, st));
    NB_CUDA(cudaStreamSynchronize(st));
    for (size_t k = 0; k < c; k++) total += ctx->h_red[k];
  }
  *result = total;
  return NB_OK;
}

int kinetic_energy(nb_context* ctx, const nb_particles* d, double* ke, cudaStream_t st) {
  const size_t n = d->count;
  *ke = 0.0;
  if (!n) return NB_OK;
  size_t grid = (n + 255) / 256;
  const size_t cap = (size_t)ctx->sm_count * 8;
  if (grid > cap) grid = cap;
  if (int rc = ctx->red.ensure(sizeof(double) * grid)) return rc;
  kinetic_kernel<<<(unsigned)grid, 256, 0, st>>>(*d, n, ctx->red.as<double>());
  NB_LAUNCH_CHECK(ctx);
  return sum_partials(ctx, grid, ke, st);
}

int potential_energy(nb_context* ctx, const nb_particles* d, float G, float eps, double* pe, cudaStream_t st) {
  const size_t n = d->count;
  *pe = 0.0;
  if (!n) return NB_OK;
  const size_t gi = (n + PE_T - 1) / PE_T;
  const size_t jtiles = (n + PE_JT - 1) / PE_JT;
  // enough j-splits to fill the machine when there are few target tiles
  size_t splits = ((size_t)ctx->sm_count * 8 + gi - 1) / gi;
  if (splits > jtiles) splits = jtiles;
  if (splits > 65535) splits = 65535;
  if (splits < 1) splits = 1;
  const int jchunk_tiles = (int)((jtiles + splits - 1) / splits);
  const size_t gy = (jtiles + jchunk_tiles - 1) / jchunk_tiles;
  if (int rc = ctx->red.ensure(sizeof(double) * gi * gy)) return rc;
  dim3 grid((unsigned)gi, (unsigned)gy);
  potential_kernel<<<grid, PE_T, 0, st>>>(*d, n, eps * eps, jchunk_tiles, ctx->red.as<double>());
  NB_LAUNCH_CHECK(ctx);
  double s = 0.0;
  if (int rc = sum_partials(ctx, gi * gy, &s, st)) return rc;
  *pe = -0.5 * (double)G * s;  // pe -= G*m_i*m_j/r over j>i (integrator.cu:97-102)
  return NB_OK;
}

// ---------------------------------------------------------------------------------------------
// conserved-quantity diagnostics: one pass over the bodies, fp64 sums of
//   M, M*x_com (3), P = sum m v (3), L = sum m x cross v (3), KE  -> NB_DIAG_MOMENTS values per CTA,
// summed on the host in CTA order (deterministic).  The reference has no equivalent; its tests check
// these quantities on the host (tests/test_integrator.cpp:51-84 for KE).
// ---------------------------------------------------------------------------------------------
constexpr int NB_DIAG_MOMENTS = 11;
__global__ void __launch_bounds__(256) moments_kernel(nb_particles d, size_t n, double* __restrict__ out) {
  __shared__ double sh[8];
  double a[NB_DIAG_MOMENTS];
#pragma unroll
  for (int q = 0; q < NB_DIAG_MOMENTS; q++) a[q] = 0.0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double m = d.mass[i];
    const double x = d.pos_x[i], y = d.pos_y[i], z = d.pos_z[i];
    const double vx = d.vel_x[i], vy = d.vel_y[i], vz = d.vel_z[i];
    a[0] += m;
    a[1] += m * x; a[2] += m * y; a[3] += m * z;
    a[4] += m * vx; a[5] += m * vy; a[6] += m * vz;
    a[7] += m * (y * vz - z * vy); a[8] += m * (z * vx - x * vz); a[9] += m * (x * vy - y * vx);
    a[10] += 0.5 * m * (vx * vx + vy * vy + vz * vz);
  }
#pragma unroll
  for (int q = 0; q < NB_DIAG_MOMENTS; q++) {
    const double t = block_sum(a[q], sh);
    if (threadIdx.x == 0) out[(size_t)blockIdx.x * NB_DIAG_MOMENTS + q] = t;
    __syncthreads();  // sh is reused by the next moment
  }
}

int moments(nb_context* ctx, const nb_particles* d, double* out11, cudaStream_t st) {
  const size_t n = d->count;
  for (int q = 0; q < NB_DIAG_MOMENTS; q++) out11[q] = 0.0;
  if (!n) return NB_OK;
  size_t grid = (n + 255) / 256;
  const size_t cap = (size_t)ctx->sm_count * 8;
  if (grid > cap) grid = cap;
  if (int rc = ctx->red.ensure(sizeof(double) * grid * NB_DIAG_MOMENTS)) return rc;
  moments_kernel<<<(unsigned)grid, 256, 0, st>>>(*d, n, ctx->red.as<double>());
  NB_LAUNCH_CHECK(ctx);
  if (!ctx->h_red) NB_CUDA(cudaMallocHost(&ctx->h_red, sizeof(double) * 65536));
  NB_CUDA(cudaMemcpyAsync(ctx->h_red, ctx->red.as<double>(), sizeof(double) * grid * NB_DIAG_MOMENTS,
                          cudaMemcpyDeviceToHost, st));  // grid <= 8 * SMs: fits the 65536-double staging buffer
  NB_CUDA(cudaStreamSynchronize(st));
  for (size_t b = 0; b < grid; b++)
    for (int q = 0; q < NB_DIAG_MOMENTS; q++) out11[q] += ctx->h_red[b * NB_DIAG_MOMENTS + q];
  return NB_OK;
}

}  // namespace nb
