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

// src/cuda/integrator.cu:81-119 re-tiled.  The reference visits every pair j > i from one thread per i (global
// loads, triangular imbalance, fp32 sum).  Here the pair set is cut into 512 x 512 tile pairs of the upper
// triangle (source tile >= target tile): CTA (bi, by) takes target tile bi against up to `chunk` source tiles
// starting `by * chunk` tiles right of the diagonal.  Each thread holds two targets as one packed fp32x2 pair,
// the source is a warp-broadcast LDS.128, and the pair term m_j / sqrt(r^2 + eps^2) costs 3 FADD2 + 3 FFMA2 +
// 2 MUFU.RSQ + 1 FMUL2 + 1 FADD2 per two pairs.  Off-diagonal tiles need no index test (and with eps > 0 no
// r = 0 test); the diagonal tile keeps j > i.  Pair terms are fp32, a tile's 512 terms are summed in four
// interleaved fp32 accumulators, everything above that in fp64.
constexpr int PE_T = 256;
constexpr int PE_TILE = 2 * PE_T;
template <bool EPS0>
__global__ void __launch_bounds__(PE_T) potential_kernel(nb_particles d, size_t n, float eps2, int n_tiles, int chunk,
                                                         double* __restrict__ out) {
  __shared__ __align__(16) float4 tile[PE_TILE];
  __shared__ double sh[8];
  const int tid = threadIdx.x;
  const int bi = blockIdx.x;
  const int jt0 = bi + (int)blockIdx.y * chunk;
  double* slot = out + (size_t)blockIdx.y * gridDim.x + blockIdx.x;
  if (jt0 >= n_tiles) {  // nothing right of the diagonal left for this CTA
    if (tid == 0) *slot = 0.0;
    return;
  }
  const int jt1 = min(jt0 + chunk, n_tiles);
  const size_t i0 = (size_t)bi * PE_TILE + tid, i1 = i0 + PE_T;
  const bool live0 = i0 < n, live1 = i1 < n;
  const f2 xi = pk(live0 ? d.pos_x[i0] : 0.f, live1 ? d.pos_x[i1] : 0.f);
  const f2 yi = pk(live0 ? d.pos_y[i0] : 0.f, live1 ? d.pos_y[i1] : 0.f);
  const f2 zi = pk(live0 ? d.pos_z[i0] : 0.f, live1 ? d.pos_z[i1] : 0.f);
  const f2 e2 = pk(eps2, eps2);
  const f2 zero2 = pk(0.f, 0.f);
  double acc0 = 0.0, acc1 = 0.0;
  for (int jt = jt0; jt < jt1; jt++) {
    const size_t jb = (size_t)jt * PE_TILE;
#pragma unroll
    for (int h = 0; h < 2; h++) {  // padding sources carry zero mass
      const size_t j = jb + tid + h * PE_T;
      tile[tid + h * PE_T] = (j < n) ? make_float4(d.pos_x[j], d.pos_y[j], d.pos_z[j], d.mass[j]) : make_float4(0, 0, 0, 0);
    }
    __syncthreads();
    f2 s[4] = {zero2, zero2, zero2, zero2};
    if (EPS0 || jt == bi) {
      // diagonal tile: keep j > i only; eps == 0: a coincident pair (r = 0) contributes nothing instead of inf
      const bool diag = (jt == bi);
      const int k_lo0 = diag ? tid + 1 : 0, k_lo1 = diag ? tid + PE_T + 1 : 0;  // first source index kept per target
#pragma unroll 4
      for (int k = 0; k < PE_TILE; k++) {
        const float4 p = tile[k];
        const f2 dx = sub2(pk(p.x, p.x), xi), dy = sub2(pk(p.y, p.y), yi), dz = sub2(pk(p.z, p.z), zi);
        const f2 r2 = fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, e2)));
        float r0, r1;
        upk(r2, r0, r1);
        float t0 = p.w * rsq(r0), t1 = p.w * rsq(r1);
        t0 = (k >= k_lo0 && r0 > 0.f) ? t0 : 0.f;
        t1 = (k >= k_lo1 && r1 > 0.f) ? t1 : 0.f;
        s[k & 3] = add2(s[k & 3], pk(t0, t1));
      }
    } else {
#pragma unroll 16
      for (int k = 0; k < PE_TILE; k++) {
        const float4 p = tile[k];
        const f2 dx = sub2(pk(p.x, p.x), xi), dy = sub2(pk(p.y, p.y), yi), dz = sub2(pk(p.z, p.z), zi);
        const f2 r2 = fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, e2)));
        float r0, r1;
        upk(r2, r0, r1);
        s[k & 3] = add2(s[k & 3], mul2(pk(p.w, p.w), pk(rsq(r0), rsq(r1))));
      }
    }
    float a0, a1, b0, b1;
    upk(add2(s[0], s[1]), a0, a1);
    upk(add2(s[2], s[3]), b0, b1);
    acc0 += (double)a0 + (double)b0;
    acc1 += (double)a1 + (double)b1;
    __syncthreads();
  }
  const double mine = (live0 ? acc0 * (double)d.mass[i0] : 0.0) + (live1 ? acc1 * (double)d.mass[i1] : 0.0);
  const double t = block_sum(mine, sh);
  if (tid == 0) *slot = t;
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
  if (n < 2) return NB_OK;
  const long long nt = (long long)((n + PE_TILE - 1) / PE_TILE);
  // about 128 CTAs' worth of tile pairs per SM, so that the triangle's ragged rows still balance
  const long long pairs = nt * (nt + 1) / 2;
  long long chunk = pairs / ((long long)ctx->sm_count * 128);
  if (chunk < 1) chunk = 1;
  long long gy = (nt + chunk - 1) / chunk;
  if (gy > 65535) {
    gy = 65535;
    chunk = (nt + gy - 1) / gy;
    gy = (nt + chunk - 1) / chunk;
  }
  if (int rc = ctx->red.ensure(sizeof(double) * (size_t)nt * (size_t)gy)) return rc;
  dim3 grid((unsigned)nt, (unsigned)gy);
  const float eps2 = eps * eps;
  if (eps2 >= 1.17549435e-38f) potential_kernel<false><<<grid, PE_T, 0, st>>>(*d, n, eps2, (int)nt, (int)chunk, ctx->red.as<double>());
  else potential_kernel<true><<<grid, PE_T, 0, st>>>(*d, n, eps2, (int)nt, (int)chunk, ctx->red.as<double>());
  NB_LAUNCH_CHECK(ctx);
  double s = 0.0;
  if (int rc = sum_partials(ctx, (size_t)nt * (size_t)gy, &s, st)) return rc;
  *pe = -(double)G * s;  // pe -= G*m_i*m_j/r over j > i (integrator.cu:97-102)
  return NB_OK;
}

// Potential energy of a target-sharded run ("energy diagnostic: same sharding, all-reduce of two
// doubles"): this rank's targets -- n_tgt bodies at tgt_off in the replicated packed source array {x,y,z,G*m} --
// against ALL sources.  Every ordered pair with i local is visited (no triangle: the other half of a cross-rank
// pair belongs to the other rank), so the all-reduced sum is halved by the caller.  Same pair arithmetic as
// potential_kernel; the self pair is excluded by global index in the tiles that overlap the target tile.
template <bool EPS0>
__global__ void __launch_bounds__(PE_T) potential_rect_kernel(const float4* __restrict__ src, size_t n_src,
                                                              size_t tgt_off, size_t n_tgt,
                                                              const float* __restrict__ mass_local, float eps2,
                                                              int n_src_tiles, int chunk, double* __restrict__ out) {
  __shared__ __align__(16) float4 tile[PE_TILE];
  __shared__ double sh[8];
  const int tid = threadIdx.x;
  const int jt0 = (int)blockIdx.y * chunk, jt1 = min(jt0 + chunk, n_src_tiles);
  const size_t l0 = (size_t)blockIdx.x * PE_TILE + tid, l1 = l0 + PE_T;  // local target indices
  const bool live0 = l0 < n_tgt, live1 = l1 < n_tgt;
  const float4 t0 = live0 ? src[tgt_off + l0] : make_float4(0, 0, 0, 0);
  const float4 t1 = live1 ? src[tgt_off + l1] : make_float4(0, 0, 0, 0);
  const f2 xi = pk(t0.x, t1.x), yi = pk(t0.y, t1.y), zi = pk(t0.z, t1.z);
  const f2 e2 = pk(eps2, eps2), zero2 = pk(0.f, 0.f);
  const size_t g_lo = tgt_off + (size_t)blockIdx.x * PE_TILE, g_hi = g_lo + PE_TILE;  // this CTA's targets, global
  double acc0 = 0.0, acc1 = 0.0;
  for (int jt = jt0; jt < jt1; jt++) {
    const size_t jb = (size_t)jt * PE_TILE;
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const size_t j = jb + tid + h * PE_T;
      tile[tid + h * PE_T] = (j < n_src) ? src[j] : make_float4(0, 0, 0, 0);
    }
    __syncthreads();
    f2 s[4] = {zero2, zero2, zero2, zero2};
    if (EPS0 || (jb < g_hi && jb + PE_TILE > g_lo)) {  // tile holds some of this CTA's own bodies
      const long long self0 = (long long)(tgt_off + l0) - (long long)jb, self1 = (long long)(tgt_off + l1) - (long long)jb;
#pragma unroll 4
      for (int k = 0; k < PE_TILE; k++) {
        const float4 p = tile[k];
        const f2 dx = sub2(pk(p.x, p.x), xi), dy = sub2(pk(p.y, p.y), yi), dz = sub2(pk(p.z, p.z), zi);
        const f2 r2 = fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, e2)));
        float r0, r1;
        upk(r2, r0, r1);
        float u0 = p.w * rsq(r0), u1 = p.w * rsq(r1);
        u0 = (k != self0 && r0 > 0.f) ? u0 : 0.f;
        u1 = (k != self1 && r1 > 0.f) ? u1 : 0.f;
        s[k & 3] = add2(s[k & 3], pk(u0, u1));
      }
    } else {
#pragma unroll 16
      for (int k = 0; k < PE_TILE; k++) {
        const float4 p = tile[k];
        const f2 dx = sub2(pk(p.x, p.x), xi), dy = sub2(pk(p.y, p.y), yi), dz = sub2(pk(p.z, p.z), zi);
        const f2 r2 = fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, e2)));
        float r0, r1;
        upk(r2, r0, r1);
        s[k & 3] = add2(s[k & 3], mul2(pk(p.w, p.w), pk(rsq(r0), rsq(r1))));
      }
    }
    float a0, a1, b0, b1;
    upk(add2(s[0], s[1]), a0, a1);
    upk(add2(s[2], s[3]), b0, b1);
    acc0 += (double)a0 + (double)b0;
    acc1 += (double)a1 + (double)b1;
    __syncthreads();
  }
  const double mine = (live0 ? acc0 * (double)mass_local[l0] : 0.0) + (live1 ? acc1 * (double)mass_local[l1] : 0.0);
  const double t = block_sum(mine, sh);
  if (tid == 0) out[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = t;
}

// sum over local targets i of m_i * sum over all sources j != i of (G m_j) / sqrt(r^2 + eps^2)
int potential_partial(nb_context* ctx, const float4* src, size_t n_src, size_t tgt_off, size_t n_tgt,
                      const float* mass_local, float eps, double* partial, cudaStream_t st) {
  *partial = 0.0;
  if (!n_tgt || !n_src) return NB_OK;
  const long long gi = (long long)((n_tgt + PE_TILE - 1) / PE_TILE);
  const long long nt = (long long)((n_src + PE_TILE - 1) / PE_TILE);
  long long chunk = gi * nt / ((long long)ctx->sm_count * 128);
  if (chunk < 1) chunk = 1;
  if (chunk > nt) chunk = nt;
  long long gy = (nt + chunk - 1) / chunk;
  if (gy > 65535) {
    gy = 65535;
    chunk = (nt + gy - 1) / gy;
    gy = (nt + chunk - 1) / chunk;
  }
  if (int rc = ctx->red.ensure(sizeof(double) * (size_t)gi * (size_t)gy)) return rc;
  dim3 grid((unsigned)gi, (unsigned)gy);
  const float eps2 = eps * eps;
  if (eps2 >= 1.17549435e-38f)  // rsqrt.approx.ftz flushes a denormal eps^2 to 0
    potential_rect_kernel<false><<<grid, PE_T, 0, st>>>(src, n_src, tgt_off, n_tgt, mass_local, eps2, (int)nt, (int)chunk,
                                                        ctx->red.as<double>());
  else
    potential_rect_kernel<true><<<grid, PE_T, 0, st>>>(src, n_src, tgt_off, n_tgt, mass_local, eps2, (int)nt, (int)chunk,
                                                       ctx->red.as<double>());
  NB_LAUNCH_CHECK(ctx);
  return sum_partials(ctx, (size_t)gi * (size_t)gy, partial, st);
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
