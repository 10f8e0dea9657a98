// direct.cu -- direct N^2 force evaluation for sm_100a and the fused Velocity-Verlet step.
//
// Replaces computeForcesDirectKernel / launchDirectForceKernel (reference
// src/cuda/force_direct.cu:10-98) and, fused around it, the three integrator kernels
// (src/cuda/integrator.cu:11-48).  Not a port: the reference runs one target per thread with four
// scalar SoA shared tiles, two branches per pair and an IEEE-guarded rsqrtf.  Here:
//   * sources are packed once per evaluation into float4 {x, y, z, G*m} (16 B/body);
//   * each thread owns I targets held as I/2 packed fp32x2 register pairs and runs the pair body
//     with Blackwell FFMA2/FADD2/FMUL2 (12 packed instructions + 2 MUFU.RSQ per two pairs), the
//     source being a warp-broadcast LDS.128;
//   * source tiles are staged by 1-D TMA bulk copies (cp.async.bulk + mbarrier), double buffered;
//   * work is split "stream-K": the (target tile x source tile) unit space is cut into gridDim.x
//     equal contiguous runs, so all 148 SMs finish together for any N; every run writes fp32
//     partials, and a reduce kernel adds them in a fixed order (deterministic) and applies the kick;
//   * accumulation is two-level: plain fp32 inside a source tile, Kahan-compensated across tiles.
#include "direct.cuh"

#include <cstdlib>

namespace nb {

// ---------------------------------------------------------------------------------------------
// pack: SoA -> float4 {x,y,z,G*m}; optionally fused with store-acc + drift.
// ---------------------------------------------------------------------------------------------
template <bool DRIFT>
__global__ void __launch_bounds__(256) pack_kernel(nb_particles d, float4* __restrict__ src, size_t n, size_t n_pad,
                                                   float G, float dt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad) return;
  float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n) {
    float x = d.pos_x[i], y = d.pos_y[i], z = d.pos_z[i];
    if (DRIFT) {
      // src/cuda/integrator.cu:39-48 then :11-21, same rounding sequence:
      // dt2_half = 0.5f*dt*dt ; pos += vel*dt + acc*dt2_half (no FMA contraction)
      const float dt2_half = __fmul_rn(__fmul_rn(0.5f, dt), dt);
      const float ax = d.acc_x[i], ay = d.acc_y[i], az = d.acc_z[i];
      d.acc_old_x[i] = ax;
      d.acc_old_y[i] = ay;
      d.acc_old_z[i] = az;
      x = __fadd_rn(x, __fadd_rn(__fmul_rn(d.vel_x[i], dt), __fmul_rn(ax, dt2_half)));
      y = __fadd_rn(y, __fadd_rn(__fmul_rn(d.vel_y[i], dt), __fmul_rn(ay, dt2_half)));
      z = __fadd_rn(z, __fadd_rn(__fmul_rn(d.vel_z[i], dt), __fmul_rn(az, dt2_half)));
      d.pos_x[i] = x;
      d.pos_y[i] = y;
      d.pos_z[i] = z;
    }
    o = make_float4(x, y, z, __fmul_rn(G, d.mass[i]));  // f = G*m_j*inv3, force_direct.cu:69
  }
  src[i] = o;
}

// pack fused with the all-gather (multi-GPU peer exchange): the packed float4 of every local body is
// stored straight into slot `slot_off` of EVERY rank's replicated source buffer -- the local one and,
// through NVLink/NVSwitch peer mappings, the R-1 remote ones -- so no separate collective runs.  The
// last CTA to finish raises this rank's step flag on every peer (release at system scope); the
// consumer there is peer_flags_wait_kernel, enqueued between the local and the remote force pass.
template <bool DRIFT>
__global__ void __launch_bounds__(256) pack_push_kernel(nb_particles d, PeerBufs pb, size_t slot_off, size_t n,
                                                        size_t n_pad, float G, float dt, unsigned step,
                                                        int flag_index, unsigned* __restrict__ done_counter) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_pad) {
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n) {
      float x = d.pos_x[i], y = d.pos_y[i], z = d.pos_z[i];
      if (DRIFT) {  // same rounding sequence as pack_kernel<true> (integrator.cu:39-48, :11-21)
        const float dt2_half = __fmul_rn(__fmul_rn(0.5f, dt), dt);
        const float ax = d.acc_x[i], ay = d.acc_y[i], az = d.acc_z[i];
        d.acc_old_x[i] = ax;
        d.acc_old_y[i] = ay;
        d.acc_old_z[i] = az;
        x = __fadd_rn(x, __fadd_rn(__fmul_rn(d.vel_x[i], dt), __fmul_rn(ax, dt2_half)));
        y = __fadd_rn(y, __fadd_rn(__fmul_rn(d.vel_y[i], dt), __fmul_rn(ay, dt2_half)));
        z = __fadd_rn(z, __fadd_rn(__fmul_rn(d.vel_z[i], dt), __fmul_rn(az, dt2_half)));
        d.pos_x[i] = x;
        d.pos_y[i] = y;
        d.pos_z[i] = z;
      }
      o = make_float4(x, y, z, __fmul_rn(G, d.mass[i]));
    }
#pragma unroll
    for (int r = 0; r < NB_MAX_PEERS; r++)
      if (r < pb.nranks) pb.buf[r][slot_off + i] = o;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(done_counter, 1u);
    if (prev == gridDim.x - 1) {  // every CTA's stores are fenced: publish
      *done_counter = 0;
      __threadfence_system();
      for (int r = 0; r < pb.nranks; r++)
        if (r != pb.self)
          asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pb.flags[r] + flag_index), "r"(step) : "memory");
    }
  }
}

int launch_pack_push(nb_context* ctx, const nb_particles* d, const PeerBufs& pb, size_t slot_off, size_t n,
                     size_t n_pad, float G, float dt, bool drift, unsigned step, int flag_index, unsigned* done_counter,
                     cudaStream_t st) {
  const unsigned grid = (unsigned)((n_pad + 255) / 256);
  if (grid == 0) return NB_OK;
  if (drift) pack_push_kernel<true><<<grid, 256, 0, st>>>(*d, pb, slot_off, n, n_pad, G, dt, step, flag_index, done_counter);
  else pack_push_kernel<false><<<grid, 256, 0, st>>>(*d, pb, slot_off, n, n_pad, G, dt, step, flag_index, done_counter);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

// One thread per peer spins (bounded) until that peer has raised `step` in this rank's flag block.
// Runs between the local-source pass and the remote-source pass, so the hot force kernel itself
// carries no waiting code (an in-kernel gate cost 5 % of the step through register pressure).
__global__ void peer_flags_wait_kernel(const unsigned* __restrict__ flags, int flag_base, unsigned step, int nranks,
                                       int self, int* __restrict__ timeout_flag) {
  const int r = threadIdx.x;
  if (r >= nranks || r == self) return;
  const unsigned* f = flags + flag_base + r;
  unsigned long long t0 = 0;
  for (unsigned spins = 0;; spins++) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
    if ((int)(v - step) >= 0) break;
    if ((spins & 1023u) == 1023u) {  // a dead peer must not hang the GPU: give up after ~20 s and say so
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 20000000000ull) {
        *timeout_flag = 1;
        break;
      }
    }
  }
}

int launch_peer_wait(nb_context* ctx, const unsigned* flags, int flag_base, unsigned step, int nranks, int self,
                     int* timeout_flag, cudaStream_t st) {
  peer_flags_wait_kernel<<<1, 32, 0, st>>>(flags, flag_base, step, nranks, self, timeout_flag);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

// ---------------------------------------------------------------------------------------------
// force kernel
// ---------------------------------------------------------------------------------------------
template <int T, int I, int JT, bool EPS0>
__global__ void __launch_bounds__(T) direct_force_kernel(const float4* __restrict__ src,
                                                          const float4* __restrict__ tgt,
                                                          float* __restrict__ partial, long long U, int UJ,
                                                          int ja0, int na, int jb0, float eps2,
                                                          unsigned long long slot_base) {
  constexpr int P = I / 2;
  constexpr int IT = T * I;
  constexpr uint32_t TILE_BYTES = JT * sizeof(float4);
  static_assert(I % 2 == 0, "targets per thread must be even (packed fp32x2)");
  __shared__ __align__(128) float4 tile[2][JT];
  __shared__ __align__(8) uint64_t full[2];

  const int tid = threadIdx.x;
  const long long u0 = (U * (long long)blockIdx.x) / gridDim.x;
  const long long u1 = (U * (long long)(blockIdx.x + 1)) / gridDim.x;
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
  }
  __syncthreads();
#ifdef NB_DIRECT_PAD  // experiment: shift the hot loop's code address by NB_DIRECT_PAD instructions
#pragma unroll
  for (int k = 0; k < NB_DIRECT_PAD; k++) asm volatile("nanosleep.u32 0;");
#endif
  if (u0 >= u1) return;

  int itile = (int)(u0 / UJ);
  int jt = (int)(u0 - (long long)itile * UJ);
  auto jtile_ptr = [&](int j) -> const float4* {
    const int g = (j < na) ? (ja0 + j) : (jb0 + (j - na));
    return src + (size_t)g * JT;
  };
  if (tid == 0) {
    mbar_expect_tx(&full[0], TILE_BYTES);
    tma_load_1d(&tile[0][0], jtile_ptr(jt), TILE_BYTES, &full[0]);
  }

  f2 xi[P], yi[P], zi[P];
  f2 sx[P], sy[P], sz[P];  // Kahan sums across tiles
  f2 cx[P], cy[P], cz[P];  // Kahan compensation
  const f2 zero2 = pk(0.f, 0.f);
  const f2 e2 = pk(eps2, eps2);

  auto load_targets = [&](int t) {
    const float4* base = tgt + (size_t)t * IT + tid;
#pragma unroll
    for (int p = 0; p < P; p++) {
      const float4 a = __ldg(base + (2 * p) * T), b = __ldg(base + (2 * p + 1) * T);
      xi[p] = pk(a.x, b.x);
      yi[p] = pk(a.y, b.y);
      zi[p] = pk(a.z, b.z);
      sx[p] = sy[p] = sz[p] = zero2;
      cx[p] = cy[p] = cz[p] = zero2;
    }
  };
  load_targets(itile);

  const int nunits = (int)(u1 - u0);
  for (int n = 0; n < nunits; n++) {
    const int st = n & 1;
    if (n + 1 < nunits && tid == 0) {
      const int jn = (jt + 1 == UJ) ? 0 : jt + 1;
      mbar_expect_tx(&full[st ^ 1], TILE_BYTES);
      tma_load_1d(&tile[st ^ 1][0], jtile_ptr(jn), TILE_BYTES, &full[st ^ 1]);
    }
    mbar_wait(&full[st], (uint32_t)((n >> 1) & 1));

    f2 ax[P], ay[P], az[P];
#pragma unroll
    for (int p = 0; p < P; p++) ax[p] = ay[p] = az[p] = zero2;
    const float4* tj = &tile[st][0];
#pragma unroll 8
    for (int j = 0; j < JT; j++) {
      const float4 s = tj[j];  // warp-broadcast LDS.128
      const f2 px = pk(s.x, s.x), py = pk(s.y, s.y), pz = pk(s.z, s.z), pm = pk(s.w, s.w);
#pragma unroll
      for (int p = 0; p < P; p++) {
        const f2 dx = sub2(px, xi[p]), dy = sub2(py, yi[p]), dz = sub2(pz, zi[p]);
        f2 r2 = fma2(dx, dx, e2);
        r2 = fma2(dy, dy, r2);
        r2 = fma2(dz, dz, r2);
        float r0, r1;
        upk(r2, r0, r1);
        float i0 = rsq(r0), i1 = rsq(r1);
        if (EPS0) {  // eps == 0: the self pair (and exact coincidences) must contribute 0, not NaN
          i0 = (r0 > 0.f) ? i0 : 0.f;
          i1 = (r1 > 0.f) ? i1 : 0.f;
        }
        const f2 ri = pk(i0, i1);
        const f2 w = mul2(mul2(ri, ri), mul2(pm, ri));  // (G m_j) / r^3
        ax[p] = fma2(w, dx, ax[p]);
        ay[p] = fma2(w, dy, ay[p]);
        az[p] = fma2(w, dz, az[p]);
      }
    }
    // fold the tile's sums into the compensated running sums
#pragma unroll
    for (int p = 0; p < P; p++) {
      f2 y, t;
      y = sub2(ax[p], cx[p]); t = add2(sx[p], y); cx[p] = sub2(sub2(t, sx[p]), y); sx[p] = t;
      y = sub2(ay[p], cy[p]); t = add2(sy[p], y); cy[p] = sub2(sub2(t, sy[p]), y); sy[p] = t;
      y = sub2(az[p], cz[p]); t = add2(sz[p], y); cz[p] = sub2(sub2(t, sz[p]), y); sz[p] = t;
    }
    __syncthreads();  // everyone is done with tile[st] before it is refilled two units later

    ++jt;
    const bool last = (n + 1 == nunits);
    if (jt == UJ || last) {
      float* out = partial + ((size_t)slot_base + blockIdx.x + itile) * (size_t)(3 * IT) + tid;
#pragma unroll
      for (int p = 0; p < P; p++) {
        float a, b;
        upk(sx[p], a, b); out[(2 * p) * T] = a; out[(2 * p + 1) * T] = b;
        upk(sy[p], a, b); out[IT + (2 * p) * T] = a; out[IT + (2 * p + 1) * T] = b;
        upk(sz[p], a, b); out[2 * IT + (2 * p) * T] = a; out[2 * IT + (2 * p + 1) * T] = b;
      }
      if (jt == UJ) {
        jt = 0;
        ++itile;
      }
      if (!last) load_targets(itile);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// reduce partials in a fixed order (+ optional fused kick, src/cuda/integrator.cu:24-36)
// ---------------------------------------------------------------------------------------------
struct ReduceArgs {
  nb_pass pass[2];
  int npass;
};

__device__ __forceinline__ void pass_range(const nb_pass& ps, int t, int& c0, int& c1) {
  const long long A = (long long)t * ps.UJ, B = A + ps.UJ;
  long long f = ((A + 1) * ps.G + ps.U - 1) / ps.U - 1;
  long long l = (B * ps.G + ps.U - 1) / ps.U - 1;
  if (f > ps.G - 1) f = ps.G - 1;
  if (l > ps.G - 1) l = ps.G - 1;
  c0 = (int)f;
  c1 = (int)l;
}

template <int IT, bool KICK>
__global__ void __launch_bounds__(256) reduce_kernel(const float* __restrict__ partial, ReduceArgs ra, nb_particles d,
                                                     size_t n, float dt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int t = (int)(i / IT), r = (int)(i % IT);
  float ax = 0.f, ay = 0.f, az = 0.f;
  for (int ps = 0; ps < ra.npass; ps++) {
    int c0, c1;
    pass_range(ra.pass[ps], t, c0, c1);
    for (int c = c0; c <= c1; c++) {
      const float* p = partial + (ra.pass[ps].slot_base + (size_t)c + (size_t)t) * (size_t)(3 * IT) + r;
      ax += p[0];
      ay += p[IT];
      az += p[2 * IT];
    }
  }
  d.acc_x[i] = ax;
  d.acc_y[i] = ay;
  d.acc_z[i] = az;
  if (KICK) {
    const float dt_half = __fmul_rn(0.5f, dt);
    d.vel_x[i] = __fadd_rn(d.vel_x[i], __fmul_rn(__fadd_rn(d.acc_old_x[i], ax), dt_half));
    d.vel_y[i] = __fadd_rn(d.vel_y[i], __fmul_rn(__fadd_rn(d.acc_old_y[i], ay), dt_half));
    d.vel_z[i] = __fadd_rn(d.vel_z[i], __fmul_rn(__fadd_rn(d.acc_old_z[i], az), dt_half));
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static const DirectCfg kCfgBig = {256, 2, 1024};
static const DirectCfg kCfgSmall = {128, 2, 256};
static const size_t kSmallN = 32768;

const DirectCfg& direct_cfg(size_t n_total_sources) { return n_total_sources < kSmallN ? kCfgSmall : kCfgBig; }

template <int T, int I, int JT>
static int occupancy(bool eps0) {
  int occ = 0;
  if (eps0)
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, direct_force_kernel<T, I, JT, true>, T, 0);
  else
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, direct_force_kernel<T, I, JT, false>, T, 0);
  return occ > 0 ? occ : 1;
}

// Launch one stream-K pass.  Source tiles [ja0, ja0+na) then [jb0, jb0+nbt) of `src` (tile = JT
// bodies) against n_itiles target tiles of `tgt`.
int launch_direct_pass(nb_context* ctx, const DirectCfg& cfg, const float4* src, const float4* tgt, int n_itiles,
                       int ja0, int na, int jb0, int nbt, float eps2, int pass_index, cudaStream_t st) {
  const int UJ = na + nbt;
  nb_pass& ps = ctx->pass[pass_index];
  ps.UJ = UJ;
  ps.n_itiles = n_itiles;
  ps.U = (long long)n_itiles * UJ;
  ps.slot_base = (pass_index == 0) ? 0 : ctx->pass[0].slot_base + (size_t)ctx->pass[0].G + ctx->pass[0].n_itiles;
  if (ps.U == 0) {
    ps.G = 0;
    return NB_OK;
  }
  const bool big = (cfg.T == kCfgBig.T);
  const bool eps0 = !(eps2 > 0.f);
  int& occ = ctx->direct_occ[(big ? 0 : 2) + (eps0 ? 1 : 0)];
  if (occ == 0) occ = big ? occupancy<256, 2, 1024>(eps0) : occupancy<128, 2, 256>(eps0);
  int use_occ = occ;
  if (const char* e = getenv("NB_DIRECT_OCC")) {  // tuning knob: resident CTAs per SM
    const int v = atoi(e);
    if (v > 0 && v < occ) use_occ = v;
  }
  long long g = (long long)use_occ * ctx->sm_count;
  if (g > ps.U) g = ps.U;
  ps.G = (int)g;
  // pass 0 reserves room for a second pass too: the buffer must not move while pass 0 runs
  const size_t slots = (pass_index == 0) ? 2 * ((size_t)occ * ctx->sm_count + n_itiles)
                                         : ps.slot_base + (size_t)ps.G + n_itiles;
  if (int rc = ctx->partial.ensure(slots * 3 * cfg.IT() * sizeof(float))) return rc;
  float* part = ctx->partial.as<float>();
#define NB_LAUNCH_DIRECT(T_, I_, JT_, E_)                                                             \
  direct_force_kernel<T_, I_, JT_, E_><<<ps.G, T_, 0, st>>>(src, tgt, part, ps.U, UJ, ja0, na, jb0, eps2, \
                                                            (unsigned long long)ps.slot_base)
  ProfScope prof(ctx, st);
  if (big) {
    if (eps0) NB_LAUNCH_DIRECT(256, 2, 1024, true);
    else NB_LAUNCH_DIRECT(256, 2, 1024, false);
  } else {
    if (eps0) NB_LAUNCH_DIRECT(128, 2, 256, true);
    else NB_LAUNCH_DIRECT(128, 2, 256, false);
  }
#undef NB_LAUNCH_DIRECT
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

int launch_reduce(nb_context* ctx, const DirectCfg& cfg, const nb_particles* d, size_t n, float dt, bool kick,
                  cudaStream_t st) {
  ReduceArgs ra;
  ra.npass = 0;
  for (int p = 0; p < ctx->n_pass; p++)
    if (ctx->pass[p].U > 0) ra.pass[ra.npass++] = ctx->pass[p];
  const unsigned grid = (unsigned)((n + 255) / 256);
  const float* part = ctx->partial.as<float>();
  const bool big = (cfg.T == kCfgBig.T);
  if (big) {
    if (kick) reduce_kernel<512, true><<<grid, 256, 0, st>>>(part, ra, *d, n, dt);
    else reduce_kernel<512, false><<<grid, 256, 0, st>>>(part, ra, *d, n, dt);
  } else {
    if (kick) reduce_kernel<256, true><<<grid, 256, 0, st>>>(part, ra, *d, n, dt);
    else reduce_kernel<256, false><<<grid, 256, 0, st>>>(part, ra, *d, n, dt);
  }
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

static size_t padded_count(size_t n, const DirectCfg& cfg) {
  const size_t it = cfg.IT(), jt = cfg.JT;
  const size_t a = (n + it - 1) / it * it, b = (n + jt - 1) / jt * jt;
  return a > b ? a : b;
}

int launch_pack(nb_context* ctx, const nb_particles* d, float4* src, size_t n, size_t n_pad, float G, float dt,
                bool drift, cudaStream_t st) {
  const unsigned grid = (unsigned)((n_pad + 255) / 256);
  if (grid == 0) return NB_OK;
  if (drift) pack_kernel<true><<<grid, 256, 0, st>>>(*d, src, n, n_pad, G, dt);
  else pack_kernel<false><<<grid, 256, 0, st>>>(*d, src, n, n_pad, G, dt);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

// forces (and, if step, the whole Verlet step) for a single GPU
int direct_run(nb_context* ctx, const nb_particles* d, float dt, float G, float eps2, bool step, cudaStream_t st) {
  const size_t n = d->count;
  if (n == 0) return NB_OK;
  if (n > 0x7fffffffull / 2) {
    set_error("direct N^2: particle count %zu too large", n);
    return NB_ERR_INVALID;
  }
  const DirectCfg& cfg = direct_cfg(n);
  const size_t n_pad = padded_count(n, cfg);
  if (int rc = ctx->src.ensure(n_pad * sizeof(float4))) return rc;
  float4* src = ctx->src.as<float4>();
  if (int rc = launch_pack(ctx, d, src, n, n_pad, G, dt, step, st)) return rc;
  const int n_itiles = (int)((n + cfg.IT() - 1) / cfg.IT());
  const int n_jt = (int)((n + cfg.JT - 1) / cfg.JT);
  ctx->n_pass = 1;
  if (int rc = launch_direct_pass(ctx, cfg, src, src, n_itiles, 0, n_jt, 0, 0, eps2, 0, st)) return rc;
  return launch_reduce(ctx, cfg, d, n, dt, step, st);
}

}  // namespace nb
