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
//   * the (target tile x source tile) unit space is cut into chunks that a persistent grid (6 CTAs per
//     SM) claims from a global counter, so all 148 SMs and all resident CTAs finish together for any
//     N; chunk c writes fp32 partials to slot c, and a reduce kernel adds a tile's chunks in chunk
//     order (deterministic, no float atomics) and applies the kick;
//   * accumulation is three-level: plain fp32 inside a source tile, Kahan-compensated across the
//     tiles of a chunk and across the chunks of a target tile.
#include "direct.cuh"

#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <type_traits>

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
// build-time tuning knobs (A/B builds through tools/ab_direct.sh)
#ifndef NB_DIRECT_THREADS_PER_SM
#define NB_DIRECT_THREADS_PER_SM 1536  // 6 CTAs of 256 threads at 40 registers (5 at 48: 404.8 ms per 1 Mi step, 6: 387.7)
#endif
#ifndef NB_DIRECT_UNROLL
#define NB_DIRECT_UNROLL 64  // sources per trip of the pair loop (8: 440.5 ms per 1 Mi step, 32: 430.9, 64: 429.4)
#endif
// Tried and dropped (profiles/r01_direct_variants_ab.log): not keeping the warps of a CTA in lockstep --
// the producer's warp alone waits on a named barrier before refilling a buffer, the others bar.arrive and
// block only on missing data -- 430.5 ms against 429.8 ms with the plain __syncthreads per tile.
#ifndef NB_DIRECT_BIG_T
#define NB_DIRECT_BIG_T 256  // threads per CTA of the large-N configuration
#endif
#ifndef NB_DIRECT_BIG_JT
#define NB_DIRECT_BIG_JT 1024  // sources per shared-memory tile of the large-N configuration
#endif
#ifndef NB_DIRECT_BIG_I
#define NB_DIRECT_BIG_I 2  // targets per thread of the large-N configuration (even: packed pairs)
#endif
#define NB_PRAGMA_(x) _Pragma(#x)
#define NB_PRAGMA_UNROLL(n) NB_PRAGMA_(unroll n)

#ifdef NB_DIRECT_TRACE  // experiment: where and when each CTA of the force kernel runs
__device__ unsigned long long nb_direct_trace_buf[3 * 4096];
#endif

// Work distribution: the (target tile x source tile) unit space is cut into n_chunks CHUNKS, chunk c =
// target tile c / C, source tiles [jc*UJ/C, (jc+1)*UJ/C) with jc = c % C, and CTAs claim chunks from a
// global counter (work[0]) until none are left.  A static equal split is not balanced on this machine:
// with 5 CTAs resident per SM the warp schedulers favour some CTAs, equal shares finished between 151
// and 457 ms of a 457 ms launch (profiles/r01_direct_cta_trace.log) and the SM ran under-occupied for
// most of it.  Which CTA computes a chunk does not affect its result, and chunk c always lands in
// partial slot c, so the sums stay reproducible.  The last CTA to leave re-arms the counters (work[1]
// counts leavers), which keeps the launch replayable from a CUDA graph.
// Round 2 tried to cut the per-chunk partial slots (6 KB each, 350 MB at any large N with 64 chunks per CTA) without
// lengthening the tail of the launch; all measured on one box against this kernel (profiles/r02_direct_same_box_ab.jsonl,
// r02_direct_chunk_plans_ab.jsonl; ms per 1 Mi step, this kernel: 387.0):
//   * 32 equal chunks per CTA (half the slots): 389.0 -- the tail is one chunk long;
//   * rows cut into chunks of descending length, claimed class by class (long chunks of all rows first): 394.1, and
//     equal chunks in that order 391.0 -- with row-major claiming the CTAs in flight cover ~32 target tiles x ALL source
//     ranges, the whole packed source array in use at once; class-major makes all 888 CTAs stream the same few tiles;
//   * long chunks for most rows and short ones for the last rows, row-major (a third of the slots, same tail): 392.4 --
//     and 391.0 / 390.2 with twice / four times as many chunks, i.e. not the schedule but the kernel: the extra
//     chunk-to-range arithmetic changed ptxas' schedule of the one-mass pair loop (the general loop stayed at 428-430).
// Kept: this kernel and 64 chunks per CTA; NB_DIRECT_CHUNKS=32 halves the scratch for +0.5 %.
template <int T, int I, int JT, bool EPS0>
__global__ void __launch_bounds__(T, NB_DIRECT_THREADS_PER_SM / T) direct_force_kernel(const float4* __restrict__ src,
                                                          const float4* __restrict__ tgt,
                                                          float* __restrict__ partial, unsigned* __restrict__ work,
                                                          int n_chunks, int C, int UJ, int ja0, int na, int jb0,
                                                          float eps2, unsigned long long slot_base) {
  constexpr int P = I / 2;
  constexpr int IT = T * I;
  constexpr uint32_t TILE_BYTES = JT * sizeof(float4);
  static_assert(I % 2 == 0, "targets per thread must be even (packed fp32x2)");
  static_assert(JT % T == 0, "the one-mass test reads JT / T sources per thread");
  __shared__ __align__(128) float4 tile[2][JT];
  __shared__ __align__(8) uint64_t full[2];
  __shared__ int claimed[2];

  const int tid = threadIdx.x;
  auto jtile_ptr = [&](int j) -> const float4* {
    const int g = (j < na) ? (ja0 + j) : (jb0 + (j - na));
    return src + (size_t)g * JT;
  };
  auto first_jtile = [&](int c) -> int { return (int)(((long long)(c % C) * UJ) / C); };
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
    const int c = (int)atomicAdd(&work[0], 1u);
    claimed[0] = c;
    if (c < n_chunks) {
      mbar_expect_tx(&full[0], TILE_BYTES);
      tma_load_1d(&tile[0][0], jtile_ptr(first_jtile(c)), TILE_BYTES, &full[0]);
    }
  }
  __syncthreads();
#ifdef NB_DIRECT_TRACE
  if (tid == 0 && blockIdx.x < 4096) {
    unsigned sm;
    unsigned long long t;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    nb_direct_trace_buf[3 * blockIdx.x] = sm;
    nb_direct_trace_buf[3 * blockIdx.x + 1] = t;
  }
#endif

  f2 xi[P], yi[P], zi[P];
  f2 sx[P], sy[P], sz[P];  // Kahan sums across the tiles of a chunk
  f2 cx[P], cy[P], cz[P];  // Kahan compensation
  const f2 zero2 = pk(0.f, 0.f);
  const f2 e2 = pk(eps2, eps2);

  int chunk = claimed[0];
  unsigned n = 0;  // tiles this CTA has consumed: buffer and mbarrier parity
  for (unsigned k = 0; chunk < n_chunks; k++) {
    const int itile = chunk / C, jc = chunk - itile * C;
    int j = (int)(((long long)jc * UJ) / C);
    const int jend = (int)(((long long)(jc + 1) * UJ) / C);
    {
      const float4* base = tgt + (size_t)itile * IT + tid;
#pragma unroll
      for (int p = 0; p < P; p++) {
        // scalar loads: each lands in its half of the packed pair (a 128-bit load pins a register quad
        // and the re-pairing moves end up inside the pair loop)
        const float* a = reinterpret_cast<const float*>(base + (2 * p) * T);
        const float* b = reinterpret_cast<const float*>(base + (2 * p + 1) * T);
        xi[p] = pk(__ldg(a), __ldg(b));
        yi[p] = pk(__ldg(a + 1), __ldg(b + 1));
        zi[p] = pk(__ldg(a + 2), __ldg(b + 2));
        sx[p] = sy[p] = sz[p] = zero2;
        cx[p] = cy[p] = cz[p] = zero2;
      }
    }
    for (; j < jend; j++, n++) {
      const unsigned st = n & 1u;
      if (tid == 0) {
        if (j + 1 < jend) {
          mbar_expect_tx(&full[st ^ 1], TILE_BYTES);
          tma_load_1d(&tile[st ^ 1][0], jtile_ptr(j + 1), TILE_BYTES, &full[st ^ 1]);
        } else {  // last tile of this chunk: claim the next chunk now and prefetch its first tile
          const int c = (int)atomicAdd(&work[0], 1u);
          claimed[(k + 1) & 1u] = c;
          if (c < n_chunks) {
            mbar_expect_tx(&full[st ^ 1], TILE_BYTES);
            tma_load_1d(&tile[st ^ 1][0], jtile_ptr(first_jtile(c)), TILE_BYTES, &full[st ^ 1]);
          }
        }
      }
      mbar_wait(&full[st], (n >> 1) & 1u);

      f2 ax[P], ay[P], az[P];
#pragma unroll
      for (int p = 0; p < P; p++) ax[p] = ay[p] = az[p] = zero2;
      const float4* tj = &tile[st][0];
      // Sources of one mass: when every body of the tile carries the same G*m (bit pattern), the factor
      // leaves the pair loop -- w = 1/r^3 costs two packed multiplies instead of three -- and is applied
      // once to the tile's sums.  Decided per tile from the data itself (all the reference's initial
      // conditions are equal-mass, particle_init.cu:290-357; mixed tiles, e.g. the zero-mass padding of
      // the last tile, take the general loop), so it holds for every caller of this kernel.
      const float gm0 = tj[0].w;
      bool mixed = false;
#pragma unroll
      for (int q = 0; q < JT / T; q++) mixed |= (__float_as_uint(tj[q * T + tid].w) != __float_as_uint(gm0));
      const bool one_mass = !__syncthreads_or(mixed);
      auto pair_loop = [&](auto uniform_tag) {
        constexpr bool ONE_MASS = decltype(uniform_tag)::value;
        NB_PRAGMA_UNROLL(NB_DIRECT_UNROLL)
        for (int jj = 0; jj < JT; jj++) {
          const float4 s = tj[jj];  // warp-broadcast LDS.128
          const f2 px = pk(s.x, s.x), py = pk(s.y, s.y), pz = pk(s.z, s.z);
#pragma unroll
          for (int p = 0; p < P; p++) {
            const f2 dx = sub2(px, xi[p]), dy = sub2(py, yi[p]), dz = sub2(pz, zi[p]);
            f2 r2 = fma2(dx, dx, e2);
            r2 = fma2(dy, dy, r2);
            r2 = fma2(dz, dz, r2);
            float r0, r1;
            upk(r2, r0, r1);
            float i0 = rsq(r0), i1 = rsq(r1);
            if (EPS0) {  // eps ~ 0: the self pair (and exact coincidences) must contribute 0, not NaN
              i0 = (r0 > eps2) ? i0 : 0.f;
              i1 = (r1 > eps2) ? i1 : 0.f;
            }
            const f2 ri = pk(i0, i1);
            f2 w;
            if (ONE_MASS) w = mul2(mul2(ri, ri), ri);                      // 1 / r^3
            else w = mul2(mul2(ri, ri), mul2(pk(s.w, s.w), ri));           // (G m_j) / r^3
            ax[p] = fma2(w, dx, ax[p]);
            ay[p] = fma2(w, dy, ay[p]);
            az[p] = fma2(w, dz, az[p]);
          }
        }
      };
      if (one_mass) {
        pair_loop(std::true_type{});
        const f2 gm = pk(gm0, gm0);
#pragma unroll
        for (int p = 0; p < P; p++) {
          ax[p] = mul2(ax[p], gm);
          ay[p] = mul2(ay[p], gm);
          az[p] = mul2(az[p], gm);
        }
      } else {
        pair_loop(std::false_type{});
      }
      // fold the tile's sums into the compensated running sums
#pragma unroll
      for (int p = 0; p < P; p++) {
        f2 y, t;
        y = sub2(ax[p], cx[p]); t = add2(sx[p], y); cx[p] = sub2(sub2(t, sx[p]), y); sx[p] = t;
        y = sub2(ay[p], cy[p]); t = add2(sy[p], y); cy[p] = sub2(sub2(t, sy[p]), y); sy[p] = t;
        y = sub2(az[p], cz[p]); t = add2(sz[p], y); cz[p] = sub2(sub2(t, sz[p]), y); sz[p] = t;
      }
      __syncthreads();  // everyone is done with tile[st] before it is refilled two tiles later
    }
    float* out = partial + ((size_t)slot_base + (size_t)chunk) * (size_t)(3 * IT) + tid;
#pragma unroll
    for (int p = 0; p < P; p++) {
      float a, b;
      upk(sx[p], a, b); out[(2 * p) * T] = a; out[(2 * p + 1) * T] = b;
      upk(sy[p], a, b); out[IT + (2 * p) * T] = a; out[IT + (2 * p + 1) * T] = b;
      upk(sz[p], a, b); out[2 * IT + (2 * p) * T] = a; out[2 * IT + (2 * p + 1) * T] = b;
    }
    chunk = claimed[(k + 1) & 1u];
  }
  if (tid == 0) {
#ifdef NB_DIRECT_TRACE
    if (blockIdx.x < 4096) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      nb_direct_trace_buf[3 * blockIdx.x + 2] = t;
    }
#endif
    // every CTA makes exactly one failing claim before it gets here, so once all have left nobody
    // touches work[0] again: re-arm both counters for the next launch
    if (atomicAdd(&work[1], 1u) == gridDim.x - 1) {
      work[0] = 0u;
      work[1] = 0u;
    }
  }
}

#ifdef NB_DIRECT_TRACE
extern "C" int nb_debug_direct_trace(unsigned long long* out, int n_ctas) {
  cudaDeviceSynchronize();
  return (int)cudaMemcpyFromSymbol(out, nb_direct_trace_buf, sizeof(unsigned long long) * 3 * n_ctas);
}
#endif

// ---------------------------------------------------------------------------------------------
// reduce partials in a fixed order (+ optional fused kick, src/cuda/integrator.cu:24-36)
// ---------------------------------------------------------------------------------------------
struct ReduceArgs {
  nb_pass pass[2];
  int npass;
};

template <int IT, bool KICK>
__global__ void __launch_bounds__(256) reduce_kernel(const float* __restrict__ partial, ReduceArgs ra, nb_particles d,
                                                     size_t n, float dt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int t = (int)(i / IT), r = (int)(i % IT);
  // chunk sums of this body's target tile in chunk order, compensated (each chunk sum already is)
  float ax = 0.f, ay = 0.f, az = 0.f, ex = 0.f, ey = 0.f, ez = 0.f;
  for (int ps = 0; ps < ra.npass; ps++) {
    const int C = ra.pass[ps].C;
    const float* p = partial + (ra.pass[ps].slot_base + (size_t)t * C) * (size_t)(3 * IT) + r;
    for (int c = 0; c < C; c++, p += 3 * IT) {
      float y, s;
      y = __fsub_rn(p[0], ex); s = __fadd_rn(ax, y); ex = __fsub_rn(__fsub_rn(s, ax), y); ax = s;
      y = __fsub_rn(p[IT], ey); s = __fadd_rn(ay, y); ey = __fsub_rn(__fsub_rn(s, ay), y); ay = s;
      y = __fsub_rn(p[2 * IT], ez); s = __fadd_rn(az, y); ez = __fsub_rn(__fsub_rn(s, az), y); az = s;
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
static const DirectCfg kCfgBig = {NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT};
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

// Chunks per target tile: enough chunks that every CTA claims about `per_cta` of them (the tail of a
// launch is at most one chunk long), never more than one chunk per source tile.
static int chunks_per_tile(int n_itiles, int UJ, long long ctas) {
  long long per_cta = 64;
  if (const char* e = getenv("NB_DIRECT_CHUNKS")) {  // tuning knob: chunks per CTA
    const long long v = atoll(e);
    if (v > 0) per_cta = v;
  }
  long long c = (per_cta * ctas + n_itiles - 1) / n_itiles;
  if (c > UJ) c = UJ;
  if (c < 1) c = 1;
  return (int)c;
}

// Launch one pass.  Source tiles [ja0, ja0+na) then [jb0, jb0+nbt) of `src` (tile = JT bodies) against
// n_itiles target tiles of `tgt`.
int launch_direct_pass(nb_context* ctx, const DirectCfg& cfg, const float4* src, const float4* tgt, int n_itiles,
                       int ja0, int na, int jb0, int nbt, float eps2, int pass_index, cudaStream_t st) {
  const int UJ = na + nbt;
  nb_pass& ps = ctx->pass[pass_index];
  ps.UJ = UJ;
  ps.n_itiles = n_itiles;
  ps.U = (long long)n_itiles * UJ;
  ps.slot_base = (pass_index == 0) ? 0 : ctx->pass[0].slot_base + (size_t)ctx->pass[0].chunks;
  if (ps.U == 0) {
    ps.G = 0;
    ps.C = 0;
    ps.chunks = 0;
    return NB_OK;
  }
  const bool big = (&cfg == &kCfgBig);
  const bool eps0 = !(eps2 >= NB_EPS2_GUARD);
  int& occ = ctx->direct_occ[(big ? 0 : 2) + (eps0 ? 1 : 0)];
  if (occ == 0) occ = big ? occupancy<NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT>(eps0) : occupancy<128, 2, 256>(eps0);
  int use_occ = occ;
  if (const char* e = getenv("NB_DIRECT_OCC")) {  // tuning knob: resident CTAs per SM
    const int v = atoi(e);
    if (v > 0 && v < occ) use_occ = v;
  }
  const long long ctas = (long long)use_occ * ctx->sm_count;
  ps.C = chunks_per_tile(n_itiles, UJ, ctas);
  ps.chunks = (long long)n_itiles * ps.C;
  if (ps.chunks > 0x7fffffffll - 65536) {
    set_error("direct N^2: %lld work chunks exceed the kernel's 32-bit chunk counter", ps.chunks);
    return NB_ERR_INVALID;
  }
  ps.G = (int)(ctas < ps.chunks ? ctas : ps.chunks);
  // pass 0 reserves room for the remote-source pass of a sharded step too ((R-1) x as many source tiles,
  // hence at most (R-1) x as many chunks): the buffer must not move while pass 0 runs
  size_t slots = ps.slot_base + (size_t)ps.chunks;
  if (pass_index == 0 && ctx->mg_nranks > 1) slots += (size_t)ps.chunks * (size_t)(ctx->mg_nranks - 1);
  if (int rc = ctx->partial.ensure(slots * 3 * cfg.IT() * sizeof(float))) return rc;
  float* part = ctx->partial.as<float>();
  unsigned* work = ctx->direct_work.as<unsigned>() + 2 * pass_index;
#define NB_LAUNCH_DIRECT(T_, I_, JT_, E_)                                                                     \
  direct_force_kernel<T_, I_, JT_, E_><<<ps.G, T_, 0, st>>>(src, tgt, part, work, (int)ps.chunks, ps.C, UJ, ja0, na, \
                                                            jb0, eps2, (unsigned long long)ps.slot_base)
  ProfScope prof(ctx, st);
  if (big) {
    if (eps0) NB_LAUNCH_DIRECT(NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT, true);
    else NB_LAUNCH_DIRECT(NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT, false);
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
  const bool big = (&cfg == &kCfgBig);
  if (big) {
    if (kick) reduce_kernel<NB_DIRECT_BIG_I * NB_DIRECT_BIG_T, true><<<grid, 256, 0, st>>>(part, ra, *d, n, dt);
    else reduce_kernel<NB_DIRECT_BIG_I * NB_DIRECT_BIG_T, false><<<grid, 256, 0, st>>>(part, ra, *d, n, dt);
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

// ---------------------------------------------------------------------------------------------
// nb_direct_forces_static: direct N^2 forces with NO device allocation at call time.
//
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same
// three kernels as nb_direct_forces -- pack, direct_force_kernel, reduce -- run with their scratch (chunk
// counters, packed sources, per-chunk partial sums) in ONE file-scope __device__ array instead of the
// context's growable buffers.  Every GPU that loads the library holds the array until the process exits.
// All calls share it: calls that overlap on the GPU (two streams, two contexts, two host threads on the same
// device) race.  The array keeps the previous call's contents; the only part that has to start at zero, the
// chunk counters, is cleared on the caller's stream at the top of every call.
// ---------------------------------------------------------------------------------------------
constexpr size_t kStaticMaxBodies = NB_DIRECT_STATIC_MAX_BODIES;
constexpr size_t kStaticHeadBytes = 256;                                  // chunk counters {claimed, leavers}
constexpr size_t kStaticSrcBytes = kStaticMaxBodies * sizeof(float4);     // packed {x, y, z, G m}: 1 MiB
constexpr size_t kStaticPartialBytes = (size_t)16 << 20;                  // per-chunk partial sums: 16 MiB
constexpr size_t kStaticBytes = kStaticHeadBytes + kStaticSrcBytes + kStaticPartialBytes;
static __device__ __align__(256) unsigned char nb_direct_static_scratch[kStaticBytes];


namespace nb {
namespace {
struct StaticSetup {
  bool ready = false;
  unsigned char* base = nullptr;  // DEVICE address of nb_direct_static_scratch on this device
  int occ[4] = {0, 0, 0, 0};      // resident CTAs per SM: {big, big eps0, small, small eps0}
};
constexpr int kMaxDevices = 64;
StaticSetup g_static_setup[kMaxDevices];
std::mutex g_static_mutex;

// once per device, with that device current: in host code the array's name is a host-side stand-in, the
// kernels need the address cudaGetSymbolAddress returns for THIS device
int static_setup(int device, StaticSetup** out) {
  if (device < 0 || device >= kMaxDevices) {
    set_error("nb_direct_forces_static: device ordinal %d out of range", device);
    return NB_ERR_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_static_mutex);
  StaticSetup& s = g_static_setup[device];
  if (!s.ready) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, nb_direct_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(nb_direct_static_scratch)", __FILE__, __LINE__);
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255u)) {
      set_error("nb_direct_forces_static: static scratch address %p is null or not 256-byte aligned", p);
      return NB_ERR_INTERNAL;
    }
    s.base = static_cast<unsigned char*>(p);
    s.occ[0] = occupancy<NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT>(false);
    s.occ[1] = occupancy<NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT>(true);
    s.occ[2] = occupancy<128, 2, 256>(false);
    s.occ[3] = occupancy<128, 2, 256>(true);
    s.ready = true;
  }
  *out = &s;
  return NB_OK;
}
}  // namespace

}  // namespace nb

extern "C" int nb_direct_forces_static(nb_context* ctx, const nb_particles* d, float G, float eps2, nb_stream s) {
  using namespace nb;
  if (!ctx) {
    set_error("null context");
    return NB_ERR_INVALID;
  }
  if (!d) {
    set_error("null particle data");
    return NB_ERR_INVALID;
  }
  const size_t n = d->count;
  if (n == 0) return NB_OK;
  if (n > kStaticMaxBodies) {
    set_error("nb_direct_forces_static: %zu bodies exceed the static scratch's maximum of %zu "
              "(NB_DIRECT_STATIC_MAX_BODIES); use nb_direct_forces, which allocates",
              n, kStaticMaxBodies);
    return NB_ERR_INVALID;
  }
  if (!d->pos_x || !d->pos_y || !d->pos_z || !d->mass || !d->acc_x || !d->acc_y || !d->acc_z) {
    set_error("particle data has null position/mass/acceleration arrays");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  StaticSetup* su = nullptr;
  if (int rc = static_setup(ctx->device, &su)) return rc;
  cudaStream_t st = (cudaStream_t)s;

  unsigned* work = reinterpret_cast<unsigned*>(su->base);
  float4* src = reinterpret_cast<float4*>(su->base + kStaticHeadBytes);
  float* part = reinterpret_cast<float*>(su->base + kStaticHeadBytes + kStaticSrcBytes);

  // same configuration, padding and chunk plan as direct_run / launch_direct_pass ...
  const DirectCfg& cfg = direct_cfg(n);
  const bool big = (&cfg == &kCfgBig);
  const bool eps0 = !(eps2 >= NB_EPS2_GUARD);
  const size_t n_pad = padded_count(n, cfg);  // <= kStaticMaxBodies: the maximum is a multiple of every tile size
  const int n_itiles = (int)((n + cfg.IT() - 1) / cfg.IT());
  const int UJ = (int)((n + cfg.JT - 1) / cfg.JT);
  const long long ctas = (long long)su->occ[(big ? 0 : 2) + (eps0 ? 1 : 0)] * ctx->sm_count;
  int C = chunks_per_tile(n_itiles, UJ, ctas);
  // ... except that the partial slots have to fit the fixed array: fewer, longer chunks per target tile when they
  // would not (n > 16384).  Chunk sums are then added in a different grouping than nb_direct_forces' -- the same
  // pair terms, equal to rounding; at n <= 16384 the plan, and so every bit of the result, is the same.
  const size_t slot_bytes = (size_t)3 * cfg.IT() * sizeof(float);
  const long long c_fit = (long long)(kStaticPartialBytes / slot_bytes) / n_itiles;
  if (C > c_fit) C = (int)c_fit;
  if (C < 1 || n_pad * sizeof(float4) > kStaticSrcBytes) {
    set_error("nb_direct_forces_static: internal sizing error (n=%zu, n_pad=%zu, C=%d)", n, n_pad, C);
    return NB_ERR_INTERNAL;
  }
  const long long chunks = (long long)n_itiles * C;
  const int grid = (int)(ctas < chunks ? ctas : chunks);

  // the chunk counters must start at zero (a launch re-arms them on its way out, but an earlier call that failed
  // part-way would leave them set); a memset node on the caller's stream, no allocation
  NB_CUDA(cudaMemsetAsync(work, 0, kStaticHeadBytes, st));
  if (int rc = launch_pack(ctx, d, src, n, n_pad, G, 0.f, false, st)) return rc;
#define NB_LAUNCH_DIRECT_STATIC(T_, I_, JT_, E_)                                                                \
  direct_force_kernel<T_, I_, JT_, E_><<<grid, T_, 0, st>>>(src, src, part, work, (int)chunks, C, UJ, 0, UJ, 0, eps2, 0ull)
  if (big) {
    if (eps0) NB_LAUNCH_DIRECT_STATIC(NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT, true);
    else NB_LAUNCH_DIRECT_STATIC(NB_DIRECT_BIG_T, NB_DIRECT_BIG_I, NB_DIRECT_BIG_JT, false);
  } else {
    if (eps0) NB_LAUNCH_DIRECT_STATIC(128, 2, 256, true);
    else NB_LAUNCH_DIRECT_STATIC(128, 2, 256, false);
  }
#undef NB_LAUNCH_DIRECT_STATIC
  NB_LAUNCH_CHECK(ctx);
  ReduceArgs ra;
  ra.npass = 1;
  ra.pass[0] = nb_pass{};
  ra.pass[0].C = C;
  ra.pass[0].slot_base = 0;
  const unsigned rgrid = (unsigned)((n + 255) / 256);
  if (big) reduce_kernel<NB_DIRECT_BIG_I * NB_DIRECT_BIG_T, false><<<rgrid, 256, 0, st>>>(part, ra, *d, n, 0.f);
  else reduce_kernel<256, false><<<rgrid, 256, 0, st>>>(part, ra, *d, n, 0.f);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}
