// sh.cu -- spatial-hash (uniform grid) short-range forces.
//
// Replaces SpatialHashGrid::build / computeForces (reference src/cuda/force_spatial_hash.cu:235-316):
// the reference counts with N global atomics, scans with Thrust, scatters with N more atomics (so
// the order inside a cell is racy) and gathers neighbours through an index indirection.  Here the
// cell ids -- same arithmetic, bit-exact (:28-49, :226-246) -- are the keys of the stable radix
// sort (sort.cu), cell_start/cell_end are read off the sorted keys (identical to the reference's
// exclusive scan, empty cells included), bodies are physically reordered into cell order as packed
// float4, and the 27-cell sweep (:83-152) walks 9 contiguous x-runs of that array.  Order inside a
// cell is ascending body index, so results are reproducible.  The force kernel gives every thread TWO
// targets of the same cell (one packed fp32x2 pair) so that each candidate is loaded once for two
// interactions: pairs are counted from each cell's first slot (a cell with an odd count ends in a
// single), and the build lists the pair leaders' slots (flag + scan over the cell-ordered bodies).
#include <cmath>
#include <new>
#include <cstring>

#include "common.cuh"

namespace nb {
int bbox_device(nb_context* ctx, const nb_particles* d, cudaStream_t st);
int sort_pairs_db(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, uint32_t* keys_alt, int32_t* vals_alt, size_t n,
                  int end_bit, bool identity_vals, uint32_t** out_keys, int32_t** out_vals, cudaStream_t st,
                  const uint32_t* copy_keys = nullptr, const int32_t* copy_vals = nullptr);
int launch_scan_tiles(nb_context* ctx, unsigned long long* tile_sum, int ntiles, unsigned long long* total,
                      cudaStream_t st);  // render.cu
}  // namespace nb

#ifndef NB_SH_SATMASK
#define NB_SH_SATMASK 0  // build-time knob (experiment): cutoff applied as a saturated FFMA mask instead of compare + select
#endif
#ifndef NB_SH_UNROLL
#define NB_SH_UNROLL 4  // candidates per trip of the sweep (build-time tuning knob)
#endif
#define NB_SH_PRAGMA_(x) _Pragma(#x)
#define NB_SH_PRAGMA_UNROLL(n) NB_SH_PRAGMA_(unroll n)

struct SHMisc {  // device + pinned host mirror
  float bbox[6];  // padded
  int dims[3];
  int too_large;
  long long total;
  unsigned long long pairs;
  unsigned long long n_pairs;  // target pairs = sum over cells of ceil(count / 2)
  int n_gaps;                  // long runs of empty cells left to sh_fill_gaps_kernel
  int pad;
};

struct SH {
  size_t n = 0;
  float cell_size = 0.f;
  int total_cells = 0;
  int dims[3] = {0, 0, 0};
  float bbox[6] = {0, 0, 0, 0, 0, 0};
  nb::DevBuf particle_cell, keys, sorted_idx, keys_alt, idx_alt, packed, spos, pair_slot, tile_sum, cell_start, cell_end, misc;
  nb::DevBuf gaps;
  uint32_t* keys_sorted = nullptr;   // where the last build's sorted cell ids / body indices live
  int32_t* idx_sorted = nullptr;
  SHMisc* h_misc = nullptr;
  bool built = false;
  float stats_cutoff2 = -1.f;  // cutoff^2 of the last force evaluation (< 0: none yet)
  bool stats_valid = false;    // misc->pairs holds the count of that evaluation
};

namespace nb {

__device__ __forceinline__ float sh_ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// :226-232 pad, :244-246 dims = (int)ceilf((max-min)/cell)+1, :249-254 limit
__global__ void sh_prepare_kernel(const int* __restrict__ bb_ord, float cell, SHMisc* __restrict__ m) {
  if (threadIdx.x != 0) return;
  float bb[6];
  for (int a = 0; a < 3; a++) {
    bb[a] = __fsub_rn(sh_ord2f(bb_ord[a]), 0.001f);
    bb[3 + a] = __fadd_rn(sh_ord2f(bb_ord[3 + a]), 0.001f);
  }
  long long total = 1;
  for (int a = 0; a < 3; a++) {
    m->bbox[a] = bb[a];
    m->bbox[3 + a] = bb[3 + a];
    const int dgrid = (int)ceilf(__fdiv_rn(__fsub_rn(bb[3 + a], bb[a]), cell)) + 1;
    m->dims[a] = dgrid;
    total *= (long long)dgrid;
  }
  m->total = total;
  m->too_large = total > 100000000LL ? 1 : 0;
  m->pairs = 0ull;
  m->n_pairs = 0ull;
  m->n_gaps = 0;
}

// :28-49
__global__ void __launch_bounds__(256) sh_assign_kernel(const float* __restrict__ px, const float* __restrict__ py,
                                                        const float* __restrict__ pz,
                                                        const float* __restrict__ mass, size_t n, float cell,
                                                        const SHMisc* __restrict__ m, int32_t* __restrict__ pcell,
                                                        uint32_t* __restrict__ keys, float4* __restrict__ packed) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int nx = m->dims[0], ny = m->dims[1], nz = m->dims[2];
  const float rx = px[i], ry = py[i], rz = pz[i];
  packed[i] = make_float4(rx, ry, rz, mass[i]);  // body order; gathered into cell order after the sort
  const float x = __fsub_rn(rx, m->bbox[0]), y = __fsub_rn(ry, m->bbox[1]), z = __fsub_rn(rz, m->bbox[2]);
  int cx = (int)floorf(__fdiv_rn(x, cell)), cy = (int)floorf(__fdiv_rn(y, cell)), cz = (int)floorf(__fdiv_rn(z, cell));
  cx = max(0, min(cx, nx - 1));
  cy = max(0, min(cy, ny - 1));
  cz = max(0, min(cz, nz - 1));
  const int c = cx + cy * nx + cz * nx * ny;
  pcell[i] = c;
  keys[i] = (uint32_t)c;  // the body indices (identity) are generated by the first radix pass
}

// After the sort, one sweep over the sorted (cell, body) pairs does both remaining jobs:
//   cell_start[c] = #bodies with cell < c ; cell_end[c] = #bodies with cell <= c  (:285-288, :61-68),
//   read off the places where the sorted key changes (empty cells in between included);
//   spos[k] = packed[sorted body k]: bodies physically in cell order.
// The thread at a key change fills the ranges of the empty cells before it itself when there are few
// (a dense grid: ~1 per change); a long run of empty cells (few bodies on a fine grid -- up to 1e8 cells are
// allowed) goes to a gap list that sh_fill_gaps_kernel fills with whole thread blocks.
constexpr int SH_GAP_INLINE = 16;
struct SHGap {
  int c0, c1;  // cells [c0, c1): start = end = k (no body), except cell_start of c1 / cell_end of c0 - 1 handled inline
  int k;
  int pad;
};
__global__ void __launch_bounds__(256) sh_ranges_gather_kernel(const uint32_t* __restrict__ keys, size_t n,
                                                               int total_cells, int32_t* __restrict__ cell_start,
                                                               int32_t* __restrict__ cell_end,
                                                               const int32_t* __restrict__ sidx,
                                                               const float4* __restrict__ packed,
                                                               float4* __restrict__ spos, SHGap* __restrict__ gaps,
                                                               int* __restrict__ ngaps) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n) return;
  if (k < n) spos[k] = __ldg(packed + sidx[k]);
  const int prev = (k == 0) ? -1 : (int)keys[k - 1];
  const int cur = (k == n) ? total_cells : (int)keys[k];
  if (prev == cur) return;
  // cells prev+1 .. cur start at k; cells prev .. cur-1 end at k
  if (cur < total_cells) cell_start[cur] = (int32_t)k;
  if (prev >= 0) cell_end[prev] = (int32_t)k;
  const int e0 = prev + 1, e1 = cur;  // the empty cells in between: start = end = k
  if (e1 - e0 <= SH_GAP_INLINE) {
    for (int c = e0; c < e1; c++) {
      cell_start[c] = (int32_t)k;
      cell_end[c] = (int32_t)k;
    }
  } else {
    const int g = atomicAdd(ngaps, 1);
    gaps[g] = SHGap{e0, e1, (int)k, 0};
  }
}

__global__ void __launch_bounds__(256) sh_fill_gaps_kernel(const SHGap* __restrict__ gaps, const int* __restrict__ ngaps,
                                                           int32_t* __restrict__ cell_start,
                                                           int32_t* __restrict__ cell_end) {
  const int ng = *ngaps;
  for (int g = blockIdx.x; g < ng; g += gridDim.x) {
    const SHGap gp = gaps[g];
    for (int c = gp.c0 + threadIdx.x; c < gp.c1; c += blockDim.x) {
      cell_start[c] = gp.k;
      cell_end[c] = gp.k;
    }
  }
}

// Target pairs of the force kernel: slot k leads a pair when it sits an even number of slots after its cell's
// first slot.  pair_slot[q] = slot of the q-th leader, in slot order: flag, tile sums, scan of the tile sums
// (launch_scan_tiles), then in-tile scan + scatter.  Tiles of 2048 slots, 8 consecutive slots per thread.
constexpr int SH_TILE = 2048;
__device__ __forceinline__ unsigned sh_lead_flags(const uint32_t* __restrict__ keys,
                                                  const int32_t* __restrict__ cell_start, size_t n, size_t base,
                                                  unsigned f[8]) {
  unsigned mine = 0;
#pragma unroll
  for (int e = 0; e < 8; e++) {
    const size_t k = base + e;
    f[e] = (k < n && (((unsigned)k - (unsigned)cell_start[keys[k]]) & 1u) == 0u) ? 1u : 0u;
    mine += f[e];
  }
  return mine;
}

__global__ void __launch_bounds__(256) sh_pair_sums_kernel(const uint32_t* __restrict__ keys,
                                                           const int32_t* __restrict__ cell_start, size_t n,
                                                           unsigned long long* __restrict__ tile_sum) {
  __shared__ unsigned ws[8];
  unsigned f[8];
  unsigned s = sh_lead_flags(keys, cell_start, n, (size_t)blockIdx.x * SH_TILE + (size_t)threadIdx.x * 8, f);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned t = 0;
    for (int w = 0; w < 8; w++) t += ws[w];
    tile_sum[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256) sh_pair_fill_kernel(const uint32_t* __restrict__ keys,
                                                           const int32_t* __restrict__ cell_start, size_t n,
                                                           const unsigned long long* __restrict__ tile_off,
                                                           int32_t* __restrict__ pair_slot) {
  __shared__ unsigned ws[8];
  const size_t base = (size_t)blockIdx.x * SH_TILE + (size_t)threadIdx.x * 8;
  unsigned f[8];
  const unsigned mine = sh_lead_flags(keys, cell_start, n, base, f);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  unsigned x = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[w] = x;
  __syncthreads();
  unsigned before = x - mine;  // leaders before this thread inside its warp
  for (int v = 0; v < w; v++) before += ws[v];
  size_t q = (size_t)tile_off[blockIdx.x] + before;
#pragma unroll
  for (int e = 0; e < 8; e++)
    if (f[e]) pair_slot[q++] = (int32_t)(base + e);
}

// :83-152.  Thread q takes the q-th target pair: the cell-ordered slots (k0, k0 + 1), k0 = pair_slot[q], as one
// packed fp32x2 pair when both lie in the same cell, else slot k0 alone (last body of a cell with an odd count).
// The 27 neighbour cells are 9 contiguous x-runs of the cell-ordered array; every candidate is one LDG.128,
// broadcast into both halves: un-softened r^2 with the reference's unfused mul/add sequence (:128-130), cutoff
// test per half, softened monopole, weight forced to 0 outside the cutoff.  A body is inside its own cell's
// run: its own term is G m * rsqrt(eps^2)^3 * 0 = 0 exactly for eps > 0 and is zeroed by a select for eps = 0
// (as are exact coincidences).  Candidates are accumulated in run order, i.e. cell order z, y, x then slot
// (:104-150).
#ifdef NB_SH_MIN_BLOCKS  // build-time tuning knob: resident CTAs per SM asked of ptxas (register cap)
#define NB_SH_BOUNDS __launch_bounds__(128, NB_SH_MIN_BLOCKS)
#else
#define NB_SH_BOUNDS __launch_bounds__(128)
#endif
// FORCE / COUNT: the timed force evaluation does not count the in-cutoff pairs (four integer instructions per
// candidate, each paid at FP32-pipe rates on this SM: DESIGN.md 3.5); nb_sh_stats counts them on demand with the
// COUNT-only instantiation, same loop, same test.
template <bool EPS0, bool FORCE, bool COUNT>
__global__ void NB_SH_BOUNDS sh_force_kernel(size_t n, const float4* __restrict__ spos,
                                                       const uint32_t* __restrict__ keys,
                                                       const int32_t* __restrict__ sidx,
                                                       const int32_t* __restrict__ pair_slot,
                                                       const int32_t* __restrict__ cell_start,
                                                       const int32_t* __restrict__ cell_end,
                                                       const SHMisc* __restrict__ m, float cell, float cutoff2, float G,
                                                       float eps2, float* __restrict__ ax, float* __restrict__ ay,
                                                       float* __restrict__ az, unsigned long long* __restrict__ pairs) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned cnt = 0;
  size_t k0 = n;
  bool two = false;
  if (t < (size_t)m->n_pairs) {
    k0 = (size_t)pair_slot[t];
    two = (k0 + 1 < n) && (keys[k0 + 1] == keys[k0]);
  }
  if (k0 < n) {
    const int nx = m->dims[0], ny = m->dims[1], nz = m->dims[2];
    // scalar loads: each lands in its half of a packed pair (see direct.cu)
    const float* p0 = reinterpret_cast<const float*>(spos + k0);
    const float* p1 = reinterpret_cast<const float*>(spos + (two ? k0 + 1 : k0));
    const float mex = __ldg(p0), mey = __ldg(p0 + 1), mez = __ldg(p0 + 2);
    const f2 xi = pk(mex, __ldg(p1)), yi = pk(mey, __ldg(p1 + 1)), zi = pk(mez, __ldg(p1 + 2));
    // own cell recomputed WITHOUT clamp (:100-102); the pair partner has the same cell
    const int cx = (int)floorf(__fdiv_rn(__fsub_rn(mex, m->bbox[0]), cell));
    const int cy = (int)floorf(__fdiv_rn(__fsub_rn(mey, m->bbox[1]), cell));
    const int cz = (int)floorf(__fdiv_rn(__fsub_rn(mez, m->bbox[2]), cell));
    const f2 e2 = pk(eps2, eps2);
    f2 sx = pk(0.f, 0.f), sy = sx, sz = sx;
#if NB_SH_SATMASK
    const float sat_big = __fdividef(1.2676506e30f, fmaxf(cutoff2, 1e-30f)), sat_c = sat_big * cutoff2;
#endif
    for (int dz = -1; dz <= 1; dz++) {
      const int z = cz + dz;
      if (z < 0 || z >= nz) continue;
      for (int dy = -1; dy <= 1; dy++) {
        const int y = cy + dy;
        if (y < 0 || y >= ny) continue;
        // the in-grid part of x-1..x+1 is one contiguous run of the cell-ordered array
        const int x0 = max(cx - 1, 0), x1 = min(cx + 1, nx - 1);
        if (x0 > x1) continue;
        const int row = y * nx + z * nx * ny;
        const int j0 = cell_start[row + x0], j1 = cell_end[row + x1];
        NB_SH_PRAGMA_UNROLL(NB_SH_UNROLL)
        for (int jj = j0; jj < j1; jj++) {
          const float4 c = __ldg(spos + jj);
          const f2 ddx = sub2(pk(c.x, c.x), xi), ddy = sub2(pk(c.y, c.y), yi), ddz = sub2(pk(c.z, c.z), zi);
          // dx*dx + dy*dy + dz*dz (:131) as the reference's binary evaluates it (nvcc contracts it to
          // FMUL dy*dy, FFMA dx*dx + ., FFMA dz*dz + . -- SASS of oracle/_ref): the cutoff decision below is
          // bit-identical to the reference's
          const f2 d2 = fma2(ddz, ddz, fma2(ddx, ddx, mul2(ddy, ddy)));
          float r0, r1;
          upk(d2, r0, r1);
          const bool in0 = (r0 < cutoff2), in1 = (r1 < cutoff2);  // un-softened test (:131-135)
#if NB_SH_SATMASK
          const float r2a = r0, r2b = r1;
#endif
          if (COUNT) cnt += (in0 ? 1u : 0u) + ((in1 && two) ? 1u : 0u);
          if (FORCE) {
            upk(add2(d2, e2), r0, r1);
            float i0 = rsq(r0), i1 = rsq(r1);
            if (EPS0) {
              i0 = (r0 > eps2) ? i0 : 0.f;
              i1 = (r1 > eps2) ? i1 : 0.f;
            }
#if NB_SH_SATMASK
            // experiment: the cutoff as a saturated multiply-add on the FP32 pipe (mask = 1 inside, 0 outside or on
            // the cutoff; BIG = 2^100 / cutoff^2, so one ulp of r^2 below the cutoff already saturates) instead of
            // two compares and two selects on the integer pipe
            const float k0_ = __saturatef(fmaf(-sat_big, r2a, sat_c)), k1_ = __saturatef(fmaf(-sat_big, r2b, sat_c));
            const f2 inv = mul2(pk(i0, i1), pk(k0_, k1_));
#else
            i0 = in0 ? i0 : 0.f;
            i1 = in1 ? i1 : 0.f;
            const f2 inv = pk(i0, i1);
#endif
            const float gm = G * c.w;
            const f2 f = mul2(pk(gm, gm), mul2(mul2(inv, inv), inv));
            sx = fma2(f, ddx, sx);
            sy = fma2(f, ddy, sy);
            sz = fma2(f, ddz, sz);
          }
        }
      }
    }
    if (FORCE) {
      float a, b;
      const int i = sidx[k0];
      const int i2 = two ? sidx[k0 + 1] : i;
      upk(sx, a, b); ax[i] = a; if (two) ax[i2] = b;
      upk(sy, a, b); ay[i] = a; if (two) ay[i2] = b;
      upk(sz, a, b); az[i] = a; if (two) az[i2] = b;
    }
    if (COUNT && cutoff2 > 0.f) cnt -= two ? 2u : 1u;  // the bodies themselves (inside their own run; r = 0 passes the test unless cutoff = 0)
  }
  if (COUNT) {  // one atomic per warp
    unsigned long long c = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(pairs, c);
  }
}

static SH* sh_get(nb_context* ctx) {
  if (!ctx->sh) ctx->sh = new (std::nothrow) SH();
  return ctx->sh;
}

void sh_destroy(nb_context* ctx) {
  if (!ctx->sh) return;
  SH* s = ctx->sh;
  s->keys_alt.release(); s->idx_alt.release(); s->gaps.release();
  s->particle_cell.release(); s->keys.release(); s->sorted_idx.release(); s->packed.release(); s->spos.release(); s->pair_slot.release(); s->tile_sum.release();
  s->cell_start.release(); s->cell_end.release(); s->misc.release();
  if (s->h_misc) cudaFreeHost(s->h_misc);
  delete s;
  ctx->sh = nullptr;
}

int sh_build(nb_context* ctx, const nb_particles* d, float cell, cudaStream_t st) {
  SH* s = sh_get(ctx);
  if (!s) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  const size_t n = d->count;
  s->built = false;
  s->stats_valid = false;
  s->stats_cutoff2 = -1.f;
  if (!(cell > 0.f) || std::isnan(cell) || std::isinf(cell)) {
    set_error("Spatial hash cell size must be positive and finite");
    return NB_ERR_INVALID;
  }
  if (n == 0) return NB_OK;
  if (n > 0x7ffffff0ull) {
    set_error("spatial hash: particle count %zu too large", n);
    return NB_ERR_INVALID;
  }
  s->n = n;
  s->cell_size = cell;
  if (int rc = s->particle_cell.ensure(n * 4)) return rc;
  if (int rc = s->keys.ensure(n * 4)) return rc;
  if (int rc = s->sorted_idx.ensure(n * 4)) return rc;
  if (int rc = s->keys_alt.ensure(n * 4)) return rc;
  if (int rc = s->idx_alt.ensure(n * 4)) return rc;
  if (int rc = s->spos.ensure(n * sizeof(float4))) return rc;
  if (int rc = s->pair_slot.ensure((n + 1) * 4)) return rc;
  if (int rc = s->tile_sum.ensure(((n + SH_TILE - 1) / SH_TILE) * sizeof(unsigned long long))) return rc;
  if (int rc = s->packed.ensure(n * sizeof(float4))) return rc;
  if (int rc = s->misc.ensure(sizeof(SHMisc))) return rc;
  if (!s->h_misc) NB_CUDA(cudaMallocHost(&s->h_misc, sizeof(SHMisc)));
  SHMisc* misc = s->misc.as<SHMisc>();
  {
    NB_STAGE(ctx, st, "sh.bbox");
    if (int rc = bbox_device(ctx, d, st)) return rc;
    sh_prepare_kernel<<<1, 32, 0, st>>>(ctx->bbox_dev.as<int>(), cell, misc);
    NB_LAUNCH_CHECK(ctx);
  }
  // the grid size decides the allocation: one small readback (the reference does 12 here)
  NB_CUDA(cudaMemcpyAsync(s->h_misc, misc, sizeof(SHMisc), cudaMemcpyDeviceToHost, st));
  NB_CUDA(cudaStreamSynchronize(st));
  if (s->h_misc->too_large) {
    set_error("Spatial hash grid too large: reduce cell_size or bounding box");  // :252-254
    return NB_ERR_GRID_TOO_LARGE;
  }
  s->total_cells = (int)s->h_misc->total;
  memcpy(s->dims, s->h_misc->dims, sizeof s->dims);
  memcpy(s->bbox, s->h_misc->bbox, sizeof s->bbox);
  if (int rc = s->cell_start.ensure((size_t)s->total_cells * 4)) return rc;
  if (int rc = s->cell_end.ensure((size_t)s->total_cells * 4)) return rc;
  {  // long runs of empty cells: one per key change at most, each longer than SH_GAP_INLINE cells
    size_t max_gaps = (size_t)s->total_cells / (SH_GAP_INLINE + 1) + 1;
    if (max_gaps > n + 1) max_gaps = n + 1;
    if (int rc = s->gaps.ensure(max_gaps * sizeof(SHGap))) return rc;
  }
  const unsigned gn = (unsigned)((n + 255) / 256);
  {
    NB_STAGE(ctx, st, "sh.assign");
    sh_assign_kernel<<<gn, 256, 0, st>>>(d->pos_x, d->pos_y, d->pos_z, d->mass, n, cell, misc,
                                         s->particle_cell.as<int32_t>(), s->keys.as<uint32_t>(),
                                         s->packed.as<float4>());
    NB_LAUNCH_CHECK(ctx);
  }
  int bits = 1;
  while (bits < 32 && (1ll << bits) < (long long)s->total_cells) bits++;
  {
    NB_STAGE(ctx, st, "sh.sort");
    if (int rc = sort_pairs_db(ctx, s->keys.as<uint32_t>(), s->sorted_idx.as<int32_t>(), s->keys_alt.as<uint32_t>(),
                               s->idx_alt.as<int32_t>(), n, bits, true, &s->keys_sorted, &s->idx_sorted, st))
      return rc;
  }
  {
    NB_STAGE(ctx, st, "sh.ranges_gather");
    sh_ranges_gather_kernel<<<(unsigned)((n + 1 + 255) / 256), 256, 0, st>>>(
        s->keys_sorted, n, s->total_cells, s->cell_start.as<int32_t>(), s->cell_end.as<int32_t>(), s->idx_sorted,
        s->packed.as<float4>(), s->spos.as<float4>(), s->gaps.as<SHGap>(), &misc->n_gaps);
    NB_LAUNCH_CHECK(ctx);
    sh_fill_gaps_kernel<<<ctx->sm_count * 4, 256, 0, st>>>(s->gaps.as<SHGap>(), &misc->n_gaps,
                                                           s->cell_start.as<int32_t>(), s->cell_end.as<int32_t>());
    NB_LAUNCH_CHECK(ctx);
  }
  NB_STAGE(ctx, st, "sh.pair_list");
  const int ntiles = (int)((n + SH_TILE - 1) / SH_TILE);
  unsigned long long* tile_sum = s->tile_sum.as<unsigned long long>();
  sh_pair_sums_kernel<<<ntiles, 256, 0, st>>>(s->keys_sorted, s->cell_start.as<int32_t>(), n, tile_sum);
  NB_LAUNCH_CHECK(ctx);
  if (int rc = launch_scan_tiles(ctx, tile_sum, ntiles, &misc->n_pairs, st)) return rc;
  sh_pair_fill_kernel<<<ntiles, 256, 0, st>>>(s->keys_sorted, s->cell_start.as<int32_t>(), n, tile_sum,
                                              s->pair_slot.as<int32_t>());
  NB_LAUNCH_CHECK(ctx);
  s->built = true;
  return NB_OK;
}

int sh_forces(nb_context* ctx, const nb_particles* d, float cutoff, float G, float eps, cudaStream_t st) {
  SH* s = sh_get(ctx);
  if (!s) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  const size_t n = d->count;
  if (n == 0) return NB_OK;
  if (!s->built || s->n != n) {
    set_error("spatial hash: forces requested without a grid built for these particles");
    return NB_ERR_INVALID;
  }
  SHMisc* misc = s->misc.as<SHMisc>();
  s->stats_cutoff2 = cutoff * cutoff;  // nb_sh_stats counts the pairs of this evaluation on demand
  s->stats_valid = false;
  const float eps2 = eps * eps;
  ProfScope prof(ctx, st);
  NB_STAGE(ctx, st, "sh.force");
  // pairs = sum over cells of ceil(count / 2) <= (n + non-empty cells) / 2; threads past the device-side count exit
  size_t nonempty_bound = n;
  if ((size_t)s->total_cells < nonempty_bound) nonempty_bound = (size_t)s->total_cells;
  const size_t threads = (n + nonempty_bound + 1) / 2;
#define NB_SH_LAUNCH(E_)                                                                                             \
  sh_force_kernel<E_, true, false><<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(                              \
      n, s->spos.as<float4>(), s->keys_sorted, s->idx_sorted, s->pair_slot.as<int32_t>(),                            \
      s->cell_start.as<int32_t>(), s->cell_end.as<int32_t>(), misc, s->cell_size, cutoff * cutoff, G, eps2,         \
      d->acc_x, d->acc_y, d->acc_z, &misc->pairs)
  if (eps2 >= NB_EPS2_GUARD) NB_SH_LAUNCH(false);
  else NB_SH_LAUNCH(true);
#undef NB_SH_LAUNCH
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_sh_build(nb_context* ctx, const nb_particles* d, float cell_size, nb_stream s) {
  if (!ctx || !d || (d->count && (!d->pos_x || !d->pos_y || !d->pos_z || !d->mass))) {
    set_error("spatial hash build: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  return sh_build(ctx, d, cell_size, (cudaStream_t)s);
}

int nb_sh_forces(nb_context* ctx, const nb_particles* d, float cutoff, float G, float eps, nb_stream s) {
  if (!ctx || !d || (d->count && (!d->acc_x || !d->acc_y || !d->acc_z))) {
    set_error("spatial hash forces: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  return sh_forces(ctx, d, cutoff, G, eps, (cudaStream_t)s);
}

int nb_sh_stats(nb_context* ctx, uint64_t* pairs) {
  if (!ctx || !ctx->sh || !ctx->sh->built || !pairs) {
    set_error("spatial hash: no grid built");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  NB_CUDA(cudaDeviceSynchronize());
  SH* s = ctx->sh;
  SHMisc* misc = s->misc.as<SHMisc>();
  if (s->stats_cutoff2 < 0.f) {  // no force evaluation on this grid yet
    *pairs = 0;
    return NB_OK;
  }
  if (!s->stats_valid) {
    // counted on demand: the same sweep and the same test as the force evaluation, without the force arithmetic
    // (the cell-ordered bodies of the last build are still in place)
    const size_t n = s->n;
    NB_CUDA(cudaMemset(&misc->pairs, 0, sizeof(unsigned long long)));
    size_t nonempty_bound = n;
    if ((size_t)s->total_cells < nonempty_bound) nonempty_bound = (size_t)s->total_cells;
    const size_t threads = (n + nonempty_bound + 1) / 2;
    sh_force_kernel<false, false, true><<<(unsigned)((threads + 127) / 128), 128>>>(
        n, s->spos.as<float4>(), s->keys_sorted, s->idx_sorted, s->pair_slot.as<int32_t>(),
        s->cell_start.as<int32_t>(), s->cell_end.as<int32_t>(), misc, s->cell_size, s->stats_cutoff2, 0.f, 0.f,
        nullptr, nullptr, nullptr, &misc->pairs);
    NB_LAUNCH_CHECK(ctx);
    NB_CUDA(cudaDeviceSynchronize());
    s->stats_valid = true;
  }
  NB_CUDA(cudaMemcpy(pairs, &misc->pairs, 8, cudaMemcpyDeviceToHost));
  return NB_OK;
}

int nb_sh_total_cells(nb_context* ctx, int* total_cells) {
  if (!ctx || !ctx->sh || !ctx->sh->built || !total_cells) {
    set_error("spatial hash: no grid built");
    return NB_ERR_INVALID;
  }
  *total_cells = ctx->sh->total_cells;
  return NB_OK;
}

int nb_sh_export(nb_context* ctx, int32_t dims[3], float bbox_padded[6], int32_t* particle_cell, int32_t* cell_start,
                 int32_t* cell_end, int32_t* sorted_indices) {
  if (!ctx || !ctx->sh || !ctx->sh->built) {
    set_error("spatial hash export: no grid built");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  NB_CUDA(cudaDeviceSynchronize());
  SH* s = ctx->sh;
  if (dims) memcpy(dims, s->dims, sizeof s->dims);
  if (bbox_padded) memcpy(bbox_padded, s->bbox, sizeof s->bbox);
  if (particle_cell) NB_CUDA(cudaMemcpy(particle_cell, s->particle_cell.p, s->n * 4, cudaMemcpyDeviceToHost));
  if (cell_start) NB_CUDA(cudaMemcpy(cell_start, s->cell_start.p, (size_t)s->total_cells * 4, cudaMemcpyDeviceToHost));
  if (cell_end) NB_CUDA(cudaMemcpy(cell_end, s->cell_end.p, (size_t)s->total_cells * 4, cudaMemcpyDeviceToHost));
  if (sorted_indices) NB_CUDA(cudaMemcpy(sorted_indices, s->idx_sorted, s->n * 4, cudaMemcpyDeviceToHost));
  return NB_OK;
}

}  // extern "C"
