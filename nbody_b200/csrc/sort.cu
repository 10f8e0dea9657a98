// sort.cu -- bounding box reduction and a CUB-free stable LSD radix sort of (u32 key, i32 value)
// pairs.  Replaces computeBoundingBoxKernel + its 6 malloc / 12 memcpy wrapper
// (reference src/cuda/force_barnes_hut.cu:66-110, 218-258) and thrust::sort_by_key (:276-280).
// Both are HBM-bound: bbox reads 12 B/body; the sort reads the keys once for all digit histograms
// and then moves 16 B/pair per 8-bit pass.
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace nb {

// ---- bounding box --------------------------------------------------------------------------------
// floats are reduced through an order-preserving int encoding so one atomicMin/atomicMax per block
// per component suffices (the reference uses CAS loops on floats).
__device__ __forceinline__ int f2ord(float f) {
  const int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__host__ __device__ __forceinline__ float ord2f(int i) {
  const int j = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
  return __int_as_float(j);
#else
  float f;
  memcpy(&f, &j, 4);
  return f;
#endif
}

__global__ void bbox_init_kernel(int* bb) {
  if (threadIdx.x < 3) bb[threadIdx.x] = 0x7fffffff;            // +max
  else if (threadIdx.x < 6) bb[threadIdx.x] = (int)0x80000000;  // -max
}

__global__ void __launch_bounds__(256) bbox_kernel(const float* __restrict__ px, const float* __restrict__ py,
                                                   const float* __restrict__ pz, size_t n, int* __restrict__ bb) {
  float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float x = px[i], y = py[i], z = pz[i];
    mn[0] = fminf(mn[0], x); mx[0] = fmaxf(mx[0], x);
    mn[1] = fminf(mn[1], y); mx[1] = fmaxf(mx[1], y);
    mn[2] = fminf(mn[2], z); mx[2] = fmaxf(mx[2], z);
  }
  __shared__ float sh[6][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const float lo = warp_min(mn[a]), hi = warp_max(mx[a]);
    if (lane == 0) {
      sh[a][w] = lo;
      sh[3 + a][w] = hi;
    }
  }
  __syncthreads();
  if (w == 0) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      float lo = lane < 8 ? sh[a][lane] : 3.0e38f, hi = lane < 8 ? sh[3 + a][lane] : -3.0e38f;
      lo = warp_min(lo);
      hi = warp_max(hi);
      if (lane == 0) {
        atomicMin(&bb[a], f2ord(lo));
        atomicMax(&bb[3 + a], f2ord(hi));
      }
    }
  }
}

// leaves 6 order-encoded ints in ctx->bbox_dev (min xyz, max xyz)
int bbox_device(nb_context* ctx, const nb_particles* d, cudaStream_t st) {
  if (int rc = ctx->bbox_dev.ensure(64)) return rc;
  int* bb = ctx->bbox_dev.as<int>();
  bbox_init_kernel<<<1, 32, 0, st>>>(bb);
  NB_LAUNCH_CHECK(ctx);
  const size_t n = d->count;
  if (n) {
    size_t grid = (n + 255) / 256;
    const size_t cap = (size_t)ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    bbox_kernel<<<(unsigned)grid, 256, 0, st>>>(d->pos_x, d->pos_y, d->pos_z, n, bb);
    NB_LAUNCH_CHECK(ctx);
  }
  return NB_OK;
}

int bbox_to_host(nb_context* ctx, float out[6], cudaStream_t st) {
  if (!ctx->h_bbox) NB_CUDA(cudaMallocHost(&ctx->h_bbox, 64));
  NB_CUDA(cudaMemcpyAsync(ctx->h_bbox, ctx->bbox_dev.p, 6 * sizeof(int), cudaMemcpyDeviceToHost, st));
  NB_CUDA(cudaStreamSynchronize(st));
  const int* hi = reinterpret_cast<const int*>(ctx->h_bbox);
  for (int a = 0; a < 6; a++) out[a] = ord2f(hi[a]);
  return NB_OK;
}

// ---- radix sort ("onesweep": one histogram read + one read/write sweep per 8-bit digit) ---------------
// Stable LSD.  rs_hist_all reads the keys once and builds the digit histograms of every pass.  Each
// pass is ONE kernel: a block takes the next 4096/8192-key tile (ticket order), ranks its keys per
// digit (each warp owns 512 consecutive keys and ranks them 32 at a time with __match_any_sync, so
// equal digits keep input order), publishes its per-digit counts and resolves the counts of all
// earlier tiles by decoupled look-back over a per-tile status word (2 flag bits + 30-bit count, one
// thread per digit), then moves keys and values through shared memory in digit order so that every
// global store is a run of consecutive addresses.  Algorithmic traffic: 4 + 16 x passes B/pair
// (68 B for 30-bit Morton keys) against 80 B for the histogram/scan/scatter-per-pass design it replaces.
constexpr int RS_ITEMS = 16;
constexpr int RS_WARP_KEYS = RS_ITEMS * 32;  // 512 consecutive keys per warp
constexpr uint32_t RS_FLAG_AGG = 1u << 30, RS_FLAG_INC = 2u << 30, RS_VAL_MASK = (1u << 30) - 1u;
constexpr int RS_MAX_PASSES = 4;

// hist[p][d] += number of keys whose pass-p digit is d
__global__ void __launch_bounds__(512) rs_hist_all_kernel(const uint32_t* __restrict__ keys, size_t n, int passes,
                                                          int end_bit, uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[RS_MAX_PASSES][256];
  for (int k = threadIdx.x; k < RS_MAX_PASSES * 256; k += blockDim.x) (&sh[0][0])[k] = 0;
  __syncthreads();
  const uint32_t last_mask = (1u << (end_bit - 8 * (passes - 1))) - 1u;
  const size_t n4 = n / 4;
  const uint4* k4 = reinterpret_cast<const uint4*>(keys);
  auto add = [&](uint32_t k) {
#pragma unroll
    for (int p = 0; p < RS_MAX_PASSES; p++) {
      if (p < passes) {
        uint32_t d = (k >> (8 * p)) & 0xffu;
        if (p == passes - 1) d &= last_mask;
        atomicAdd(&sh[p][d], 1u);
      }
    }
  };
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = k4[i];
    add(v.x); add(v.y); add(v.z); add(v.w);
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n - n4 * 4)) add(keys[n4 * 4 + threadIdx.x]);
  __syncthreads();
  for (int k = threadIdx.x; k < passes * 256; k += blockDim.x) {
    const uint32_t c = (&sh[0][0])[k];
    if (c) atomicAdd(&hist[k], c);
  }
}

__device__ __forceinline__ uint32_t ld_status(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_status(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// one pass.  status: tiles x 256 words, zeroed; ticket: zeroed; hist: this pass's 256 global digit counts
template <int WARPS, bool MATCH>
__global__ void __launch_bounds__(WARPS * 32, 1024 / (WARPS * 32)) rs_onesweep_kernel(const uint32_t* __restrict__ keys_in,
                                                                 const int32_t* __restrict__ vals_in,
                                                                 uint32_t* __restrict__ keys_out,
                                                                 int32_t* __restrict__ vals_out, size_t n, int shift,
                                                                 uint32_t mask, const uint32_t* __restrict__ hist,
                                                                 uint32_t* __restrict__ status,
                                                                 uint32_t* __restrict__ ticket) {
  constexpr int THREADS = WARPS * 32;
  constexpr int TILE = WARPS * RS_WARP_KEYS;
  __shared__ uint32_t cnt[WARPS][256];  // per-warp digit counts, then exclusive offsets across warps
  __shared__ uint32_t lstart[256];      // first slot of the digit inside the tile's digit-ordered buffer
  __shared__ uint32_t gbase[256];       // global position of the tile's first key of the digit
  extern __shared__ __align__(16) uint32_t rs_buf[];  // keys then values of the tile in digit order
  uint32_t* buf_k = rs_buf;
  uint32_t* buf_v = rs_buf + TILE;
  __shared__ uint32_t wsum[8];
  __shared__ uint32_t tile_s;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) tile_s = atomicAdd(ticket, 1u);
  for (int k = tid; k < WARPS * 256; k += THREADS) (&cnt[0][0])[k] = 0;
  __syncthreads();
  const uint32_t tile = tile_s;
  const size_t tbase = (size_t)tile * TILE;
  const uint32_t tcount = (uint32_t)((n - tbase) < (size_t)TILE ? (n - tbase) : (size_t)TILE);

  // ---- load + rank inside the warp's 512 consecutive keys -----------------------------------------
  // peers = lanes holding the same digit, from 8 independent ballots (pipelined over the 16 keys; the
  // MATCH instruction serialises over the distinct values in the warp).  The lowest peer adds the
  // group's size to the warp's digit counter with one shared-memory atomic -- a warp's atomics retire in
  // issue order, so key k sees the counts of keys < k without any warp synchronisation -- and hands the
  // old value to its peers by shuffle.
  const size_t wbase = tbase + (size_t)w * RS_WARP_KEYS;
  uint32_t key[RS_ITEMS];
  uint32_t pos[RS_ITEMS];  // rank among the warp's keys of the same digit
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const size_t i = wbase + (size_t)k * 32 + lane;
    key[k] = (i < n) ? keys_in[i] : 0xffffffffu;
  }
  const uint32_t full_valid = (wbase + RS_WARP_KEYS <= n) ? 1u : 0u;
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const size_t i = wbase + (size_t)k * 32 + lane;
    const bool ok = full_valid || (i < n);
    const uint32_t dgt = (key[k] >> shift) & mask;
    uint32_t peers;
    if (MATCH) {
      peers = __match_any_sync(0xffffffffu, ok ? dgt : 0x100u);  // padding forms its own group
    } else {
      peers = __ballot_sync(0xffffffffu, ok);  // padding lanes are nobody's peer
#pragma unroll
      for (int b = 0; b < 8; b++) {
        if ((mask >> b) & 1u) {  // warp-uniform: a 6-bit last pass needs 6 ballots
          const uint32_t v = __ballot_sync(0xffffffffu, (dgt >> b) & 1u);
          peers &= ((dgt >> b) & 1u) ? v : ~v;
        }
      }
    }
    const int leader = __ffs(peers) - 1;
    uint32_t before = 0;
    if (ok && lane == leader) before = atomicAdd(&cnt[w][dgt], (uint32_t)__popc(peers));
    before = __shfl_sync(0xffffffffu, before, leader < 0 ? 0 : leader);
    pos[k] = before + __popc(peers & lt);
  }
  __syncthreads();

  // ---- per digit: offsets across warps, tile count; publish, look back -----------------------------
  uint32_t my_total = 0;
  if (tid < 256) {
#pragma unroll
    for (int k = 0; k < WARPS; k++) {
      const uint32_t c = cnt[k][tid];
      cnt[k][tid] = my_total;
      my_total += c;
    }
    uint32_t* mine = status + (size_t)tile * 256 + tid;
    if (tile == 0) st_status(mine, RS_FLAG_INC | my_total);
    else st_status(mine, RS_FLAG_AGG | my_total);
  }
  // exclusive scan of the 256 tile counts (slot of each digit in buf) and of the global histogram
  {
    uint32_t x = (tid < 256) ? my_total : 0u, g = (tid < 256) ? hist[tid] : 0u;
    const uint32_t x0 = x, g0 = g;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o), h = __shfl_up_sync(0xffffffffu, g, o);
      if (lane >= o) { x += y; g += h; }
    }
    __shared__ uint32_t gsum[8];
    if (tid < 256 && lane == 31) { wsum[w] = x; gsum[w] = g; }
    __syncthreads();
    if (tid < 256) {
      uint32_t px = 0, pg = 0;
      for (int k = 0; k < w; k++) { px += wsum[k]; pg += gsum[k]; }
      lstart[tid] = px + x - x0;
      uint32_t excl = 0;  // keys of this digit in earlier tiles
      if (tile > 0) {
        const uint32_t* look = status + (size_t)(tile - 1) * 256 + tid;
        for (;;) {
          const uint32_t s = ld_status(look);
          if (s & RS_FLAG_INC) { excl += s & RS_VAL_MASK; break; }
          if (s & RS_FLAG_AGG) { excl += s & RS_VAL_MASK; look -= 256; }
        }
        st_status(status + (size_t)tile * 256 + tid, RS_FLAG_INC | (excl + my_total));
      }
      gbase[tid] = pg + g - g0 + excl;
    }
  }
  __syncthreads();

  // ---- registers -> buf (digit order) -> global (runs of consecutive addresses) ---------------------------
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const size_t i = wbase + (size_t)k * 32 + lane;
    const uint32_t dgt = (key[k] >> shift) & mask;
    if (i < n) {
      const uint32_t slot = lstart[dgt] + cnt[w][dgt] + pos[k];
      buf_k[slot] = key[k];
      buf_v[slot] = (uint32_t)vals_in[i];
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const uint32_t p = (uint32_t)k * THREADS + tid;
    if (p < tcount) {
      const uint32_t kk = buf_k[p];
      const uint32_t dgt = (kk >> shift) & mask;
      const uint32_t g = gbase[dgt] + (p - lstart[dgt]);
      keys_out[g] = kk;
      vals_out[g] = (int32_t)buf_v[p];
    }
  }
}

// sorts in place (result in d_keys/d_vals); bits [0, end_bit) are significant
int sort_pairs(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, size_t n, int end_bit, cudaStream_t st) {
  if (n < 2) return NB_OK;
  if (end_bit < 1) end_bit = 1;
  if (end_bit > 32) end_bit = 32;
  if (n > (size_t)RS_VAL_MASK) {
    set_error("sort: too many elements");
    return NB_ERR_INVALID;
  }
  static bool attr_set_dev[64] = {};  // function attributes are per device: a process may hold a context on several
  bool& attr_set = attr_set_dev[ctx->device & 63];
  if (!attr_set) {
    NB_CUDA(cudaFuncSetAttribute(rs_onesweep_kernel<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * RS_WARP_KEYS * 8));
    NB_CUDA(cudaFuncSetAttribute(rs_onesweep_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16 * RS_WARP_KEYS * 8));
    NB_CUDA(cudaFuncSetAttribute(rs_onesweep_kernel<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * RS_WARP_KEYS * 8));
    NB_CUDA(cudaFuncSetAttribute(rs_onesweep_kernel<16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16 * RS_WARP_KEYS * 8));
    attr_set = true;
  }
  const char* e = getenv("NB_SORT_WARPS");  // tuning knob: 8 (4096-key tiles) or 16 (8192-key tiles)
  const int warps_cfg = (e && atoi(e) == 8) ? 8 : 16;
  const char* me = getenv("NB_SORT_MATCH");  // tuning knob: 1 = MATCH.ANY instruction, 0 = eight ballots
  const bool use_match = me && atoi(me) == 1;
  const int warps = warps_cfg;
  const int tile = warps * RS_WARP_KEYS;
  const size_t smem = (size_t)tile * 8;
  const int passes = (end_bit + 7) / 8;
  const size_t ntiles = (n + tile - 1) / tile;
  if (int rc = ctx->sort_keys2.ensure(n * sizeof(uint32_t))) return rc;
  if (int rc = ctx->sort_vals2.ensure(n * sizeof(int32_t))) return rc;
  // scratch: [passes][256] histograms, [passes] tickets (padded to 64 words), [passes][ntiles][256] status
  const size_t head = (size_t)RS_MAX_PASSES * 256 + 64;
  const size_t words = head + (size_t)passes * ntiles * 256;
  if (int rc = ctx->sort_hist.ensure(words * sizeof(uint32_t))) return rc;
  uint32_t* hist = ctx->sort_hist.as<uint32_t>();
  uint32_t* ticket = hist + RS_MAX_PASSES * 256;
  uint32_t* status = hist + head;
  NB_CUDA(cudaMemsetAsync(hist, 0, words * sizeof(uint32_t), st));
  {
    size_t g = (n / 4 + 511) / 512;
    const size_t cap = (size_t)ctx->sm_count * 4;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    rs_hist_all_kernel<<<(unsigned)g, 512, 0, st>>>(d_keys, n, passes, end_bit, hist);
    NB_LAUNCH_CHECK(ctx);
  }
  uint32_t *ka = d_keys, *kb = ctx->sort_keys2.as<uint32_t>();
  int32_t *va = d_vals, *vb = ctx->sort_vals2.as<int32_t>();
  for (int p = 0; p < passes; p++) {
    const int shift = p * 8;
    const int bits = (end_bit - shift) < 8 ? (end_bit - shift) : 8;
    const uint32_t mask = (1u << bits) - 1u;
    uint32_t* stp = status + (size_t)p * ntiles * 256;
#define NB_RS_LAUNCH(W_, M_)                                                                                   \
  rs_onesweep_kernel<W_, M_><<<(unsigned)ntiles, W_ * 32, smem, st>>>(ka, va, kb, vb, n, shift, mask, hist + p * 256, \
                                                                      stp, ticket + p)
    if (warps == 8) {
      if (use_match) NB_RS_LAUNCH(8, true);
      else NB_RS_LAUNCH(8, false);
    } else {
      if (use_match) NB_RS_LAUNCH(16, true);
      else NB_RS_LAUNCH(16, false);
    }
#undef NB_RS_LAUNCH
    NB_LAUNCH_CHECK(ctx);
    uint32_t* tk = ka; ka = kb; kb = tk;
    int32_t* tv = va; va = vb; vb = tv;
  }
  if (ka != d_keys) {
    NB_CUDA(cudaMemcpyAsync(d_keys, ka, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    NB_CUDA(cudaMemcpyAsync(d_vals, va, n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  }
  return NB_OK;
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_sort_pairs(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, size_t n, int end_bit, nb_stream s) {
  if (!ctx || (n && (!d_keys || !d_vals))) {
    set_error("sort: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  return sort_pairs(ctx, d_keys, d_vals, n, end_bit, (cudaStream_t)s);
}

int nb_bounding_box(nb_context* ctx, const nb_particles* d, float out[6], nb_stream s) {
  if (!ctx || !d || !out || (d->count && (!d->pos_x || !d->pos_y || !d->pos_z))) {
    set_error("bounding box: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  if (int rc = bbox_device(ctx, d, (cudaStream_t)s)) return rc;
  return bbox_to_host(ctx, out, (cudaStream_t)s);
}

}  // extern "C"
