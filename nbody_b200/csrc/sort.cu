// sort.cu -- bounding box reduction and a CUB-free stable LSD radix sort of (u32 key, i32 value)
// pairs.  Replaces computeBoundingBoxKernel + its 6 malloc / 12 memcpy wrapper
// (reference src/cuda/force_barnes_hut.cu:66-110, 218-258) and thrust::sort_by_key (:276-280).
// Both are HBM-bound: bbox reads 12 B/body; the sort moves 20 B/body per 8-bit pass
// (histogram pass reads the keys, scatter pass reads and writes key + value).
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

// ---- radix sort -------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 keys per block
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_WARP_KEYS = RS_TILE / RS_WARPS;  // 512 consecutive keys per warp

// hist[digit * nblocks + block] = number of keys of `digit` in the block's tile
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t* __restrict__ keys, size_t n, int shift,
                                                             uint32_t mask, uint32_t* __restrict__ hist, int nblocks) {
  __shared__ uint32_t sh[256];
  sh[threadIdx.x] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * RS_TILE;
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const size_t i = base + (size_t)k * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&sh[(keys[i] >> shift) & mask], 1u);
  }
  __syncthreads();
  hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = sh[threadIdx.x];
}

// one block per digit: exclusive scan of its row in place, row total to totals[digit]
__global__ void __launch_bounds__(1024) rs_rowscan_kernel(uint32_t* __restrict__ hist, int nblocks,
                                                          uint32_t* __restrict__ totals) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t carry_s;
  uint32_t* row = hist + (size_t)blockIdx.x * nblocks;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int b0 = 0; b0 < nblocks; b0 += 1024) {
    const int i = b0 + threadIdx.x;
    const uint32_t v = i < nblocks ? row[i] : 0u;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    if (w == 0) {
      uint32_t s = wsum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += y;
      }
      wsum[lane] = s;  // inclusive over warps
    }
    __syncthreads();
    const uint32_t carry = carry_s;
    const uint32_t excl = carry + (w ? wsum[w - 1] : 0u) + x - v;
    if (i < nblocks) row[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + wsum[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[blockIdx.x] = carry_s;
}

// stable scatter: warp w of a block owns 512 consecutive keys and ranks them 32 at a time with
// __match_any_sync; ranks are made block-wide by a scan over the per-warp digit counts.
__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(const uint32_t* __restrict__ keys_in,
                                                                const int32_t* __restrict__ vals_in,
                                                                uint32_t* __restrict__ keys_out,
                                                                int32_t* __restrict__ vals_out, size_t n, int shift,
                                                                uint32_t mask, const uint32_t* __restrict__ hist,
                                                                const uint32_t* __restrict__ totals, int nblocks) {
  __shared__ uint32_t cnt[RS_WARPS][256];
  __shared__ uint32_t base[256];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int k = threadIdx.x; k < RS_WARPS * 256; k += RS_THREADS) (&cnt[0][0])[k] = 0;
  // digit base = exclusive scan of the 256 totals (+ this block's offset inside the digit)
  {
    const uint32_t v = totals[threadIdx.x];
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    __shared__ uint32_t ws[8];
    if (lane == 31) ws[w] = x;
    __syncthreads();
    uint32_t pre = 0;
    for (int k = 0; k < w; k++) pre += ws[k];
    base[threadIdx.x] = pre + x - v + hist[(size_t)threadIdx.x * nblocks + blockIdx.x];
  }
  __syncthreads();

  const size_t wbase = (size_t)blockIdx.x * RS_TILE + (size_t)w * RS_WARP_KEYS;
  uint32_t key[RS_ITEMS];
  int32_t val[RS_ITEMS];
  uint32_t rank[RS_ITEMS];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const size_t i = wbase + (size_t)k * 32 + lane;
    const bool ok = i < n;
    key[k] = ok ? keys_in[i] : 0xffffffffu;
    val[k] = ok ? vals_in[i] : 0;
    const uint32_t dgt = (key[k] >> shift) & mask;
    const uint32_t peers = __match_any_sync(0xffffffffu, dgt);
    const uint32_t before = cnt[w][dgt];
    rank[k] = before + __popc(peers & lt);
    __syncwarp();
    if ((peers & lt) == 0) cnt[w][dgt] = before + __popc(peers);  // lowest peer lane updates the count
    __syncwarp();
  }
  __syncthreads();
  // per digit: turn per-warp counts into exclusive offsets across warps
  {
    uint32_t run = 0;
#pragma unroll
    for (int k = 0; k < RS_WARPS; k++) {
      const uint32_t c = cnt[k][threadIdx.x];
      cnt[k][threadIdx.x] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const size_t i = wbase + (size_t)k * 32 + lane;
    if (i < n) {
      const uint32_t dgt = (key[k] >> shift) & mask;
      const size_t pos = (size_t)base[dgt] + cnt[w][dgt] + rank[k];
      keys_out[pos] = key[k];
      vals_out[pos] = val[k];
    }
  }
}

// sorts in place (result in d_keys/d_vals); bits [0, end_bit) are significant
int sort_pairs(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, size_t n, int end_bit, cudaStream_t st) {
  if (n < 2) return NB_OK;
  if (end_bit < 1) end_bit = 1;
  if (end_bit > 32) end_bit = 32;
  if (n > 0xfffffff0ull) {
    set_error("sort: too many elements");
    return NB_ERR_INVALID;
  }
  const int passes = (end_bit + 7) / 8;
  const int nblocks = (int)((n + RS_TILE - 1) / RS_TILE);
  if (int rc = ctx->sort_keys2.ensure(n * sizeof(uint32_t))) return rc;
  if (int rc = ctx->sort_vals2.ensure(n * sizeof(int32_t))) return rc;
  if (int rc = ctx->sort_hist.ensure(((size_t)nblocks * 256 + 256) * sizeof(uint32_t))) return rc;
  uint32_t* hist = ctx->sort_hist.as<uint32_t>();
  uint32_t* totals = hist + (size_t)nblocks * 256;
  uint32_t *ka = d_keys, *kb = ctx->sort_keys2.as<uint32_t>();
  int32_t *va = d_vals, *vb = ctx->sort_vals2.as<int32_t>();
  for (int p = 0; p < passes; p++) {
    const int shift = p * 8;
    const int bits = (end_bit - shift) < 8 ? (end_bit - shift) : 8;
    const uint32_t mask = (1u << bits) - 1u;
    rs_hist_kernel<<<nblocks, RS_THREADS, 0, st>>>(ka, n, shift, mask, hist, nblocks);
    NB_LAUNCH_CHECK(ctx);
    rs_rowscan_kernel<<<256, 1024, 0, st>>>(hist, nblocks, totals);
    NB_LAUNCH_CHECK(ctx);
    rs_scatter_kernel<<<nblocks, RS_THREADS, 0, st>>>(ka, va, kb, vb, n, shift, mask, hist, totals, nblocks);
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
