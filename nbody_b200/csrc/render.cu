// render.cu -- headless point-splat renderer: the reference's GL point-sprite pipeline
// (src/render/renderer.cpp:16-77 shaders, :110-113 blend state, :184-224 draw call) restated as
// CUDA kernels writing a device RGBA8 framebuffer.  No window, no swapchain, no GL interop.
//
// GL semantics kept, per body (= one GL_POINTS vertex, drawn in body order):
//   viewPos = view * (p, 1); clip = projection * viewPos; culled unless -w <= x,y,z <= w;
//   size = clamp(pointSize / max(-viewPos.z, 0.1), size_min, size_max);  depth = -viewPos.z;
//   window centre (xw, yw) = ((ndc + 1) / 2) * (W, H), origin bottom-left;
//   fragments: every pixel whose centre lies inside the size x size square around (xw, yw);
//   gl_PointCoord = 0.5 + (pixel centre - window centre) / size (t flipped: upper-left origin);
//   fragment shader: discard if |coord - 0.5| > 0.5; colour ramp by mode; alpha = 1 - smoothstep(.4,.5,d);
//   blend: dst = src * a + dst * (1 - a) on all four channels, in body order, fp32, over the clear
//   colour; one round-to-nearest conversion to unorm8 at the end.
// Blending is order dependent, so the fragments are sorted by pixel with the STABLE radix sort of
// sort.cu (key = pixel, value = body; fragments are emitted in body order, so each pixel's run is in
// draw order) and one thread resolves each pixel: it walks back from the end of the run to the last
// fully opaque fragment (a = 1 replaces dst exactly) and blends forward from there.
// Every fp32 operation is an explicit round-to-nearest intrinsic in the order oracle/render.c spells
// out, so the framebuffer is bit-exact with the oracle.
#include <cstring>
#include <new>

#include "common.cuh"

namespace nb {
int sort_pairs(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, size_t n, int end_bit, cudaStream_t st);
}

struct RenderState {
  nb::DevBuf splat;    // per body: float4 {xw, yw, size, shade t} ; size < 0 = culled
  nb::DevBuf count;    // per body: fragments (u32), then exclusive offsets
  nb::DevBuf block;    // scan scratch
  nb::DevBuf keys, vals;
  unsigned long long* h_total = nullptr;  // pinned
  unsigned long long fragments = 0;
};

namespace nb {

__device__ __forceinline__ float clamp01(float t) { return fminf(fmaxf(t, 0.0f), 1.0f); }

// rows of a column-major 4x4 (glm layout: m[col*4 + row]) times (x, y, z, 1), left to right
__device__ __forceinline__ float mat_row(const float* m, int r, float x, float y, float z, float w) {
  return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[r], x), __fmul_rn(m[4 + r], y)), __fmul_rn(m[8 + r], z)),
                   __fmul_rn(m[12 + r], w));
}

// first and one-past-last pixel index whose centre (i + 0.5) lies in [c - h, c + h), clipped to [0, extent)
__device__ __forceinline__ void pixel_span(float c, float h, int extent, int& lo, int& hi) {
  const float a = ceilf(__fsub_rn(__fsub_rn(c, h), 0.5f)), b = ceilf(__fsub_rn(__fadd_rn(c, h), 0.5f));
  lo = (int)fmaxf(a, 0.0f);
  hi = (int)fminf(b, (float)extent);
  if (hi < lo) hi = lo;
}

// vertex stage + fragment counting
__global__ void __launch_bounds__(256) splat_project_kernel(nb_particles d, size_t n, nb_render_params p,
                                                            float4* __restrict__ splat, uint32_t* __restrict__ count) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = d.pos_x[i], y = d.pos_y[i], z = d.pos_z[i];
  const float vx = mat_row(p.view, 0, x, y, z, 1.0f), vy = mat_row(p.view, 1, x, y, z, 1.0f),
              vz = mat_row(p.view, 2, x, y, z, 1.0f), vw = mat_row(p.view, 3, x, y, z, 1.0f);
  const float cx = mat_row(p.proj, 0, vx, vy, vz, vw), cy = mat_row(p.proj, 1, vx, vy, vz, vw),
              cz = mat_row(p.proj, 2, vx, vy, vz, vw), cw = mat_row(p.proj, 3, vx, vy, vz, vw);
  float4 s = make_float4(0.f, 0.f, -1.f, 0.f);
  uint32_t cnt = 0;
  const bool inside = (cx >= -cw && cx <= cw) && (cy >= -cw && cy <= cw) && (cz >= -cw && cz <= cw) && cw > 0.0f;
  if (inside) {
    const float depth = -vz;
    float size = __fdiv_rn(p.point_size, fmaxf(depth, 0.1f));
    size = fminf(fmaxf(size, p.point_size_min), p.point_size_max);
    const float xw = __fmul_rn(__fmul_rn(__fadd_rn(__fdiv_rn(cx, cw), 1.0f), 0.5f), (float)p.width);
    const float yw = __fmul_rn(__fmul_rn(__fadd_rn(__fdiv_rn(cy, cw), 1.0f), 0.5f), (float)p.height);
    float t;
    if (p.color_mode == 1) {  // velocity = length(aVel)
      const float ux = d.vel_x[i], uy = d.vel_y[i], uz = d.vel_z[i];
      const float sp = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(ux, ux), __fmul_rn(uy, uy)), __fmul_rn(uz, uz)));
      t = clamp01(__fdiv_rn(sp, p.max_velocity));
    } else {
      t = clamp01(__fdiv_rn(depth, p.max_depth));
    }
    const float h = __fmul_rn(size, 0.5f);
    int x0, x1, y0, y1;
    pixel_span(xw, h, p.width, x0, x1);
    pixel_span(yw, h, p.height, y0, y1);
    cnt = (uint32_t)(x1 - x0) * (uint32_t)(y1 - y0);
    s = make_float4(xw, yw, size, t);
  }
  splat[i] = s;
  count[i] = cnt;
}

// ---- exclusive scan of the per-body fragment counts (3 small kernels; u64 total) -----------------------
constexpr int SCAN_TILE = 2048;  // 256 threads x 8

__global__ void __launch_bounds__(256) scan_tile_sums_kernel(const uint32_t* __restrict__ v, size_t n,
                                                             unsigned long long* __restrict__ tile_sum) {
  __shared__ unsigned long long ws[8];
  const size_t base = (size_t)blockIdx.x * SCAN_TILE;
  unsigned long long s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const size_t i = base + (size_t)k * 256 + threadIdx.x;
    if (i < n) s += v[i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int k = 0; k < 8; k++) t += ws[k];
    tile_sum[blockIdx.x] = t;
  }
}

// single block: exclusive scan of the tile sums in place, grand total to *total
__global__ void __launch_bounds__(1024) scan_tiles_kernel(unsigned long long* __restrict__ tile_sum, int ntiles,
                                                          unsigned long long* __restrict__ total) {
  __shared__ unsigned long long ws[32];
  __shared__ unsigned long long carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int b0 = 0; b0 < ntiles; b0 += 1024) {
    const int i = b0 + threadIdx.x;
    const unsigned long long v = i < ntiles ? tile_sum[i] : 0ull;
    unsigned long long x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) ws[w] = x;
    __syncthreads();
    if (w == 0) {
      unsigned long long s = ws[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += y;
      }
      ws[lane] = s;
    }
    __syncthreads();
    const unsigned long long carry = carry_s;
    if (i < ntiles) tile_sum[i] = carry + (w ? ws[w - 1] : 0ull) + x - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + ws[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry_s;
}

// shared with sh.cu: in-place exclusive scan of ntiles 64-bit tile sums, grand total to *total (device memory)
int launch_scan_tiles(nb_context* ctx, unsigned long long* tile_sum, int ntiles, unsigned long long* total,
                      cudaStream_t st) {
  scan_tiles_kernel<<<1, 1024, 0, st>>>(tile_sum, ntiles, total);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

// fragments of body i, emitted in body order: key = pixel (row-major, origin bottom-left), value = body
__global__ void __launch_bounds__(256) splat_emit_kernel(const float4* __restrict__ splat,
                                                         const uint32_t* __restrict__ count, size_t n,
                                                         const unsigned long long* __restrict__ tile_off, int width,
                                                         int height, uint32_t* __restrict__ keys,
                                                         int32_t* __restrict__ vals) {
  __shared__ unsigned long long ws[8];
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * 8;  // 8 consecutive bodies per thread
  uint32_t c[8];
  unsigned long long mine = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    c[k] = (base + k < n) ? count[base + k] : 0u;
    mine += c[k];
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  unsigned long long x = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[w] = x;
  __syncthreads();
  unsigned long long off = tile_off[blockIdx.x] + x - mine;
  for (int k = 0; k < w; k++) off += ws[k];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    if (c[k] == 0) continue;
    const float4 s = splat[base + k];
    const float h = __fmul_rn(s.z, 0.5f);
    int x0, x1, y0, y1;
    pixel_span(s.x, h, width, x0, x1);
    pixel_span(s.y, h, height, y0, y1);
    unsigned long long o = off;
    for (int py = y0; py < y1; py++)
      for (int px = x0; px < x1; px++) {
        keys[o] = (uint32_t)(py * width + px);
        vals[o] = (int32_t)(base + k);
        o++;
      }
    off += c[k];
  }
}

// fragment stage for (pixel, body): false = discarded; else src colour and alpha
__device__ __forceinline__ bool splat_shade(const float4 s, int px, int py, int mode, float& r, float& g, float& b,
                                            float& a) {
  // gl_PointCoord - 0.5 (the t flip does not change the length)
  const float u = __fdiv_rn(__fsub_rn(__fadd_rn((float)px, 0.5f), s.x), s.z);
  const float v = __fdiv_rn(__fsub_rn(__fadd_rn((float)py, 0.5f), s.y), s.z);
  const float dist = __fsqrt_rn(__fadd_rn(__fmul_rn(u, u), __fmul_rn(v, v)));
  if (dist > 0.5f) return false;
  float lo[3], hi[3];
  if (mode == 0) { lo[0] = 1.0f; lo[1] = 0.5f; lo[2] = 0.0f; hi[0] = 0.0f; hi[1] = 0.5f; hi[2] = 1.0f; }
  else if (mode == 1) { lo[0] = 0.0f; lo[1] = 0.0f; lo[2] = 1.0f; hi[0] = 1.0f; hi[1] = 0.0f; hi[2] = 0.0f; }
  else { lo[0] = 0.2f; lo[1] = 0.2f; lo[2] = 0.8f; hi[0] = 1.0f; hi[1] = 1.0f; hi[2] = 0.2f; }
  const float t = s.w, omt = __fsub_rn(1.0f, t);  // mix(x, y, t) = x * (1 - t) + y * t
  r = __fadd_rn(__fmul_rn(lo[0], omt), __fmul_rn(hi[0], t));
  g = __fadd_rn(__fmul_rn(lo[1], omt), __fmul_rn(hi[1], t));
  b = __fadd_rn(__fmul_rn(lo[2], omt), __fmul_rn(hi[2], t));
  // 1 - smoothstep(0.4, 0.5, dist)
  const float q = clamp01(__fdiv_rn(__fsub_rn(dist, 0.4f), __fsub_rn(0.5f, 0.4f)));
  a = __fsub_rn(1.0f, __fmul_rn(__fmul_rn(q, q), __fsub_rn(3.0f, __fmul_rn(2.0f, q))));
  return true;
}

__device__ __forceinline__ unsigned to_unorm8(float c) {
  return (unsigned)__float2int_rn(__fmul_rn(clamp01(c), 255.0f));
}

__global__ void __launch_bounds__(256) splat_clear_kernel(uint32_t* __restrict__ fb, size_t pixels, uint32_t rgba) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < pixels) fb[i] = rgba;
}

// one thread per sorted fragment; the thread holding the LAST fragment of a pixel resolves that pixel
__global__ void __launch_bounds__(256) splat_resolve_kernel(const uint32_t* __restrict__ keys,
                                                            const int32_t* __restrict__ vals, size_t nf,
                                                            const float4* __restrict__ splat, nb_render_params p,
                                                            uint32_t* __restrict__ fb) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nf) return;
  const uint32_t pix = keys[k];
  if (k + 1 < nf && keys[k + 1] == pix) return;
  const int px = (int)(pix % (uint32_t)p.width), py = (int)(pix / (uint32_t)p.width);
  // walk back to the last fully opaque fragment (alpha == 1 replaces the destination exactly)
  size_t first = k;
  for (;;) {
    float r, g, b, a;
    const bool live = splat_shade(splat[vals[first]], px, py, p.color_mode, r, g, b, a);
    if (live && a == 1.0f) break;
    if (first == 0 || keys[first - 1] != pix) {
      first = (size_t)-1;  // no opaque fragment: start from the clear colour, blend the whole run
      break;
    }
    first--;
  }
  float R = p.clear_color[0], G = p.clear_color[1], B = p.clear_color[2], A = p.clear_color[3];
  size_t j;
  if (first == (size_t)-1) {
    j = k;
    while (j > 0 && keys[j - 1] == pix) j--;
  } else {
    j = first;
  }
  for (; j <= k; j++) {
    float r, g, b, a;
    if (!splat_shade(splat[vals[j]], px, py, p.color_mode, r, g, b, a)) continue;
    const float oma = __fsub_rn(1.0f, a);  // GL_SRC_ALPHA, GL_ONE_MINUS_SRC_ALPHA on all four channels
    R = __fadd_rn(__fmul_rn(r, a), __fmul_rn(R, oma));
    G = __fadd_rn(__fmul_rn(g, a), __fmul_rn(G, oma));
    B = __fadd_rn(__fmul_rn(b, a), __fmul_rn(B, oma));
    A = __fadd_rn(__fmul_rn(a, a), __fmul_rn(A, oma));
  }
  fb[pix] = to_unorm8(R) | (to_unorm8(G) << 8) | (to_unorm8(B) << 16) | (to_unorm8(A) << 24);
}

// K16 (src/render/cuda_gl_interop.cpp:7-28): SoA -> interleaved xyz, what a GL vertex buffer holds
__global__ void __launch_bounds__(256) interleave3_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                          const float* __restrict__ z, size_t n,
                                                          float* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[3 * i] = x[i];
  out[3 * i + 1] = y[i];
  out[3 * i + 2] = z[i];
}

}  // namespace nb

namespace nb {

static RenderState* render_get(nb_context* ctx) {
  if (!ctx->render) ctx->render = new (std::nothrow) RenderState();
  return ctx->render;
}

void render_destroy(nb_context* ctx) {
  if (!ctx->render) return;
  RenderState* r = ctx->render;
  r->splat.release(); r->count.release(); r->block.release(); r->keys.release(); r->vals.release();
  if (r->h_total) cudaFreeHost(r->h_total);
  delete r;
  ctx->render = nullptr;
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_render_splats(nb_context* ctx, const nb_particles* d, const nb_render_params* p, uint8_t* d_rgba, nb_stream s) {
  if (!ctx || !d || !p || !d_rgba || (d->count && (!d->pos_x || !d->pos_y || !d->pos_z))) {
    set_error("render: null argument");
    return NB_ERR_INVALID;
  }
  if (p->width < 1 || p->height < 1 || (long long)p->width * p->height > (1ll << 28) || p->color_mode < 0 ||
      p->color_mode > 2 || !(p->point_size_min > 0.f) || !(p->point_size_max >= p->point_size_min) ||
      !(p->max_depth > 0.f) || !(p->max_velocity > 0.f)) {
    set_error("render: bad parameters (size %dx%d, colour mode %d, point size range [%g, %g])", p->width, p->height,
              p->color_mode, p->point_size_min, p->point_size_max);
    return NB_ERR_INVALID;
  }
  if (p->color_mode == 1 && d->count && (!d->vel_x || !d->vel_y || !d->vel_z)) {
    set_error("render: velocity colour mode needs velocities");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)s;
  RenderState* r = render_get(ctx);
  if (!r) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  const size_t n = d->count, pixels = (size_t)p->width * p->height;
  auto u8 = [](float c) { return (uint32_t)lrintf(fminf(fmaxf(c, 0.f), 1.f) * 255.0f); };
  const uint32_t clear = u8(p->clear_color[0]) | (u8(p->clear_color[1]) << 8) | (u8(p->clear_color[2]) << 16) |
                         (u8(p->clear_color[3]) << 24);
  splat_clear_kernel<<<(unsigned)((pixels + 255) / 256), 256, 0, st>>>(reinterpret_cast<uint32_t*>(d_rgba), pixels, clear);
  NB_LAUNCH_CHECK(ctx);
  r->fragments = 0;
  if (n == 0) return NB_OK;
  const int ntiles = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
  if (int rc = r->splat.ensure(n * sizeof(float4))) return rc;
  if (int rc = r->count.ensure(n * sizeof(uint32_t))) return rc;
  if (int rc = r->block.ensure(((size_t)ntiles + 1) * sizeof(unsigned long long))) return rc;
  if (!r->h_total) NB_CUDA(cudaMallocHost(&r->h_total, sizeof(unsigned long long)));
  unsigned long long* tile_sum = r->block.as<unsigned long long>();
  unsigned long long* total = tile_sum + ntiles;
  const unsigned gn = (unsigned)((n + 255) / 256);
  splat_project_kernel<<<gn, 256, 0, st>>>(*d, n, *p, r->splat.as<float4>(), r->count.as<uint32_t>());
  NB_LAUNCH_CHECK(ctx);
  scan_tile_sums_kernel<<<ntiles, 256, 0, st>>>(r->count.as<uint32_t>(), n, tile_sum);
  NB_LAUNCH_CHECK(ctx);
  scan_tiles_kernel<<<1, 1024, 0, st>>>(tile_sum, ntiles, total);
  NB_LAUNCH_CHECK(ctx);
  // the fragment count decides two allocations: one 8-byte readback
  NB_CUDA(cudaMemcpyAsync(r->h_total, total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  NB_CUDA(cudaStreamSynchronize(st));
  const unsigned long long nf = *r->h_total;
  r->fragments = nf;
  if (nf == 0) return NB_OK;
  if (nf >= (1ull << 30)) {
    set_error("render: %llu fragments exceed the 2^30 the fragment sort handles (lower point_size_max)", nf);
    return NB_ERR_INVALID;
  }
  if (int rc = r->keys.ensure(nf * sizeof(uint32_t))) return rc;
  if (int rc = r->vals.ensure(nf * sizeof(int32_t))) return rc;
  splat_emit_kernel<<<ntiles, 256, 0, st>>>(r->splat.as<float4>(), r->count.as<uint32_t>(), n, tile_sum, p->width,
                                            p->height, r->keys.as<uint32_t>(), r->vals.as<int32_t>());
  NB_LAUNCH_CHECK(ctx);
  int bits = 1;
  while (bits < 32 && (1ull << bits) < pixels) bits++;
  if (int rc = sort_pairs(ctx, r->keys.as<uint32_t>(), r->vals.as<int32_t>(), (size_t)nf, bits, st)) return rc;
  splat_resolve_kernel<<<(unsigned)((nf + 255) / 256), 256, 0, st>>>(r->keys.as<uint32_t>(), r->vals.as<int32_t>(),
                                                                     (size_t)nf, r->splat.as<float4>(), *p,
                                                                     reinterpret_cast<uint32_t*>(d_rgba));
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

int nb_render_stats(nb_context* ctx, uint64_t* fragments) {
  if (!ctx || !ctx->render || !fragments) {
    set_error("render stats: nothing rendered");
    return NB_ERR_INVALID;
  }
  *fragments = ctx->render->fragments;
  return NB_OK;
}

int nb_interop_pack(nb_context* ctx, const nb_particles* d, float* d_positions_xyz, float* d_velocities_xyz, nb_stream s) {
  if (!ctx || !d || (!d_positions_xyz && !d_velocities_xyz)) {
    set_error("interop pack: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  const size_t n = d->count;
  if (n == 0) return NB_OK;
  const unsigned gn = (unsigned)((n + 255) / 256);
  if (d_positions_xyz) {
    interleave3_kernel<<<gn, 256, 0, (cudaStream_t)s>>>(d->pos_x, d->pos_y, d->pos_z, n, d_positions_xyz);
    NB_LAUNCH_CHECK(ctx);
  }
  if (d_velocities_xyz) {
    interleave3_kernel<<<gn, 256, 0, (cudaStream_t)s>>>(d->vel_x, d->vel_y, d->vel_z, n, d_velocities_xyz);
    NB_LAUNCH_CHECK(ctx);
  }
  return NB_OK;
}

}  // extern "C"
