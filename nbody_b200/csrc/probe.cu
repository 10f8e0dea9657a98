// probe.cu -- in-run measurement of the FP32 pipe rates the direct-N^2 roofline is quoted against
// (bench.py `roofline.peak_measured_*`): dependent-free chains of scalar FFMA or packed FFMA2
// (fma.rn.f32x2) for some tens of milliseconds, timed with CUDA events on the caller's stream.  Measured
// by the run that prints it, not copied from an earlier capture.
#include "common.cuh"

namespace nb {

template <int CH, bool PACKED>
__global__ void __launch_bounds__(256) fp32_probe_kernel(float* __restrict__ out, float a, float b, int iters) {
  if (PACKED) {
    f2 v[CH];
    const f2 a2 = pk(a, a * 1.0001f), b2 = pk(b, b * 1.0001f);
#pragma unroll
    for (int c = 0; c < CH; c++) v[c] = pk(threadIdx.x * 1e-3f + c, threadIdx.x * 2e-3f + c);
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int c = 0; c < CH; c++) v[c] = fma2(v[c], a2, b2);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CH; c++) {
      float x, y;
      upk(v[c], x, y);
      s += x + y;
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
  } else {
    float v[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) v[c] = threadIdx.x * 1e-3f + c;
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int c = 0; c < CH; c++) v[c] = fmaf(v[c], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CH; c++) s += v[c];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
  }
}

}  // namespace nb

using namespace nb;

extern "C" int nb_probe_fp32(nb_context* ctx, int packed, double* tflops, double* ms_out, nb_stream s) {
  if (!ctx || !tflops) {
    set_error("fp32 probe: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)s;
  constexpr int CH = 16;
  const int grid = ctx->sm_count * 8, threads = 256;
  if (int rc = ctx->red.ensure((size_t)grid * threads * sizeof(float))) return rc;
  float* out = ctx->red.as<float>();
  const int iters = 16384;  // x CH fma per thread: ~20-40 ms on a B200
  cudaEvent_t e0, e1;
  NB_CUDA(cudaEventCreate(&e0));
  NB_CUDA(cudaEventCreate(&e1));
  float best = 0.f;
  for (int rep = 0; rep < 3; rep++) {  // the first repetition warms the clocks up; best of the rest
    NB_CUDA(cudaEventRecord(e0, st));
    if (packed) fp32_probe_kernel<CH, true><<<grid, threads, 0, st>>>(out, 0.999f, 0.001f, iters);
    else fp32_probe_kernel<CH, false><<<grid, threads, 0, st>>>(out, 0.999f, 0.001f, iters);
    ctx->launches++;
    NB_CUDA(cudaEventRecord(e1, st));
    NB_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    NB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && (best == 0.f || ms < best)) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  NB_CUDA(cudaGetLastError());
  const double flops = 2.0 * (packed ? 2.0 : 1.0) * CH * (double)iters * (double)grid * threads;
  *tflops = flops / (best * 1e-3) / 1e12;
  if (ms_out) *ms_out = best;
  return NB_OK;
}
