// Follow-up to ubench.cu: what limits the packed fp32x2 instructions on B200 (sm_100a)?  FFMA2 alone reaches 86 % of
// the FP32 pipe in ubench.cu against 96 % for scalar FFMA; this file varies the operand pattern (distinct / repeated /
// broadcast sources) and the opcode (add, mul, fma).  Standalone: nvcc -gencode arch=compute_100a,code=sm_100a -O3.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA %s at line %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

typedef unsigned long long f2;
__device__ __forceinline__ f2 pk(float a, float b) { f2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(f2 v, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { f2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { f2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// MODE 0: v = v*a + b (two loop-invariant sources)    1: v = v*v + v (one register pair for all three sources)
//      2: v = v*a + v                                 3: v = v + a        4: v = v * a
//      5: v[c] = v[c]*w[c] + u[c] (three distinct, all changing)   6: v = fma(v, bcast(s), b) with s a scalar register
template <int MODE, int CH>
__global__ void __launch_bounds__(256) k(float* out, float a, float b, int iters) {
  f2 v[CH], w[CH], u[CH];
  const f2 a2 = pk(a, a * 1.0001f), b2 = pk(b, b * 1.0001f);
#pragma unroll
  for (int c = 0; c < CH; c++) {
    v[c] = pk(threadIdx.x * 1e-3f + c, threadIdx.x * 2e-3f + c);
    w[c] = pk(1.0f + c * 1e-3f, 1.0f - c * 1e-3f);
    u[c] = pk(c * 1e-4f, -c * 1e-4f);
  }
  float s = a;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CH; c++) {
      if (MODE == 0) v[c] = fma2(v[c], a2, b2);
      if (MODE == 1) v[c] = fma2(v[c], v[c], v[c]);
      if (MODE == 2) v[c] = fma2(v[c], a2, v[c]);
      if (MODE == 3) v[c] = add2(v[c], a2);
      if (MODE == 4) v[c] = mul2(v[c], a2);
      if (MODE == 5) { v[c] = fma2(v[c], w[c], u[c]); }
      if (MODE == 6) v[c] = fma2(v[c], pk(s, s), b2);
    }
    if (MODE == 5) {  // keep w and u live and changing without adding packed work to the count
#pragma unroll
      for (int c = 0; c < CH; c += 8) { w[c] ^= (f2)it; }
    }
  }
  float acc = 0;
#pragma unroll
  for (int c = 0; c < CH; c++) { float x, y; upk(v[c], x, y); acc += x + y; upk(w[c], x, y); acc += x; upk(u[c], x, y); acc += y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <typename F>
float time_ms(F f, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
  int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const int sms = pr.multiProcessorCount, iters = 4096;
  float* out; CK(cudaMalloc(&out, sizeof(float) * sms * 8 * 256));
  printf("%s SMs=%d clock=%d kHz; packed lane-ops per clk per SM (128 = one per FP32 lane per clk)\n", pr.name, sms, clk_khz);
#define RUN(MODE, CH, BPS, NAME)                                                                      \
  {                                                                                                   \
    float ms = time_ms([&] { k<MODE, CH><<<sms * BPS, 256>>>(out, 1.0001f, 0.5f, iters); }, 5);       \
    double ops = 2.0 * CH * iters * 256.0 * sms * BPS;                                                \
    printf("%-44s chains=%2d blocks/SM=%d: %.3f ms  %.1f lanes/clk/SM\n", NAME, CH, BPS, ms,           \
           ops / (ms * 1e-3) / sms / (clk_khz * 1e3));                                                \
  }
  RUN(0, 16, 4, "FFMA2 v = v*a + b (invariant a, b)")
  RUN(0, 8, 4, "FFMA2 v = v*a + b (invariant a, b)")
  RUN(0, 16, 8, "FFMA2 v = v*a + b (invariant a, b)")
  RUN(1, 16, 4, "FFMA2 v = v*v + v (one source pair)")
  RUN(2, 16, 4, "FFMA2 v = v*a + v")
  RUN(5, 16, 4, "FFMA2 v = v*w + u (three distinct pairs)")
  RUN(6, 16, 4, "FFMA2 v = v*bcast(s) + b")
  RUN(3, 16, 4, "FADD2 v = v + a")
  RUN(4, 16, 4, "FMUL2 v = v * a")
  RUN(3, 16, 8, "FADD2 v = v + a")
  RUN(4, 16, 8, "FMUL2 v = v * a")
  return 0;
}
