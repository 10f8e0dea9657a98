// Do packed fp32x2 instructions leave issue slots free for other pipes?  (B200, sm_100a)
// A warp stream of NF independent FFMA2 chains, optionally interleaved with NI independent integer (ALU pipe)
// instructions or NS scalar FFMA per trip.  If FFMA2 holds the scheduler's issue port for both of its pipe cycles,
// every added instruction lengthens the loop; if it holds only the pipe, integer work rides along for free.
// Used for DESIGN.md sec. 3.5 (what bounds the Barnes-Hut traversal).  nvcc -O3 -arch=sm_100a -o ubench3 ubench3.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint64_t pk(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}

template <int NF, int NI, int NS, int NM>
__global__ void __launch_bounds__(256) k_mix(float* out, float a, float b, unsigned m, int iters) {
  uint64_t v[NF > 0 ? NF : 1];
  unsigned q[NI > 0 ? NI : 1];
  float s[NS > 0 ? NS : 1];
  float r[NM > 0 ? NM : 1];
  const uint64_t a2 = pk(a, a * 1.0001f), b2 = pk(b, b * 1.0001f);
#pragma unroll
  for (int c = 0; c < NF; c++) v[c] = pk(threadIdx.x * 1e-3f + c, threadIdx.x * 2e-3f + c);
#pragma unroll
  for (int c = 0; c < NI; c++) q[c] = threadIdx.x * 7u + c;
#pragma unroll
  for (int c = 0; c < NS; c++) s[c] = threadIdx.x * 3e-3f + c;
#pragma unroll
  for (int c = 0; c < NM; c++) r[c] = threadIdx.x * 3e-3f + c + 1.f;
  constexpr int M1 = NF > NI ? NF : NI, M2 = NS > NM ? NS : NM, MAXN = M1 > M2 ? M1 : M2;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < MAXN; c++) {
      if (c < NF) v[c] = fma2(v[c], a2, b2);
      if (c < NI) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(q[c]) : "r"(m), "r"(q[(c + 1) % (NI > 0 ? NI : 1)]));
      if (c < NS) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(s[c]) : "f"(a), "f"(b));
      if (c < NM) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(r[c]));
    }
  }
  float acc = 0.f;
#pragma unroll
  for (int c = 0; c < NF; c++) {
    float x, y;
    upk(v[c], x, y);
    acc += x + y;
  }
#pragma unroll
  for (int c = 0; c < NI; c++) acc += (float)q[c];
#pragma unroll
  for (int c = 0; c < NS; c++) acc += s[c];
#pragma unroll
  for (int c = 0; c < NM; c++) acc += r[c];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int NF, int NI, int NS, int NM>
static void run(const char* name, float* out, int sms) {
  const int iters = 8192, grid = sms * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0);
    k_mix<NF, NI, NS, NM><<<grid, 256>>>(out, 0.999f, 0.001f, 0x5a5a5a5au, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep && ms < best) best = ms;
  }
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  // warp-instructions per scheduler: grid*8 warps / (sms*4) schedulers = 16 warps per scheduler
  const double trips = 16.0 * iters;
  const double cyc = best * 1e-3 * khz * 1e3;  // at the nominal boost clock (upper bound on the cycles)
  printf("{\"kernel\": \"%s\", \"ffma2\": %d, \"lop3\": %d, \"ffma\": %d, \"mufu\": %d, \"ms\": %.4f, \"cycles_per_trip_per_scheduler_at_max_clock\": %.2f}\n",
         name, NF, NI, NS, NM, best, cyc / trips);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  float* out;
  cudaMalloc(&out, (size_t)p.multiProcessorCount * 8 * 256 * sizeof(float));
  run<12, 0, 0, 0>("ffma2 x12", out, p.multiProcessorCount);
  run<12, 3, 0, 0>("ffma2 x12 + lop3 x3", out, p.multiProcessorCount);
  run<12, 6, 0, 0>("ffma2 x12 + lop3 x6", out, p.multiProcessorCount);
  run<12, 12, 0, 0>("ffma2 x12 + lop3 x12", out, p.multiProcessorCount);
  run<12, 0, 0, 2>("ffma2 x12 + mufu x2", out, p.multiProcessorCount);
  run<12, 6, 0, 2>("ffma2 x12 + mufu x2 + lop3 x6", out, p.multiProcessorCount);
  run<12, 12, 0, 2>("ffma2 x12 + mufu x2 + lop3 x12", out, p.multiProcessorCount);
  run<0, 12, 0, 0>("lop3 x12", out, p.multiProcessorCount);
  run<0, 0, 12, 0>("ffma x12", out, p.multiProcessorCount);
  run<0, 0, 24, 0>("ffma x24", out, p.multiProcessorCount);
  run<0, 12, 24, 0>("ffma x24 + lop3 x12", out, p.multiProcessorCount);
  run<0, 0, 0, 2>("mufu x2", out, p.multiProcessorCount);
  return 0;
}
