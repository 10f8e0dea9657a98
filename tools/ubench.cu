// Pipe-throughput microbenchmarks for B200 (sm_100a): FFMA, FFMA2 (fma.rn.f32x2),
// MUFU.RSQ, and quick variants of the direct-N^2 pair body.  Used to pick the
// shape of the production kernel in nbody_b200/csrc/direct.cu; the numbers it
// prints are recorded in profiles/.  Standalone: nvcc -o ubench ubench.cu.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      fprintf(stderr, "CUDA %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

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
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float rsq(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// ---- raw pipes ---------------------------------------------------------------
template <int CH>
__global__ void k_ffma(float* out, float a, float b, int iters) {
  float v[CH];
#pragma unroll
  for (int c = 0; c < CH; c++) v[c] = threadIdx.x * 1e-3f + c;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CH; c++) v[c] = fmaf(v[c], a, b);
  }
  float s = 0;
#pragma unroll
  for (int c = 0; c < CH; c++) s += v[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CH>
__global__ void k_ffma2(float* out, float a, float b, int iters) {
  uint64_t v[CH];
  uint64_t a2 = pk(a, a * 1.0001f), b2 = pk(b, b * 1.0001f);
#pragma unroll
  for (int c = 0; c < CH; c++) v[c] = pk(threadIdx.x * 1e-3f + c, threadIdx.x * 2e-3f + c);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CH; c++) v[c] = fma2(v[c], a2, b2);
  }
  float s = 0;
#pragma unroll
  for (int c = 0; c < CH; c++) {
    float x, y;
    upk(v[c], x, y);
    s += x + y;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CH>
__global__ void k_mufu(float* out, int iters) {
  float v[CH];
#pragma unroll
  for (int c = 0; c < CH; c++) v[c] = 1.0f + threadIdx.x * 1e-3f + c;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CH; c++) v[c] = rsq(v[c]);
  }
  float s = 0;
#pragma unroll
  for (int c = 0; c < CH; c++) s += v[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FFMA2 + MUFU mix at the pair-body ratio: 6 FFMA2 : 1 MUFU per "pair".
template <int CH>
__global__ void k_mix2(float* out, float a, float b, int iters) {
  uint64_t v[CH];
  float m[CH];
  uint64_t a2 = pk(a, a * 1.0001f), b2 = pk(b, b * 1.0001f);
#pragma unroll
  for (int c = 0; c < CH; c++) {
    v[c] = pk(threadIdx.x * 1e-3f + c, threadIdx.x * 2e-3f + c);
    m[c] = 1.0f + c;
  }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CH; c++) {
#pragma unroll
      for (int r = 0; r < 6; r++) v[c] = fma2(v[c], a2, b2);
      m[c] = rsq(m[c]);
    }
  }
  float s = 0;
#pragma unroll
  for (int c = 0; c < CH; c++) {
    float x, y;
    upk(v[c], x, y);
    s += x + y + m[c];
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FFMA2 : FFMA mixes -- does scalar FFMA fill what FFMA2 leaves of the FP32 pipe?
template <int R2, int R1>
__global__ void k_mixfs(float* out, float a, float b, int iters) {
  uint64_t v[R2 > 0 ? R2 : 1];
  float w[R1 > 0 ? R1 : 1];
  uint64_t a2 = pk(a, a * 1.0001f), b2 = pk(b, b * 1.0001f);
#pragma unroll
  for (int c = 0; c < R2; c++) v[c] = pk(threadIdx.x * 1e-3f + c, threadIdx.x * 2e-3f + c);
#pragma unroll
  for (int c = 0; c < R1; c++) w[c] = threadIdx.x * 3e-3f + c;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < (R2 > R1 ? R2 : R1); c++) {
      if (c < R2) v[c] = fma2(v[c], a2, b2);
      if (c < R1) w[c] = fmaf(w[c], a, b);
    }
  }
  float s = 0;
#pragma unroll
  for (int c = 0; c < R2; c++) {
    float x, y;
    upk(v[c], x, y);
    s += x + y;
  }
#pragma unroll
  for (int c = 0; c < R1; c++) s += w[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- direct pair body variants ----------------------------------------------
// I targets per thread, T threads, tile of T sources (float4) per sync.
template <int I, int T>
__global__ void __launch_bounds__(T) k_direct_scalar(const float4* __restrict__ src, float* __restrict__ acc,
                                                     int n, float eps2) {
  __shared__ float4 tile[T];
  float xi[I], yi[I], zi[I], ax[I], ay[I], az[I];
  const int base = blockIdx.x * (I * T) + threadIdx.x;
#pragma unroll
  for (int k = 0; k < I; k++) {
    float4 p = src[base + k * T];
    xi[k] = p.x; yi[k] = p.y; zi[k] = p.z;
    ax[k] = ay[k] = az[k] = 0.f;
  }
  for (int t = 0; t < n; t += T) {
    tile[threadIdx.x] = src[t + threadIdx.x];
    __syncthreads();
#pragma unroll 8
    for (int j = 0; j < T; j++) {
      float4 p = tile[j];
#pragma unroll
      for (int k = 0; k < I; k++) {
        float dx = p.x - xi[k], dy = p.y - yi[k], dz = p.z - zi[k];
        float r2 = fmaf(dx, dx, eps2);
        r2 = fmaf(dy, dy, r2);
        r2 = fmaf(dz, dz, r2);
        float ri = rsq(r2);
        float s = (ri * ri) * (p.w * ri);
        ax[k] = fmaf(s, dx, ax[k]);
        ay[k] = fmaf(s, dy, ay[k]);
        az[k] = fmaf(s, dz, az[k]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < I; k++) {
    acc[base + k * T] = ax[k];
    acc[n + base + k * T] = ay[k];
    acc[2 * n + base + k * T] = az[k];
  }
}

// Packed: I must be even; I/2 packed target pairs per thread, source broadcast.
template <int I, int T>
__global__ void __launch_bounds__(T) k_direct_packed(const float4* __restrict__ src, float* __restrict__ acc,
                                                     int n, float eps2) {
  __shared__ float4 tile[T];
  constexpr int P = I / 2;
  uint64_t xi[P], yi[P], zi[P], ax[P], ay[P], az[P];
  const int base = blockIdx.x * (I * T) + threadIdx.x;
#pragma unroll
  for (int k = 0; k < P; k++) {
    float4 p = src[base + (2 * k) * T], q = src[base + (2 * k + 1) * T];
    xi[k] = pk(p.x, q.x); yi[k] = pk(p.y, q.y); zi[k] = pk(p.z, q.z);
    ax[k] = ay[k] = az[k] = pk(0.f, 0.f);
  }
  const uint64_t e2 = pk(eps2, eps2);
  for (int t = 0; t < n; t += T) {
    tile[threadIdx.x] = src[t + threadIdx.x];
    __syncthreads();
#pragma unroll 8
    for (int j = 0; j < T; j++) {
      float4 p = tile[j];
      uint64_t px = pk(p.x, p.x), py = pk(p.y, p.y), pz = pk(p.z, p.z), pm = pk(p.w, p.w);
#pragma unroll
      for (int k = 0; k < P; k++) {
        uint64_t dx = sub2(px, xi[k]), dy = sub2(py, yi[k]), dz = sub2(pz, zi[k]);
        uint64_t r2 = fma2(dx, dx, e2);
        r2 = fma2(dy, dy, r2);
        r2 = fma2(dz, dz, r2);
        float r0, r1;
        upk(r2, r0, r1);
        uint64_t ri = pk(rsq(r0), rsq(r1));
        uint64_t s = mul2(mul2(ri, ri), mul2(pm, ri));
        ax[k] = fma2(s, dx, ax[k]);
        ay[k] = fma2(s, dy, ay[k]);
        az[k] = fma2(s, dz, az[k]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < P; k++) {
    float a, b;
    upk(ax[k], a, b); acc[base + 2 * k * T] = a; acc[base + (2 * k + 1) * T] = b;
    upk(ay[k], a, b); acc[n + base + 2 * k * T] = a; acc[n + base + (2 * k + 1) * T] = b;
    upk(az[k], a, b); acc[2 * n + base + 2 * k * T] = a; acc[2 * n + base + (2 * k + 1) * T] = b;
  }
}

template <typename F>
static float time_ms(F&& launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  launch();
  launch();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int r = 0; r < reps; r++) launch();
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  CK(cudaGetLastError());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return ms / reps;
}

int main() {
  cudaDeviceProp pr;
  CK(cudaGetDeviceProperties(&pr, 0));
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  printf("device: %s sm_%d%d SMs=%d clock=%d kHz smem/SM=%zu smem/blk(optin)=%zu regs/SM=%d L2=%d MB mem=%.1f GB\n",
         pr.name, pr.major, pr.minor, pr.multiProcessorCount, clk_khz, pr.sharedMemPerMultiprocessor,
         pr.sharedMemPerBlockOptin, pr.regsPerMultiprocessor, pr.l2CacheSize >> 20, pr.totalGlobalMem / 1e9);
  const int sms = pr.multiProcessorCount;
  float* out;
  CK(cudaMalloc(&out, sizeof(float) * sms * 8 * 1024));

  // raw pipes: threads 256, blocks sms*8 (2048 thr/SM), iters big
  const int iters = 4096;
  {
    constexpr int CH = 16;
    for (int bps : {2, 4, 8}) {
      float ms = time_ms([&] { k_ffma<CH><<<sms * bps, 256>>>(out, 1.0001f, 0.5f, iters); }, 5);
      double fl = 2.0 * CH * iters * 256.0 * sms * bps;
      printf("FFMA   blocks/SM=%d: %.3f ms  %.2f TFLOP/s  (%.1f FMA lanes/clk/SM @%d kHz nominal)\n", bps, ms,
             fl / ms / 1e9, fl / 2 / (ms * 1e-3) / sms / (clk_khz * 1e3), clk_khz);
    }
    for (int bps : {2, 4, 8}) {
      float ms = time_ms([&] { k_ffma2<CH><<<sms * bps, 256>>>(out, 1.0001f, 0.5f, iters); }, 5);
      double fl = 4.0 * CH * iters * 256.0 * sms * bps;
      printf("FFMA2  blocks/SM=%d: %.3f ms  %.2f TFLOP/s  (%.1f FMA lanes/clk/SM)\n", bps, ms, fl / ms / 1e9,
             fl / 2 / (ms * 1e-3) / sms / (clk_khz * 1e3));
    }
    for (int bps : {2, 4, 8}) {
      float ms = time_ms([&] { k_mufu<8><<<sms * bps, 256>>>(out, iters); }, 5);
      double ops = 8.0 * iters * 256.0 * sms * bps;
      printf("MUFU   blocks/SM=%d: %.3f ms  %.2f Gop/s  (%.2f lanes/clk/SM)\n", bps, ms, ops / ms / 1e6,
             ops / (ms * 1e-3) / sms / (clk_khz * 1e3));
    }
    for (int bps : {2, 4, 8}) {
      float ms = time_ms([&] { k_mix2<8><<<sms * bps, 256>>>(out, 1.0001f, 0.5f, iters); }, 5);
      double fl = 4.0 * 6 * 8 * iters * 256.0 * sms * bps;
      printf("MIX 6xFFMA2+1xMUFU blocks/SM=%d: %.3f ms  %.2f TFLOP/s fma  (%.1f FMA lanes/clk/SM)\n", bps, ms,
             fl / ms / 1e9, fl / 2 / (ms * 1e-3) / sms / (clk_khz * 1e3));
    }
  }

#define RUN_MIX(R2, R1)                                                                                   \
  {                                                                                                       \
    float ms = time_ms([&] { k_mixfs<R2, R1><<<sms * 4, 256>>>(out, 1.0001f, 0.5f, iters); }, 5);        \
    double fl = (4.0 * R2 + 2.0 * R1) * iters * 256.0 * sms * 4;                                          \
    printf("MIX %2d FFMA2 : %2d FFMA  %.3f ms  %.2f TFLOP/s  (%.1f FMA lanes/clk/SM)\n", R2, R1, ms,     \
           fl / ms / 1e9, fl / 2 / (ms * 1e-3) / sms / (clk_khz * 1e3));                                  \
  }
  RUN_MIX(8, 0) RUN_MIX(8, 2) RUN_MIX(8, 4) RUN_MIX(8, 8) RUN_MIX(4, 8) RUN_MIX(2, 8) RUN_MIX(0, 8) RUN_MIX(0, 16)
  RUN_MIX(12, 4) RUN_MIX(6, 6)
  if (getenv("UBENCH_PIPES_ONLY")) return 0;
  // direct variants
  const int n = sms * 2048;  // 303104 on 148 SMs
  std::vector<float4> h(n);
  srand(1);
  for (int i = 0; i < n; i++)
    h[i] = make_float4(rand() / (float)RAND_MAX, rand() / (float)RAND_MAX, rand() / (float)RAND_MAX, 1.0f / n);
  float4* src;
  float* acc;
  CK(cudaMalloc(&src, sizeof(float4) * n));
  CK(cudaMalloc(&acc, sizeof(float) * 3 * n));
  CK(cudaMemcpy(src, h.data(), sizeof(float4) * n, cudaMemcpyHostToDevice));
  const double pairs = (double)n * n;
  auto report = [&](const char* name, int I, int T, float ms) {
    std::vector<float> a(8);
    CK(cudaMemcpy(a.data(), acc, sizeof(float) * 8, cudaMemcpyDeviceToHost));
    printf("%-8s I=%d T=%3d grid=%5d: %8.3f ms  %7.1f Gint/s  %5.2f TFLOP/s(20)  clk/pair/SMSP=%.2f  acc0=%.6e\n", name, I,
           T, n / (I * T), ms, pairs / ms / 1e6, pairs * 20 / ms / 1e9,
           (ms * 1e-3) * (clk_khz * 1e3) * sms * 4 * 32 / pairs, a[0]);
  };
#define RUN_S(I, T) report("scalar", I, T, time_ms([&] { k_direct_scalar<I, T><<<n / (I * T), T>>>(src, acc, n, 1e-4f); }, 3))
#define RUN_P(I, T) report("packed", I, T, time_ms([&] { k_direct_packed<I, T><<<n / (I * T), T>>>(src, acc, n, 1e-4f); }, 3))
  RUN_S(1, 256);
  RUN_S(2, 256);
  RUN_S(4, 256);
  RUN_S(4, 128);
  RUN_S(8, 128);
  RUN_P(2, 256);
  RUN_P(4, 256);
  RUN_P(4, 128);
  RUN_P(8, 128);
  RUN_P(8, 256);
  RUN_P(16, 128);
  printf("done\n");
  return 0;
}
