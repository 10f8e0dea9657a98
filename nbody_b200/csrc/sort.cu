// sort.cu -- bounding box reduction and a CUB-free stable LSD radix sort of (u32 key, i32 value)
// pairs.  Replaces computeBoundingBoxKernel + its 6 malloc / 12 memcpy wrapper
// (reference src/cuda/force_barnes_hut.cu:66-110, 218-258) and thrust::sort_by_key (:276-280).
// Both are HBM-side kernels: bbox reads 12 B/body; the sort reads the keys once for all digit histograms
// and then moves 16 B/pair per pass, with digits of up to 11 bits (3 passes for 30-bit Morton keys, 2 for
// the 20/21-bit cell ids of the spatial hash and the renderer's pixel ids).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <type_traits>

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

// ---- radix sort ("onesweep": one histogram read + one read/write sweep per digit of up to 11 bits) ------
// Stable LSD.  rs_hist_kernel reads the keys once and builds the digit histograms of every pass,
// rs_scan_kernel turns them into global digit offsets.  Each pass is ONE kernel: a block takes the next
// 8192-key tile (ticket order), ranks its keys per digit -- each warp owns 512 consecutive keys and ranks
// them 32 at a time: peers (lanes with the same digit) come from one ballot per digit bit, the lowest
// peer adds the group size to the warp's digit counter with one shared-memory atomic and hands the old
// value to its peers by shuffle, so equal digits keep input order -- publishes its per-digit counts and
// resolves the counts of all earlier tiles by decoupled look-back over a per-tile status word (2 flag bits
// + 30-bit count), then moves keys and values through shared memory in digit order so that global stores
// are runs of consecutive addresses.
//
// Round 1 used 8-bit digits (4 passes for 30 bits, 0.60 ms for 16 Mi pairs = 29 % of the measured HBM
// bandwidth) and was bound by integer issue, not by DRAM (ncu: ALU pipe 58 %, DRAM 18 %): most of the
// ~140 thread-instructions per key and pass are per-pass work (loads, slot computation, two trips
// through shared memory, 64-bit addressing), only ~4 per digit BIT are ranking.  So: fewer, wider passes
// (the ranking cost per key bit stays, everything else drops by the pass count), per-warp digit counters
// packed two to a 32-bit word (a warp holds 512 keys, a count fits 16 bits) and aliased with the
// digit-ordered staging buffer, which keeps two 512-thread CTAs resident per SM with 2048 bins, and 32-bit
// index arithmetic inside a tile.  Algorithmic traffic: 4 + 16 x passes B/pair (52 B for 30-bit keys); a
// sort whose values are the identity permutation (tree and grid builds) generates them in the first pass
// instead of reading them (-4 B), and internal callers take the result from whichever buffer the last pass
// wrote (no copy back; nb_sort_pairs, which sorts in place, copies after an odd number of passes).
constexpr int RS_ITEMS = 16;
constexpr int RS_WARPS = 16;
constexpr int RS_THREADS = RS_WARPS * 32;
constexpr int RS_WARP_KEYS = RS_ITEMS * 32;      // 512 consecutive keys per warp
constexpr int RS_TILE = RS_WARPS * RS_WARP_KEYS;  // 8192 keys per tile
constexpr uint32_t RS_FLAG_AGG = 1u << 30, RS_FLAG_INC = 2u << 30, RS_VAL_MASK = (1u << 30) - 1u;
constexpr int RS_MAX_PASSES = 3;   // ceil(32 / 11)
constexpr int RS_MAX_BITS = 11;
constexpr int RS_MAX_BINS = 1 << RS_MAX_BITS;
#ifndef NB_RS_LOOK
#define NB_RS_LOOK 1  // build-time knob: 1 -> 0.463 ms, 2 -> 0.470, 4 -> 0.485 for the 16 Mi Morton sort (profiles/r02_bh_tree_tile_shapes_ab_v2.jsonl)
#endif
constexpr int RS_LOOK = NB_RS_LOOK;  // predecessors examined per look-back round trip
#ifndef NB_RS_MIN_BLOCKS
#define NB_RS_MIN_BLOCKS 2  // build-time knob: resident CTAs per SM the pass kernel is compiled for (3 = 40 registers with spills: 0.507 ms against 0.488)
#endif
#ifndef NB_RS_MATCH
#define NB_RS_MATCH 0  // build-time knob: rank with match.any.sync instead of one ballot per digit bit
#endif
#ifndef NB_RS_BATCH
#define NB_RS_BATCH 16  // build-time knob: keys a thread holds in registers at a time while ranking / staging (8: 0.488 ms, 16: 0.483 for 16 Mi x 30 bits in place)
#endif
constexpr int RS_BATCH = NB_RS_BATCH;
static_assert(RS_ITEMS % RS_BATCH == 0, "the batch must divide the keys per thread");

struct RsPlan {
  int passes;
  int shift[RS_MAX_PASSES];
  int bits[RS_MAX_PASSES];
};

// digits of (almost) equal width, at most 11 bits each, low digit first
static RsPlan rs_plan(int end_bit) {
  RsPlan p{};
  p.passes = (end_bit + RS_MAX_BITS - 1) / RS_MAX_BITS;
  int done = 0;
  for (int k = 0; k < p.passes; k++) {
    const int left = p.passes - k;
    const int b = (end_bit - done + left - 1) / left;
    p.shift[k] = done;
    p.bits[k] = b;
    done += b;
  }
  return p;
}

// hist[p][d] += number of keys whose pass-p digit is d.  COPY: the pairs are also copied to (keys_to, vals_to) on
// the way -- an in-place sort with an odd number of passes starts from the scratch copy, so that its last pass
// lands in the caller's arrays (12 B per pair more here instead of a 16 B per pair copy back at the end).
template <bool COPY>
__global__ void __launch_bounds__(512) rs_hist_kernel(const uint32_t* __restrict__ keys, size_t n, RsPlan plan,
                                                      uint32_t* __restrict__ hist, const int32_t* __restrict__ vals,
                                                      uint32_t* __restrict__ keys_to, int32_t* __restrict__ vals_to) {
  extern __shared__ uint32_t rs_hsh[];  // [passes][RS_MAX_BINS]
  const int total = plan.passes * RS_MAX_BINS;
  for (int k = threadIdx.x; k < total; k += blockDim.x) rs_hsh[k] = 0;
  __syncthreads();
  const size_t n4 = n / 4;
  const uint4* k4 = reinterpret_cast<const uint4*>(keys);
  auto add = [&](uint32_t k) {
#pragma unroll
    for (int p = 0; p < RS_MAX_PASSES; p++)
      if (p < plan.passes) atomicAdd(&rs_hsh[p * RS_MAX_BINS + ((k >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u))], 1u);
  };
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = k4[i];
    if (COPY) {
      reinterpret_cast<uint4*>(keys_to)[i] = v;
      reinterpret_cast<int4*>(vals_to)[i] = reinterpret_cast<const int4*>(vals)[i];
    }
    add(v.x); add(v.y); add(v.z); add(v.w);
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n - n4 * 4)) {
    const size_t i = n4 * 4 + threadIdx.x;
    const uint32_t k = keys[i];
    if (COPY) {
      keys_to[i] = k;
      vals_to[i] = vals[i];
    }
    add(k);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < total; k += blockDim.x) {
    const uint32_t c = rs_hsh[k];
    if (c) atomicAdd(&hist[k], c);
  }
}

// in place: hist[p][d] -> number of keys whose pass-p digit is < d.  One block per pass, 2 bins per thread.
__global__ void __launch_bounds__(RS_MAX_BINS / 2) rs_scan_kernel(uint32_t* __restrict__ hist) {
  __shared__ uint32_t ws[32];
  uint32_t* h = hist + (size_t)blockIdx.x * RS_MAX_BINS;
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  const uint32_t a = h[2 * t], b = h[2 * t + 1];
  uint32_t x = a + b;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[w] = x;
  __syncthreads();
  if (w == 0) {
    uint32_t v = ws[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += y;
    }
    ws[lane] = v;
  }
  __syncthreads();
  const uint32_t excl = x - (a + b) + (w ? ws[w - 1] : 0u);
  h[2 * t] = excl;
  h[2 * t + 1] = excl + a;
}

__device__ __forceinline__ unsigned long long ld_status2(const uint32_t* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_status2(uint32_t* p, uint32_t lo, uint32_t hi) {
  asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(lo), "r"(hi) : "memory");
}

// One pass.  RB: digit capacity of this instantiation (bins = 2^RB sizes the shared memory and RB ballots rank a
// key -- a run-time ballot count costs a branch per ballot), `bits` <= RB the digit width of this pass (the
// ballots of the unused high bits see zeros and change nothing).  gexcl: global exclusive digit offsets; status: tiles x 2^bits words,
// zeroed; ticket: zeroed.  IDENT: the values are the identity permutation (value = index) and not read.
template <int RB, bool IDENT>
__global__ void __launch_bounds__(RS_THREADS, NB_RS_MIN_BLOCKS) rs_pass_kernel(const uint32_t* __restrict__ keys_in,
                                                                 const int32_t* __restrict__ vals_in,
                                                                 uint32_t* __restrict__ keys_out,
                                                                 int32_t* __restrict__ vals_out, size_t n, int shift,
                                                                 int bits, const uint32_t* __restrict__ gexcl,
                                                                 uint32_t* __restrict__ status,
                                                                 uint32_t* __restrict__ ticket,
                                                                 int* __restrict__ err_flag) {
  constexpr int BINS = 1 << RB;
  constexpr int CW = BINS / 2;  // packed counter words per warp: digits 2j (low half) and 2j + 1 (high half)
  constexpr int CNT_WORDS = RS_WARPS * CW;
  constexpr int STAGE_WORDS = 2 * RS_TILE;
  constexpr int UNION_WORDS = CNT_WORDS > STAGE_WORDS ? CNT_WORDS : STAGE_WORDS;
  extern __shared__ __align__(16) uint32_t rs_sm[];
  uint32_t* cntw = rs_sm;                 // [RS_WARPS][CW], live until every key knows its slot
  uint32_t* buf_k = rs_sm;                // then: keys and values of the tile in digit order
  uint32_t* buf_v = rs_sm + RS_TILE;
  uint32_t* lstart = rs_sm + UNION_WORDS;  // [BINS] first slot of the digit inside the tile
  uint32_t* gbase2 = lstart + BINS;        // [BINS] global position of the digit's first key of this tile - lstart
  __shared__ uint32_t wsum[RS_WARPS];
  __shared__ uint32_t tile_s;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t mask = (1u << bits) - 1u;
  if (tid == 0) tile_s = atomicAdd(ticket, 1u);
  {
    uint4* z = reinterpret_cast<uint4*>(cntw);
    for (int k = tid; k < CNT_WORDS / 4; k += RS_THREADS) z[k] = make_uint4(0u, 0u, 0u, 0u);
  }
  __syncthreads();
  const uint32_t tile = tile_s;
  const size_t tbase = (size_t)tile * RS_TILE;
  const uint32_t tcount = (uint32_t)((n - tbase) < (size_t)RS_TILE ? (n - tbase) : (size_t)RS_TILE);

  // ---- load + rank inside the warp's 512 consecutive keys -----------------------------------------
  const uint32_t wofs = (uint32_t)w * RS_WARP_KEYS + lane;  // offset of this lane's first key inside the tile
  // Registers: a thread never holds its 16 keys at once.  It ranks them RS_BATCH at a time and keeps, per key, only
  // pd = (digit << 16) | rank (rank < 512, later the slot < 8192: the low half never carries into the digit); the
  // keys are read again (from L2: this CTA loaded them microseconds ago) when they move to the staging buffer.
  // 16 live registers across the look-back instead of 32.
  uint32_t pd[RS_ITEMS];
  const uint32_t lt = (1u << lane) - 1u;
  uint32_t* mycnt = cntw + w * CW;
  const bool full = (uint32_t)w * RS_WARP_KEYS + RS_WARP_KEYS <= tcount;  // warp-uniform
  const uint32_t* kin = keys_in + tbase + wofs;
  // ranking runs at the top level of the kernel (no enclosing branch: the ballots need no divergence checks);
  // a partial last tile is handled by one extra ballot per key (padding lanes are nobody's peer)
#pragma unroll
  for (int h = 0; h < RS_ITEMS; h += RS_BATCH) {
    uint32_t key[RS_BATCH];
    if (full) {
#pragma unroll
      for (int k = 0; k < RS_BATCH; k++) key[k] = kin[(h + k) * 32];
    } else {
#pragma unroll
      for (int k = 0; k < RS_BATCH; k++) key[k] = (wofs + (h + k) * 32 < tcount) ? kin[(h + k) * 32] : 0xffffffffu;
    }
#pragma unroll
    for (int k = 0; k < RS_BATCH; k++) {
      const bool ok = wofs + (h + k) * 32 < tcount;
      const uint32_t dgt = (key[k] >> shift) & mask;
      uint32_t peers = __ballot_sync(0xffffffffu, ok);
#if NB_RS_MATCH
      peers &= __match_any_sync(0xffffffffu, dgt);  // experiment: MATCH.ANY instead of one ballot per digit bit
#else
#pragma unroll
      for (int b = 0; b < RB; b++) {
        // v = lanes whose bit b is set, x = 0 if mine is set else ~0: peers &= ~(v ^ x) ... written as v ^ ~x below.
        // PTX on purpose: the compiler's own sequence tests the bit twice and shifts first (6 instructions per bit).
        uint32_t v, x;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            ".reg .b32 t;\n"
            "and.b32 t, %2, %3;\n"
            "setp.ne.u32 p, t, 0;\n"
            "vote.sync.ballot.b32 %0, p, 0xffffffff;\n"
            "selp.b32 %1, 0, 0xffffffff, p;\n"
            "}"
            : "=r"(v), "=r"(x)
            : "r"(dgt), "r"(1u << b));
        peers &= v ^ x;
      }
#endif
      const int leader = max(__ffs(peers) - 1, 0);
      const uint32_t sh16 = (dgt & 1u) * 16u;
      uint32_t old = 0;
      // a warp's shared-memory atomics retire in issue order: key k sees the counts of keys < k
      if (ok && lane == leader) old = atomicAdd(&mycnt[dgt >> 1], (uint32_t)__popc(peers) << sh16);
      old = __shfl_sync(0xffffffffu, old, leader);
      pd[h + k] = (dgt << 16) | (((old >> sh16) & 0xffffu) + __popc(peers & lt));
    }
  }
  __syncthreads();

  // ---- per digit: offsets across warps, tile count; publish, look back -----------------------------
  // thread t owns the packed words [t * WPT, (t + 1) * WPT) = digits [2 t WPT, 2 (t + 1) WPT)
  constexpr int WPT = (CW + RS_THREADS - 1) / RS_THREADS;
  const int nwords = (int)((mask + 1u + 1u) >> 1);  // words in use (a 1-bit digit still uses one word)
  uint32_t tot[WPT];                     // packed tile counts of my words
  uint32_t mysum = 0;
#pragma unroll
  for (int j = 0; j < WPT; j++) {
    const int wd = tid * WPT + j;
    uint32_t run = 0;
    if (wd < nwords) {
#pragma unroll
      for (int k = 0; k < RS_WARPS; k++) {
        const uint32_t c = cntw[k * CW + wd];
        cntw[k * CW + wd] = run;  // exclusive over warps; the halves cannot carry (<= 8192 keys per tile)
        run += c;
      }
    }
    tot[j] = run;
    mysum += (run & 0xffffu) + (run >> 16);
  }
  uint32_t x = mysum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) wsum[w] = x;
  __syncthreads();
  {
    uint32_t before = x - mysum;
    for (int k = 0; k < w; k++) before += wsum[k];
#pragma unroll
    for (int j = 0; j < WPT; j++) {
      const int wd = tid * WPT + j;
      if (wd < nwords) {
        const uint32_t t0 = tot[j] & 0xffffu, t1 = tot[j] >> 16;
        const uint32_t l0 = before, l1 = before + t0;
        before += t0 + t1;
        lstart[2 * wd] = l0;
        lstart[2 * wd + 1] = l1;
        uint32_t e0 = 0, e1 = 0;  // keys of these digits in earlier tiles
        uint32_t* mine = status + ((size_t)tile << bits) + 2 * wd;
        if (bits == 0 || tile == 0) {
          st_status2(mine, RS_FLAG_INC | t0, RS_FLAG_INC | t1);
        } else {
          st_status2(mine, RS_FLAG_AGG | t0, RS_FLAG_AGG | t1);
          // Decoupled look-back, RS_LOOK predecessors per round trip: the status words of tiles t-1 .. t-RS_LOOK are
          // loaded together and consumed in order (aggregate: add and go on; inclusive: add and stop; not yet
          // published: poll again from there).  One predecessor per trip left every tile waiting ~13 dependent L2
          // round trips in this phase (ncu: a quarter of the pass's instructions and stall samples were this loop).
          const size_t stride = (size_t)1 << bits;
          const uint32_t* look = mine - stride;
          uint32_t left = tile;  // predecessors not yet consumed (tile 0 always publishes an inclusive count)
          unsigned long long t_start = 0;
          for (unsigned spins = 1;; spins++) {
            unsigned long long s2[RS_LOOK];
#pragma unroll
            for (int q = 0; q < RS_LOOK; q++) s2[q] = ((uint32_t)q < left) ? ld_status2(look - (size_t)q * stride) : 0ull;
            bool finished = false;
            int consumed = 0;
#pragma unroll
            for (int q = 0; q < RS_LOOK; q++) {
              // both words of a pair are always published together: the flags of the low word decide
              const uint32_t s_lo = (uint32_t)s2[q], s_hi = (uint32_t)(s2[q] >> 32);
              if (!finished && consumed == q && (s_lo & (RS_FLAG_INC | RS_FLAG_AGG))) {
                e0 += s_lo & RS_VAL_MASK;
                e1 += s_hi & RS_VAL_MASK;
                consumed++;
                finished = (s_lo & RS_FLAG_INC) != 0u;
              }
            }
            if (finished) break;
            look -= (size_t)consumed * stride;
            left -= (uint32_t)consumed;
            if ((spins & 1023u) == 0u) {  // a faulted predecessor must not hang the device: give up after ~2 s
              unsigned long long now;
              asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
              if (t_start == 0) t_start = now;
              else if (now - t_start > 2000000000ull) {
                *err_flag = 1;
                break;
              }
            }
          }
          st_status2(mine, RS_FLAG_INC | (e0 + t0), RS_FLAG_INC | (e1 + t1));
        }
        gbase2[2 * wd] = gexcl[2 * wd] + e0 - l0;
        if (2 * wd + 1 <= (int)mask) gbase2[2 * wd + 1] = gexcl[2 * wd + 1] + e1 - l1;
      }
    }
  }
  // every key's slot in the digit-ordered staging buffer (the last use of the per-warp counters)
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const uint32_t dgt = pd[k] >> 16;
    pd[k] += ((mycnt[dgt >> 1] >> ((dgt & 1u) * 16u)) & 0xffffu);
  }
  __syncthreads();  // lstart complete, counters dead: the staging buffer may overwrite them
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) pd[k] = lstart[pd[k] >> 16] + (pd[k] & 0xffffu);
  // ---- global (again) -> staging buffer (digit order) -> global (runs of consecutive addresses) ------
  {
    const int32_t* vin = IDENT ? nullptr : vals_in + tbase + wofs;
#pragma unroll
    for (int h = 0; h < RS_ITEMS; h += RS_BATCH) {
      uint32_t key[RS_BATCH], val[RS_BATCH];
#pragma unroll
      for (int k = 0; k < RS_BATCH; k++) {
        const bool ok = full || wofs + (h + k) * 32 < tcount;
        key[k] = ok ? kin[(h + k) * 32] : 0u;
        if (IDENT) val[k] = (uint32_t)(tbase + wofs + (h + k) * 32);
        else val[k] = ok ? (uint32_t)vin[(h + k) * 32] : 0u;
      }
#pragma unroll
      for (int k = 0; k < RS_BATCH; k++) {
        if (full || wofs + (h + k) * 32 < tcount) {
          buf_k[pd[h + k]] = key[k];
          buf_v[pd[h + k]] = val[k];
        }
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const uint32_t p = (uint32_t)k * RS_THREADS + tid;
    if (p < tcount) {
      const uint32_t kk = buf_k[p];
      const uint32_t g = gbase2[(kk >> shift) & mask] + p;
      keys_out[g] = kk;
      vals_out[g] = (int32_t)buf_v[p];
    }
  }
}

template <int RB>
static constexpr size_t rs_smem_bytes() {
  constexpr size_t cnt = (size_t)RS_WARPS * ((1 << RB) / 2), stage = 2 * (size_t)RS_TILE;
  return ((cnt > stage ? cnt : stage) + 2 * ((size_t)1 << RB)) * sizeof(uint32_t);
}

// per-device kernel attributes (opt-in shared memory), set once when a context is created on the device
int sort_init_device(nb_context* ctx) {
  (void)ctx;
#define NB_RS_ATTR(RB_)                                                                                    \
  NB_CUDA(cudaFuncSetAttribute(rs_pass_kernel<RB_, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                               (int)rs_smem_bytes<RB_>()));                                               \
  NB_CUDA(cudaFuncSetAttribute(rs_pass_kernel<RB_, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                               (int)rs_smem_bytes<RB_>()))
  NB_RS_ATTR(4);
  NB_RS_ATTR(8);
  NB_RS_ATTR(10);
  NB_RS_ATTR(11);
#undef NB_RS_ATTR
  return NB_OK;
}

void sort_destroy(nb_context* ctx) {
  if (ctx->h_sort_err) cudaFreeHost(ctx->h_sort_err);
  ctx->h_sort_err = nullptr;
}

int sort_poll_error(nb_context* ctx) {
  if (ctx->h_sort_err && *ctx->h_sort_err) {
    *ctx->h_sort_err = 0;
    set_error("radix sort: a tile waited more than 2 s for its predecessors (device fault?); result invalid");
    return NB_ERR_INTERNAL;
  }
  return NB_OK;
}

// Sorts (keys, vals) by bits [0, end_bit) of the keys.  The passes ping-pong between (d_keys, d_vals) and the
// caller's second pair of buffers (keys_alt, vals_alt); *out_keys / *out_vals say where the result is (no copy).
// identity_vals: vals need not be initialised -- the first pass writes value = index.  copy_keys / copy_vals
// (optional): the pairs are there, not in (d_keys, d_vals) yet; the histogram sweep copies them over.
int sort_pairs_db(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, uint32_t* keys_alt, int32_t* vals_alt, size_t n,
                  int end_bit, bool identity_vals, uint32_t** out_keys, int32_t** out_vals, cudaStream_t st,
                  const uint32_t* copy_keys = nullptr, const int32_t* copy_vals = nullptr) {
  *out_keys = d_keys;
  *out_vals = d_vals;
  if (end_bit < 1) end_bit = 1;
  if (end_bit > 32) end_bit = 32;
  if (n > (size_t)RS_VAL_MASK) {
    set_error("sort: too many elements");
    return NB_ERR_INVALID;
  }
  if (n == 0) return NB_OK;
  if (!ctx->h_sort_err) {
    NB_CUDA(cudaHostAlloc(&ctx->h_sort_err, sizeof(int), cudaHostAllocMapped));
    *ctx->h_sort_err = 0;
    NB_CUDA(cudaHostGetDevicePointer(&ctx->d_sort_err, ctx->h_sort_err, 0));
  }
  if (int rc = sort_poll_error(ctx)) return rc;
  const RsPlan plan = rs_plan(end_bit);
  const size_t ntiles = (n + RS_TILE - 1) / RS_TILE;
  // scratch: [passes][2048] histograms, [passes] tickets (padded to 64 words), per pass [ntiles][2^bits] status
  const size_t head = (size_t)RS_MAX_PASSES * RS_MAX_BINS + 64;
  size_t words = head;
  for (int p = 0; p < plan.passes; p++) words += ntiles << plan.bits[p];
  if (int rc = ctx->sort_hist.ensure(words * sizeof(uint32_t))) return rc;
  uint32_t* hist = ctx->sort_hist.as<uint32_t>();
  uint32_t* ticket = hist + RS_MAX_PASSES * RS_MAX_BINS;
  uint32_t* status = hist + head;
  NB_CUDA(cudaMemsetAsync(hist, 0, words * sizeof(uint32_t), st));
  {
    size_t g = (n / 4 + 511) / 512;
    const size_t cap = (size_t)ctx->sm_count * 4;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    const size_t hsm = (size_t)plan.passes * RS_MAX_BINS * sizeof(uint32_t);
    const bool aligned = (((uintptr_t)d_keys | (uintptr_t)d_vals | (uintptr_t)copy_keys | (uintptr_t)copy_vals) & 15u) == 0;
    if (copy_keys && aligned) {
      rs_hist_kernel<true><<<(unsigned)g, 512, hsm, st>>>(copy_keys, n, plan, hist, copy_vals, d_keys, d_vals);
    } else {
      if (copy_keys) {  // odd alignment: plain copies, then the histogram
        NB_CUDA(cudaMemcpyAsync(d_keys, copy_keys, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        NB_CUDA(cudaMemcpyAsync(d_vals, copy_vals, n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
      }
      rs_hist_kernel<false><<<(unsigned)g, 512, hsm, st>>>(d_keys, n, plan, hist, nullptr, nullptr, nullptr);
    }
    NB_LAUNCH_CHECK(ctx);
    rs_scan_kernel<<<plan.passes, RS_MAX_BINS / 2, 0, st>>>(hist);
    NB_LAUNCH_CHECK(ctx);
  }
  uint32_t *ka = d_keys, *kb = keys_alt;
  int32_t *va = d_vals, *vb = vals_alt;
  for (int p = 0; p < plan.passes; p++) {
    const bool ident = identity_vals && p == 0;
#define NB_RS_LAUNCH(RB_, ID_)                                                                                        \
  rs_pass_kernel<RB_, ID_><<<(unsigned)ntiles, RS_THREADS, rs_smem_bytes<RB_>(), st>>>(                               \
      ka, va, kb, vb, n, plan.shift[p], plan.bits[p], hist + p * RS_MAX_BINS, status, ticket + p, ctx->d_sort_err)
    if (plan.bits[p] <= 4) {
      if (ident) NB_RS_LAUNCH(4, true);
      else NB_RS_LAUNCH(4, false);
    } else if (plan.bits[p] <= 8) {
      if (ident) NB_RS_LAUNCH(8, true);
      else NB_RS_LAUNCH(8, false);
    } else if (plan.bits[p] <= 10) {
      if (ident) NB_RS_LAUNCH(10, true);
      else NB_RS_LAUNCH(10, false);
    } else {
      if (ident) NB_RS_LAUNCH(11, true);
      else NB_RS_LAUNCH(11, false);
    }
#undef NB_RS_LAUNCH
    NB_LAUNCH_CHECK(ctx);
    status += ntiles << plan.bits[p];
    uint32_t* tk = ka; ka = kb; kb = tk;
    int32_t* tv = va; va = vb; vb = tv;
  }
  *out_keys = ka;
  *out_vals = va;
  return NB_OK;
}

// sorts in place (result in d_keys/d_vals); bits [0, end_bit) are significant
int sort_pairs(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, size_t n, int end_bit, cudaStream_t st) {
  if (n < 2) return NB_OK;
  uint32_t* ko;
  int32_t* vo;
  if (int rc = ctx->sort_keys2.ensure(n * sizeof(uint32_t))) return rc;
  if (int rc = ctx->sort_vals2.ensure(n * sizeof(int32_t))) return rc;
  uint32_t* k2 = ctx->sort_keys2.as<uint32_t>();
  int32_t* v2 = ctx->sort_vals2.as<int32_t>();
  if (end_bit < 1) end_bit = 1;
  if (end_bit > 32) end_bit = 32;
  if (rs_plan(end_bit).passes % 2) {
    // odd number of passes: start from the scratch copy (made by the histogram sweep) so that the last pass
    // writes the caller's arrays
    if (int rc = sort_pairs_db(ctx, k2, v2, d_keys, d_vals, n, end_bit, false, &ko, &vo, st, d_keys, d_vals)) return rc;
  } else {
    if (int rc = sort_pairs_db(ctx, d_keys, d_vals, k2, v2, n, end_bit, false, &ko, &vo, st)) return rc;
  }
  if (ko != d_keys) {  // not reached with the plans above
    NB_CUDA(cudaMemcpyAsync(d_keys, ko, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    NB_CUDA(cudaMemcpyAsync(d_vals, vo, n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
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

int nb_sort_passes(int end_bit) {
  if (end_bit < 1) end_bit = 1;
  if (end_bit > 32) end_bit = 32;
  return rs_plan(end_bit).passes;
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
