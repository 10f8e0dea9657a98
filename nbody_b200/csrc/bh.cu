// bh.cu -- Barnes-Hut on the GPU, rebuilt every step: bbox -> Morton keys -> radix sort ->
// Karras LBVH -> bottom-up centres of mass -> warp-cooperative traversal.
//
// Replaces BarnesHutTree::build / computeForces (reference src/cuda/force_barnes_hut.cu:282-289,
// 488-498).  The reference copies everything to the host, inserts bodies one by one into an octree
// on one CPU thread (and, as shipped, never splits a leaf: SURVEY.md sec. 0) and traverses it with a
// per-thread 1 KiB local-memory stack.  Here the integer front half (bbox, keys, stable sort) is
// bit-exact with the reference's arithmetic (:23-38, :113-127, :260-280), the tree is a binary radix
// tree over those keys (Karras 2012), and the traversal keeps the reference's acceptance test
// (is-leaf or size^2/(d^2+eps^2) < theta^2, :167) and monopole formula (:175-181), with "size" the
// side of the Morton cell the node lives in (== the reference's octree cell 2*half_size at every
// third level for a cubic box).  Bodies walk the tree in groups (maximal subtrees of <= 64 leaves,
// one warp each): a node is used only if it passes the test at the group's bounding box, which
// implies it passes for every body of the group, so each body's force is at least as accurate as
// with its own per-body test.
#include <cstdlib>
#include <new>
#include <cstring>

#include "bh.cuh"
#include "direct.cuh"

namespace nb {

int bbox_device(nb_context* ctx, const nb_particles* d, cudaStream_t st);
int sort_pairs(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, size_t n, int end_bit, cudaStream_t st);

}  // namespace nb

struct BH {
  size_t n = 0;
  nb::DevBuf keys, perm, packed, node_com, meta, parent, flags, misc;  // misc: size2 table[32] + counter + bbox floats
  nb::DevBuf range, groups;  // int2 {first,last} per internal node; int2 {first,count} per traversal group
  float h_bbox[6] = {0, 0, 0, 0, 0, 0};
  int* h_flag = nullptr;  // host-mapped: traversal stack overflow
  int* d_flag = nullptr;
  int group = 64;         // max bodies per traversal group / warp (32 or 64; NB_BH_GROUP overrides)
  int super = 512;        // max leaves of the subtree a group is cut from (NB_BH_SUPER overrides; >= group)
  int occ = 0;
  int stack_cap = 2048;   // entries per warp traversal stack (NB_BH_STACK overrides)
  bool built = false;
};

namespace nb {

__device__ __forceinline__ float ord2f_dev(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// src/cuda/force_barnes_hut.cu:23-29
__device__ __forceinline__ unsigned expand_bits(unsigned v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

// misc layout (floats): [0..31] size2 table by prefix length, [32..37] bbox, [40..41] u64 counter,
// [44] int number of traversal groups, [45] int work-queue cursor of the traversal kernel,
// [46] int deepest traversal stack of the last force evaluation
__global__ void bh_prepare_kernel(const int* __restrict__ bb_ord, float* __restrict__ misc) {
  const int t = threadIdx.x;
  __shared__ float bb[6];
  if (t < 6) {
    bb[t] = ord2f_dev(bb_ord[t]);
    misc[32 + t] = bb[t];
  }
  if (t == 6) *reinterpret_cast<unsigned long long*>(misc + 40) = 0ull;
  if (t == 7) reinterpret_cast<int*>(misc)[44] = 0;  // number of traversal groups
  if (t == 8) reinterpret_cast<int*>(misc)[46] = 0;  // deepest traversal stack seen
  __syncthreads();
  if (t <= 30) {
    // side of the Morton cell at prefix length L: x is split first, then y, then z
    // (x is bit 2 of each key triple, :37); extents as in :265-267
    const float sx = __fadd_rn(__fsub_rn(bb[3], bb[0]), 0.001f), sy = __fadd_rn(__fsub_rn(bb[4], bb[1]), 0.001f),
                sz = __fadd_rn(__fsub_rn(bb[5], bb[2]), 0.001f);
    const float ex = ldexpf(sx, -((t + 2) / 3)), ey = ldexpf(sy, -((t + 1) / 3)), ez = ldexpf(sz, -(t / 3));
    const float s = fmaxf(ex, fmaxf(ey, ez));
    misc[t] = __fmul_rn(s, s);
  }
  if (t == 31) misc[31] = 0.f;
}

// :260-274 (host part: size = max-min+0.001f, inv = 1.0f/size) and :113-127 (device part)
__global__ void __launch_bounds__(256) bh_morton_kernel(const float* __restrict__ px, const float* __restrict__ py,
                                                        const float* __restrict__ pz,
                                                        const float* __restrict__ mass, size_t n,
                                                        const float* __restrict__ misc, uint32_t* __restrict__ keys,
                                                        int32_t* __restrict__ perm, float4* __restrict__ packed) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = px[i], y = py[i], z = pz[i];
  // {x,y,z,m} in body order: the leaf gather after the sort is then ONE random 16-B read per body
  // instead of four random 4-B reads (each of which costs a whole DRAM sector)
  packed[i] = make_float4(x, y, z, mass[i]);
  const float mnx = misc[32], mny = misc[33], mnz = misc[34];
  const float ix = __fdiv_rn(1.0f, __fadd_rn(__fsub_rn(misc[35], mnx), 0.001f));
  const float iy = __fdiv_rn(1.0f, __fadd_rn(__fsub_rn(misc[36], mny), 0.001f));
  const float iz = __fdiv_rn(1.0f, __fadd_rn(__fsub_rn(misc[37], mnz), 0.001f));
  const float nx = __fmul_rn(__fsub_rn(x, mnx), ix);
  const float ny = __fmul_rn(__fsub_rn(y, mny), iy);
  const float nz = __fmul_rn(__fsub_rn(z, mnz), iz);
  // static_cast<unsigned>(x * 1024.0f) clamped to [0, 1023] (:31-35); cvt.rzi.u32.f32 saturates
  const unsigned ux = min(__float2uint_rz(__fmul_rn(nx, 1024.0f)), 1023u);
  const unsigned uy = min(__float2uint_rz(__fmul_rn(ny, 1024.0f)), 1023u);
  const unsigned uz = min(__float2uint_rz(__fmul_rn(nz, 1024.0f)), 1023u);
  keys[i] = (expand_bits(ux) << 2) | (expand_bits(uy) << 1) | expand_bits(uz);
  perm[i] = (int32_t)i;
}

// leaves of the tree = bodies in sorted order, packed {x,y,z,m}
__global__ void __launch_bounds__(256) bh_gather_kernel(const float4* __restrict__ packed, size_t n,
                                                        const int32_t* __restrict__ perm,
                                                        float4* __restrict__ leaves) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  leaves[k] = __ldg(packed + perm[k]);
}

__device__ __forceinline__ int bh_delta(const uint32_t* __restrict__ k, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint32_t a = k[i], b = k[j];
  if (a == b) return 32 + __clz((unsigned)i ^ (unsigned)j);
  return __clz(a ^ b);
}

// Karras 2012, sec. 4: one thread per internal node.  meta[i] = {left, right, prefix, 0}.
__global__ void __launch_bounds__(256) bh_karras_kernel(const uint32_t* __restrict__ keys, int n,
                                                        int4* __restrict__ meta, int32_t* __restrict__ parent,
                                                        int2* __restrict__ range) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int leaf0 = n - 1;
  const int d = (bh_delta(keys, n, i, i + 1) - bh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  const int dmin = bh_delta(keys, n, i, i - d);
  int lmax = 2;
  while (bh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (bh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = bh_delta(keys, n, i, j);
  int s = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (bh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  const int gamma = i + s * d + (d < 0 ? d : 0);
  const int lo = min(i, j), hi = max(i, j);
  const int lc = (lo == gamma) ? leaf0 + gamma : gamma;
  const int rc = (hi == gamma + 1) ? leaf0 + gamma + 1 : gamma + 1;
  int p = dnode - 2;
  if (p > 30) p = 30;
  meta[i] = make_int4(lc, rc, p, 0);
  range[i] = make_int2(lo, hi);
  parent[lc] = i;
  parent[rc] = i;
  if (i == 0) parent[0] = -1;
}

// bottom-up centres of mass: the second thread to reach a node combines its two children.
// c = (ca*ma + cb*mb) / (ma+mb), a = left child, b = right child: order-independent of arrival.
__global__ void __launch_bounds__(256) bh_com_kernel(int n, const int4* __restrict__ meta,
                                                     const int32_t* __restrict__ parent, int* __restrict__ flags,
                                                     float4* node_com) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int p = parent[n - 1 + k];
  while (p >= 0) {
    __threadfence();
    if (atomicAdd(&flags[p], 1) == 0) return;
    const int4 m = meta[p];
    const float4 a = __ldcg(node_com + m.x), b = __ldcg(node_com + m.y);
    const float ms = __fadd_rn(a.w, b.w);
    float4 c;
    if (ms > 0.f) {
      c.x = __fdiv_rn(__fadd_rn(__fmul_rn(a.x, a.w), __fmul_rn(b.x, b.w)), ms);
      c.y = __fdiv_rn(__fadd_rn(__fmul_rn(a.y, a.w), __fmul_rn(b.y, b.w)), ms);
      c.z = __fdiv_rn(__fadd_rn(__fmul_rn(a.z, a.w), __fmul_rn(b.z, b.w)), ms);
    } else {
      c.x = a.x; c.y = a.y; c.z = a.z;
    }
    c.w = ms;
    __stcg(node_com + p, c);
    p = parent[p];
  }
}

// The same bottom-up pass with the lower levels in shared memory.  A CTA owns COM_TILE consecutive leaves; the
// internal nodes whose leaf range lies inside that tile (all but ~1 in COM_TILE of them; node i always lies inside
// its own range, so their indices fall in the tile too) have their topology staged in shared memory, arrive on
// shared-memory counters and hand their values up through shared memory; only nodes that span tiles go through the
// global flags as above.  Every node is still combine(left child, right child) with the same operations, so the
// result is bit-identical to bh_com_kernel's.
constexpr int COM_TILE = 512;
__device__ __forceinline__ float4 bh_combine(const float4& a, const float4& b) {
  const float ms = __fadd_rn(a.w, b.w);
  float4 c;
  if (ms > 0.f) {
    c.x = __fdiv_rn(__fadd_rn(__fmul_rn(a.x, a.w), __fmul_rn(b.x, b.w)), ms);
    c.y = __fdiv_rn(__fadd_rn(__fmul_rn(a.y, a.w), __fmul_rn(b.y, b.w)), ms);
    c.z = __fdiv_rn(__fadd_rn(__fmul_rn(a.z, a.w), __fmul_rn(b.z, b.w)), ms);
  } else {
    c.x = a.x; c.y = a.y; c.z = a.z;
  }
  c.w = ms;
  return c;
}
__global__ void __launch_bounds__(COM_TILE) bh_com_tile_kernel(int n, const int4* __restrict__ meta,
                                                               const int32_t* __restrict__ parent,
                                                               const int2* __restrict__ range, int* __restrict__ flags,
                                                               float4* node_com) {
  __shared__ int4 s_meta[COM_TILE];
  __shared__ int2 s_range[COM_TILE];
  __shared__ int s_parent[COM_TILE];
  __shared__ int s_cnt[COM_TILE];
  __shared__ float4 s_com[COM_TILE];
  const int tid = threadIdx.x;
  const int t0 = blockIdx.x * COM_TILE, t1 = min(t0 + COM_TILE, n);  // this CTA's leaves [t0, t1)
  const int leaf0 = n - 1;
  if (t0 + tid < n - 1) {
    s_meta[tid] = meta[t0 + tid];
    s_range[tid] = range[t0 + tid];
    s_parent[tid] = parent[t0 + tid];
  }
  s_cnt[tid] = 0;
  __syncthreads();
  const int k = t0 + tid;
  if (k >= n) return;
  int node = leaf0 + k;
  int p = parent[node];
  float4 cur = __ldcg(node_com + node);
  while (p >= 0) {  // levels inside the tile
    const int q = p - t0;
    if (q < 0 || q >= COM_TILE) break;
    const int2 r = s_range[q];
    if (r.x < t0 || r.y >= t1) break;
    __threadfence_block();
    if (atomicAdd(&s_cnt[q], 1) == 0) return;  // the sibling's thread finishes this node
    __threadfence_block();  // the sibling's value was published before its arrival: read it after ours
    const int4 m = s_meta[q];
    const bool left = (m.x == node);
    const int sib = left ? m.y : m.x;
    const float4 sv = (sib >= leaf0) ? __ldcg(node_com + sib) : s_com[sib - t0];
    cur = left ? bh_combine(cur, sv) : bh_combine(sv, cur);
    s_com[q] = cur;
    __stcg(node_com + p, cur);
    node = p;
    p = s_parent[q];
  }
  while (p >= 0) {  // nodes spanning tiles: as bh_com_kernel
    __threadfence();
    if (atomicAdd(&flags[p], 1) == 0) return;
    const int4 m = meta[p];
    const float4 c = bh_combine(__ldcg(node_com + m.x), __ldcg(node_com + m.y));
    __stcg(node_com + p, c);
    p = parent[p];
  }
}

// Traversal groups.  A "supergroup" is a maximal subtree with at most `smax` leaves: its bodies share
// a Morton prefix, so it is as compact as its cell (a fixed run of consecutive sorted bodies can
// straddle distant cells in sparse regions).  Each supergroup of c bodies is cut into
// ceil(c / gmax) runs of consecutive bodies of (almost) equal length <= gmax, run i starting at
// floor(i * c / runs): with smax == gmax the groups are the subtrees themselves (mean fill 69 % of
// gmax), with smax = 8 * gmax the runs are 89-100 % full and still confined to one cell.
// One thread per leaf climbs to its supergroup's root; the leaf that starts a run publishes it.
__global__ void __launch_bounds__(256) bh_groups_kernel(int n, int smax, int gmax, const int32_t* __restrict__ parent,
                                                        const int2* __restrict__ range, int2* __restrict__ groups,
                                                        int* __restrict__ ngroups) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int first = k, count = 1;
  int p = parent[n - 1 + k];
  while (p >= 0) {
    const int2 r = range[p];
    if (r.y - r.x + 1 > smax) break;
    first = r.x;
    count = r.y - r.x + 1;
    p = parent[p];
  }
  const int runs = (count + gmax - 1) / gmax;
  const int r = k - first;
  const int i = (int)(((long long)r * runs + count - 1) / count);  // first run starting at or after r
  if (i < runs && (int)(((long long)i * count) / runs) == r) {
    const int next = (int)(((long long)(i + 1) * count) / runs);
    groups[atomicAdd(ngroups, 1)] = make_int2(k, next - r);
  }
}

// ---- traversal ---------------------------------------------------------------------------------------
// One warp owns 32*BPL Morton-consecutive bodies and a stack of node indices in shared memory.
// Each round the warp pops up to 32 nodes, ONE PER LANE: every lane fetches its node and tests it
// against the group's bounding box (dmin = distance from the node's centre of mass to the box, so
// dmin <= every body's own distance and a node the group accepts is a node every body would accept
// with the reference's per-body test, :160-172).  __ballot_sync + popc compact the accepted nodes
// into an interaction list and the children of the rejected ones back onto the stack.  Whenever
// the list holds 32 entries they are evaluated: entry e is a warp-broadcast LDS.128 and every lane
// adds the monopole term (:175-181) for its own bodies.  Order is deterministic (lane order), and
// oracle/lbvh.c restates it step for step.
#ifndef NB_BH_UNROLL
#define NB_BH_UNROLL 32  // list entries per trip of the evaluation loop (8: 33.3 ms traversal at 16 Mi, 32: 32.6)
#endif
#define NB_BH_PRAGMA_(x) _Pragma(#x)
#define NB_BH_PRAGMA_UNROLL(n) NB_BH_PRAGMA_(unroll n)
constexpr int BH_WARPS = 4;
constexpr int BH_LIST = 64;

// SHARDED (multi-GPU, mgpu.cu): every rank holds the same tree; this rank traverses only the groups it
// is dealt (by the group's first sorted body, so the deal is identical on every rank although the
// order of the group array is not) and stores each body's acceleration, in SORTED order, into the
// exchange buffer of EVERY rank over NVLink peer mappings -- contiguous runs of up to 64 float4 per
// group and peer -- so the all-gather of accelerations is the traversal kernel's own epilogue.  The
// last CTA to finish raises this rank's step flag on every peer.
template <int BPL, bool EPS0, bool SHARDED>
__global__ void __launch_bounds__(BH_WARPS * 32) bh_force_kernel(int n, const float4* __restrict__ node_com,
                                                                  const int4* __restrict__ meta,
                                                                  const int32_t* __restrict__ perm,
                                                                  const float* __restrict__ misc, float theta2, float G,
                                                                  float eps2, float* __restrict__ ax,
                                                                  float* __restrict__ ay, float* __restrict__ az,
                                                                  unsigned long long* __restrict__ counter,
                                                                  int* __restrict__ overflow,
                                                                  const int2* __restrict__ groups,
                                                                  const int* __restrict__ ngroups,
                                                                  int* __restrict__ cursor, int stack_cap,
                                                                  int* __restrict__ max_stack, BhShard sh) {
  extern __shared__ __align__(16) unsigned char bh_smem[];
  __shared__ float size2_tab[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x < 32) size2_tab[threadIdx.x] = misc[threadIdx.x];
  __syncthreads();
  const size_t smem_per_warp = (size_t)stack_cap * 4 + BH_LIST * 16;
  float4* list = reinterpret_cast<float4*>(bh_smem + (size_t)w * smem_per_warp);
  int* stack = reinterpret_cast<int*>(bh_smem + (size_t)w * smem_per_warp + BH_LIST * 16);
  int deepest = 0;
  const int leaf0 = n - 1;
  const unsigned lt = (1u << lane) - 1u;
  const int total_groups = *ngroups;
  unsigned long long visits_bodies = 0;

  for (;;) {  // persistent warp: pull traversal groups from a work queue until it is empty
  int g = 0;
  if (lane == 0) g = atomicAdd(cursor, 1);
  g = __shfl_sync(0xffffffffu, g, 0);
  if (g >= total_groups) break;
  const int2 grp = groups[g];
  const int k0 = grp.x, gcnt = grp.y;
  if (SHARDED && (((unsigned)k0 * 2654435761u) >> 20) % (unsigned)sh.nranks != (unsigned)sh.rank) continue;  // dealt to another rank

  // scalar loads on purpose: a float4 load pins x,y,z to one aligned register quad, and ptxas then
  // re-packs the (body 0, body 1) pairs with MOVs for every list entry (4.5 MOVs per entry measured)
  struct { float x, y, z; } me[BPL];
  float sx[BPL], sy[BPL], sz[BPL];
  float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
#pragma unroll
  for (int b = 0; b < BPL; b++) {
    me[b].x = me[b].y = me[b].z = 0.f;
    if (b * 32 + lane < gcnt) {
      const float* q = reinterpret_cast<const float*>(node_com + leaf0 + k0 + b * 32 + lane);
      me[b].x = __ldg(q); me[b].y = __ldg(q + 1); me[b].z = __ldg(q + 2);
      lo[0] = fminf(lo[0], me[b].x); hi[0] = fmaxf(hi[0], me[b].x);
      lo[1] = fminf(lo[1], me[b].y); hi[1] = fmaxf(hi[1], me[b].y);
      lo[2] = fminf(lo[2], me[b].z); hi[2] = fmaxf(hi[2], me[b].z);
    }
    sx[b] = sy[b] = sz[b] = 0.f;
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
    lo[a] = warp_min(lo[a]);
    hi[a] = warp_max(hi[a]);
  }
  const f2 e2 = pk(eps2, eps2);
  const f2 xi2 = pk(me[0].x, me[BPL - 1].x), yi2 = pk(me[0].y, me[BPL - 1].y), zi2 = pk(me[0].z, me[BPL - 1].z);
  f2 sx2 = pk(0.f, 0.f), sy2 = sx2, sz2 = sx2;

  int sp = 1, nl = 0;
  unsigned visits = 0;
  if (lane == 0) stack[0] = 0;
  __syncwarp();
  for (;;) {
    const int m = min(sp, 32);
    const bool have = lane < m;
    const int ni = have ? stack[sp - m + lane] : 0;
    sp -= m;
    __syncwarp();
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    int4 mt = make_int4(0, 0, 31, 0);
    const bool leaf = ni >= leaf0;
    if (have) {  // both fetches issue together (no dependence of the second on the first's data)
      c = node_com[ni];
      if (!leaf) mt = meta[ni];
    }
    const bool valid = have && c.w != 0.f;  // total_mass == 0 -> skipped (:157-158)
    bool acc = leaf;
    if (valid && !leaf) {
      const float ddx = fmaxf(fmaxf(lo[0] - c.x, c.x - hi[0]), 0.f);
      const float ddy = fmaxf(fmaxf(lo[1] - c.y, c.y - hi[1]), 0.f);
      const float ddz = fmaxf(fmaxf(lo[2] - c.z, c.z - hi[2]), 0.f);
      const float d2 = fmaf(ddz, ddz, fmaf(ddy, ddy, fmaf(ddx, ddx, eps2)));
      acc = size2_tab[mt.z] < theta2 * d2;  // size2 / dist2 < theta2 (:167)
    }
    const bool is_acc = valid && acc, is_rej = valid && !acc;
    const unsigned amask = __ballot_sync(0xffffffffu, is_acc), rmask = __ballot_sync(0xffffffffu, is_rej);
    if (is_acc) list[nl + __popc(amask & lt)] = make_float4(c.x, c.y, c.z, G * c.w);  // G*M once per node
    const int npush = 2 * __popc(rmask);
    deepest = max(deepest, sp + npush);
    if (sp + npush > stack_cap) {  // never silently drop subtrees (the reference does, :186-194)
      if (lane == 0) *overflow = 1;
      break;
    }
    if (is_rej) {
      const int pos = sp + 2 * __popc(rmask & lt);
      stack[pos] = mt.x;
      stack[pos + 1] = mt.y;
    }
    nl += __popc(amask);
    sp += npush;
    __syncwarp();
    const bool done = (sp == 0);
    while (nl >= 32 || (done && nl > 0)) {
      const int ne = min(nl, 32);
      auto body = [&](int e) {
        const float4 p = list[e];  // warp-broadcast LDS.128; p.w = G * M
        const float gm = p.w;
        if constexpr (BPL == 2) {
          // the lane's two bodies as one packed fp32x2 pair (FADD2/FFMA2/FMUL2, as in direct.cu)
          const f2 dx = sub2(pk(p.x, p.x), xi2), dy = sub2(pk(p.y, p.y), yi2), dz = sub2(pk(p.z, p.z), zi2);
          const f2 d2 = fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, e2)));
          float r0, r1;
          upk(d2, r0, r1);
          float i0 = rsq(r0), i1 = rsq(r1);
          if (EPS0) {  // own leaf / exact coincidence at eps = 0
            i0 = (r0 > 0.f) ? i0 : 0.f;
            i1 = (r1 > 0.f) ? i1 : 0.f;
          }
          const f2 ri = pk(i0, i1);
          const f2 f = mul2(pk(gm, gm), mul2(mul2(ri, ri), ri));
          sx2 = fma2(f, dx, sx2);
          sy2 = fma2(f, dy, sy2);
          sz2 = fma2(f, dz, sz2);
        } else {
#pragma unroll
          for (int b = 0; b < BPL; b++) {
            const float dx = p.x - me[b].x, dy = p.y - me[b].y, dz = p.z - me[b].z;
            const float d2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps2)));
            float inv = rsq(d2);
            if (EPS0) inv = (d2 > 0.f) ? inv : 0.f;  // own leaf / exact coincidence at eps = 0
            const float f = gm * (inv * inv * inv);
            sx[b] = fmaf(f, dx, sx[b]);
            sy[b] = fmaf(f, dy, sy[b]);
            sz[b] = fmaf(f, dz, sz[b]);
          }
        }
      };
      if (ne == 32) {
        NB_BH_PRAGMA_UNROLL(NB_BH_UNROLL)
        for (int e = 0; e < 32; e++) body(e);
      } else {
        for (int e = 0; e < ne; e++) body(e);
      }
      visits += ne;
      __syncwarp();
      const int rest = nl - ne;  // < 32
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
      if (lane < rest) t = list[ne + lane];
      __syncwarp();
      if (lane < rest) list[lane] = t;
      __syncwarp();
      nl = rest;
    }
    if (done) break;
  }
  if constexpr (BPL == 2) {
    upk(sx2, sx[0], sx[1]);
    upk(sy2, sy[0], sy[1]);
    upk(sz2, sz[0], sz[1]);
  }
#pragma unroll
  for (int b = 0; b < BPL; b++) {
    if (b * 32 + lane < gcnt) {
      if (SHARDED) {
        const float4 v = make_float4(sx[b], sy[b], sz[b], 0.f);
#pragma unroll
        for (int r = 0; r < 8; r++)
          if (r < sh.nranks) sh.out[r][k0 + b * 32 + lane] = v;
      } else {
        const int i = perm[k0 + b * 32 + lane];
        ax[i] = sx[b];
        ay[i] = sy[b];
        az[i] = sz[b];
      }
    }
  }
  visits_bodies += (unsigned long long)visits * (unsigned long long)gcnt;
  __syncwarp();
  }  // next group
  if (lane == 0 && visits_bodies) {
    atomicAdd(counter, visits_bodies);
    atomicMax(max_stack, deepest);
  }
  if (SHARDED) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned prev = atomicAdd(sh.done_counter, 1u);
      if (prev == gridDim.x - 1) {  // every CTA's peer stores are fenced: publish
        *sh.done_counter = 0;
        __threadfence_system();
        for (int r = 0; r < sh.nranks; r++)
          if (r != sh.rank)
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(sh.flags[r] + sh.flag_index), "r"(sh.step)
                         : "memory");
      }
    }
  }
}

// accelerations arrive in sorted order; hand them back to body order, optionally with the kick
// (src/cuda/integrator.cu:24-36, same rounding sequence as integrate.cu)
template <bool KICK>
__global__ void __launch_bounds__(256) bh_shard_finish_kernel(const float4* __restrict__ acc_sorted,
                                                              const int32_t* __restrict__ perm, nb_particles d,
                                                              size_t n, float dt) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int i = perm[k];
  const float4 a = acc_sorted[k];
  d.acc_x[i] = a.x;
  d.acc_y[i] = a.y;
  d.acc_z[i] = a.z;
  if (KICK) {
    const float dt_half = __fmul_rn(0.5f, dt);
    d.vel_x[i] = __fadd_rn(d.vel_x[i], __fmul_rn(__fadd_rn(d.acc_old_x[i], a.x), dt_half));
    d.vel_y[i] = __fadd_rn(d.vel_y[i], __fmul_rn(__fadd_rn(d.acc_old_y[i], a.y), dt_half));
    d.vel_z[i] = __fadd_rn(d.vel_z[i], __fmul_rn(__fadd_rn(d.acc_old_z[i], a.z), dt_half));
  }
}

static BH* bh_get(nb_context* ctx) {
  if (!ctx->bh) ctx->bh = new (std::nothrow) BH();
  return ctx->bh;
}

void bh_destroy(nb_context* ctx) {
  if (!ctx->bh) return;
  BH* b = ctx->bh;
  b->keys.release(); b->perm.release(); b->packed.release(); b->node_com.release(); b->meta.release();
  b->parent.release(); b->flags.release(); b->misc.release(); b->range.release(); b->groups.release();
  if (b->h_flag) cudaFreeHost(b->h_flag);
  delete b;
  ctx->bh = nullptr;
}

int bh_build(nb_context* ctx, const nb_particles* d, cudaStream_t st) {
  BH* b = bh_get(ctx);
  if (!b) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  const size_t n = d->count;
  b->built = false;
  if (n == 0) return NB_OK;
  if (n > 0x3fffffffull) {
    set_error("Barnes-Hut: particle count %zu too large", n);
    return NB_ERR_INVALID;
  }
  b->n = n;
  if (int rc = b->keys.ensure(n * 4)) return rc;
  if (int rc = b->perm.ensure(n * 4)) return rc;
  if (int rc = b->packed.ensure(n * sizeof(float4))) return rc;
  if (int rc = b->node_com.ensure((2 * n) * sizeof(float4))) return rc;
  if (int rc = b->meta.ensure(n * sizeof(int4))) return rc;
  if (int rc = b->parent.ensure(2 * n * 4)) return rc;
  if (int rc = b->flags.ensure(n * 4)) return rc;
  if (int rc = b->misc.ensure(64 * 4)) return rc;
  if (int rc = b->range.ensure(n * sizeof(int2))) return rc;
  if (int rc = b->groups.ensure(n * sizeof(int2))) return rc;
  {
    const char* e = getenv("NB_BH_GROUP");
    const int g = (e && atoi(e) == 32) ? 32 : 64;
    if (g != b->group) b->occ = 0;
    b->group = g;
    if (const char* ce = getenv("NB_BH_STACK")) {
      int c = atoi(ce);
      c = c < 512 ? 512 : (c > 2816 ? 2816 : c);  // 4 warps x (cap x 4 + 1 KiB) <= 48 KiB
      if (c != b->stack_cap) b->occ = 0;
      b->stack_cap = c;
    }
    const char* se = getenv("NB_BH_SUPER");
    int sg = se ? atoi(se) : 8 * g;
    if (sg < g) sg = g;
    b->super = sg;
  }
  float* misc = b->misc.as<float>();
  {
    NB_STAGE(ctx, st, "bh.bbox");
    if (int rc = bbox_device(ctx, d, st)) return rc;
    bh_prepare_kernel<<<1, 32, 0, st>>>(ctx->bbox_dev.as<int>(), misc);
    NB_LAUNCH_CHECK(ctx);
  }
  const unsigned gn = (unsigned)((n + 255) / 256);
  {
    NB_STAGE(ctx, st, "bh.morton");
    bh_morton_kernel<<<gn, 256, 0, st>>>(d->pos_x, d->pos_y, d->pos_z, d->mass, n, misc, b->keys.as<uint32_t>(),
                                         b->perm.as<int32_t>(), b->packed.as<float4>());
    NB_LAUNCH_CHECK(ctx);
  }
  {
    NB_STAGE(ctx, st, "bh.sort");
    if (int rc = sort_pairs(ctx, b->keys.as<uint32_t>(), b->perm.as<int32_t>(), n, 30, st)) return rc;
  }
  float4* com = b->node_com.as<float4>();
  {
    NB_STAGE(ctx, st, "bh.gather");
    bh_gather_kernel<<<gn, 256, 0, st>>>(b->packed.as<float4>(), n, b->perm.as<int32_t>(), com + (n - 1));
    NB_LAUNCH_CHECK(ctx);
  }
  NB_CUDA(cudaMemsetAsync(b->parent.p, 0xff, sizeof(int32_t), st));  // parent[0] = -1 (also when n == 1)
  if (n > 1) {
    {
      NB_STAGE(ctx, st, "bh.karras");
      NB_CUDA(cudaMemsetAsync(b->flags.p, 0, (n - 1) * 4, st));
      const unsigned gi = (unsigned)((n - 1 + 255) / 256);
      bh_karras_kernel<<<gi, 256, 0, st>>>(b->keys.as<uint32_t>(), (int)n, b->meta.as<int4>(),
                                           b->parent.as<int32_t>(), b->range.as<int2>());
      NB_LAUNCH_CHECK(ctx);
    }
    NB_STAGE(ctx, st, "bh.com");
    if (getenv("NB_BH_COM_GLOBAL"))  // the all-global version, kept for A/B
      bh_com_kernel<<<gn, 256, 0, st>>>((int)n, b->meta.as<int4>(), b->parent.as<int32_t>(), b->flags.as<int>(), com);
    else
      bh_com_tile_kernel<<<(unsigned)((n + COM_TILE - 1) / COM_TILE), COM_TILE, 0, st>>>(
          (int)n, b->meta.as<int4>(), b->parent.as<int32_t>(), b->range.as<int2>(), b->flags.as<int>(), com);
    NB_LAUNCH_CHECK(ctx);
  }
  {
    NB_STAGE(ctx, st, "bh.groups");
    bh_groups_kernel<<<gn, 256, 0, st>>>((int)n, b->super, b->group, b->parent.as<int32_t>(), b->range.as<int2>(),
                                         b->groups.as<int2>(), reinterpret_cast<int*>(misc) + 44);
    NB_LAUNCH_CHECK(ctx);
  }
  b->built = true;
  return NB_OK;
}

static int bh_check_overflow(BH* b) {
  if (b->h_flag && *b->h_flag) {
    *b->h_flag = 0;
    set_error("Barnes-Hut traversal stack overflow (tree deeper than the %d-entry warp stack allows)", b->stack_cap);
    return NB_ERR_INTERNAL;
  }
  return NB_OK;
}

// host-visible error state of launches that have already run (graph replays included): c_api.cu polls it
int bh_poll_error(nb_context* ctx) { return ctx->bh ? bh_check_overflow(ctx->bh) : NB_OK; }

int bh_forces(nb_context* ctx, const nb_particles* d, float theta, float G, float eps, cudaStream_t st,
              const BhShard* shard) {
  BH* b = bh_get(ctx);
  if (!b) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  const size_t n = d->count;
  if (n == 0) return NB_OK;
  if (!b->built || b->n != n) {
    set_error("Barnes-Hut: forces requested without a tree built for these particles");
    return NB_ERR_INVALID;
  }
  if (!b->h_flag) {
    NB_CUDA(cudaHostAlloc(&b->h_flag, sizeof(int), cudaHostAllocMapped));
    *b->h_flag = 0;
    NB_CUDA(cudaHostGetDevicePointer(&b->d_flag, b->h_flag, 0));
  }
  if (int rc = bh_check_overflow(b)) return rc;  // reported by an earlier (asynchronous) launch
  float* misc = b->misc.as<float>();
  int* cursor = reinterpret_cast<int*>(misc) + 45;
  NB_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int), st));
  const size_t smem = (size_t)BH_WARPS * ((size_t)b->stack_cap * 4 + BH_LIST * 16);
  const float eps2 = eps * eps;
  const bool eps0 = !(eps2 > 0.f);
  if (!b->occ) {
    int o = 0;
    if (b->group == 64)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, bh_force_kernel<2, false, false>, BH_WARPS * 32, smem);
    else
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, bh_force_kernel<1, false, false>, BH_WARPS * 32, smem);
    b->occ = o > 0 ? o : 1;
  }
  long long max_groups = (long long)n;  // upper bound (the real count lives on the device)
  long long want = (max_groups + BH_WARPS - 1) / BH_WARPS;
  long long cap = (long long)b->occ * ctx->sm_count;
  const unsigned grid = (unsigned)(want < cap ? want : cap);
  ProfScope prof(ctx, st);
  NB_STAGE(ctx, st, "bh.traversal");
  BhShard none{};
  const BhShard shv = shard ? *shard : none;
#define NB_BH_LAUNCH_S(BPL_, E_, S_)                                                                               \
  bh_force_kernel<BPL_, E_, S_><<<grid, BH_WARPS * 32, smem, st>>>(                                                \
      (int)n, b->node_com.as<float4>(), b->meta.as<int4>(), b->perm.as<int32_t>(), misc, theta * theta, G, eps2,  \
      d->acc_x, d->acc_y, d->acc_z, reinterpret_cast<unsigned long long*>(misc + 40), b->d_flag,                   \
      b->groups.as<int2>(), reinterpret_cast<const int*>(misc) + 44, cursor, b->stack_cap,                        \
      reinterpret_cast<int*>(misc) + 46, shv)
#define NB_BH_LAUNCH(BPL_, E_)               \
  do {                                       \
    if (shard) NB_BH_LAUNCH_S(BPL_, E_, true); \
    else NB_BH_LAUNCH_S(BPL_, E_, false);    \
  } while (0)
  if (b->group == 64) {
    if (eps0) NB_BH_LAUNCH(2, true);
    else NB_BH_LAUNCH(2, false);
  } else {
    if (eps0) NB_BH_LAUNCH(1, true);
    else NB_BH_LAUNCH(1, false);
  }
#undef NB_BH_LAUNCH_S
#undef NB_BH_LAUNCH
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

int bh_shard_wait(nb_context* ctx, const unsigned* flags, int flag_base, unsigned step, int nranks, int self,
                  int* timeout_flag, cudaStream_t st) {
  return launch_peer_wait(ctx, flags, flag_base, step, nranks, self, timeout_flag, st);
}

int bh_shard_finish(nb_context* ctx, const float4* acc_sorted, const nb_particles* d, float dt, bool kick,
                    cudaStream_t st) {
  BH* b = bh_get(ctx);
  if (!b) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  const size_t n = d->count;
  if (n == 0) return NB_OK;
  if (!b->built || b->n != n) {
    set_error("Barnes-Hut: sharded finish without a tree built for these particles");
    return NB_ERR_INVALID;
  }
  const unsigned gn = (unsigned)((n + 255) / 256);
  if (kick) bh_shard_finish_kernel<true><<<gn, 256, 0, st>>>(acc_sorted, b->perm.as<int32_t>(), *d, n, dt);
  else bh_shard_finish_kernel<false><<<gn, 256, 0, st>>>(acc_sorted, b->perm.as<int32_t>(), *d, n, dt);
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_bh_build(nb_context* ctx, const nb_particles* d, nb_stream s) {
  if (!ctx || !d || (d->count && (!d->pos_x || !d->pos_y || !d->pos_z || !d->mass))) {
    set_error("Barnes-Hut build: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  return bh_build(ctx, d, (cudaStream_t)s);
}

int nb_bh_forces(nb_context* ctx, const nb_particles* d, float theta, float G, float eps, nb_stream s) {
  if (!ctx || !d || (d->count && (!d->acc_x || !d->acc_y || !d->acc_z))) {
    set_error("Barnes-Hut forces: null argument");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  return bh_forces(ctx, d, theta, G, eps, (cudaStream_t)s, nullptr);
}

int nb_bh_export(nb_context* ctx, uint32_t* keys, int32_t* perm, int32_t* left, int32_t* right, int32_t* parent,
                 int32_t* prefix, float* node_com, float* bbox) {
  if (!ctx || !ctx->bh || !ctx->bh->built) {
    set_error("Barnes-Hut export: no tree built");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  NB_CUDA(cudaDeviceSynchronize());
  BH* b = ctx->bh;
  const size_t n = b->n;
  if (keys) NB_CUDA(cudaMemcpy(keys, b->keys.p, n * 4, cudaMemcpyDeviceToHost));
  if (perm) NB_CUDA(cudaMemcpy(perm, b->perm.p, n * 4, cudaMemcpyDeviceToHost));
  if (parent) NB_CUDA(cudaMemcpy(parent, b->parent.p, (2 * n - 1) * 4, cudaMemcpyDeviceToHost));
  if (node_com) NB_CUDA(cudaMemcpy(node_com, b->node_com.p, (2 * n - 1) * sizeof(float4), cudaMemcpyDeviceToHost));
  if ((left || right || prefix) && n > 1) {
    int4* h = (int4*)malloc((n - 1) * sizeof(int4));
    if (!h) {
      set_error("out of host memory");
      return NB_ERR_RESOURCE;
    }
    cudaError_t e = cudaMemcpy(h, b->meta.p, (n - 1) * sizeof(int4), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
      free(h);
      return cuda_fail(e, "cudaMemcpy", __FILE__, __LINE__);
    }
    for (size_t i = 0; i + 1 < n; i++) {
      if (left) left[i] = h[i].x;
      if (right) right[i] = h[i].y;
      if (prefix) prefix[i] = h[i].z;
    }
    free(h);
  }
  if (bbox) NB_CUDA(cudaMemcpy(bbox, b->misc.as<float>() + 32, 6 * 4, cudaMemcpyDeviceToHost));
  return NB_OK;
}

int nb_bh_stats(nb_context* ctx, uint64_t* interactions, int* node_count) {
  if (!ctx || !ctx->bh || !ctx->bh->built) {
    set_error("Barnes-Hut stats: no tree built");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  NB_CUDA(cudaDeviceSynchronize());
  if (int rc = bh_check_overflow(ctx->bh)) return rc;
  if (interactions) NB_CUDA(cudaMemcpy(interactions, ctx->bh->misc.as<float>() + 40, 8, cudaMemcpyDeviceToHost));
  if (node_count) *node_count = (int)(2 * ctx->bh->n - 1);
  return NB_OK;
}

int nb_bh_max_stack(nb_context* ctx, int* deepest, int* capacity) {
  if (!ctx || !ctx->bh || !ctx->bh->built) {
    set_error("Barnes-Hut stats: no tree built");
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  NB_CUDA(cudaDeviceSynchronize());
  if (deepest) NB_CUDA(cudaMemcpy(deepest, ctx->bh->misc.as<int>() + 46, 4, cudaMemcpyDeviceToHost));
  if (capacity) *capacity = ctx->bh->stack_cap;
  return NB_OK;
}

}  // extern "C"
