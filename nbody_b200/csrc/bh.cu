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
int launch_scan_tiles(nb_context* ctx, unsigned long long* tile_sum, int ntiles, unsigned long long* total,
                      cudaStream_t st);  // render.cu
int sort_pairs_db(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, uint32_t* keys_alt, int32_t* vals_alt, size_t n,
                  int end_bit, bool identity_vals, uint32_t** out_keys, int32_t** out_vals, cudaStream_t st,
                  const uint32_t* copy_keys = nullptr, const int32_t* copy_vals = nullptr);

}  // namespace nb

struct BH {
  size_t n = 0;
  // keys/perm: two buffers each, the radix passes ping-pong between them (the result stays in BH-owned memory, so
  // a later sort on the same context cannot clobber the permutation the traversal scatters through)
  nb::DevBuf keys, perm, keys_alt, perm_alt, packed, node_com, meta, misc;  // misc: size2 table[32] + counter + bbox floats
  nb::DevBuf roots;   // int2 {first, last} of the subtrees the tiles hand to the top pass
  nb::DevBuf supers;  // int4 {node, first, last, -} per supergroup
  nb::DevBuf glen;    // group length at the group's first leaf, compacted into `groups` in leaf order
  nb::DevBuf gtiles;  // per-block counts / offsets of that compaction
  int hybrid = 1;     // NB_BH_HYBRID=0: equal runs only (A/B)
  int pipeline = 1;   // NB_BH_PIPELINE=0: fetch the next round's nodes after the list evaluation (A/B)
  nb::DevBuf slot;    // int per split position: meeting point of the two children of a node that spans tiles
  nb::DevBuf parent;  // filled on export only
  nb::DevBuf inv;     // inverse sort permutation (sharded finish only)
  nb::DevBuf groups;  // int2 {first,count} per traversal group
  nb::DevBuf pairs;   // 3 x 16 B per internal node: both children's {centre of mass, mass}, {ids, size^2}; + 1 for the root
  int walk_pairs = 1; // NB_BH_WALK=nodes: round-1/2 walk, one node per stack entry (A/B)
  uint32_t* keys_sorted = nullptr;  // where the last build's sorted keys / permutation live
  int32_t* perm_sorted = nullptr;
  float h_bbox[6] = {0, 0, 0, 0, 0, 0};
  int* h_flag = nullptr;  // host-mapped: traversal stack overflow
  int* d_flag = nullptr;
  int group = 64;         // max bodies per traversal group / warp (32 or 64; NB_BH_GROUP overrides)
  int super = 512;        // max leaves of the subtree a group is cut from (NB_BH_SUPER overrides; >= group)
  int occ = 0;
  int stack_cap = 1280;   // entries per warp traversal stack (NB_BH_STACK overrides): deepest seen 440 (16 Mi Plummer);
                          // 2048 -> 5 CTAs per SM and 32.2 ms for that traversal, 1024 -> 8 CTAs and 31.65 ms
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
// [46] int deepest traversal stack of the last force evaluation, [47] int subtrees left to bh_tree_top_kernel,
// [48] int supergroups recorded, [50..51] u64 number of groups (the compaction's total)
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
  if (t == 9) reinterpret_cast<int*>(misc)[47] = 0;  // subtrees handed to the top pass of the tree build
  if (t == 10) reinterpret_cast<int*>(misc)[48] = 0;  // supergroups recorded by the tree passes
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
                                                        float4* __restrict__ packed) {
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
  // the identity permutation is not written: the first radix pass generates value = index itself
}

// ---- tree build: ONE kernel for leaf gather + radix-tree topology + centres of mass + traversal groups -----
// Round 1 built the tree in four kernels -- leaf gather, Karras' per-node binary searches, a bottom-up
// centre-of-mass pass (one thread per leaf climbing, second arriver combines) and a group pass -- 1.84 ms of
// the 2.6 ms build at 16 Mi bodies, all of them bound by instruction issue at low lane occupancy (the climbing
// passes lose half their lanes per level, ncu: 22 % / 22 % / 17 % of the HBM bandwidth).  This kernel builds the
// same tree bottom-up instead (Apetrei 2014): a complete subtree [l, r] is a left child iff
// delta(r) > delta(l-1) (delta(k) = common prefix of sorted keys k, k+1, with Karras' index tie-break for equal
// keys); two adjacent complete subtrees [l, g], [g+1, r] are siblings iff delta(g) exceeds both outer deltas.
// Node numbering stays Karras' (a left child is node r, a right child node l, the root node 0; children of a
// node split at g are g and g+1, or leaves), so left/right/prefix and the centres of mass are bit-identical to
// the per-node searches and to oracle/lbvh.c.  Schedule, per CTA tile of 4096 consecutive leaves in shared
// memory:
//   phase 0  leaves: one random 16-B read per body through the sort permutation (the round-1 gather kernel),
//            stored coalesced as leaf records and kept in shared memory; delta of every adjacent pair;
//   phase 1  every thread merges inside its OWN 16 consecutive leaves with the sequential stack algorithm (the
//            stack is implicit: a current subtree [a, b] keeps its value at slot a and its ends at slots a
//            and b), one action -- merge the top two subtrees or push the next leaf -- per trip, so all lanes
//            work until their own leaves run out: ~5 warp-instructions per leaf where the climbing passes
//            spent ~37;
//   phase 2  log2(256) rounds: thread j joins blocks 2j and 2j+1 of the previous round across their common
//            boundary (only subtrees touching the boundary can newly merge), dense thread indices;
// and then bh_tree_top_kernel: the few subtrees whose sibling lies in another tile (the two spines of a tile,
// ~0.6 % of the nodes) climb with one global exchange per level, one thread per subtree.
// The thread that completes a node writes its record, and emits the traversal groups of a child that is a
// maximal subtree of <= smax leaves (the round-1 group pass).
#ifndef NB_TB_THREADS
#define NB_TB_THREADS 128  // build-time knob: leaves per tile / 16.  Small tiles on purpose: the phases of a tile are
                          // latency chains (random gather, then serial merges along the tile's spines), so the SM
                          // needs many independent tiles in different phases: 256 threads (2 tiles per SM, all in the
                          // same phase) built 16 Mi leaves in 1.79 ms, 128 threads in 1.49 ms
#endif
#ifndef NB_TB_K
#define NB_TB_K 4  // build-time knob: consecutive leaves per thread in phase 1 (a power of two)
#endif
constexpr int TB_THREADS = NB_TB_THREADS;
constexpr int TB_K = NB_TB_K;                   // consecutive leaves per thread in phase 1
constexpr int TB_TILE = TB_THREADS * TB_K;      // leaves per CTA
constexpr int TB_PADDED = TB_TILE + TB_TILE / TB_K + 2;
constexpr int TB_MAX_ROOTS = 128;               // subtrees left per tile: two spines of <= 62 levels each
// thread j's leaf i sits at (TB_K + 1) j + i: an odd stride, no shared-memory bank conflicts in phase 1
__device__ __forceinline__ int tb_pad(int q) { return q + q / TB_K; }

#ifndef NB_BH_SCALAR
#define NB_BH_SCALAR 0  // build-time knob (experiment): scalar FFMA pair body for two bodies per lane instead of packed fp32x2
#endif
#ifndef NB_BH_COM_RCP
#define NB_BH_COM_RCP 0  // build-time knob (experiment, changes the last bit of the centres of mass: not the oracle's formula)
#endif
__device__ __forceinline__ float4 bh_combine(const float4& a, const float4& b) {
  const float ms = __fadd_rn(a.w, b.w);
  float4 c;
  if (ms > 0.f) {
#if NB_BH_COM_RCP
    const float inv = __frcp_rn(ms);  // experiment: one IEEE reciprocal and three multiplies instead of three divides
    c.x = __fmul_rn(__fadd_rn(__fmul_rn(a.x, a.w), __fmul_rn(b.x, b.w)), inv);
    c.y = __fmul_rn(__fadd_rn(__fmul_rn(a.y, a.w), __fmul_rn(b.y, b.w)), inv);
    c.z = __fmul_rn(__fadd_rn(__fmul_rn(a.z, a.w), __fmul_rn(b.z, b.w)), inv);
#else
    c.x = __fdiv_rn(__fadd_rn(__fmul_rn(a.x, a.w), __fmul_rn(b.x, b.w)), ms);
    c.y = __fdiv_rn(__fadd_rn(__fmul_rn(a.y, a.w), __fmul_rn(b.y, b.w)), ms);
    c.z = __fdiv_rn(__fadd_rn(__fmul_rn(a.z, a.w), __fmul_rn(b.z, b.w)), ms);
#endif
  } else {
    c.x = a.x; c.y = a.y; c.z = a.z;
  }
  c.w = ms;
  return c;
}

// common prefix of sorted keys k and k+1 (Karras 2012 sec. 4: equal keys fall back to the positions); -1 outside
__device__ __forceinline__ int bh_delta_adj(const uint32_t* __restrict__ keys, int n, int k) {
  if (k < 0 || k >= n - 1) return -1;
  const uint32_t a = keys[k], b = keys[k + 1];
  if (a == b) return 32 + __clz((unsigned)k ^ (unsigned)(k + 1));
  return __clz(a ^ b);
}

#ifndef NB_BH_SEPARATE_GATHER
#define NB_BH_SEPARATE_GATHER 0  // build-time knob: 1 = the leaves are gathered by a kernel of their own before the tile pass
#endif
// leaves of the tree = bodies in sorted order, packed {x,y,z,m}: one random 16-B read per body
__global__ void __launch_bounds__(256) bh_gather_kernel(const float4* __restrict__ packed, size_t n,
                                                        const int32_t* __restrict__ perm,
                                                        float4* __restrict__ leaves) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  leaves[k] = __ldg(packed + perm[k]);
}

struct TreeOut {
  float4* node_com;  // [2n-1]: internal nodes 0..n-2, leaf k of the sorted order at n-1+k
  int4* meta;        // [n-1] {left, right, prefix, 0}
  int2* groups;
  int* ngroups;
  unsigned char* glen;  // [n rounded up to GC_TILE] group length at the group's first leaf, else 0
  int n, leaf0, smax, gmax;
  int hybrid;  // 1: descend the tree inside a supergroup while that costs no extra group (default); 0: equal runs only
  int2* roots;  // subtrees whose sibling lies in another tile: {first, last} leaf, joined by bh_tree_top_kernel
  int* nroots;
  int4* supers;  // supergroups {node, first leaf, last leaf, -}: cut into traversal groups by bh_groups_kernel
  int* nsupers;
  int* err;      // host-mapped: a capacity of the build or the traversal was exceeded (reported, never dropped)
};

// Traversal groups of a supergroup (a maximal subtree of <= smax leaves, node `id` with leaves [cl, cr]).
// Wanted: as few groups as possible (a warp costs the same whether its 64 lane-slots are full or not) that are as
// compact as possible (a group opens every node its bounding box is close to).  ceil(c / gmax) groups is the
// minimum; cutting the c bodies into that many runs of (almost) equal length, run i starting at
// floor(i * c / runs), achieves it but ignores the tree, following the tree all the way down to subtrees of
// <= gmax leaves is compact but needs ~45 % more groups (mean fill 69 %).  So: descend the tree as long as that
// costs no extra group -- ceil(a / gmax) + ceil(b / gmax) == ceil(c / gmax) for the children's leaf counts a, b --
// and cut equal runs where it would (3296 instead of 3429 list entries x lane-slots per body on a 262 144-body
// Plummer sphere, same number of groups, same error: oracle/lbvh.c counts it on the CPU).  The subtree's nodes are
// complete when this runs; the ones written by other thread blocks are read through L2.
// A group is recorded as its length at its first leaf (glen[], zero elsewhere); bh_group_fill_kernel compacts
// the marks into the group list in LEAF ORDER.  Appending with an atomic counter instead left the list in whatever
// order the supergroups were finished, and the traversal -- whose warps take groups in list order, so that
// neighbouring warps walk neighbouring parts of the tree and share its nodes in L2 -- ran 10 % slower.
__device__ __forceinline__ void tb_emit_runs(int cl, int c, int runs, const TreeOut& o) {
  for (int i = 0; i < runs; i++) {
    const int s0 = (int)(((long long)i * c) / runs), s1 = (int)(((long long)(i + 1) * c) / runs);
    o.glen[cl + s0] = (unsigned char)(s1 - s0);
  }
}
// The tree passes only RECORD the supergroups (node, first and last leaf); bh_groups_kernel cuts them into groups
// afterwards, one thread per supergroup: the descent reads nodes through L2 (~1 us per level) and used to run
// inside the tile pass, where it stalled a whole tile each time.
__device__ __forceinline__ void tb_emit(int id, int cl, int cr, const TreeOut& o) {
  o.supers[atomicAdd(o.nsupers, 1)] = make_int4(id, cl, cr, 0);
}
__global__ void __launch_bounds__(128) bh_groups_kernel(TreeOut o) {
  const int nsup = *o.nsupers;
  constexpr int CAP = 32;
  for (int sidx = blockIdx.x * blockDim.x + threadIdx.x; sidx < nsup; sidx += gridDim.x * blockDim.x) {
    const int4 sg = o.supers[sidx];
    int sid[CAP], sl[CAP], sr[CAP];
    int sp = 1;
    sid[0] = sg.x; sl[0] = sg.y; sr[0] = sg.z;
    while (sp > 0) {
      sp--;
      const int ni = sid[sp], l = sl[sp], r = sr[sp];
      const int c = r - l + 1;
      const int need = (c + o.gmax - 1) / o.gmax;
      if (need == 1) {
        tb_emit_runs(l, c, 1, o);
        continue;
      }
      const int4 m = __ldcg(o.meta + ni);
      const int g = (m.x >= o.leaf0) ? m.x - o.leaf0 : m.x;  // the left child ends at the split
      const int a = g - l + 1, b = r - g;
      if (o.hybrid && (a + o.gmax - 1) / o.gmax + (b + o.gmax - 1) / o.gmax == need && sp + 2 <= CAP) {
        sid[sp] = m.y; sl[sp] = g + 1; sr[sp] = r;
        sp++;
        sid[sp] = m.x; sl[sp] = l; sr[sp] = g;
        sp++;
      } else {
        tb_emit_runs(l, c, need, o);
      }
    }
  }
}

constexpr int GC_TILE = 4096;  // leaves per block of the group compaction: 256 threads x 16 marks
__device__ __forceinline__ unsigned gc_count16(const uint4& v) {
  return (__popc(__vcmpne4(v.x, 0u)) + __popc(__vcmpne4(v.y, 0u)) + __popc(__vcmpne4(v.z, 0u)) +
          __popc(__vcmpne4(v.w, 0u))) >> 3;
}
__global__ void __launch_bounds__(256) bh_group_count_kernel(const unsigned char* __restrict__ glen,
                                                             unsigned long long* __restrict__ tile_sum) {
  __shared__ unsigned ws[8];
  const uint4 v = reinterpret_cast<const uint4*>(glen)[(size_t)blockIdx.x * 256 + threadIdx.x];
  unsigned c = gc_count16(v);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned t = 0;
    for (int w = 0; w < 8; w++) t += ws[w];
    tile_sum[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(256) bh_group_fill_kernel(const unsigned char* __restrict__ glen,
                                                            const unsigned long long* __restrict__ tile_off,
                                                            const unsigned long long* __restrict__ total,
                                                            int2* __restrict__ groups, int* __restrict__ ngroups) {
  __shared__ unsigned ws[8];
  const size_t base = ((size_t)blockIdx.x * 256 + threadIdx.x) * 16;
  const uint4 v = reinterpret_cast<const uint4*>(glen)[(size_t)blockIdx.x * 256 + threadIdx.x];
  const unsigned mine = gc_count16(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  unsigned x = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[w] = x;
  __syncthreads();
  unsigned before = x - mine;
  for (int k = 0; k < w; k++) before += ws[k];
  size_t at = (size_t)tile_off[blockIdx.x] + before;
  const unsigned words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int q = 0; q < 16; q++) {
    const unsigned len = (words[q >> 2] >> ((q & 3) * 8)) & 0xffu;
    if (len) groups[at++] = make_int2((int)(base + q), (int)len);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *ngroups = (int)*total;
}

// Completes node [l, r] split at g (all global leaf indices) from its children's values; dL = delta(l-1),
// dR = delta(r), dS = delta(g).  Writes the node's record and emits the groups of supergroup children.
__device__ __forceinline__ float4 tb_form(int l, int g, int r, int dL, int dR, int dS, const float4& a, const float4& b,
                                          const TreeOut& o) {
  const float4 c = bh_combine(a, b);
  const int id = (dR > dL) ? r : l;  // a left child is numbered by its last leaf, a right child (and the root) by its first
  const int lc = (l == g) ? o.leaf0 + g : g;
  const int rc = (r == g + 1) ? o.leaf0 + g + 1 : g + 1;
  int p = dS - 2;  // key bits start at bit 29
  if (p > 30) p = 30;
  o.meta[id] = make_int4(lc, rc, p, 0);
  o.node_com[id] = c;
  if (r - l + 1 > o.smax) {
    if (g - l + 1 <= o.smax) tb_emit(lc, l, g, o);
    if (r - g <= o.smax) tb_emit(rc, g + 1, r, o);
  } else if (l == 0 && r == o.n - 1) {
    tb_emit(id, l, r, o);  // the whole tree is one supergroup
  }
  return c;
}

#ifndef NB_TB_MIN_BLOCKS
#define NB_TB_MIN_BLOCKS (1024 / NB_TB_THREADS)  // build-time knob: resident tiles per SM asked of ptxas (register cap)
#endif
__global__ void __launch_bounds__(TB_THREADS, NB_TB_MIN_BLOCKS) bh_tree_kernel(const uint32_t* __restrict__ keys,
                                                                 const int32_t* __restrict__ perm,
                                                                 const float4* __restrict__ packed, TreeOut o) {
  extern __shared__ __align__(16) unsigned char tb_smem[];
  float4* s_com = reinterpret_cast<float4*>(tb_smem);            // value of the current subtree starting here
  int* s_d = reinterpret_cast<int*>(s_com + TB_PADDED);          // s_d[q]: delta between local leaves q-1 and q
  short* s_lo = reinterpret_cast<short*>(s_d + TB_PADDED);       // s_lo[b]: first leaf of the current subtree ending at b
  short* s_hi = s_lo + TB_PADDED;                                // s_hi[a]: last leaf of the current subtree starting at a
  __shared__ int s_roots[TB_MAX_ROOTS];
  __shared__ int s_nroots, s_base;
  const int tid = threadIdx.x;
  const int n = o.n;
  const int t0 = blockIdx.x * TB_TILE;
  const int tcount = min(TB_TILE, n - t0);

  // ---- phase 0: leaves and deltas ---------------------------------------------------------------------
  {
    int idx[TB_K];
#pragma unroll
    for (int i = 0; i < TB_K; i++) {
      const int q = tid + i * TB_THREADS;
      idx[i] = (q < tcount && !NB_BH_SEPARATE_GATHER) ? perm[t0 + q] : -1;
    }
#pragma unroll
    for (int i = 0; i < TB_K; i++) {
      const int q = tid + i * TB_THREADS;
      if (q < tcount) {
#if NB_BH_SEPARATE_GATHER
        const float4 v = o.node_com[o.leaf0 + t0 + q];
        s_com[tb_pad(q)] = v;
#else
        const float4 v = __ldg(packed + idx[i]);
        s_com[tb_pad(q)] = v;
        o.node_com[o.leaf0 + t0 + q] = v;
#endif
        s_lo[tb_pad(q)] = (short)q;
        s_hi[tb_pad(q)] = (short)q;
        s_d[tb_pad(q + 1)] = bh_delta_adj(keys, n, t0 + q);
      }
    }
    if (tid == 0) s_d[0] = bh_delta_adj(keys, n, t0 - 1);
  }
  __syncthreads();
  if (n == 1) {  // a single leaf: no internal node, one group
    if (tid == 0) tb_emit(0, 0, 0, o);
    return;
  }

  // ---- phase 1: every thread merges inside its own TB_K consecutive leaves ---------------------------------
  {
    const int blk0 = tid * TB_K, blk1 = min(blk0 + TB_K, tcount);
    int q = blk0;  // the top subtree ends at leaf q
    int a_top, g, a, dS, dL, dR;
    // can the top subtree [a_top, q] and the one below it [a, g] be joined?
    auto siblings = [&]() -> bool {
      a_top = s_lo[tb_pad(q)];
      if (a_top <= blk0) return false;
      g = a_top - 1;
      a = s_lo[tb_pad(g)];
      dS = s_d[tb_pad(g + 1)];
      dL = s_d[tb_pad(a)];
      dR = s_d[tb_pad(q + 1)];
      return dS > dL && dS > dR;
    };
    // One trip = at most one push (cheap) and one merge (expensive, ONE site): a warp's lanes alternate between the
    // two at different times, and a trip that did only one of them would cost every lane the other as well.
    while (q < blk1) {
      bool can = siblings();
      if (!can) {  // push the next leaf (its one-leaf subtree is already in place) and look again
        if (++q >= blk1) break;
        can = siblings();
      }
      if (can) {
        const float4 c = tb_form(t0 + a, t0 + g, t0 + q, dL, dR, dS, s_com[tb_pad(a)], s_com[tb_pad(a_top)], o);
        s_com[tb_pad(a)] = c;
        s_hi[tb_pad(a)] = (short)q;
        s_lo[tb_pad(q)] = (short)a;
      }
    }
  }
  __syncthreads();

  // ---- phase 2: join neighbouring blocks across their common boundary, doubling the block size ---------------
  for (int half = TB_K; half < TB_TILE; half <<= 1) {
    const int p = (2 * tid + 1) * half;  // boundary between local leaves p-1 and p
    if (p < tcount) {
      const int lo_lim = p - half, hi_lim = min(p + half, tcount);
      // start from the subtree that ends at the boundary; every trip joins it with its left or its right neighbour
      // if the two are siblings (ONE merge site for both directions: the lanes of a warp stay together)
      int a = s_lo[tb_pad(p - 1)], b = p - 1;
      int dL = s_d[tb_pad(a)], dR = s_d[tb_pad(p)];
      float4 c = s_com[tb_pad(a)];
      bool changed = false;
      for (;;) {
        bool left = false, right = false;
        int o2 = 0, d2 = 0;
        if (a > lo_lim) {
          o2 = s_lo[tb_pad(a - 1)];
          d2 = s_d[tb_pad(o2)];
          left = dL > d2 && dL > dR;
        }
        if (!left && b + 1 < hi_lim) {
          o2 = s_hi[tb_pad(b + 1)];
          d2 = s_d[tb_pad(o2 + 1)];
          right = dR > dL && dR > d2;
        }
        if (!left && !right) break;
        const float4 other = s_com[tb_pad(left ? o2 : b + 1)];
        const int nl = left ? o2 : a, ng = left ? a - 1 : b, nr = left ? b : o2;
        const int ndL = left ? d2 : dL, ndR = left ? dR : d2, dS = left ? dL : dR;
        c = tb_form(t0 + nl, t0 + ng, t0 + nr, ndL, ndR, dS, left ? other : c, left ? c : other, o);
        a = nl; b = nr; dL = ndL; dR = ndR;
        changed = true;
      }
      if (changed) {
        s_com[tb_pad(a)] = c;
        s_hi[tb_pad(a)] = (short)b;
        s_lo[tb_pad(b)] = (short)a;
      }
    }
    __syncthreads();
  }
  if (gridDim.x == 1) return;  // the tile holds the whole tree: its root is complete

  // ---- the subtrees left over (the tile's two spines: their siblings lie in other tiles) go to the top pass -----
  if (tid == 0) {
    int nr = 0;
    int q = 0;
    for (; q < tcount && nr < TB_MAX_ROOTS; q = s_hi[tb_pad(q)] + 1) s_roots[nr++] = q;
    if (q < tcount) *o.err = 2;  // more left-over subtrees than two spines of 62 levels can have: cannot happen
    s_nroots = nr;
    s_base = atomicAdd(o.nroots, nr);
  }
  __syncthreads();
  for (int i = tid; i < s_nroots; i += TB_THREADS) {
    const int a = s_roots[i];
    o.roots[s_base + i] = make_int2(t0 + a, t0 + s_hi[tb_pad(a)]);
  }
}

// Top of the tree: every subtree a tile could not finish climbs through global memory, one thread each.  A
// subtree [l, r] is a left child iff delta(r) > delta(l-1); the two children of a node meet at its split g:
// the first to arrive leaves its far end in slot[g] and stops, the second takes it, reads the sibling's record
// and completes the node.  This used to be the tail of the tile kernel, where each of the ~20 dependent global
// round trips of a spine held a whole 100 KB tile resident (1.9 ms for 16 Mi leaves, 14 of every 15 issue
// slots stalled); as a kernel of its own it is ~100 000 independent threads deep.
__global__ void __launch_bounds__(256) bh_tree_top_kernel(const uint32_t* __restrict__ keys, int* __restrict__ slot,
                                                          TreeOut o) {
  const int n = o.n;
  const int nroots = *o.nroots;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nroots; i += gridDim.x * blockDim.x) {
  const int2 rt = o.roots[i];
  int l = rt.x, r = rt.y;
  int dL = bh_delta_adj(keys, n, l - 1), dR = bh_delta_adj(keys, n, r);
  // the subtree's own record: a leaf, or node r (left child) / node l (right child)
  float4 cur = __ldcg(o.node_com + ((l == r) ? o.leaf0 + l : ((dR > dL) ? r : l)));
  while (!(l == 0 && r == n - 1)) {
    const bool is_left = dR > dL;
    const int g = is_left ? r : l - 1;  // split of the parent: where the two children meet
    __threadfence();                    // this subtree's record is visible before its arrival
    const int old = atomicExch(&slot[g], (is_left ? l : r) + 1);
    if (old == 0) break;  // first to arrive: the sibling's thread completes the parent
    slot[g] = 0;          // both children have arrived: the slot is clean for the next build
    __threadfence();
    const int far = old - 1;
    const int dS = is_left ? dR : dL;
    if (is_left) {
      const float4 sib = __ldcg(o.node_com + ((far == g + 1) ? o.leaf0 + g + 1 : g + 1));
      r = far;
      dR = bh_delta_adj(keys, n, r);
      cur = tb_form(l, g, r, dL, dR, dS, cur, sib, o);
    } else {
      const float4 sib = __ldcg(o.node_com + ((far == g) ? o.leaf0 + g : g));
      l = far;
      dL = bh_delta_adj(keys, n, l - 1);
      cur = tb_form(l, g, r, dL, dR, dS, sib, cur, o);
    }
  }
  }  // next subtree
}

// parent[] is needed by nb_bh_export only (parity tests): derived from the children on demand
__global__ void __launch_bounds__(256) bh_parent_kernel(int n, const int4* __restrict__ meta, int32_t* __restrict__ parent) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) parent[0] = -1;
  if (i >= n - 1) return;
  const int4 m = meta[i];
  parent[m.x] = i;
  parent[m.y] = i;
}

// ---- child-pair records for the traversal ---------------------------------------------------------------
// The walk (bh_force_pairs_kernel) keeps only nodes that have to be OPENED on its stack and tests both children of an
// opened node in one go; for that it wants both children's centres of mass, ids and cell sizes in one 48-byte record
// per internal node, instead of one 32-byte fetch per child after a trip through the stack:
//   rec[3 i + 0] = centre of mass + mass of the left child, rec[3 i + 1] = of the right child,
//   rec[3 i + 2] = {left id, right id, size^2 of the left child's cell, of the right child's} (a leaf's is 0).
// Record n - 1 (one past the internal nodes) is a virtual parent of the root -- {root, a massless dummy} -- where
// every walk starts, so the root needs no special case (n == 1: the root is the only leaf).
__global__ void __launch_bounds__(256) bh_pairs_kernel(int n, const float4* __restrict__ node_com,
                                                       const int4* __restrict__ meta, const float* __restrict__ misc,
                                                       float4* __restrict__ rec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int leaf0 = n - 1;
  if (i > leaf0) return;
  float4 a, b = make_float4(0.f, 0.f, 0.f, 0.f);
  int4 m;
  if (i == leaf0) {  // virtual parent of the root
    a = __ldg(node_com);
    m = make_int4(0, 0, __float_as_int(n > 1 ? misc[__ldg(&meta[0].z)] : 0.f), 0);
  } else {
    const int4 me = __ldg(meta + i);
    a = __ldg(node_com + me.x);
    b = __ldg(node_com + me.y);
    const float sa = me.x >= leaf0 ? 0.f : misc[__ldg(&meta[me.x].z)];
    const float sb = me.y >= leaf0 ? 0.f : misc[__ldg(&meta[me.y].z)];
    m = make_int4(me.x, me.y, __float_as_int(sa), __float_as_int(sb));
  }
  rec[3 * (size_t)i] = a;
  rec[3 * (size_t)i + 1] = b;
  rec[3 * (size_t)i + 2] = make_float4(__int_as_float(m.x), __int_as_float(m.y), __int_as_float(m.z), __int_as_float(m.w));
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
#ifndef NB_BH_MIN_BLOCKS
#define NB_BH_MIN_BLOCKS 6  // build-time knob: resident CTAs per SM the pair-record traversal is compiled for (register cap)
#endif
constexpr int BH_WARPS = 4;
constexpr int BH_LIST = 64;
constexpr int BH_STAGE = 32 * 32;  // bytes per warp: one {centre of mass, node record} per lane

// What a warp holds for its traversal group, shared by both walks: the lane's BPL bodies (as one packed fp32x2 pair when
// BPL == 2), their running sums, and the group's bounding box.
template <int BPL, bool EPS0>
struct BhGroup {
  struct { float x, y, z; } me[BPL];
  float sx[BPL], sy[BPL], sz[BPL];
  float lo[3], hi[3];
  f2 xi2, yi2, zi2, sx2, sy2, sz2, e2;
  float eps2;

  // scalar loads on purpose: a float4 load pins x,y,z to one aligned register quad, and ptxas then re-packs the
  // (body 0, body 1) pairs with MOVs for every list entry (4.5 MOVs per entry measured)
  __device__ __forceinline__ void load(const float4* __restrict__ leaves, int k0, int gcnt, int lane, float eps2_) {
    eps2 = eps2_;
#pragma unroll
    for (int a = 0; a < 3; a++) {
      lo[a] = 3.0e38f;
      hi[a] = -3.0e38f;
    }
#pragma unroll
    for (int b = 0; b < BPL; b++) {
      me[b].x = me[b].y = me[b].z = 0.f;
      if (b * 32 + lane < gcnt) {
        const float* q = reinterpret_cast<const float*>(leaves + k0 + b * 32 + lane);
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
    e2 = pk(eps2, eps2);
    xi2 = pk(me[0].x, me[BPL - 1].x);
    yi2 = pk(me[0].y, me[BPL - 1].y);
    zi2 = pk(me[0].z, me[BPL - 1].z);
    sx2 = sy2 = sz2 = pk(0.f, 0.f);
  }

  // distance^2 from a centre of mass to the group's box, softened: the denominator of the acceptance test.  dmin <=
  // every body's own distance, so a node the group accepts is a node every body would accept with the reference's
  // per-body test (:160-172).
  __device__ __forceinline__ float box_d2(const float4& c) const {
    const float ddx = fmaxf(fmaxf(lo[0] - c.x, c.x - hi[0]), 0.f);
    const float ddy = fmaxf(fmaxf(lo[1] - c.y, c.y - hi[1]), 0.f);
    const float ddz = fmaxf(fmaxf(lo[2] - c.z, c.z - hi[2]), 0.f);
    return fmaf(ddz, ddz, fmaf(ddy, ddy, fmaf(ddx, ddx, eps2)));
  }

  // the monopole term (:175-181) of one list entry {x, y, z, G * M} for the lane's bodies
  __device__ __forceinline__ void add(const float4 p) {
    const float gm = p.w;
    if constexpr (BPL == 2 && !NB_BH_SCALAR) {
      // the lane's two bodies as one packed fp32x2 pair (FADD2/FFMA2/FMUL2, as in direct.cu)
      const f2 dx = sub2(pk(p.x, p.x), xi2), dy = sub2(pk(p.y, p.y), yi2), dz = sub2(pk(p.z, p.z), zi2);
      const f2 d2 = fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, e2)));
      float r0, r1;
      upk(d2, r0, r1);
      float i0 = rsq(r0), i1 = rsq(r1);
      if (EPS0) {  // own leaf / exact coincidence at eps ~ 0
        i0 = (r0 > eps2) ? i0 : 0.f;
        i1 = (r1 > eps2) ? i1 : 0.f;
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
        if (EPS0) inv = (d2 > eps2) ? inv : 0.f;  // own leaf / exact coincidence at eps ~ 0
        const float f = gm * (inv * inv * inv);
        sx[b] = fmaf(f, dx, sx[b]);
        sy[b] = fmaf(f, dy, sy[b]);
        sz[b] = fmaf(f, dz, sz[b]);
      }
    }
  }

  // ne <= 32 list entries, each a warp-broadcast LDS.128
  __device__ __forceinline__ void evaluate(const float4* lcur, int ne) {
    if (ne == 32) {
      NB_BH_PRAGMA_UNROLL(NB_BH_UNROLL)
      for (int e = 0; e < 32; e++) add(lcur[e]);
    } else {
      for (int e = 0; e < ne; e++) add(lcur[e]);
    }
  }

  // accelerations out: body order through the sort permutation, or (SHARDED) sorted order into every rank's
  // exchange buffer
  template <bool SHARDED>
  __device__ __forceinline__ void store(int k0, int gcnt, int lane, const int32_t* __restrict__ perm,
                                        float* __restrict__ ax, float* __restrict__ ay, float* __restrict__ az,
                                        const BhShard& sh) {
    if constexpr (BPL == 2 && !NB_BH_SCALAR) {
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
  }
};

// SHARDED epilogue of a traversal kernel: the last CTA to finish raises this rank's step flag on every peer
__device__ __forceinline__ void bh_shard_publish(const BhShard& sh) {
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
                                                                  int* __restrict__ max_stack, int pipeline,
                                                                  BhShard sh) {
  extern __shared__ __align__(16) unsigned char bh_smem[];
  __shared__ float size2_tab[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x < 32) size2_tab[threadIdx.x] = misc[threadIdx.x];
  __syncthreads();
  const size_t smem_per_warp = (size_t)stack_cap * 4 + BH_LIST * 16 + BH_STAGE;
  float4* list = reinterpret_cast<float4*>(bh_smem + (size_t)w * smem_per_warp);
  float4* stage_c = list + BH_LIST;                              // next round's nodes, one per lane (cp.async)
  int4* stage_m = reinterpret_cast<int4*>(stage_c + 32);
  int* stack = reinterpret_cast<int*>(bh_smem + (size_t)w * smem_per_warp + BH_LIST * 16 + BH_STAGE);
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

  BhGroup<BPL, EPS0> bodies;
  bodies.load(node_com + leaf0, k0, gcnt, lane, eps2);

  int sp = 1, nl = 0, lbase = 0;
  unsigned visits = 0;
  if (lane == 0) stack[0] = 0;
  __syncwarp();
  // The node fetches are software-pipelined: the next round's nodes are popped and their fetches issued BEFORE the
  // current interaction list is evaluated, so the (L2 / DRAM) latency of the walk hides under the pair arithmetic.
  // The fetches are cp.async copies into a per-lane staging slot in shared memory: no destination registers, hence
  // no scoreboard wait until the data is needed.  (With plain loads ptxas put a MOV of the node's mass right behind
  // the load -- the wait for the data BEFORE the evaluation: 11.5 % of the kernel's stall samples on that one MOV.)
  int ni;
  bool have;
  float4 c;
  int4 mt;
  auto pop_and_fetch = [&]() {
    const int m = min(sp, 32);
    have = lane < m;
    ni = have ? stack[sp - m + lane] : 0;  // lanes without a node fetch the root (ignored)
    sp -= m;
    __syncwarp();
    const int mi = max(min(ni, leaf0 - 1), 0);  // leaves have no record: some internal node's is fetched and ignored
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage_c + lane)), "l"(node_com + ni) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage_m + lane)), "l"(meta + mi) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  pop_and_fetch();
  for (;;) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    c = stage_c[lane];
    mt = stage_m[lane];
    const bool leaf = ni >= leaf0;
    const bool valid = have && c.w != 0.f;  // total_mass == 0 -> skipped (:157-158)
    bool acc = leaf;
    if (valid && !leaf) acc = size2_tab[mt.z] < theta2 * bodies.box_d2(c);  // size2 / dist2 < theta2 (:167)
    const bool is_acc = valid && acc, is_rej = valid && !acc;
    const unsigned amask = __ballot_sync(0xffffffffu, is_acc), rmask = __ballot_sync(0xffffffffu, is_rej);
    // the list is a ring of two 32-entry halves: a batch is one half, what is left over already sits in the other
    if (is_acc) list[(lbase + nl + __popc(amask & lt)) & (BH_LIST - 1)] = make_float4(c.x, c.y, c.z, G * c.w);  // G*M once per node
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
    if (!done && pipeline) pop_and_fetch();  // in flight while the list is evaluated
    while (nl >= 32 || (done && nl > 0)) {
      const int ne = min(nl, 32);
      bodies.evaluate(list + lbase, ne);
      visits += ne;
      __syncwarp();  // every lane has read the batch before later entries wrap into its half
      if (ne == 32) lbase ^= 32;  // what is left (< 32 entries) already sits at the start of the other half
      nl -= ne;
    }
    if (done) break;
    if (!pipeline) pop_and_fetch();
  }
  bodies.template store<SHARDED>(k0, gcnt, lane, perm, ax, ay, az, sh);
  visits_bodies += (unsigned long long)visits * (unsigned long long)gcnt;
  __syncwarp();
  }  // next group
  if (lane == 0 && visits_bodies) {
    atomicAdd(counter, visits_bodies);
    atomicMax(max_stack, deepest);
  }
  if (SHARDED) bh_shard_publish(sh);
}

// ---- traversal over child-pair records (the default) ----------------------------------------------------------
// Same groups, same acceptance test per node, hence the same interaction SET as bh_force_kernel; what changes is the
// bookkeeping.  bh_force_kernel pushes every child of an opened node and pops it again just to test it: two pops, two
// 32-byte fetches and two trips through the ballot/compaction code per accepted entry.  The kernel is bound by
// instruction issue (FFMA2 holds the issue port for both its cycles and every integer instruction of the walk adds to
// it: profiles/r02_issue_slots_ubench.jsonl), so the walk's instructions are paid in full.  Here the stack holds only
// nodes that must be OPENED; a lane pops one, fetches its 48-byte pair record (bh_pairs_kernel) and tests BOTH
// children: accepted ones go to the interaction list, rejected internal ones onto the stack.  Half the pops and
// rounds, four ballots per round instead of two per child.  The list is a ring of four 32-entry batches (a round adds
// up to 64 entries to fewer than 32 left over).  List order: lane order, a lane's left child before its right one;
// oracle/lbvh.c restates it step for step.
constexpr int BH_PLIST = 128;
constexpr int BH_PSTAGE = 32 * 48;  // bytes per warp: one pair record per lane

template <int BPL, bool EPS0, bool SHARDED>
__global__ void __launch_bounds__(BH_WARPS * 32, NB_BH_MIN_BLOCKS)
    bh_force_pairs_kernel(int n, const float4* __restrict__ node_com, const float4* __restrict__ rec,
                          const int32_t* __restrict__ perm, float theta2, float G, float eps2, float* __restrict__ ax,
                          float* __restrict__ ay, float* __restrict__ az, unsigned long long* __restrict__ counter,
                          int* __restrict__ overflow, const int2* __restrict__ groups, const int* __restrict__ ngroups,
                          int* __restrict__ cursor, int stack_cap, int* __restrict__ max_stack, BhShard sh) {
  extern __shared__ __align__(16) unsigned char bh_smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const size_t smem_per_warp = (size_t)stack_cap * 4 + BH_PLIST * 16 + BH_PSTAGE;
  float4* list = reinterpret_cast<float4*>(bh_smem + (size_t)w * smem_per_warp);
  float4* stage_a = list + BH_PLIST;  // next round's records, one per lane (cp.async): left child, right child, ids
  float4* stage_b = stage_a + 32;
  float4* stage_m = stage_b + 32;
  int* stack = reinterpret_cast<int*>(stage_m + 32);
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

  BhGroup<BPL, EPS0> bodies;
  bodies.load(node_com + leaf0, k0, gcnt, lane, eps2);

  int sp = 1, nl = 0, lbase = 0;
  unsigned visits = 0;
  if (lane == 0) stack[0] = leaf0;  // the virtual parent of the root
  __syncwarp();
  int popped = 0;
  auto pop_and_fetch = [&]() {
    const int m = min(sp, 32);
    popped = m;
    const int ni = (lane < m) ? stack[sp - m + lane] : leaf0;  // lanes without a node fetch the virtual record (ignored)
    sp -= m;
    __syncwarp();
    const float4* r = rec + 3 * (size_t)ni;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage_a + lane)), "l"(r) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage_b + lane)), "l"(r + 1) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage_m + lane)), "l"(r + 2) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  pop_and_fetch();
  for (;;) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    const float4 ca = stage_a[lane], cb = stage_b[lane];
    const float4 mf = stage_m[lane];
    const int ia = __float_as_int(mf.x), ib = __float_as_int(mf.y);
    const bool have = lane < popped;
    // a child is used for the whole group iff it is a leaf or size^2 < theta^2 * d^2 (:167); total mass 0 -> skipped
    // (:157-158).  Both tests run unconditionally: no branches around them.
    const bool va = have && ca.w != 0.f, vb = have && cb.w != 0.f;
    const bool ta = (ia >= leaf0) | (mf.z < theta2 * bodies.box_d2(ca));
    const bool tb = (ib >= leaf0) | (mf.w < theta2 * bodies.box_d2(cb));
    const bool acc_a = va && ta, acc_b = vb && tb, rej_a = va && !ta, rej_b = vb && !tb;
    const unsigned ma = __ballot_sync(0xffffffffu, acc_a), mb = __ballot_sync(0xffffffffu, acc_b);
    const unsigned ra = __ballot_sync(0xffffffffu, rej_a), rb = __ballot_sync(0xffffffffu, rej_b);
    const int lpos = lbase + nl + __popc(ma & lt) + __popc(mb & lt);
    if (acc_a) list[lpos & (BH_PLIST - 1)] = make_float4(ca.x, ca.y, ca.z, G * ca.w);  // G*M once per node
    if (acc_b) list[(lpos + (acc_a ? 1 : 0)) & (BH_PLIST - 1)] = make_float4(cb.x, cb.y, cb.z, G * cb.w);
    const int npush = __popc(ra) + __popc(rb);
    deepest = max(deepest, sp + npush);
    if (sp + npush > stack_cap) {  // never silently drop subtrees (the reference does, :186-194)
      if (lane == 0) *overflow = 1;
      break;
    }
    const int spos = sp + __popc(ra & lt) + __popc(rb & lt);
    if (rej_a) stack[spos] = ia;
    if (rej_b) stack[spos + (rej_a ? 1 : 0)] = ib;
    nl += __popc(ma) + __popc(mb);
    sp += npush;
    __syncwarp();
    const bool done = (sp == 0);
    if (!done) pop_and_fetch();  // in flight while the list is evaluated
    while (nl >= 32 || (done && nl > 0)) {
      const int ne = min(nl, 32);
      bodies.evaluate(list + lbase, ne);
      visits += ne;
      __syncwarp();  // every lane has read the batch before later entries wrap into it
      if (ne == 32) lbase = (lbase + 32) & (BH_PLIST - 1);
      nl -= ne;
    }
    if (done) break;
  }
  bodies.template store<SHARDED>(k0, gcnt, lane, perm, ax, ay, az, sh);
  visits_bodies += (unsigned long long)visits * (unsigned long long)gcnt;
  __syncwarp();
  }  // next group
  if (lane == 0 && visits_bodies) {
    atomicAdd(counter, visits_bodies);
    atomicMax(max_stack, deepest);
  }
  if (SHARDED) bh_shard_publish(sh);
}

// accelerations arrive in sorted order; hand them back to body order, optionally with the kick
// (src/cuda/integrator.cu:24-36, same rounding sequence as integrate.cu).  One thread per BODY: its arrays (acc,
// vel, acc_old: nine of them) are read and written coalesced, and the acceleration is ONE random 16-B read through the
// inverse permutation.  Round 1 ran one thread per sorted slot and scattered through perm: twelve random 4-B accesses
// per body, ~4.5 ms for 16 Mi bodies on every rank of a sharded step.
__global__ void __launch_bounds__(256) bh_inverse_perm_kernel(const int32_t* __restrict__ perm, size_t n,
                                                              int32_t* __restrict__ inv) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) inv[perm[k]] = (int32_t)k;
}
template <bool KICK>
__global__ void __launch_bounds__(256) bh_shard_finish_kernel(const float4* __restrict__ acc_sorted,
                                                              const int32_t* __restrict__ inv, nb_particles d,
                                                              size_t n, float dt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 a = __ldg(acc_sorted + inv[i]);
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

constexpr size_t kTreeSmem = (size_t)TB_PADDED * (sizeof(float4) + sizeof(int) + 2 * sizeof(short));

// per-device kernel attribute (opt-in shared memory), set when a context is created on the device
int bh_init_device(nb_context* ctx) {
  (void)ctx;
  NB_CUDA(cudaFuncSetAttribute(bh_tree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTreeSmem));
  return NB_OK;
}

static BH* bh_get(nb_context* ctx) {
  if (!ctx->bh) ctx->bh = new (std::nothrow) BH();
  return ctx->bh;
}

void bh_destroy(nb_context* ctx) {
  if (!ctx->bh) return;
  BH* b = ctx->bh;
  b->keys.release(); b->perm.release(); b->keys_alt.release(); b->perm_alt.release(); b->packed.release();
  b->node_com.release(); b->meta.release(); b->parent.release(); b->slot.release(); b->misc.release();
  b->roots.release();
  b->inv.release();
  b->supers.release();
  b->glen.release();
  b->gtiles.release();
  b->groups.release();
  b->pairs.release();
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
  if (!b->h_flag) {
    NB_CUDA(cudaHostAlloc(&b->h_flag, sizeof(int), cudaHostAllocMapped));
    *b->h_flag = 0;
    NB_CUDA(cudaHostGetDevicePointer(&b->d_flag, b->h_flag, 0));
  }
  if (int rc = b->keys.ensure(n * 4)) return rc;
  if (int rc = b->perm.ensure(n * 4)) return rc;
  if (int rc = b->keys_alt.ensure(n * 4)) return rc;
  if (int rc = b->perm_alt.ensure(n * 4)) return rc;
  if (int rc = b->packed.ensure(n * sizeof(float4))) return rc;
  if (int rc = b->node_com.ensure((2 * n) * sizeof(float4))) return rc;
  if (int rc = b->meta.ensure(n * sizeof(int4))) return rc;
  if (int rc = b->misc.ensure(64 * 4)) return rc;
  if (int rc = b->groups.ensure(n * sizeof(int2))) return rc;
  if (int rc = b->roots.ensure(((n + TB_TILE - 1) / TB_TILE) * TB_MAX_ROOTS * sizeof(int2))) return rc;
  // supergroups: children (<= smax leaves) of the nodes with more: those nodes number < 2 n / (smax + 1)
  if (b->slot.cap < n * 4) {
    // meeting slots of the tree build: zero when idle (the second child to arrive cleans its slot), so they
    // are cleared only when (re)allocated
    if (int rc = b->slot.ensure(n * 4)) return rc;
    NB_CUDA(cudaMemsetAsync(b->slot.p, 0, b->slot.cap, st));
  }
  {
    const char* e = getenv("NB_BH_GROUP");
    const int g = (e && atoi(e) == 32) ? 32 : 64;
    if (g != b->group) b->occ = 0;
    b->group = g;
    {
      // One walk for single-GPU and sharded steps alike: a sharded step stays bit-identical to the single-GPU one.
      // (On 8 ranks the 0.24 ms record pass per 16 Mi bodies, paid by every rank, slightly exceeds the ~4 % it saves
      // on an eighth of the traversal; NB_BH_WALK=nodes selects the other walk.)
      int wp = 1;
      if (const char* we = getenv("NB_BH_WALK")) wp = strcmp(we, "nodes") != 0;
      if (wp != b->walk_pairs) b->occ = 0;
      b->walk_pairs = wp;
    }
    {
      // 48 KiB of dynamic shared memory without opt-in: 4 warps x (cap x 4 B + list and staging: 2 KiB node walk,
      // 3.5 KiB pair walk)
      const int cmax = b->walk_pairs ? 2176 : 2528;  // the node walk also holds a 128-byte static table
      int c = b->stack_cap;
      if (const char* ce = getenv("NB_BH_STACK")) c = atoi(ce);
      c = c < 512 ? 512 : (c > cmax ? cmax : c);
      if (c != b->stack_cap) b->occ = 0;
      b->stack_cap = c;
    }
    if (const char* he = getenv("NB_BH_HYBRID")) b->hybrid = atoi(he) != 0;
    if (const char* pe = getenv("NB_BH_PIPELINE")) b->pipeline = atoi(pe) != 0;
    const char* se = getenv("NB_BH_SUPER");
    int sg = se ? atoi(se) : 8 * g;
    if (sg < g) sg = g;
    b->super = sg;
  }
  if (b->walk_pairs)
    if (int rc = b->pairs.ensure(n * 3 * sizeof(float4))) return rc;
  if (int rc = b->supers.ensure((4 * n / ((size_t)b->super + 1) + 16) * sizeof(int4))) return rc;
  const size_t gc_tiles = (n + GC_TILE - 1) / GC_TILE;
  if (int rc = b->glen.ensure(gc_tiles * GC_TILE)) return rc;
  if (int rc = b->gtiles.ensure(gc_tiles * sizeof(unsigned long long))) return rc;
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
                                         b->packed.as<float4>());
    NB_LAUNCH_CHECK(ctx);
  }
  {
    NB_STAGE(ctx, st, "bh.sort");
    if (int rc = sort_pairs_db(ctx, b->keys.as<uint32_t>(), b->perm.as<int32_t>(), b->keys_alt.as<uint32_t>(),
                               b->perm_alt.as<int32_t>(), n, 30, true, &b->keys_sorted, &b->perm_sorted, st))
      return rc;
  }
  {
    NB_STAGE(ctx, st, "bh.tree");
    TreeOut o;
    o.node_com = b->node_com.as<float4>();
    o.meta = b->meta.as<int4>();
    o.groups = b->groups.as<int2>();
    o.ngroups = reinterpret_cast<int*>(misc) + 44;
    o.n = (int)n;
    o.leaf0 = (int)n - 1;
    o.smax = b->super;
    o.gmax = b->group;
    o.hybrid = b->hybrid;
    o.roots = b->roots.as<int2>();
    o.nroots = reinterpret_cast<int*>(misc) + 47;
    o.glen = b->glen.as<unsigned char>();
    NB_CUDA(cudaMemsetAsync(o.glen, 0, gc_tiles * GC_TILE, st));
    o.err = b->d_flag;
    o.supers = b->supers.as<int4>();
    o.nsupers = reinterpret_cast<int*>(misc) + 48;
    const unsigned tiles = (unsigned)((n + TB_TILE - 1) / TB_TILE);
#if NB_BH_SEPARATE_GATHER
    bh_gather_kernel<<<gn, 256, 0, st>>>(b->packed.as<float4>(), n, b->perm_sorted, o.node_com + o.leaf0);
    NB_LAUNCH_CHECK(ctx);
#endif
    bh_tree_kernel<<<tiles, TB_THREADS, kTreeSmem, st>>>(b->keys_sorted, b->perm_sorted, b->packed.as<float4>(), o);
    NB_LAUNCH_CHECK(ctx);
    if (tiles > 1) {  // the subtrees the tiles left over (at most 2 x 62 each: one per level of each spine)
      bh_tree_top_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(b->keys_sorted, b->slot.as<int>(), o);
      NB_LAUNCH_CHECK(ctx);
    }
    bh_groups_kernel<<<ctx->sm_count * 8, 128, 0, st>>>(o);
    NB_LAUNCH_CHECK(ctx);
    unsigned long long* gt = b->gtiles.as<unsigned long long>();
    unsigned long long* gtotal = reinterpret_cast<unsigned long long*>(misc + 50);
    bh_group_count_kernel<<<(unsigned)gc_tiles, 256, 0, st>>>(o.glen, gt);
    NB_LAUNCH_CHECK(ctx);
    if (int rc = launch_scan_tiles(ctx, gt, (int)gc_tiles, gtotal, st)) return rc;
    bh_group_fill_kernel<<<(unsigned)gc_tiles, 256, 0, st>>>(o.glen, gt, gtotal, o.groups, o.ngroups);
    NB_LAUNCH_CHECK(ctx);
  }
  if (b->walk_pairs) {
    NB_STAGE(ctx, st, "bh.pairs");
    bh_pairs_kernel<<<gn, 256, 0, st>>>((int)n, b->node_com.as<float4>(), b->meta.as<int4>(), misc,
                                        b->pairs.as<float4>());
    NB_LAUNCH_CHECK(ctx);
  }
  b->built = true;
  return NB_OK;
}

static int bh_check_overflow(BH* b) {
  if (b->h_flag && *b->h_flag) {
    const int what = *b->h_flag;
    *b->h_flag = 0;
    if (what == 2) set_error("Barnes-Hut tree build: a tile left more subtrees than its list holds");
    else set_error("Barnes-Hut traversal stack overflow (tree deeper than the %d-entry warp stack allows; NB_BH_STACK raises it)", b->stack_cap);
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
  const size_t smem = b->walk_pairs
                          ? (size_t)BH_WARPS * ((size_t)b->stack_cap * 4 + BH_PLIST * 16 + BH_PSTAGE)
                          : (size_t)BH_WARPS * ((size_t)b->stack_cap * 4 + BH_LIST * 16 + BH_STAGE);
  const float eps2 = eps * eps;
  const bool eps0 = !(eps2 >= NB_EPS2_GUARD);
  if (!b->occ) {
    int o = 0;
    if (b->walk_pairs && b->group == 64)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, bh_force_pairs_kernel<2, false, false>, BH_WARPS * 32, smem);
    else if (b->walk_pairs)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, bh_force_pairs_kernel<1, false, false>, BH_WARPS * 32, smem);
    else if (b->group == 64)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, bh_force_kernel<2, false, false>, BH_WARPS * 32, smem);
    else
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, bh_force_kernel<1, false, false>, BH_WARPS * 32, smem);
    b->occ = o > 0 ? o : 1;
    if (const char* oe = getenv("NB_BH_OCC")) {  // tuning knob: resident CTAs per SM of the traversal
      const int v = atoi(oe);
      if (v > 0 && v < b->occ) b->occ = v;
    }
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
      (int)n, b->node_com.as<float4>(), b->meta.as<int4>(), b->perm_sorted, misc, theta * theta, G, eps2,         \
      d->acc_x, d->acc_y, d->acc_z, reinterpret_cast<unsigned long long*>(misc + 40), b->d_flag,                   \
      b->groups.as<int2>(), reinterpret_cast<const int*>(misc) + 44, cursor, b->stack_cap,                        \
      reinterpret_cast<int*>(misc) + 46, b->pipeline, shv)
#define NB_BH_LAUNCH_P(BPL_, E_, S_)                                                                               \
  bh_force_pairs_kernel<BPL_, E_, S_><<<grid, BH_WARPS * 32, smem, st>>>(                                          \
      (int)n, b->node_com.as<float4>(), b->pairs.as<float4>(), b->perm_sorted, theta * theta, G, eps2, d->acc_x,   \
      d->acc_y, d->acc_z, reinterpret_cast<unsigned long long*>(misc + 40), b->d_flag, b->groups.as<int2>(),       \
      reinterpret_cast<const int*>(misc) + 44, cursor, b->stack_cap, reinterpret_cast<int*>(misc) + 46, shv)
#define NB_BH_LAUNCH(BPL_, E_)                                  \
  do {                                                          \
    if (b->walk_pairs) {                                        \
      if (shard) NB_BH_LAUNCH_P(BPL_, E_, true);                \
      else NB_BH_LAUNCH_P(BPL_, E_, false);                     \
    } else if (shard) NB_BH_LAUNCH_S(BPL_, E_, true);           \
    else NB_BH_LAUNCH_S(BPL_, E_, false);                       \
  } while (0)
  if (b->group == 64) {
    if (eps0) NB_BH_LAUNCH(2, true);
    else NB_BH_LAUNCH(2, false);
  } else {
    if (eps0) NB_BH_LAUNCH(1, true);
    else NB_BH_LAUNCH(1, false);
  }
#undef NB_BH_LAUNCH_S
#undef NB_BH_LAUNCH_P
#undef NB_BH_LAUNCH
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
}

int bh_shard_wait(nb_context* ctx, const unsigned* flags, int flag_base, unsigned step, int nranks, int self,
                  int* timeout_flag, cudaStream_t st) {
  return launch_peer_wait(ctx, flags, flag_base, step, nranks, self, timeout_flag, st);
}

// inverse of the sort permutation of the tree just built (enqueued behind this rank's traversal, i.e. before the
// wait for the peers' accelerations)
int bh_shard_inverse(nb_context* ctx, cudaStream_t st) {
  BH* b = bh_get(ctx);
  if (!b || !b->built) {
    set_error("Barnes-Hut: no tree built");
    return NB_ERR_INVALID;
  }
  const size_t n = b->n;
  if (int rc = b->inv.ensure(n * 4)) return rc;
  bh_inverse_perm_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(b->perm_sorted, n, b->inv.as<int32_t>());
  NB_LAUNCH_CHECK(ctx);
  return NB_OK;
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
  if (kick) bh_shard_finish_kernel<true><<<gn, 256, 0, st>>>(acc_sorted, b->inv.as<int32_t>(), *d, n, dt);
  else bh_shard_finish_kernel<false><<<gn, 256, 0, st>>>(acc_sorted, b->inv.as<int32_t>(), *d, n, dt);
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
  if (keys) NB_CUDA(cudaMemcpy(keys, b->keys_sorted, n * 4, cudaMemcpyDeviceToHost));
  if (perm) NB_CUDA(cudaMemcpy(perm, b->perm_sorted, n * 4, cudaMemcpyDeviceToHost));
  if (parent) {
    if (int rc = b->parent.ensure(2 * n * 4)) return rc;
    bh_parent_kernel<<<(unsigned)((n + 255) / 256), 256>>>((int)n, b->meta.as<int4>(), b->parent.as<int32_t>());
    NB_CUDA(cudaGetLastError());
    NB_CUDA(cudaMemcpy(parent, b->parent.p, (2 * n - 1) * 4, cudaMemcpyDeviceToHost));
  }
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
