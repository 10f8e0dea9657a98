/*
 * oracle/lbvh.c -- sequential restatement of the binary radix tree (Karras, "Maximizing
 * Parallelism in the Construction of BVHs, Octrees, and k-d Trees", HPG 2012, sec. 4) that the
 * GPU path builds from the reference's sorted Morton keys
 * (src/cuda/force_barnes_hut.cu:113-127, :276-280), plus the bottom-up centres of mass and the
 * group ("warp-cooperative") traversal.  The opening test and the monopole formula are the
 * reference's (src/cuda/force_barnes_hut.cu:160-183).  TEST INFRASTRUCTURE ONLY (see oracle.h).
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

static inline int clz32(uint32_t v) { return v ? __builtin_clz(v) : 32; }

/* delta(i,j): length of the common prefix of keys i and j; equal keys fall back to the
 * positions (Karras 2012, sec. 4, "duplicate keys"). Out of range -> -1. */
static inline int delta(const uint32_t* k, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  uint32_t a = k[i], b = k[j];
  if (a == b) return 32 + clz32((uint32_t)i ^ (uint32_t)j);
  return clz32(a ^ b);
}

void orc_lbvh_build(int n, const uint32_t* k, int32_t* left, int32_t* right, int32_t* parent,
                    int32_t* prefix, int32_t* range_first, int32_t* range_last) {
  if (n <= 0) return;
  parent[0] = -1;
  if (n == 1) return; /* single leaf = node 0, no internal nodes */
  const int leaf0 = n - 1;
  for (int i = 0; i < n - 1; i++) {
    int d = (delta(k, n, i, i + 1) - delta(k, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(k, n, i, i - d);
    int lmax = 2;
    while (delta(k, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
      if (delta(k, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(k, n, i, j);
    int s = 0, t = l;
    do {
      t = (t + 1) / 2;
      if (delta(k, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + (d < 0 ? d : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    int lc = (lo == gamma) ? leaf0 + gamma : gamma;
    int rc = (hi == gamma + 1) ? leaf0 + gamma + 1 : gamma + 1;
    left[i] = lc;
    right[i] = rc;
    parent[lc] = i;
    parent[rc] = i;
    /* key bits start at bit 29: prefix in key bits = clz - 2, capped at 30 (equal keys) */
    int p = dnode - 2;
    prefix[i] = p > 30 ? 30 : p;
    if (range_first) range_first[i] = lo;
    if (range_last) range_last[i] = hi;
  }
}

/* node ordering guarantees nothing about child > parent, so do an explicit post-order walk. */
void orc_lbvh_com(int n, const int32_t* left, const int32_t* right, const float* sx, const float* sy,
                  const float* sz, const float* sm, float* node_com) {
  if (n <= 0) return;
  const int leaf0 = n - 1;
  for (int k = 0; k < n; k++) {
    float* c = node_com + 4 * (size_t)(leaf0 + k);
    c[0] = sx[k]; c[1] = sy[k]; c[2] = sz[k]; c[3] = sm[k];
  }
  if (n == 1) return;
  /* iterative post-order */
  int32_t* stack = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)n);
  uint8_t* done = (uint8_t*)calloc((size_t)n, 1);
  int sp = 0;
  stack[sp++] = 0;
  while (sp > 0) {
    int i = stack[sp - 1];
    int l = left[i], r = right[i];
    int need = 0;
    if (l < leaf0 && !done[l]) { stack[sp++] = l; need = 1; }
    if (r < leaf0 && !done[r]) { stack[sp++] = r; need = 1; }
    if (need) continue;
    sp--;
    const float* a = node_com + 4 * (size_t)l;
    const float* b = node_com + 4 * (size_t)r;
    float* c = node_com + 4 * (size_t)i;
    /* same operation order as the GPU kernel (nbody_b200/csrc/bh.cu: bh_com_kernel):
     * m = ma + mb; c = (ca*ma + cb*mb) / m with separately rounded mul, add, div. */
    float m = a[3] + b[3];
    if (m > 0.0f) {
      c[0] = (a[0] * a[3] + b[0] * b[3]) / m;
      c[1] = (a[1] * a[3] + b[1] * b[3]) / m;
      c[2] = (a[2] * a[3] + b[2] * b[3]) / m;
    } else {
      c[0] = a[0]; c[1] = a[1]; c[2] = a[2];
    }
    c[3] = m;
    done[i] = 1;
  }
  free(stack);
  free(done);
}

/* Side of the Morton cell a node with common-prefix length L lives in: the key grid splits
 * x first, then y, then z (x is bit 2 of each triple, src/cuda/force_barnes_hut.cu:37), each
 * axis normalised by its own extent + 0.001f (:265-267); the opening test uses the largest
 * side, which reduces to the reference's octree cell side 2*half_size (:164) at L = 3k for a
 * cubic box. */
float orc_lbvh_cell_size(int L, const float bbox[6]) {
  if (L > 30) L = 30;
  if (L < 0) L = 0;
  const float sx = bbox[3] - bbox[0] + 0.001f, sy = bbox[4] - bbox[1] + 0.001f, sz = bbox[5] - bbox[2] + 0.001f;
  float ex = ldexpf(sx, -((L + 2) / 3)), ey = ldexpf(sy, -((L + 1) / 3)), ez = ldexpf(sz, -(L / 3));
  return fmaxf(ex, fmaxf(ey, ez));
}

/* Groups of a supergroup, as nbody_b200/csrc/bh.cu tb_emit forms them (mode bit 0; mode 0 = equal runs only, kept for
 * the comparison): follow the tree split whenever that costs no extra warp
 * (ceil(a / group) + ceil(b / group) == ceil(c / group)), else cut equal runs. */
static int g_hybrid_slack_from = 0; /* experiment knob (orc_set_hybrid_slack): allow ONE extra group when need >= this */
void orc_set_hybrid_slack(int from) { g_hybrid_slack_from = from; }
/* experiment knobs / counters of the walk (tools/ and tests/measure_* only) */
static int g_leaf_bucket = 0;      /* a rejected node with at most this many leaves hands over all its leaves */
static uint64_t g_walk_pops = 0;   /* nodes popped by the last orc_lbvh_forces_ex call */
static uint64_t g_walk_rounds = 0; /* rounds (pops of up to 32 entries) of that call */
void orc_set_leaf_bucket(int b) { g_leaf_bucket = b; }
static int g_walk_pairs = 1;       /* 1: restate bh_force_pairs_kernel (the default walk), 0: bh_force_kernel (NB_BH_WALK=nodes) */
void orc_set_walk_pairs(int on) { g_walk_pairs = on; }
uint64_t orc_get_walk_pops(void) { return g_walk_pops; }
uint64_t orc_get_walk_rounds(void) { return g_walk_rounds; }
static void hybrid_groups(int ni, int leaf0, int group, const int32_t* left, const int32_t* right,
                          const int32_t* range_first, const int32_t* range_last, int32_t* gfirst, int32_t* gcount,
                          int* ngroups) {
  if (ni >= leaf0) {
    gfirst[*ngroups] = ni - leaf0;
    gcount[(*ngroups)++] = 1;
    return;
  }
  const int first = range_first[ni], c = range_last[ni] - range_first[ni] + 1;
  if (c <= group) {
    gfirst[*ngroups] = first;
    gcount[(*ngroups)++] = c;
    return;
  }
  const int l = left[ni], r = right[ni];
  const int a = l >= leaf0 ? 1 : range_last[l] - range_first[l] + 1, b = c - a;
  const int need = (c + group - 1) / group;
  const int slack = (g_hybrid_slack_from > 0 && need >= g_hybrid_slack_from) ? 1 : 0;
  if ((a + group - 1) / group + (b + group - 1) / group <= need + slack) {
    hybrid_groups(l, leaf0, group, left, right, range_first, range_last, gfirst, gcount, ngroups);
    hybrid_groups(r, leaf0, group, left, right, range_first, range_last, gfirst, gcount, ngroups);
    return;
  }
  for (int i = 0; i < need; i++) {
    const int b0 = (int)(((long long)i * c) / need), b1 = (int)(((long long)(i + 1) * c) / need);
    gfirst[*ngroups] = first + b0;
    gcount[(*ngroups)++] = b1 - b0;
  }
}

void orc_lbvh_forces(int n, const int32_t* left, const int32_t* right, const int32_t* prefix,
                     const int32_t* range_first, const int32_t* range_last, const float* node_com,
                     const float* sx, const float* sy, const float* sz, const float bbox[6], float theta, float G,
                     float eps, int group, int super, float* ax, float* ay, float* az, uint64_t* interactions) {
  orc_lbvh_forces_ex(n, left, right, prefix, range_first, range_last, node_com, sx, sy, sz, bbox, theta, G, eps, group,
                     super, 1, 0.0f, ax, ay, az, interactions, 0);
}

/* mode bit 0: hybrid grouping; alpha: acceptance distance^2 = (1 - alpha) * dmin^2 + alpha * dcentre^2 (alpha = 0: the
 * conservative group-box test).  stats[0] = groups, stats[1] = sum over groups of accepted list entries. */
void orc_lbvh_forces_ex(int n, const int32_t* left, const int32_t* right, const int32_t* prefix,
                        const int32_t* range_first, const int32_t* range_last, const float* node_com,
                        const float* sx, const float* sy, const float* sz, const float bbox[6], float theta, float G,
                        float eps, int group, int super, int mode, float alpha, float* ax, float* ay, float* az,
                        uint64_t* interactions, uint64_t* stats) {
  /* Sequential restatement of nbody_b200/csrc/bh.cu: bh_groups_kernel + bh_force_pairs_kernel / bh_force_kernel.  A
   * supergroup is a maximal subtree with at most `super` (>= group) leaves (a lone leaf whose parent
   * is larger is a supergroup of one); its c bodies are cut into ceil(c / group) runs of consecutive
   * bodies, run i starting at floor(i * c / runs): those runs are the traversal groups (with
   * super == group they are the subtrees themselves).
   * A group (one warp) owns a stack of node indices.  Two walks, same acceptance test per node and so the same
   * interaction set, different order (orc_set_walk_pairs):
   *  - pairs (default, bh_force_pairs_kernel): the stack holds nodes to OPEN, starting with a virtual parent of the
   *    root.  Each round pops the top min(32, size) entries (lane 0..m-1 = bottom to top); lane l tests the left
   *    and then the right child of its node; accepted children are appended to the interaction list and rejected
   *    ones pushed, both in lane order, a lane's left child first.
   *  - nodes (bh_force_kernel, NB_BH_WALK=nodes): the stack holds nodes to TEST; each round pops up to 32, accepted
   *    ones join the list in lane order, the children (left, right) of rejected ones are pushed in lane order.
   * A node is used for the whole
   * group iff it is a leaf or size^2 < theta^2 * (dmin^2 + eps^2), dmin = distance from the node's
   * centre of mass to the group's bounding box (<= every body's own distance, so every body of the
   * group would also accept it under the reference's per-body test,
   * src/cuda/force_barnes_hut.cu:160-172); otherwise its children are pushed (left, right) in lane
   * order.  Accepted nodes are appended in lane order to an interaction list that is evaluated 32
   * entries at a time, each body accumulating in list order with the reference's monopole formula
   * (:175-181).  A term with d^2 == 0 (the body's own leaf or an exact coincidence at eps = 0)
   * contributes nothing; with eps > 0 the own leaf contributes exactly +0. */
  const float theta2 = theta * theta, eps2 = eps * eps;
  const int leaf0 = n - 1;
  if (super < group) super = group;
  float size2_of[32];
  for (int L = 0; L <= 30; L++) {
    float s = orc_lbvh_cell_size(L, bbox);
    size2_of[L] = s * s;
  }
  size2_of[31] = 0.0f;
  uint64_t total = 0, entries = 0, pops = 0, rounds = 0;
  const int bucket = g_leaf_bucket;
  /* groups: top-down, a node whose leaf range fits becomes a group */
  int32_t* gfirst = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
  int32_t* gcount = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
  int ngroups = 0;
  {
    int32_t* st = (int32_t*)malloc(sizeof(int32_t) * (size_t)(2 * n + 2));
    int sp = 0;
    st[sp++] = 0;
    while (sp > 0) {
      const int ni = st[--sp];
      if (ni >= leaf0) {
        gfirst[ngroups] = ni - leaf0;
        gcount[ngroups++] = 1;
      } else if ((mode & 1) && range_last[ni] - range_first[ni] + 1 <= super) {
        hybrid_groups(ni, leaf0, group, left, right, range_first, range_last, gfirst, gcount, &ngroups);
      } else if (range_last[ni] - range_first[ni] + 1 <= super) {
        const int first = range_first[ni], c = range_last[ni] - range_first[ni] + 1;
        const int runs = (c + group - 1) / group;
        for (int i = 0; i < runs; i++) {
          const int b0 = (int)(((long long)i * c) / runs), b1 = (int)(((long long)(i + 1) * c) / runs);
          gfirst[ngroups] = first + b0;
          gcount[ngroups++] = b1 - b0;
        }
      } else {
        st[sp++] = right[ni];
        st[sp++] = left[ni];
      }
    }
    free(st);
  }
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : total, entries, pops, rounds)
  for (int g = 0; g < ngroups; g++) {
    const int k0 = gfirst[g], cnt = gcount[g];
    float accx[64], accy[64], accz[64];
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int t = 0; t < cnt; t++) {
      accx[t] = accy[t] = accz[t] = 0.0f;
      lo[0] = fminf(lo[0], sx[k0 + t]); hi[0] = fmaxf(hi[0], sx[k0 + t]);
      lo[1] = fminf(lo[1], sy[k0 + t]); hi[1] = fmaxf(hi[1], sy[k0 + t]);
      lo[2] = fminf(lo[2], sz[k0 + t]); hi[2] = fmaxf(hi[2], sz[k0 + t]);
    }
    int32_t* stack = (int32_t*)malloc(sizeof(int32_t) * 8192);
    int32_t* list = (int32_t*)malloc(sizeof(int32_t) * (size_t)(96 + 32 * (bucket > 0 ? bucket : 0)));
    int sp = 0, nl = 0;
    stack[sp++] = 0;
    const int walk_pairs = g_walk_pairs && alpha == 0.0f && bucket == 0; /* the experiment knobs belong to the node walk */
    if (walk_pairs) stack[0] = -1; /* the virtual parent of the root (record n - 1 of bh_pairs_kernel) */
    for (;;) {
      const int m = sp < 32 ? sp : 32;
      int32_t popped[32];
      for (int l = 0; l < m; l++) popped[l] = stack[sp - m + l];
      sp -= m;
      pops += (uint64_t)m;
      rounds++;
      int32_t push[64];
      int np = 0;
      /* pair walk (bh_force_pairs_kernel): the stack holds nodes to OPEN; lane l tests both children of its node,
       * left before right; accepted children join the list and rejected ones the stack, both in lane order */
      for (int l = 0; walk_pairs && l < m; l++) {
        const int pn = popped[l];
        const int kids[2] = {pn < 0 ? 0 : left[pn], pn < 0 ? -1 : right[pn]};
        for (int q = 0; q < 2; q++) {
          const int ni = kids[q];
          if (ni < 0) continue;
          const float* c = node_com + 4 * (size_t)ni;
          if (c[3] == 0.0f) continue; /* src/cuda/force_barnes_hut.cu:157-158 */
          int accept = 1;
          if (ni < leaf0) {
            float ddx = fmaxf(fmaxf(lo[0] - c[0], c[0] - hi[0]), 0.0f);
            float ddy = fmaxf(fmaxf(lo[1] - c[1], c[1] - hi[1]), 0.0f);
            float ddz = fmaxf(fmaxf(lo[2] - c[2], c[2] - hi[2]), 0.0f);
            float d2 = fmaf(ddz, ddz, fmaf(ddy, ddy, fmaf(ddx, ddx, eps2)));
            accept = size2_of[prefix[ni]] < theta2 * d2;
          }
          if (accept) list[nl++] = ni;
          else push[np++] = ni;
        }
      }
      for (int l = 0; !walk_pairs && l < m; l++) {
        const int ni = popped[l];
        const float* c = node_com + 4 * (size_t)ni;
        if (c[3] == 0.0f) continue; /* src/cuda/force_barnes_hut.cu:157-158 */
        int accept = 1;
        if (ni < leaf0) {
          float ddx = fmaxf(fmaxf(lo[0] - c[0], c[0] - hi[0]), 0.0f);
          float ddy = fmaxf(fmaxf(lo[1] - c[1], c[1] - hi[1]), 0.0f);
          float ddz = fmaxf(fmaxf(lo[2] - c[2], c[2] - hi[2]), 0.0f);
          float d2 = fmaf(ddz, ddz, fmaf(ddy, ddy, fmaf(ddx, ddx, eps2)));
          if (alpha != 0.0f) {
            const float cx = 0.5f * (lo[0] + hi[0]) - c[0], cy = 0.5f * (lo[1] + hi[1]) - c[1],
                        cz = 0.5f * (lo[2] + hi[2]) - c[2];
            const float dc2 = cx * cx + cy * cy + cz * cz + eps2;
            d2 = (1.0f - alpha) * d2 + alpha * dc2;
          }
          accept = size2_of[prefix[ni]] < theta2 * d2;
        }
        if (accept) {
          list[nl++] = ni;
        } else if (bucket > 0 && range_last[ni] - range_first[ni] + 1 <= bucket) {
          for (int k = range_first[ni]; k <= range_last[ni]; k++)
            if (node_com[4 * (size_t)(leaf0 + k) + 3] != 0.0f) list[nl++] = leaf0 + k;
        } else {
          push[np++] = left[ni];
          push[np++] = right[ni];
        }
      }
      for (int q = 0; q < np; q++) stack[sp++] = push[q];
      const int done = (sp == 0);
      while (nl >= 32 || (done && nl > 0)) {
        const int ne = nl < 32 ? nl : 32;
        for (int e = 0; e < ne; e++) {
          const float* c = node_com + 4 * (size_t)list[e];
          const float gm = G * c[3];
          for (int t = 0; t < cnt; t++) {
            const int k = k0 + t;
            float dx = c[0] - sx[k], dy = c[1] - sy[k], dz = c[2] - sz[k];
            float d2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps2)));
            if (!(d2 > 0.0f)) continue;
            float inv = 1.0f / sqrtf(d2);
            float f = gm * (inv * inv * inv);
            accx[t] = fmaf(f, dx, accx[t]);
            accy[t] = fmaf(f, dy, accy[t]);
            accz[t] = fmaf(f, dz, accz[t]);
          }
        }
        total += (uint64_t)ne * (uint64_t)cnt;
        entries += (uint64_t)ne;
        for (int e = ne; e < nl; e++) list[e - ne] = list[e];
        nl -= ne;
      }
      if (done) break;
    }
    free(stack);
    free(list);
    for (int t = 0; t < cnt; t++) {
      ax[k0 + t] = accx[t]; ay[k0 + t] = accy[t]; az[k0 + t] = accz[t];
    }
  }
  free(gfirst);
  free(gcount);
  if (interactions) *interactions = total;
  g_walk_pops = pops;
  g_walk_rounds = rounds;
  if (stats) {
    stats[0] = (uint64_t)ngroups;
    stats[1] = entries;
  }
}
