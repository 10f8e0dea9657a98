/*
 * oracle/tree.c -- CPU restatement of the Barnes-Hut front half (bbox, Morton keys, stable
 * sort), the reference's host octree (as written and as intended), its traversal, and the
 * Karras LBVH that the GPU path builds.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* src/cuda/force_barnes_hut.cu:66-110 -- min/max are exact and order-independent. */
void orc_bbox(int n, const float* px, const float* py, const float* pz, float out[6]) {
  float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
  for (int i = 0; i < n; i++) {
    mn[0] = fminf(mn[0], px[i]); mx[0] = fmaxf(mx[0], px[i]);
    mn[1] = fminf(mn[1], py[i]); mx[1] = fmaxf(mx[1], py[i]);
    mn[2] = fminf(mn[2], pz[i]); mx[2] = fmaxf(mx[2], pz[i]);
  }
  for (int a = 0; a < 3; a++) { out[a] = mn[a]; out[3 + a] = mx[a]; }
}

/* src/cuda/force_barnes_hut.cu:23-29 */
uint32_t orc_expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

/* float -> unsigned as the device does it (cvt.rzi.u32.f32 saturates; NaN -> 0). */
static uint32_t f2u_sat(float f) {
  if (!(f > 0.0f)) return 0u;
  if (f >= 4294967296.0f) return 0xFFFFFFFFu;
  return (uint32_t)f;
}

/* src/cuda/force_barnes_hut.cu:31-38 */
uint32_t orc_morton_code(float x, float y, float z) {
  uint32_t ix = f2u_sat(x * 1024.0f), iy = f2u_sat(y * 1024.0f), iz = f2u_sat(z * 1024.0f);
  if (ix > 1023u) ix = 1023u;
  if (iy > 1023u) iy = 1023u;
  if (iz > 1023u) iz = 1023u;
  return (orc_expand_bits(ix) << 2) | (orc_expand_bits(iy) << 1) | orc_expand_bits(iz);
}

/* src/cuda/force_barnes_hut.cu:260-274 (host: size = max-min+0.001f, inv = 1.0f/size) and
 * :113-127 (device: n = (p-min)*inv, key, indices[i]=i). */
void orc_morton_keys(int n, const float* px, const float* py, const float* pz, const float bbox[6],
                     uint32_t* keys, int32_t* indices) {
  const float sx = bbox[3] - bbox[0] + 0.001f, sy = bbox[4] - bbox[1] + 0.001f, sz = bbox[5] - bbox[2] + 0.001f;
  const float ix = 1.0f / sx, iy = 1.0f / sy, iz = 1.0f / sz;
  for (int i = 0; i < n; i++) {
    float nx = (px[i] - bbox[0]) * ix;
    float ny = (py[i] - bbox[1]) * iy;
    float nz = (pz[i] - bbox[2]) * iz;
    keys[i] = orc_morton_code(nx, ny, nz);
    indices[i] = i;
  }
}

/* src/cuda/force_barnes_hut.cu:276-280: thrust::sort_by_key on u32 keys = stable LSD radix. */
void orc_sort_pairs(int n, uint32_t* keys, int32_t* vals) {
  uint32_t* k2 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
  int32_t* v2 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  uint32_t *ka = keys, *kb = k2;
  int32_t *va = vals, *vb = v2;
  for (int pass = 0; pass < 4; pass++) {
    size_t cnt[257];
    memset(cnt, 0, sizeof cnt);
    const int sh = pass * 8;
    for (int i = 0; i < n; i++) cnt[((ka[i] >> sh) & 255u) + 1]++;
    for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
    for (int i = 0; i < n; i++) {
      size_t p = cnt[(ka[i] >> sh) & 255u]++;
      kb[p] = ka[i];
      vb[p] = va[i];
    }
    uint32_t* tk = ka; ka = kb; kb = tk;
    int32_t* tv = va; va = vb; vb = tv;
  }
  /* 4 passes: result is back in keys/vals */
  free(k2);
  free(v2);
}

/* ---------------------------------------------------------------------------------------
 * Octree: src/cuda/force_barnes_hut.cu:335-431 (insertion) and :438-486 (centres of mass).
 * ------------------------------------------------------------------------------------- */
static void oct_init(orc_octnode* nd) {
  memset(nd, 0, sizeof *nd);
  for (int c = 0; c < 8; c++) nd->children[c] = -1;
  nd->particle_index = -1;
  nd->is_leaf = 1; /* include/nbody/barnes_hut_tree.hpp:20-29 default ctor */
}

static int oct_octant(const orc_octnode* nd, float x, float y, float z) {
  int o = 0; /* :365-371 */
  if (x >= nd->center[0]) o |= 1;
  if (y >= nd->center[1]) o |= 2;
  if (z >= nd->center[2]) o |= 4;
  return o;
}

static void oct_make_leaf(orc_octnode* nodes, int parent, int octant, int new_idx, int body, float x, float y,
                          float z, float m) {
  orc_octnode* ch = &nodes[new_idx];
  const orc_octnode* pa = &nodes[parent];
  oct_init(ch);
  float hs = pa->half_size * 0.5f; /* :379-383 */
  ch->center[0] = pa->center[0] + ((octant & 1) ? hs : -hs);
  ch->center[1] = pa->center[1] + ((octant & 2) ? hs : -hs);
  ch->center[2] = pa->center[2] + ((octant & 4) ? hs : -hs);
  ch->half_size = hs;
  ch->is_leaf = 1;
  ch->particle_index = body;
  ch->particle_count = 1;
  ch->total_mass = m;
  ch->com[0] = x; ch->com[1] = y; ch->com[2] = z;
  nodes[parent].children[octant] = new_idx;
}

int orc_octree_build(int n, const float* px, const float* py, const float* pz, const float* m,
                     const int32_t* sorted_idx, const float bbox[6], int as_written, orc_octnode* nodes,
                     int max_nodes, int* dropped) {
  int node_count = 1, ndropped = 0;
  orc_octnode* root = &nodes[0];
  oct_init(root);
  root->center[0] = (bbox[0] + bbox[3]) * 0.5f; /* :339-340 */
  root->center[1] = (bbox[1] + bbox[4]) * 0.5f;
  root->center[2] = (bbox[2] + bbox[5]) * 0.5f;
  float max_size = fmaxf(fmaxf(bbox[3] - bbox[0], bbox[4] - bbox[1]), bbox[5] - bbox[2]);
  root->half_size = max_size * 0.5f + 0.001f; /* :342-344 */
  root->is_leaf = 0;
  root->particle_count = n;

  for (int t = 0; t < n; t++) {
    const int idx = sorted_idx[t];
    const float x = px[idx], y = py[idx], z = pz[idx], mass = m[idx];
    int current = 0, depth = 0, placed = 0;
    if (as_written) {
      /* verbatim :363-394: the loop ends when it meets an existing leaf; nothing is inserted. */
      while (!nodes[current].is_leaf && depth < 20) {
        int o = oct_octant(&nodes[current], x, y, z);
        if (nodes[current].children[o] < 0) {
          if (node_count >= max_nodes) return -1;
          oct_make_leaf(nodes, current, o, node_count++, idx, x, y, z, mass);
          placed = 1;
          break;
        }
        current = nodes[current].children[o];
        depth++;
      }
      if (!placed && !nodes[current].is_leaf && depth >= 20) { /* :400-428 */
        int o = oct_octant(&nodes[current], x, y, z);
        if (node_count >= max_nodes) return -1;
        oct_make_leaf(nodes, current, o, node_count++, idx, x, y, z, mass);
        placed = 1;
      }
      if (!placed) ndropped++;
      continue;
    }
    /* intended algorithm: same geometry / octant rule, one body per leaf, an occupied leaf is
     * split; at the reference's depth cap (20) bodies are merged into the leaf. */
    for (;;) {
      orc_octnode* nd = &nodes[current];
      if (nd->is_leaf) {
        if (depth >= 20) {
          float tm = nd->total_mass + mass;
          if (tm > 0.0f) {
            nd->com[0] = (nd->com[0] * nd->total_mass + x * mass) / tm;
            nd->com[1] = (nd->com[1] * nd->total_mass + y * mass) / tm;
            nd->com[2] = (nd->com[2] * nd->total_mass + z * mass) / tm;
          }
          nd->total_mass = tm;
          nd->particle_count++;
          break;
        }
        /* split: push the resident body one level down */
        int ob = nd->particle_index;
        float ox = nd->com[0], oy = nd->com[1], oz = nd->com[2], om = nd->total_mass;
        nd->is_leaf = 0;
        nd->particle_index = -1;
        int oo = oct_octant(nd, ox, oy, oz);
        if (node_count >= max_nodes) return -1;
        oct_make_leaf(nodes, current, oo, node_count++, ob, ox, oy, oz, om);
        nd = &nodes[current];
      }
      int o = oct_octant(nd, x, y, z);
      if (nd->children[o] < 0) {
        if (node_count >= max_nodes) return -1;
        oct_make_leaf(nodes, current, o, node_count++, idx, x, y, z, mass);
        break;
      }
      current = nd->children[o];
      depth++;
    }
  }

  /* :452-481 reverse sweep (children always have larger indices than their parent) */
  for (int i = node_count - 1; i >= 0; i--) {
    orc_octnode* nd = &nodes[i];
    if (nd->is_leaf) {
      if (nd->particle_index >= 0 && nd->particle_count == 1) {
        int b = nd->particle_index;
        nd->com[0] = px[b]; nd->com[1] = py[b]; nd->com[2] = pz[b];
        nd->total_mass = m[b];
      }
    } else {
      float tm = 0.0f, cx = 0.0f, cy = 0.0f, cz = 0.0f;
      int cnt = 0;
      for (int c = 0; c < 8; c++) {
        if (nd->children[c] >= 0) {
          const orc_octnode* ch = &nodes[nd->children[c]];
          tm += ch->total_mass;
          cx += ch->com[0] * ch->total_mass;
          cy += ch->com[1] * ch->total_mass;
          cz += ch->com[2] * ch->total_mass;
          cnt += ch->particle_count;
        }
      }
      nd->total_mass = tm;
      if (i != 0) nd->particle_count = cnt;
      if (tm > 0.0f) {
        nd->com[0] = cx / tm; nd->com[1] = cy / tm; nd->com[2] = cz / tm;
      }
    }
  }
  if (dropped) *dropped = ndropped;
  return node_count;
}

/* src/cuda/force_barnes_hut.cu:130-201 */
void orc_octree_forces(int n, const float* px, const float* py, const float* pz, const orc_octnode* nodes,
                       int num_nodes, float theta, float G, float eps, float* ax, float* ay, float* az) {
  const float theta2 = theta * theta, eps2 = eps * eps; /* :494-495 */
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n; i++) {
    const float xi = px[i], yi = py[i], zi = pz[i];
    float sx = 0.0f, sy = 0.0f, sz = 0.0f;
    int stack[256], sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
      int ni = stack[--sp];
      if (ni < 0 || ni >= num_nodes) continue;
      const orc_octnode* nd = &nodes[ni];
      if (nd->total_mass == 0.0f) continue;
      float dx = nd->com[0] - xi, dy = nd->com[1] - yi, dz = nd->com[2] - zi;
      float dist2 = dx * dx + dy * dy + dz * dz + eps2;
      float size2 = 4.0f * nd->half_size * nd->half_size;
      if (nd->is_leaf || size2 / dist2 < theta2) {
        if (nd->particle_index != i) {
          float inv = 1.0f / sqrtf(dist2);
          float inv3 = inv * inv * inv;
          float f = G * nd->total_mass * inv3;
          sx += f * dx; sy += f * dy; sz += f * dz;
        }
      } else if (sp + 8 <= 256) {
        for (int c = 0; c < 8; c++)
          if (nd->children[c] >= 0) stack[sp++] = nd->children[c];
      }
    }
    ax[i] = sx; ay[i] = sy; az[i] = sz;
  }
}
