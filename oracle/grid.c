/*
 * oracle/grid.c -- CPU restatement of the reference's spatial-hash path
 * (src/cuda/force_spatial_hash.cu).  TEST INFRASTRUCTURE ONLY (see oracle.h).
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* src/cuda/force_spatial_hash.cu:14-17 */
void orc_cell_index(float x, float y, float z, float cell_size, int32_t out[3]) {
  out[0] = (int32_t)floorf(x / cell_size);
  out[1] = (int32_t)floorf(y / cell_size);
  out[2] = (int32_t)floorf(z / cell_size);
}

/* src/cuda/force_spatial_hash.cu:226-232 (pad +-0.001f on the host) and :244-246 (dims) */
void orc_grid_dims(float bbox[6], float cell_size, int32_t dims[3]) {
  for (int a = 0; a < 3; a++) {
    bbox[a] -= 0.001f;
    bbox[3 + a] += 0.001f;
  }
  for (int a = 0; a < 3; a++) dims[a] = (int32_t)ceilf((bbox[3 + a] - bbox[a]) / cell_size) + 1;
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* src/cuda/force_spatial_hash.cu:28-49 */
void orc_assign_cells(int n, const float* px, const float* py, const float* pz, const float bb[6],
                      float cell_size, const int32_t dims[3], int32_t* particle_cell) {
  for (int i = 0; i < n; i++) {
    float x = px[i] - bb[0], y = py[i] - bb[1], z = pz[i] - bb[2];
    int cx = (int)floorf(x / cell_size), cy = (int)floorf(y / cell_size), cz = (int)floorf(z / cell_size);
    cx = clampi(cx, 0, dims[0] - 1);
    cy = clampi(cy, 0, dims[1] - 1);
    cz = clampi(cz, 0, dims[2] - 1);
    particle_cell[i] = cx + cy * dims[0] + cz * dims[0] * dims[1];
  }
}

/* :52-58 counts, :285-288 exclusive scan, :61-68 end = start + count, :71-80 scatter.  The
 * reference's scatter order inside a cell depends on atomicAdd arrival; the deterministic
 * representative used here is ascending body index. */
void orc_cell_ranges(int n, const int32_t* particle_cell, int total_cells, int32_t* cell_start,
                     int32_t* cell_end, int32_t* sorted_indices) {
  int32_t* cnt = (int32_t*)calloc((size_t)total_cells + 1, sizeof(int32_t));
  for (int i = 0; i < n; i++) cnt[particle_cell[i]]++;
  int32_t run = 0;
  for (int c = 0; c < total_cells; c++) {
    cell_start[c] = run;
    run += cnt[c];
    cell_end[c] = run;
    cnt[c] = 0;
  }
  for (int i = 0; i < n; i++) {
    int c = particle_cell[i];
    sorted_indices[cell_start[c] + cnt[c]++] = i;
  }
  free(cnt);
}

/* src/cuda/force_spatial_hash.cu:83-152: own cell recomputed without clamp (:100-102), 27
 * neighbour cells dz,dy,dx outer->inner, out-of-grid skipped, j==i skipped, un-softened
 * r^2 < cutoff^2 test, then softened force. */
void orc_spatial_forces(int n, const float* px, const float* py, const float* pz, const float* m,
                        const int32_t* cell_start, const int32_t* cell_end, const int32_t* sorted_indices,
                        const int32_t dims[3], float cell_size, const float bb[6], float cutoff, float G,
                        float eps, float* ax, float* ay, float* az, uint64_t* pairs) {
  const float cutoff2 = cutoff * cutoff, eps2 = eps * eps; /* :313-314 */
  uint64_t total = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : total)
  for (int i = 0; i < n; i++) {
    const float xi = px[i], yi = py[i], zi = pz[i];
    float sx = 0.0f, sy = 0.0f, sz = 0.0f;
    const int cx = (int)floorf((xi - bb[0]) / cell_size);
    const int cy = (int)floorf((yi - bb[1]) / cell_size);
    const int cz = (int)floorf((zi - bb[2]) / cell_size);
    for (int dz = -1; dz <= 1; dz++)
      for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++) {
          int nx = cx + dx, ny = cy + dy, nz = cz + dz;
          if (nx < 0 || nx >= dims[0] || ny < 0 || ny >= dims[1] || nz < 0 || nz >= dims[2]) continue;
          int cell = nx + ny * dims[0] + nz * dims[0] * dims[1];
          for (int k = cell_start[cell]; k < cell_end[cell]; k++) {
            int j = sorted_indices[k];
            if (j == i) continue;
            float ddx = px[j] - xi, ddy = py[j] - yi, ddz = pz[j] - zi;
            /* :131 as the reference's BINARY evaluates it: nvcc (default -fmad=true) contracts dx*dx + dy*dy + dz*dz
             * to FMUL dy*dy, FFMA dx*dx + ., FFMA dz*dz + . (SASS of oracle/_ref/libnbody_ref.so,
             * spatialHashForceKernel); the in-cutoff decision follows that rounding */
            float dist2 = fmaf(ddz, ddz, fmaf(ddx, ddx, ddy * ddy));
            if (dist2 < cutoff2) {
              dist2 += eps2;
              float inv = 1.0f / sqrtf(dist2);
              float inv3 = inv * inv * inv;
              float f = G * m[j] * inv3;
              sx += f * ddx; sy += f * ddy; sz += f * ddz;
              total++;
            }
          }
        }
    ax[i] = sx; ay[i] = sy; az[i] = sz;
  }
  if (pairs) *pairs = total;
}
