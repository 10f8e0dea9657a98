/* oracle/render.c -- TEST INFRASTRUCTURE (see oracle.h).  CPU restatement of the reference's point-sprite
 * pipeline: vertex shader src/render/renderer.cpp:16-36, fragment shader :40-77, blend state :110-113
 * (GL_SRC_ALPHA, GL_ONE_MINUS_SRC_ALPHA, no depth test), clear colour :9-11, draw order = body order
 * (:214), plus the fixed-function parts between the two shaders that OpenGL 3.3 specifies: clip-volume
 * culling of the point's centre, viewport transform, point-size clamp, point rasterisation (pixels whose
 * centre lies inside the size x size square), gl_PointCoord.  Bodies are drawn one after the other into an
 * fp32 RGBA framebuffer that is converted to unorm8 once at the end.  Built with -ffp-contract=off so every
 * operation rounds exactly as the __f*_rn intrinsics in nbody_b200/csrc/render.cu.
 * No GL implementation is available on this box (graphics: unavailable), so this oracle is pinned by the
 * shader source and the GL specification only. */
#include <math.h>
#include <stdlib.h>

#include "oracle.h"

static float clamp01(float t) { return fminf(fmaxf(t, 0.0f), 1.0f); }

static float mat_row(const float* m, int r, float x, float y, float z, float w) {
  return ((m[r] * x + m[4 + r] * y) + m[8 + r] * z) + m[12 + r] * w;
}

static void pixel_span(float c, float h, int extent, int* lo, int* hi) {
  const float a = ceilf((c - h) - 0.5f), b = ceilf((c + h) - 0.5f);
  *lo = (int)fmaxf(a, 0.0f);
  *hi = (int)fminf(b, (float)extent);
  if (*hi < *lo) *hi = *lo;
}

static unsigned unorm8(float c) { return (unsigned)lrintf(clamp01(c) * 255.0f); }

uint64_t orc_render_splats(int n, const float* px, const float* py, const float* pz, const float* vx, const float* vy,
                           const float* vz, const float view[16], const float proj[16], float point_size,
                           float max_depth, float max_velocity, int color_mode, int width, int height,
                           float size_min, float size_max, const float clear[4], uint8_t* rgba) {
  const size_t pixels = (size_t)width * (size_t)height;
  float* fb = (float*)malloc(pixels * 4 * sizeof(float));
  for (size_t i = 0; i < pixels; i++)
    for (int c = 0; c < 4; c++) fb[4 * i + c] = clear[c];
  static const float ramp[3][2][3] = {{{1.0f, 0.5f, 0.0f}, {0.0f, 0.5f, 1.0f}},   /* depth   (:59-61) */
                                      {{0.0f, 0.0f, 1.0f}, {1.0f, 0.0f, 0.0f}},   /* velocity (:63-65) */
                                      {{0.2f, 0.2f, 0.8f}, {1.0f, 1.0f, 0.2f}}};  /* density (:67-69) */
  uint64_t fragments = 0;
  for (int i = 0; i < n; i++) {
    const float ex = mat_row(view, 0, px[i], py[i], pz[i], 1.0f), ey = mat_row(view, 1, px[i], py[i], pz[i], 1.0f),
                ez = mat_row(view, 2, px[i], py[i], pz[i], 1.0f), ew = mat_row(view, 3, px[i], py[i], pz[i], 1.0f);
    const float cx = mat_row(proj, 0, ex, ey, ez, ew), cy = mat_row(proj, 1, ex, ey, ez, ew),
                cz = mat_row(proj, 2, ex, ey, ez, ew), cw = mat_row(proj, 3, ex, ey, ez, ew);
    if (!((cx >= -cw && cx <= cw) && (cy >= -cw && cy <= cw) && (cz >= -cw && cz <= cw) && cw > 0.0f)) continue;
    const float depth = -ez;
    float size = point_size / fmaxf(depth, 0.1f);
    size = fminf(fmaxf(size, size_min), size_max);
    const float xw = ((cx / cw + 1.0f) * 0.5f) * (float)width, yw = ((cy / cw + 1.0f) * 0.5f) * (float)height;
    float t;
    if (color_mode == 1) {
      const float sp = sqrtf((vx[i] * vx[i] + vy[i] * vy[i]) + vz[i] * vz[i]);
      t = clamp01(sp / max_velocity);
    } else {
      t = clamp01(depth / max_depth);
    }
    const float omt = 1.0f - t;
    const float* lo = ramp[color_mode][0];
    const float* hi = ramp[color_mode][1];
    const float r = lo[0] * omt + hi[0] * t, g = lo[1] * omt + hi[1] * t, b = lo[2] * omt + hi[2] * t;
    const float h = size * 0.5f;
    int x0, x1, y0, y1;
    pixel_span(xw, h, width, &x0, &x1);
    pixel_span(yw, h, height, &y0, &y1);
    for (int qy = y0; qy < y1; qy++)
      for (int qx = x0; qx < x1; qx++) {
        fragments++;
        const float u = (((float)qx + 0.5f) - xw) / size, v = (((float)qy + 0.5f) - yw) / size;
        const float dist = sqrtf(u * u + v * v);
        if (dist > 0.5f) continue; /* discard */
        const float q = clamp01((dist - 0.4f) / (0.5f - 0.4f));
        const float a = 1.0f - (q * q) * (3.0f - 2.0f * q);
        const float oma = 1.0f - a;
        float* d = fb + 4 * ((size_t)qy * (size_t)width + (size_t)qx);
        d[0] = r * a + d[0] * oma;
        d[1] = g * a + d[1] * oma;
        d[2] = b * a + d[2] * oma;
        d[3] = a * a + d[3] * oma;
      }
  }
  for (size_t i = 0; i < pixels; i++)
    for (int c = 0; c < 4; c++) rgba[4 * i + c] = (uint8_t)unorm8(fb[4 * i + c]);
  free(fb);
  return fragments;
}
