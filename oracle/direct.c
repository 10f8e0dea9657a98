/*
 * oracle/direct.c -- CPU restatement of direct N^2, Velocity Verlet and energy.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Compile with -ffp-contract=off.
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

void orc_set_num_threads(int t) {
#ifdef _OPENMP
  if (t > 0) omp_set_num_threads(t);
#else
  (void)t;
#endif
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* examples/example_force_methods.cpp:44-66 -- the reference's own CPU loop.
 * dist2 = dx*dx + dy*dy + dz*dz + eps*eps (left to right), inv = 1/sqrt, inv3 = inv*inv*inv,
 * f = G*m_j*inv3, a += f*d. */
void orc_direct_f32(int n, const float* px, const float* py, const float* pz, const float* m,
                    float* ax, float* ay, float* az, float G, float eps, int i0, int i1) {
  orc_direct_f32_eps2(n, px, py, pz, m, ax, ay, az, G, eps * eps, i0, i1);
}

/* src/cuda/force_direct.cu:55-75 -- same expression with eps2 = softening^2 precomputed
 * (include/nbody/force_calculator.hpp:64-67).  rsqrtf is replaced by 1/sqrtf as in the
 * reference's CPU code (force_direct.cu:112). */
void orc_direct_f32_eps2(int n, const float* px, const float* py, const float* pz, const float* m,
                         float* ax, float* ay, float* az, float G, float eps2, int i0, int i1) {
#pragma omp parallel for schedule(static)
  for (int i = i0; i < i1; i++) {
    const float xi = px[i], yi = py[i], zi = pz[i];
    float sx = 0.0f, sy = 0.0f, sz = 0.0f;
    for (int j = 0; j < n; j++) {
      if (j == i) continue;
      float dx = px[j] - xi;
      float dy = py[j] - yi;
      float dz = pz[j] - zi;
      float dist2 = dx * dx + dy * dy + dz * dz + eps2;
      float inv_dist = 1.0f / sqrtf(dist2);
      float inv_dist3 = inv_dist * inv_dist * inv_dist;
      float f = G * m[j] * inv_dist3;
      sx += f * dx;
      sy += f * dy;
      sz += f * dz;
    }
    ax[i] = sx;
    ay[i] = sy;
    az[i] = sz;
  }
}

/* fp64 evaluation of the same formula on the same fp32 inputs: the accuracy truth. */
void orc_direct_f64(int n, const float* px, const float* py, const float* pz, const float* m,
                    double* ax, double* ay, double* az, double G, double eps2, const int* idx, int nt) {
#pragma omp parallel for schedule(dynamic, 4)
  for (int t = 0; t < nt; t++) {
    const int i = idx ? idx[t] : t;
    const double xi = px[i], yi = py[i], zi = pz[i];
    double sx = 0.0, sy = 0.0, sz = 0.0;
    for (int j = 0; j < n; j++) {
      if (j == i) continue;
      double dx = (double)px[j] - xi;
      double dy = (double)py[j] - yi;
      double dz = (double)pz[j] - zi;
      double d2 = dx * dx + dy * dy + dz * dz + eps2;
      double inv = 1.0 / sqrt(d2);
      double f = G * (double)m[j] * inv * inv * inv;
      sx += f * dx;
      sy += f * dy;
      sz += f * dz;
    }
    ax[t] = sx;
    ay[t] = sy;
    az[t] = sz;
  }
}

/* src/cuda/force_direct.cu:109-117 */
void orc_pair_force(const float p1[3], const float p2[3], float m1, float m2, float G, float eps,
                    float out[3]) {
  (void)m1;
  float rx = p2[0] - p1[0], ry = p2[1] - p1[1], rz = p2[2] - p1[2];
  float dist2 = (rx * rx + ry * ry + rz * rz) + eps * eps;
  float inv_dist = 1.0f / sqrtf(dist2);
  float inv_dist3 = inv_dist * inv_dist * inv_dist;
  float f = G * m2 * inv_dist3;
  out[0] = rx * f;
  out[1] = ry * f;
  out[2] = rz * f;
}

/* src/cuda/integrator.cu:39-48 */
void orc_store_acc(int n, const float* ax, const float* ay, const float* az, float* ox, float* oy, float* oz) {
  memcpy(ox, ax, sizeof(float) * (size_t)n);
  memcpy(oy, ay, sizeof(float) * (size_t)n);
  memcpy(oz, az, sizeof(float) * (size_t)n);
}

/* src/cuda/integrator.cu:11-21: dt2_half = 0.5f*dt*dt; pos += vel*dt + acc*dt2_half */
void orc_drift(int n, float* px, float* py, float* pz, const float* vx, const float* vy, const float* vz,
               const float* ax, const float* ay, const float* az, float dt) {
  const float dt2_half = 0.5f * dt * dt;
  for (int i = 0; i < n; i++) {
    px[i] += vx[i] * dt + ax[i] * dt2_half;
    py[i] += vy[i] * dt + ay[i] * dt2_half;
    pz[i] += vz[i] * dt + az[i] * dt2_half;
  }
}

/* src/cuda/integrator.cu:24-36: dt_half = 0.5f*dt; vel += (acc_old + acc_new)*dt_half */
void orc_kick(int n, float* vx, float* vy, float* vz, const float* ox, const float* oy, const float* oz,
              const float* ax, const float* ay, const float* az, float dt) {
  const float dt_half = 0.5f * dt;
  for (int i = 0; i < n; i++) {
    vx[i] += (ox[i] + ax[i]) * dt_half;
    vy[i] += (oy[i] + ay[i]) * dt_half;
    vz[i] += (oz[i] + az[i]) * dt_half;
  }
}

/* src/cuda/integrator.cu:224-238.  arrays[] follows include/nbody/types.hpp:234-262:
 * 0-2 pos, 3-5 vel, 6-8 acc, 9-11 acc_old, 12 mass. */
void orc_step_direct(int n, float* const a[13], float dt, float G, float eps2) {
  orc_store_acc(n, a[6], a[7], a[8], a[9], a[10], a[11]);
  orc_drift(n, a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8], dt);
  orc_direct_f32_eps2(n, a[0], a[1], a[2], a[12], a[6], a[7], a[8], G, eps2, 0, n);
  orc_kick(n, a[3], a[4], a[5], a[9], a[10], a[11], a[6], a[7], a[8], dt);
}

void orc_run_direct(int n, float* const a[13], float dt, float G, float eps2, int steps) {
  for (int s = 0; s < steps; s++) orc_step_direct(n, a, dt, G, eps2);
}

/* src/cuda/integrator.cu:59-63: ke_i = 0.5*m*(vx^2+vy^2+vz^2); fp64 accumulation. */
double orc_kinetic_f64(int n, const float* vx, const float* vy, const float* vz, const float* m) {
  double ke = 0.0;
#pragma omp parallel for reduction(+ : ke) schedule(static)
  for (int i = 0; i < n; i++) {
    double v2 = (double)vx[i] * vx[i] + (double)vy[i] * vy[i] + (double)vz[i] * vz[i];
    ke += 0.5 * (double)m[i] * v2;
  }
  return ke;
}

/* Conserved-quantity moments in fp64 (checker for nb_diagnostics; the reference's tests compute the same sums on
 * the host, tests/test_integrator.cpp:51-84): out = {M, sum m x (3), sum m v (3), sum m x cross v (3), KE}. */
void orc_moments_f64(int n, const float* px, const float* py, const float* pz, const float* vx, const float* vy,
                     const float* vz, const float* m, double* out11) {
  for (int q = 0; q < 11; q++) out11[q] = 0.0;
  for (int i = 0; i < n; i++) {
    const double mi = m[i], x = px[i], y = py[i], z = pz[i], u = vx[i], v = vy[i], w = vz[i];
    out11[0] += mi;
    out11[1] += mi * x; out11[2] += mi * y; out11[3] += mi * z;
    out11[4] += mi * u; out11[5] += mi * v; out11[6] += mi * w;
    out11[7] += mi * (y * w - z * v); out11[8] += mi * (z * u - x * w); out11[9] += mi * (x * v - y * u);
    out11[10] += 0.5 * mi * (u * u + v * v + w * w);
  }
}

/* src/cuda/integrator.cu:90-103: pe -= G*m_i*m_j / sqrt(r^2 + eps^2) over j > i; fp64. */
double orc_potential_f64(int n, const float* px, const float* py, const float* pz, const float* m,
                         double G, double eps) {
  double pe = 0.0;
  const double e2 = eps * eps;
#pragma omp parallel for reduction(+ : pe) schedule(dynamic, 16)
  for (int i = 0; i < n; i++) {
    const double xi = px[i], yi = py[i], zi = pz[i], mi = m[i];
    double s = 0.0;
    for (int j = i + 1; j < n; j++) {
      double dx = (double)px[j] - xi, dy = (double)py[j] - yi, dz = (double)pz[j] - zi;
      s += (double)m[j] / sqrt(dx * dx + dy * dy + dz * dz + e2);
    }
    pe -= G * mi * s;
  }
  return pe;
}
