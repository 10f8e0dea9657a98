/*
 * oracle.h -- CPU restatement of LessUp/n-body's force-evaluation + Velocity-Verlet path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under nbody_b200/ (the product) may include, link
 * or call this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, as the checker or as the reported CPU baseline.
 *
 * Every function cites the reference file:line it follows (paths relative to the
 * reference checkout).  Built with -ffp-contract=off so fp32 arithmetic is the same
 * sequence of IEEE operations the reference source spells out.
 *
 * Parity pinning (PINNED): the reference's own tests hold no golden vectors for this path, only
 * analytic anchors (tests/test_force_calculation.cpp:13-60, tests/test_integrator.cpp:15-84,
 * tests/test_spatial_hash.cpp:38-51); tests/test_oracle_golden.py checks the oracle against all of
 * them, against tests/golden/ref_host_helpers.json (outputs of the reference's own host-callable
 * functions) and against tests/golden/ref_gpu_n4096.npz (outputs of the reference's own CUDA kernels,
 * compiled unmodified by oracle/build_ref.sh and run on a B200 by tests/golden/make_golden.py).  Where
 * oracle/_ref is present, tests/test_gpu_vs_reference.py also runs the reference kernels live on the
 * GPU box on the same inputs.  Only oracle/render.c is unpinned against a GL implementation
 * (graphics: unavailable on this box); it is pinned by the shader source alone.
 */
#ifndef NBODY_ORACLE_H
#define NBODY_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- direct N^2 ------------------------------------------------------------------ */
/* fp32 "as the reference CPU loop": examples/example_force_methods.cpp:44-66
 * (j ascending, j==i skipped, 1/sqrt, single fp32 accumulator).  eps is the softening
 * LENGTH (the loop squares it).  Targets [i0, i1). OpenMP over targets. */
void orc_direct_f32(int n, const float* px, const float* py, const float* pz, const float* m,
                    float* ax, float* ay, float* az, float G, float eps, int i0, int i1);
/* same pair math with eps2 passed directly as the kernel does, src/cuda/force_direct.cu:55-75 */
void orc_direct_f32_eps2(int n, const float* px, const float* py, const float* pz, const float* m,
                         float* ax, float* ay, float* az, float G, float eps2, int i0, int i1);
/* fp64 truth from the same fp32 inputs; targets given by an index list (idx==NULL: 0..nt-1). */
void orc_direct_f64(int n, const float* px, const float* py, const float* pz, const float* m,
                    double* ax, double* ay, double* az, double G, double eps2, const int* idx, int nt);
/* pairwise anchor: src/cuda/force_direct.cu:109-117 (force on p1 from p2; m1 unused there too) */
void orc_pair_force(const float p1[3], const float p2[3], float m1, float m2, float G, float eps,
                    float out[3]);

/* ---- Velocity Verlet --------------------------------------------------------------- */
/* src/cuda/integrator.cu:39-48 */
void orc_store_acc(int n, const float* ax, const float* ay, const float* az, float* ox, float* oy, float* oz);
/* src/cuda/integrator.cu:11-21 */
void orc_drift(int n, float* px, float* py, float* pz, const float* vx, const float* vy, const float* vz,
               const float* ax, const float* ay, const float* az, float dt);
/* src/cuda/integrator.cu:24-36 */
void orc_kick(int n, float* vx, float* vy, float* vz, const float* ox, const float* oy, const float* oz,
              const float* ax, const float* ay, const float* az, float dt);
/* src/cuda/integrator.cu:224-238 with the direct calculator; arrays in ParticleData order. */
void orc_step_direct(int n, float* const arrays[13], float dt, float G, float eps2);
/* run `steps` steps; convenience for the CPU baseline (config 1). */
void orc_run_direct(int n, float* const arrays[13], float dt, float G, float eps2, int steps);

/* ---- energy ------------------------------------------------------------------------ */
/* src/cuda/integrator.cu:59-63 (KE) and :90-103 (PE), accumulated in fp64 */
double orc_kinetic_f64(int n, const float* vx, const float* vy, const float* vz, const float* m);
void orc_moments_f64(int n, const float* px, const float* py, const float* pz, const float* vx, const float* vy,
                     const float* vz, const float* m, double* out11);
double orc_potential_f64(int n, const float* px, const float* py, const float* pz, const float* m,
                         double G, double eps);

/* ---- bounding box / Morton / sort --------------------------------------------------- */
/* src/cuda/force_barnes_hut.cu:66-110: exact min/max. out = {minx,miny,minz,maxx,maxy,maxz} */
void orc_bbox(int n, const float* px, const float* py, const float* pz, float out[6]);
/* src/cuda/force_barnes_hut.cu:23-38 */
uint32_t orc_expand_bits(uint32_t v);
uint32_t orc_morton_code(float x, float y, float z);
/* src/cuda/force_barnes_hut.cu:260-274 + :113-127 (host +0.001f, 1.0f/size; device (p-min)*inv) */
void orc_morton_keys(int n, const float* px, const float* py, const float* pz, const float bbox[6],
                     uint32_t* keys, int32_t* indices);
/* src/cuda/force_barnes_hut.cu:276-280: ascending by key, stable (ties keep ascending index) */
void orc_sort_pairs(int n, uint32_t* keys, int32_t* vals);

/* ---- Barnes-Hut ---------------------------------------------------------------------- */
/* Octree node, same field order/size (76 B) as include/nbody/barnes_hut_tree.hpp:9-30 */
typedef struct {
  float center[3];
  float half_size;
  float com[3];
  float total_mass;
  int32_t children[8];
  int32_t particle_index;
  uint8_t is_leaf;
  uint8_t pad_[3];
  int32_t particle_count;
} orc_octnode;
/* src/cuda/force_barnes_hut.cu:335-431 + :438-486.  as_written=1: verbatim (never splits an
 * occupied leaf, bodies dropped); as_written=0: intended algorithm (occupied leaf is split,
 * depth cap 20 keeps the reference's "force insert" branch).  Returns node count, or -1 if
 * max_nodes is too small.  sorted_idx = Morton-sorted body order (insertion order). */
int orc_octree_build(int n, const float* px, const float* py, const float* pz, const float* m,
                     const int32_t* sorted_idx, const float bbox[6], int as_written,
                     orc_octnode* nodes, int max_nodes, int* dropped);
/* src/cuda/force_barnes_hut.cu:130-201 (per-body DFS, 256-entry stack, same child push order) */
void orc_octree_forces(int n, const float* px, const float* py, const float* pz, const orc_octnode* nodes,
                       int num_nodes, float theta, float G, float eps, float* ax, float* ay, float* az);

/* Karras (2012) binary radix tree over the sorted keys; ties broken by position as in the
 * paper (delta = 32 + clz(i^j) for equal keys).  Internal nodes 0..n-2, leaves n..2n-1 are
 * encoded as child index (n-1)+k for sorted position k.  prefix[i] = common-prefix length of
 * node i capped at 30.  parent[] has 2n-1 entries (internal then leaf), root parent = -1. */
void orc_lbvh_build(int n, const uint32_t* sorted_keys, int32_t* left, int32_t* right, int32_t* parent,
                    int32_t* prefix, int32_t* range_first, int32_t* range_last);
/* bottom-up mass / centre of mass for the LBVH; node_com = 4 floats (x,y,z,m) per node, 2n-1 nodes;
 * sorted body k is node (n-1)+k. */
void orc_lbvh_com(int n, const int32_t* left, const int32_t* right, const float* sx, const float* sy,
                  const float* sz, const float* sm, float* node_com);
/* cell side used by the opening test for a node with common-prefix length L */
float orc_lbvh_cell_size(int prefix_len, const float bbox[6]);
/* warp-cooperative traversal restated sequentially (see lbvh.c): a traversal group = a maximal
 * subtree with at most `group` (32 or 64) leaves; it shares a node stack popped 32 at a time; a node is used iff it is a
 * leaf or size^2 < theta^2 (dmin^2 + eps^2), dmin = distance from its COM to the group's bounding
 * box; accepted nodes are evaluated 32 at a time in list order (monopole,
 * src/cuda/force_barnes_hut.cu:175-181).  Outputs in SORTED order. *interactions = node-body
 * force evaluations. */
void orc_set_hybrid_slack(int from);
void orc_set_leaf_bucket(int b);   /* experiment knob of the walk, see lbvh.c */
uint64_t orc_get_walk_pops(void);
uint64_t orc_get_walk_rounds(void);
void orc_set_walk_pairs(int on);  /* 1 (default): restate bh_force_pairs_kernel; 0: bh_force_kernel (NB_BH_WALK=nodes) */
void orc_lbvh_forces_ex(int n, const int32_t* left, const int32_t* right, const int32_t* prefix,
                        const int32_t* range_first, const int32_t* range_last, const float* node_com,
                        const float* sx, const float* sy, const float* sz, const float bbox[6], float theta, float G,
                        float eps, int group, int super, int mode, float alpha, float* ax, float* ay, float* az,
                        uint64_t* interactions, uint64_t* stats);
void orc_lbvh_forces(int n, const int32_t* left, const int32_t* right, const int32_t* prefix,
                     const int32_t* range_first, const int32_t* range_last, const float* node_com, const float* sx, const float* sy, const float* sz,
                     const float bbox[6], float theta, float G, float eps, int group, int super, float* ax,
                     float* ay, float* az, uint64_t* interactions);

/* ---- spatial hash -------------------------------------------------------------------- */
/* src/cuda/force_spatial_hash.cu:14-17 */
void orc_cell_index(float x, float y, float z, float cell_size, int32_t out[3]);
/* src/cuda/force_spatial_hash.cu:226-246: padded bbox (in/out: raw bbox -> padded) and dims */
void orc_grid_dims(float bbox[6], float cell_size, int32_t dims[3]);
/* src/cuda/force_spatial_hash.cu:28-49 with the padded bbox */
void orc_assign_cells(int n, const float* px, const float* py, const float* pz, const float bbox_padded[6],
                      float cell_size, const int32_t dims[3], int32_t* particle_cell);
/* :52-80 + :285-288: counts -> exclusive scan -> end; sorted_indices in ascending body index inside
 * each cell (the reference's order inside a cell is racy; the set per cell is the contract). */
void orc_cell_ranges(int n, const int32_t* particle_cell, int total_cells, int32_t* cell_start,
                     int32_t* cell_end, int32_t* sorted_indices);
/* :83-152 */
void orc_spatial_forces(int n, const float* px, const float* py, const float* pz, const float* m,
                        const int32_t* cell_start, const int32_t* cell_end, const int32_t* sorted_indices,
                        const int32_t dims[3], float cell_size, const float bbox_padded[6], float cutoff,
                        float G, float eps, float* ax, float* ay, float* az, uint64_t* pairs);

/* ---- headless point-splat renderer (render.c) ------------------------------------------ */
/* src/render/renderer.cpp:16-77, :110-113, :184-224; returns the number of fragments generated */
uint64_t orc_render_splats(int n, const float* px, const float* py, const float* pz, const float* vx, const float* vy,
                           const float* vz, const float view[16], const float proj[16], float point_size,
                           float max_depth, float max_velocity, int color_mode, int width, int height,
                           float size_min, float size_max, const float clear[4], uint8_t* rgba);

int orc_num_threads(void);
void orc_set_num_threads(int t); /* overrides OMP_NUM_THREADS (torchrun sets it to 1) */

#ifdef __cplusplus
}
#endif
#endif
