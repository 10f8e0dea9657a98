/*
 * nbody_b200.h -- C ABI of the B200-native force-evaluation + Velocity-Verlet engine.
 *
 * This is the drop-in boundary for LessUp/n-body's hot path: every entry point replaces one
 * reference interface (cited as file:line, paths relative to the reference checkout).  Plain C:
 * raw device/host pointers, sizes, scalars and an opaque stream handle (cudaStream_t passed as
 * void*; NULL = legacy default stream, which is what the reference uses everywhere).  No C++
 * types and no exceptions cross this boundary; include/nbody/ *.hpp wraps it in the reference's
 * class names and turns return codes back into the reference's exception types.
 *
 * Return value: 0 on success; > 0 is a cudaError_t; < 0 is one of NB_ERR_*.  nb_last_error()
 * gives the message of the calling thread's last failure.
 *
 * There is no CPU fallback: every compute entry point needs a CUDA device of compute capability
 * 10.x and fails with NB_ERR_NO_DEVICE otherwise.
 */
#ifndef NBODY_B200_H
#define NBODY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB_OK 0
#define NB_ERR_INVALID (-1)     /* bad argument (maps to nbody::ValidationException)                 */
#define NB_ERR_NO_DEVICE (-2)   /* no sm_100 device / library built without a usable kernel image     */
#define NB_ERR_RESOURCE (-3)    /* allocation failed (maps to nbody::ResourceException)               */
#define NB_ERR_GRID_TOO_LARGE (-4) /* spatial hash > 1e8 cells: src/cuda/force_spatial_hash.cu:252-254 (std::runtime_error) */
#define NB_ERR_NCCL (-5)        /* NCCL failure in the multi-GPU path                                  */
#define NB_ERR_INTERNAL (-6)
#define NB_ERR_STALE_GRAPH (-7) /* nb_graph_launch: scratch the graph points into was reallocated; capture again */

typedef struct nb_context nb_context; /* one per device; owns all scratch; NOT thread-safe
                                         (same contract as include/nbody/force_calculator.hpp:8-20).  The scratch
                                         is shared by every entry point: keep ONE stream's work in flight per
                                         context (nb_host_* run on the context's own stream and synchronise
                                         before returning); use a second context for concurrent streams. */
typedef void* nb_stream;              /* cudaStream_t */

/* Same layout as nbody::ParticleData (include/nbody/types.hpp:234-276): 13 float* then count,
 * 112 bytes, so a `ParticleData*` can be passed as `nb_particles*` unchanged.  The 13 arrays are
 * caller-owned, separately allocated, and never freed or reallocated by this library. */
typedef struct nb_particles {
  float *pos_x, *pos_y, *pos_z;
  float *vel_x, *vel_y, *vel_z;
  float *acc_x, *acc_y, *acc_z;
  float *acc_old_x, *acc_old_y, *acc_old_z;
  float* mass;
  size_t count;
} nb_particles;

/* ---- library / context ------------------------------------------------------------------- */
int nb_version(void);
const char* nb_last_error(void);
/* device query used by validateResourceRequirements (src/utils/error_handling.cpp:25-44) */
int nb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem, int* clock_khz);
int nb_context_create(int device, size_t max_bodies_hint, nb_context** out);
int nb_context_destroy(nb_context* ctx);
int nb_context_sm_count(const nb_context* ctx);
int nb_synchronize(nb_context* ctx, nb_stream stream);
/* number of this library's kernels launched through ctx since creation (bench.py gpu_launches) */
uint64_t nb_launch_count(const nb_context* ctx);

/* ---- streams and CUDA graphs for host code without CUDA headers.  A step at small N is launch-bound;
 * run one step eagerly, capture the next on a stream from nb_stream_create, replay with
 * nb_graph_launch (named as future work by the reference,
 * openspec/changes/evolve-nbody-platform/tasks.md:24-31).  Capturable: nb_direct_forces,
 * nb_step_direct, nb_bh_build, nb_bh_forces, nb_store_accelerations, nb_update_positions,
 * nb_update_velocities.  Everything that synchronises (nb_sh_build, energies, exports, stats,
 * nb_render_splats) is not. */
int nb_device_alloc(nb_context* ctx, size_t bytes, void** out);   /* raw device buffer (e.g. a framebuffer) */
int nb_device_free(nb_context* ctx, void* p);
int nb_device_download(nb_context* ctx, void* host_dst, const void* device_src, size_t bytes, nb_stream s); /* synchronises */
int nb_device_upload(nb_context* ctx, void* device_dst, const void* host_src, size_t bytes, nb_stream s);     /* synchronises */
typedef struct nb_graph nb_graph;
int nb_stream_create(nb_context* ctx, nb_stream* out);
int nb_stream_destroy(nb_context* ctx, nb_stream s);
int nb_graph_begin_capture(nb_context* ctx, nb_stream s);
int nb_graph_end_capture(nb_context* ctx, nb_stream s, nb_graph** out);
/* A captured step holds raw pointers into context-owned scratch (packed sources, partials, tree and sort
 * buffers).  Any later call on the context that needs MORE scratch (a renderer frame with more fragments, a
 * second system with more bodies) reallocates it; nb_graph_launch then refuses with NB_ERR_STALE_GRAPH instead
 * of replaying on freed memory, and nb_graph_is_stale says so beforehand: run one eager step and re-capture.
 * nb_graph_launch and nb_synchronize also report device-side failures flagged by earlier (replayed) work:
 * Barnes-Hut traversal stack overflow, sort look-back timeout, multi-GPU exchange timeout. */
int nb_graph_launch(nb_graph* g, nb_stream s);
int nb_graph_is_stale(const nb_graph* g);
uint64_t nb_scratch_generation(const nb_context* ctx);
int nb_graph_node_count(const nb_graph* g, size_t* nodes);
int nb_graph_destroy(nb_graph* g);

/* optional device-side timing of the dominant kernels (CUDA events on the launching stream around
 * every direct-force / BH-traversal / SH-force launch).  nb_profile_read synchronises, returns the
 * summed kernel milliseconds and launch count since the last read, and resets them. */
int nb_profile_enable(nb_context* ctx, int on);
int nb_profile_read(nb_context* ctx, double* kernel_ms, uint64_t* kernel_launches);
/* per-stage device timing (bench.py `paths`): with on != 0 every stage of a path (bbox, morton, sort, tree,
 * traversal, assign, ranges, ...) is bracketed by CUDA events on the launching stream.
 * nb_profile_stages_read synchronises, writes one line "name total_ms launches\n" per stage seen since the
 * last read (first-seen order) into buf, and resets.  Not capturable into a graph. */
int nb_profile_stages_enable(nb_context* ctx, int on);
int nb_profile_stages_read(nb_context* ctx, char* buf, size_t cap);
/* device bytes held by the library's own scratch buffers (packed sources, partials, tree, grid, sort and
 * renderer buffers) in this process; the caller's 13 particle arrays are not counted */
size_t nb_scratch_bytes(const nb_context* ctx);
/* FP32 pipe rate of this device, measured now: independent chains of scalar FFMA (packed = 0) or packed
 * fma.rn.f32x2 (packed = 1) for a few tens of milliseconds, CUDA events on stream s; TFLOP/s with 2 flop per
 * FMA lane.  bench.py quotes the direct-N^2 kernel against these next to the datasheet figure. */
int nb_probe_fp32(nb_context* ctx, int packed, double* tflops, double* ms, nb_stream s);

/* ---- ParticleDataManager (src/cuda/particle_init.cu:143-283) ------------------------------ */
int nb_particles_alloc_device(nb_particles* d, size_t count); /* :143-168, acc/acc_old zeroed   */
int nb_particles_free_device(nb_particles* d);                /* :170-198                        */
int nb_particles_upload(nb_context* ctx, nb_particles* d, const nb_particles* h, nb_stream s);   /* :257-269: pos, vel, acc, mass (acc_old not copied) */
int nb_particles_download(nb_context* ctx, nb_particles* h, const nb_particles* d, nb_stream s); /* :271-283 */

/* ---- direct N^2: DirectForceCalculator::computeForces -> launchDirectForceKernel
 *      (src/cuda/force_direct.cu:103-106, 88-98, 10-85).  Overwrites acc_x/y/z only. --------- */
int nb_direct_forces(nb_context* ctx, const nb_particles* d, float G, float eps2, nb_stream s);
/* nb_direct_forces with NO device allocation at call time.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same kernels, arguments and result as nb_direct_forces; the
 * scratch (chunk counters, packed sources, per-chunk partial sums) lives in a file-scope static __device__ array
 * of 17 MiB + 256 B that every GPU loading the library holds until exit.  Limit: d->count <=
 * NB_DIRECT_STATIC_MAX_BODIES, beyond it NB_ERR_INVALID with a message and nothing launched.  Calls no cudaMalloc,
 * cudaFree or allocating library routine (the first call per device resolves the array's device address with
 * cudaGetSymbolAddress and queries kernel occupancy; nothing is allocated -- make that first call outside stream
 * capture).  All calls share the array: calls that
 * overlap on the GPU race -- keep them on one stream, or order them.  Results: bit-identical to nb_direct_forces
 * for count <= 16384; above that the partial sums of a target tile are grouped into fewer chunks to fit the array,
 * so results agree to summation rounding (same pair terms). */
#define NB_DIRECT_STATIC_MAX_BODIES 65536
int nb_direct_forces_static(nb_context* ctx, const nb_particles* d, float G, float eps2, nb_stream s);

/* ---- Integrator (src/cuda/integrator.cu) --------------------------------------------------- */
int nb_store_accelerations(nb_context* ctx, const nb_particles* d, nb_stream s);   /* :248-250, :39-48 */
int nb_update_positions(nb_context* ctx, const nb_particles* d, float dt, nb_stream s);  /* :240-242, :11-21 */
int nb_update_velocities(nb_context* ctx, const nb_particles* d, float dt, nb_stream s); /* :244-246, :24-36 */
/* Integrator::integrate with a DirectForceCalculator, fused (:224-238): store+drift+pack, force,
 * reduce+kick.  Post-state identical to the 4-call sequence. */
int nb_step_direct(nb_context* ctx, const nb_particles* d, float dt, float G, float eps2, nb_stream s);
/* Integrator::compute{Kinetic,Potential}Energy (:252-289), fp64 accumulation; synchronises. */
int nb_kinetic_energy(nb_context* ctx, const nb_particles* d, double* ke, nb_stream s);
int nb_potential_energy(nb_context* ctx, const nb_particles* d, float G, float eps, double* pe, nb_stream s);
/* Conserved-quantity diagnostics in one call (no reference equivalent -- the reference's
 * tests compute these on the host, tests/test_integrator.cpp:51-162): fp64 sums over the device arrays.
 * with_potential != 0 adds the O(N^2) potential energy and the virial ratio 2 KE / |PE|.  Synchronises. */
typedef struct nb_diagnostics_t {
  double total_mass;
  double center_of_mass[3];
  double momentum[3];          /* sum m v */
  double angular_momentum[3];  /* sum m x cross v, about the origin */
  double kinetic;
  double potential;            /* 0 unless has_potential */
  double total_energy;         /* kinetic + potential */
  double virial_ratio;         /* 2 KE / |PE|, 0 unless has_potential */
  int has_potential;
} nb_diagnostics_t;
int nb_diagnostics(nb_context* ctx, const nb_particles* d, float G, float eps, int with_potential,
                   nb_diagnostics_t* out, nb_stream s);

/* ---- host-buffer entry points (what bench.py's e2e times): host SoA in, host SoA out, H2D and
 *      D2H inside the call.  Equivalent to ParticleSystem::setState -> update x steps -> getState
 *      (src/core/particle_system.cpp:240-292, 115-127, 213-238) without re-allocation. ---------- */
int nb_host_direct_forces(nb_context* ctx, const nb_particles* h, float G, float eps2);
int nb_host_step_direct(nb_context* ctx, nb_particles* h, float dt, float G, float eps2, int steps,
                        int recompute_initial_forces);

/* ---- Barnes-Hut: BarnesHutTree::build / computeForces (src/cuda/force_barnes_hut.cu:282-289,
 *      488-498), BarnesHutCalculator::computeForces (:526-532).  Tree rebuilt on the GPU. ------- */
int nb_bh_build(nb_context* ctx, const nb_particles* d, nb_stream s);
int nb_bh_forces(nb_context* ctx, const nb_particles* d, float theta, float G, float eps, nb_stream s);
/* exports for parity tests / getTree(): host buffers, may be NULL to skip.  keys/perm = sorted
 * Morton keys and sort permutation (n each); left/right/prefix = n-1 internal nodes; parent and
 * node_com (x,y,z,m) = 2n-1 nodes; bbox = {minx,miny,minz,maxx,maxy,maxz}. */
int nb_bh_export(nb_context* ctx, uint32_t* keys, int32_t* perm, int32_t* left, int32_t* right,
                 int32_t* parent, int32_t* prefix, float* node_com, float* bbox);
int nb_bh_stats(nb_context* ctx, uint64_t* interactions, int* node_count);
/* deepest per-warp traversal stack reached by the last nb_bh_forces, and its capacity (overflow is an error, never a drop) */
int nb_bh_max_stack(nb_context* ctx, int* deepest, int* capacity);

/* ---- spatial hash: SpatialHashGrid::build / computeForces (src/cuda/force_spatial_hash.cu:235-316),
 *      SpatialHashCalculator::computeForces (:370-377). ------------------------------------------ */
int nb_sh_build(nb_context* ctx, const nb_particles* d, float cell_size, nb_stream s);
int nb_sh_forces(nb_context* ctx, const nb_particles* d, float cutoff, float G, float eps, nb_stream s);
/* copyCellDataToHost (:318-335) + getGridDims/getTotalCells; NULL to skip. */
int nb_sh_export(nb_context* ctx, int32_t dims[3], float bbox_padded[6], int32_t* particle_cell,
                 int32_t* cell_start, int32_t* cell_end, int32_t* sorted_indices);
int nb_sh_total_cells(nb_context* ctx, int* total_cells);
/* body-body pairs inside the cutoff evaluated by the last nb_sh_forces (ordered pairs, self excluded).  A measurement
 * hook: counted on demand by a second sweep over the grid of the last nb_sh_build (same loop and test as the force
 * evaluation, no force arithmetic), so the timed force kernel carries no counters; synchronises the device. */
int nb_sh_stats(nb_context* ctx, uint64_t* pairs);

/* ---- stand-alone primitives (exported for tests and reuse) ---------------------------------- */
/* stable LSD radix sort of (u32 key, i32 value) pairs on device; replaces thrust::sort_by_key
 * (src/cuda/force_barnes_hut.cu:276-280).  end_bit <= 32, begin_bit = 0. */
int nb_sort_pairs(nb_context* ctx, uint32_t* d_keys, int32_t* d_vals, size_t n, int end_bit, nb_stream s);
/* number of read/write sweeps nb_sort_pairs makes over the pairs for keys of end_bit significant bits */
int nb_sort_passes(int end_bit);
/* bbox kernel (src/cuda/force_barnes_hut.cu:218-258); out = 6 host floats; synchronises. */
int nb_bounding_box(nb_context* ctx, const nb_particles* d, float out[6], nb_stream s);

/* ---- headless point-splat renderer: the reference's GL point-sprite
 *      pipeline -- vertex + fragment shaders src/render/renderer.cpp:16-77, blend state :110-113,
 *      draw call :184-224 -- as CUDA kernels writing a device RGBA8 framebuffer (row 0 = bottom row,
 *      as glReadPixels returns it).  Matrices are column-major like glm's.  No GL anywhere. ------- */
typedef struct nb_render_params {
  float view[16], proj[16];     /* Camera::getViewMatrix / getProjectionMatrix (src/render/camera.cpp:36-44) */
  float point_size;             /* uniform pointSize (default 2), divided by max(-viewPos.z, 0.1)            */
  float max_depth, max_velocity;/* uniforms maxDepth (100), maxVelocity (10)                                  */
  int color_mode;               /* 0 depth, 1 velocity, 2 density (include/nbody/types.hpp ColorMode)        */
  int width, height;
  float point_size_min, point_size_max; /* GL's implementation point-size range; 1 and 63 by default        */
  float clear_color[4];         /* 0.02, 0.02, 0.05, 1 (renderer.cpp:9-11, :188)                             */
} nb_render_params;
int nb_render_splats(nb_context* ctx, const nb_particles* d, const nb_render_params* p, uint8_t* d_rgba, nb_stream s);
int nb_render_stats(nb_context* ctx, uint64_t* fragments);
/* copyPositionsToVBOKernel / copyVelocitiesToVBOKernel (src/render/cuda_gl_interop.cpp:7-28): SoA ->
 * interleaved xyz device buffers (either may be NULL) */
int nb_interop_pack(nb_context* ctx, const nb_particles* d, float* d_positions_xyz, float* d_velocities_xyz, nb_stream s);

/* ---- multi-GPU direct N^2 (no counterpart in the reference; target-sharded) ---
 * The caller (one process per GPU) provides the exchange: `gathered_src` is this rank's replicated
 * packed source buffer.  Flow per step: nb_mgpu_pack_local -> (all-gather of the local slice, any
 * transport, on comm stream) -> nb_mgpu_forces_local (own slice as sources, overlaps the gather) ->
 * nb_mgpu_forces_remote (all other slices, after the gather) -> nb_mgpu_finish (reduce [+kick]). */
int nb_mgpu_layout(nb_context* ctx, size_t n_total, int rank, int nranks, size_t* slice_begin,
                   size_t* slice_count, size_t* padded_slice, size_t* packed_bytes_total);
int nb_mgpu_pack_local(nb_context* ctx, const nb_particles* d_local, float* gathered_src, float dt,
                       float G, int do_drift, nb_stream s);
int nb_mgpu_forces_local(nb_context* ctx, const float* gathered_src, float eps2, nb_stream s);
int nb_mgpu_forces_remote(nb_context* ctx, const float* gathered_src, float eps2, nb_stream s);
int nb_mgpu_finish(nb_context* ctx, const nb_particles* d_local, float dt, int do_kick, nb_stream s);

/* Peer exchange (default transport of nbody_b200/dist.py): the all-gather is fused into the pack
 * kernel, which stores every packed body into all R replicated source buffers over NVLink peer
 * mappings and then raises a per-peer step flag; the local-source force pass overlaps the peers'
 * pushes, a one-warp kernel waits for the flags, the remote-source pass follows in ring order.
 * No collective-library call in the step.  Setup (collective, once, after nb_mgpu_layout):
 * every rank calls nb_mgpu_peer_alloc, the 64-byte CUDA-IPC handles are exchanged by the caller
 * (any transport), every rank calls nb_mgpu_peer_open for each other rank, then a barrier.
 * nb_mgpu_peer_step must be called the same number of times on every rank (it is the collective);
 * do_step = 0: forces only (pack without drift, reduce without kick). */
int nb_mgpu_peer_alloc(nb_context* ctx, unsigned char ipc_handle_out[64]);
int nb_mgpu_peer_open(nb_context* ctx, int peer_rank, const unsigned char ipc_handle[64]);
int nb_mgpu_peer_step(nb_context* ctx, const nb_particles* d_local, float dt, float G, float eps2, int do_step,
                      nb_stream s);
/* Energy diagnostic of a target-sharded run ("same sharding, all-reduce of two doubles"; the
 * reference is single-GPU, src/cuda/integrator.cu:252-293).  *partial = sum over this rank's bodies i of
 * m_i * sum over ALL bodies j != i of G m_j / sqrt(r^2 + eps^2), fp64; the caller all-reduces it together
 * with nb_kinetic_energy of the local slice and takes PE = -sum / 2.  gathered_src: the replicated packed
 * source array (NCCL exchange), or NULL for the peer exchange block of the last step.  Synchronises. */
int nb_mgpu_potential_partial(nb_context* ctx, const nb_particles* d_local, const float* gathered_src, float eps,
                              double* partial, nb_stream s);
/* the two halves of nb_mgpu_peer_step (pack + push + flag | local pass, flag wait, remote pass, reduce [+ kick]) */
int nb_mgpu_peer_push(nb_context* ctx, const nb_particles* d_local, float dt, float G, int do_drift, nb_stream s);
int nb_mgpu_peer_forces(nb_context* ctx, const nb_particles* d_local, float dt, float eps2, int do_kick, nb_stream s);
/* same-process peers (several contexts in one process, e.g. one host thread driving all GPUs with
 * cudaDeviceEnablePeerAccess): map by raw device pointer instead of by IPC handle */
int nb_mgpu_peer_block(nb_context* ctx, void** block, size_t* bytes);
int nb_mgpu_peer_map(nb_context* ctx, int peer_rank, void* peer_block);
int nb_mgpu_peer_close(nb_context* ctx);

/* ---- target-sharded Barnes-Hut over replicated state (named as future
 * work by the reference, openspec/changes/evolve-nbody-platform/tasks.md:24-31).  Every rank holds
 * ALL bodies and rebuilds the identical tree; a rank traverses only the groups dealt to it and its
 * traversal kernel stores the accelerations (sorted order, float4/body) into every rank's exchange
 * buffer over NVLink peer mappings, then raises its step flag; nb_mgpu_bh_finish waits for all flags,
 * scatters to body order and kicks every body, so the replicas stay bit-identical and positions are
 * never exchanged.  Setup as for the direct peer exchange: alloc on every rank, swap the 64-byte
 * handles, open (or map by pointer inside one process), barrier.  nb_mgpu_bh_step is collective. */
int nb_mgpu_bh_alloc(nb_context* ctx, size_t n_total, int rank, int nranks, unsigned char ipc_handle_out[64]);
int nb_mgpu_bh_open(nb_context* ctx, int peer_rank, const unsigned char ipc_handle[64]);
int nb_mgpu_bh_block(nb_context* ctx, void** block, size_t* bytes);
int nb_mgpu_bh_map(nb_context* ctx, int peer_rank, void* peer_block);
int nb_mgpu_bh_step(nb_context* ctx, const nb_particles* d_all, float dt, float theta, float G, float eps,
                    int do_step, nb_stream s);
int nb_mgpu_bh_traverse(nb_context* ctx, const nb_particles* d_all, float dt, float theta, float G, float eps,
                        int do_drift, nb_stream s);
int nb_mgpu_bh_finish(nb_context* ctx, const nb_particles* d_all, float dt, int do_kick, nb_stream s);
int nb_mgpu_bh_close(nb_context* ctx);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
