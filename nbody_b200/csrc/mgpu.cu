// mgpu.cu -- target-sharded direct N^2 across the GPUs of one node (no counterpart in the
// reference, which is single-GPU: SURVEY.md sec. 2.2).  Rank r owns targets
// [r*N/R, (r+1)*N/R) and every rank holds a replicated packed source array
// float4[R][padded_slice].  The exchange itself (all-gather of each rank's slice) is done by the
// caller on its own stream -- NCCL through torch.distributed in nbody_b200/dist.py -- between
// nb_mgpu_pack_local and nb_mgpu_forces_remote; the local-source pass needs no remote data and
// overlaps the gather.
#include <cstring>
#include <new>

#include "bh.cuh"
#include "direct.cuh"

using namespace nb;

extern "C" {

int nb_mgpu_layout(nb_context* ctx, size_t n_total, int rank, int nranks, size_t* slice_begin,
                   size_t* slice_count, size_t* padded_slice, size_t* packed_bytes_total) {
  if (!ctx || nranks < 1 || rank < 0 || rank >= nranks || n_total == 0) {
    set_error("mgpu layout: bad arguments");
    return NB_ERR_INVALID;
  }
  const DirectCfg& cfg = direct_cfg(n_total);
  const size_t b = n_total * (size_t)rank / nranks, e = n_total * (size_t)(rank + 1) / nranks;
  const size_t max_cnt = (n_total + nranks - 1) / nranks;
  const size_t q = (size_t)(cfg.IT() > cfg.JT ? cfg.IT() : cfg.JT);  // IT divides JT or vice versa
  const size_t pad = (max_cnt + q - 1) / q * q;
  ctx->mg_n_total = n_total;
  ctx->mg_begin = b;
  ctx->mg_count = e - b;
  ctx->mg_padded = pad;
  ctx->mg_rank = rank;
  ctx->mg_nranks = nranks;
  if (slice_begin) *slice_begin = b;
  if (slice_count) *slice_count = e - b;
  if (padded_slice) *padded_slice = pad;
  if (packed_bytes_total) *packed_bytes_total = pad * nranks * sizeof(float4);
  return NB_OK;
}

static int mg_check(nb_context* ctx, const void* p) {
  if (!ctx || !p || ctx->mg_n_total == 0) {
    set_error("mgpu: nb_mgpu_layout not called or null buffer");
    return NB_ERR_INVALID;
  }
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice", __FILE__, __LINE__);
  return NB_OK;
}

int nb_mgpu_pack_local(nb_context* ctx, const nb_particles* d_local, float* gathered_src, float dt, float G,
                       int do_drift, nb_stream s) {
  if (int rc = mg_check(ctx, gathered_src)) return rc;
  if (!d_local || d_local->count != ctx->mg_count) {
    set_error("mgpu pack: local particle count %zu does not match the slice (%zu)", d_local ? d_local->count : 0,
              ctx->mg_count);
    return NB_ERR_INVALID;
  }
  float4* mine = reinterpret_cast<float4*>(gathered_src) + (size_t)ctx->mg_rank * ctx->mg_padded;
  return launch_pack(ctx, d_local, mine, ctx->mg_count, ctx->mg_padded, G, dt, do_drift != 0, (cudaStream_t)s);
}

int nb_mgpu_forces_local(nb_context* ctx, const float* gathered_src, float eps2, nb_stream s) {
  if (int rc = mg_check(ctx, gathered_src)) return rc;
  const DirectCfg& cfg = direct_cfg(ctx->mg_n_total);
  const float4* src = reinterpret_cast<const float4*>(gathered_src);
  const float4* tgt = src + (size_t)ctx->mg_rank * ctx->mg_padded;
  const int ptiles = (int)(ctx->mg_padded / cfg.JT);
  const int n_itiles = (int)((ctx->mg_count + cfg.IT() - 1) / cfg.IT());
  ctx->n_pass = 1;
  ctx->pass[1].U = 0;
  return launch_direct_pass(ctx, cfg, src, tgt, n_itiles, ctx->mg_rank * ptiles, ptiles, 0, 0, eps2, 0,
                            (cudaStream_t)s);
}

int nb_mgpu_forces_remote(nb_context* ctx, const float* gathered_src, float eps2, nb_stream s) {
  if (int rc = mg_check(ctx, gathered_src)) return rc;
  if (ctx->n_pass < 1) {
    set_error("mgpu: forces_remote called before forces_local");
    return NB_ERR_INVALID;
  }
  const DirectCfg& cfg = direct_cfg(ctx->mg_n_total);
  const float4* src = reinterpret_cast<const float4*>(gathered_src);
  const float4* tgt = src + (size_t)ctx->mg_rank * ctx->mg_padded;
  const int ptiles = (int)(ctx->mg_padded / cfg.JT);
  const int n_itiles = (int)((ctx->mg_count + cfg.IT() - 1) / cfg.IT());
  ctx->n_pass = 2;
  // the other R-1 slices in ring order: rank+1, ..., R-1, 0, ..., rank-1
  return launch_direct_pass(ctx, cfg, src, tgt, n_itiles, (ctx->mg_rank + 1) * ptiles,
                            (ctx->mg_nranks - 1 - ctx->mg_rank) * ptiles, 0, ctx->mg_rank * ptiles, eps2, 1,
                            (cudaStream_t)s);
}

int nb_mgpu_finish(nb_context* ctx, const nb_particles* d_local, float dt, int do_kick, nb_stream s) {
  if (int rc = mg_check(ctx, d_local)) return rc;
  const DirectCfg& cfg = direct_cfg(ctx->mg_n_total);
  return launch_reduce(ctx, cfg, d_local, ctx->mg_count, dt, do_kick != 0, (cudaStream_t)s);
}

// ---- peer exchange: no collective library in the step ----------------------------------------------------
// Each rank owns ONE cudaMalloc'ed exchange block {flags[2][8] | replicated sources, parity 0 | parity 1}
// and maps every peer's block with CUDA IPC.  Per step:
//   pack_push_kernel   store-acc + drift + pack, stored into all R replicated buffers over NVLink, then
//                      this rank's step flag raised on every peer;
//   direct_force_kernel  local-source pass (overlaps the peers' pushes), peer_flags_wait_kernel (one
//                      thread per peer, acquire at system scope), remote-source pass in ring order;
//   reduce_kernel      fixed-order sum of the stream-K partials + kick.
// Buffers alternate by step parity: a rank can only run one step ahead of its slowest peer (it needs
// that peer's slice of the current step), so the slice it overwrites was consumed two steps ago.
struct PeerState {
  void* block = nullptr;           // own exchange block
  size_t block_bytes = 0;
  void* peer_block[NB_MAX_PEERS] = {nullptr};
  bool opened[NB_MAX_PEERS] = {false};
  unsigned step = 0;               // evaluations done (same on every rank: the step is collective)
  unsigned* done_counter = nullptr;
  int* h_timeout = nullptr;        // host-mapped
  int* d_timeout = nullptr;
};
static const size_t kFlagBytes = 4096;

extern "C++" {
namespace nb {
// a peer's slice / accelerations did not arrive (peer_flags_wait_kernel gave up): polled by c_api.cu after
// synchronisation and before graph replays, so the failure is reported by the call that observes it
int peer_poll_error(nb_context* ctx) {
  for (PeerState* ps : {ctx->peer, ctx->peer_bh})
    if (ps && ps->h_timeout && *ps->h_timeout) {
      set_error("multi-GPU exchange: a peer's data did not arrive within 20 s (peer process gone?)");
      return NB_ERR_INTERNAL;
    }
  return NB_OK;
}
}  // namespace nb
}  // extern "C++"

static int peer_block_close(nb_context* ctx, PeerState*& slot) {
  if (!slot) return NB_OK;
  PeerState* ps = slot;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < NB_MAX_PEERS; r++) {
    if (ps->opened[r] && ps->peer_block[r]) cudaIpcCloseMemHandle(ps->peer_block[r]);
    ps->peer_block[r] = nullptr;
    ps->opened[r] = false;
  }
  if (ps->block) cudaFree(ps->block);
  if (ps->done_counter) cudaFree(ps->done_counter);
  if (ps->h_timeout) cudaFreeHost(ps->h_timeout);
  delete ps;
  slot = nullptr;
  return NB_OK;
}

// one zeroed cudaMalloc'ed block {flags | payload}, its IPC handle, a done counter and a host-mapped
// timeout flag
static int peer_block_alloc(nb_context* ctx, PeerState*& slot, size_t payload_bytes, int self,
                            unsigned char handle_out[64]) {
  NB_CUDA(cudaSetDevice(ctx->device));
  peer_block_close(ctx, slot);
  PeerState* ps = slot = new (std::nothrow) PeerState();
  if (!ps) {
    set_error("out of host memory");
    return NB_ERR_RESOURCE;
  }
  ps->block_bytes = kFlagBytes + payload_bytes;
  NB_CUDA(cudaMalloc(&ps->block, ps->block_bytes));
  NB_CUDA(cudaMemset(ps->block, 0, ps->block_bytes));
  NB_CUDA(cudaMalloc(&ps->done_counter, sizeof(unsigned)));
  NB_CUDA(cudaMemset(ps->done_counter, 0, sizeof(unsigned)));
  NB_CUDA(cudaHostAlloc(&ps->h_timeout, sizeof(int), cudaHostAllocMapped));
  *ps->h_timeout = 0;
  NB_CUDA(cudaHostGetDevicePointer(&ps->d_timeout, ps->h_timeout, 0));
  NB_CUDA(cudaDeviceSynchronize());
  ps->step = 0;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  cudaIpcMemHandle_t h;
  NB_CUDA(cudaIpcGetMemHandle(&h, ps->block));
  memcpy(handle_out, &h, 64);
  ps->peer_block[self] = ps->block;
  return NB_OK;
}

static int peer_block_open(nb_context* ctx, PeerState* ps, int self, int peer_rank, const unsigned char handle[64]) {
  if (peer_rank == self) return NB_OK;
  NB_CUDA(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  NB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  ps->peer_block[peer_rank] = p;
  ps->opened[peer_rank] = true;
  return NB_OK;
}

static size_t peer_buf_floats4(const nb_context* ctx) { return (size_t)ctx->mg_nranks * ctx->mg_padded; }

int nb_mgpu_peer_alloc(nb_context* ctx, unsigned char handle_out[64]) {
  if (!ctx || ctx->mg_n_total == 0 || !handle_out) {
    set_error("mgpu peer: nb_mgpu_layout not called or null handle");
    return NB_ERR_INVALID;
  }
  if (ctx->mg_nranks > NB_MAX_PEERS) {
    set_error("mgpu peer: at most %d ranks", NB_MAX_PEERS);
    return NB_ERR_INVALID;
  }
  return peer_block_alloc(ctx, ctx->peer, 2 * peer_buf_floats4(ctx) * sizeof(float4), ctx->mg_rank, handle_out);
}

int nb_mgpu_peer_open(nb_context* ctx, int peer_rank, const unsigned char handle[64]) {
  if (!ctx || !ctx->peer || !ctx->peer->block || !handle || peer_rank < 0 || peer_rank >= ctx->mg_nranks) {
    set_error("mgpu peer open: bad arguments or nb_mgpu_peer_alloc not called");
    return NB_ERR_INVALID;
  }
  return peer_block_open(ctx, ctx->peer, ctx->mg_rank, peer_rank, handle);
}

int nb_mgpu_peer_close(nb_context* ctx) {
  if (!ctx) return NB_OK;
  return peer_block_close(ctx, ctx->peer);
}

int nb_mgpu_peer_block(nb_context* ctx, void** block, size_t* bytes) {
  if (!ctx || !ctx->peer || !ctx->peer->block) {
    set_error("mgpu peer: nb_mgpu_peer_alloc not called");
    return NB_ERR_INVALID;
  }
  if (block) *block = ctx->peer->block;
  if (bytes) *bytes = ctx->peer->block_bytes;
  return NB_OK;
}

int nb_mgpu_peer_map(nb_context* ctx, int peer_rank, void* peer_block) {
  if (!ctx || !ctx->peer || !ctx->peer->block || !peer_block || peer_rank < 0 || peer_rank >= ctx->mg_nranks) {
    set_error("mgpu peer map: bad arguments or nb_mgpu_peer_alloc not called");
    return NB_ERR_INVALID;
  }
  if (peer_rank != ctx->mg_rank) ctx->peer->peer_block[peer_rank] = peer_block;
  return NB_OK;
}

static int peer_prepare(nb_context* ctx, const nb_particles* d_local, const char* what) {
  if (!ctx || !ctx->peer || !ctx->peer->block || !d_local) {
    set_error("mgpu peer %s: nb_mgpu_peer_alloc not called or null particles", what);
    return NB_ERR_INVALID;
  }
  if (d_local->count != ctx->mg_count) {
    set_error("mgpu peer %s: local particle count %zu does not match the slice (%zu)", what, d_local->count,
              ctx->mg_count);
    return NB_ERR_INVALID;
  }
  if (*ctx->peer->h_timeout) {
    set_error("mgpu peer %s: a peer's slice did not arrive within 20 s (peer process gone?)", what);
    return NB_ERR_INTERNAL;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  for (int r = 0; r < ctx->mg_nranks; r++)
    if (!ctx->peer->peer_block[r]) {
      set_error("mgpu peer %s: rank %d's exchange block is not mapped (nb_mgpu_peer_open / _map)", what, r);
      return NB_ERR_INVALID;
    }
  return NB_OK;
}

static void peer_bufs(const nb_context* ctx, int parity, PeerBufs& pb) {
  const PeerState* ps = ctx->peer;
  const size_t per_parity = peer_buf_floats4(ctx);
  pb.nranks = ctx->mg_nranks;
  pb.self = ctx->mg_rank;
  for (int r = 0; r < ctx->mg_nranks; r++) {
    unsigned char* base = static_cast<unsigned char*>(ps->peer_block[r]);
    pb.flags[r] = reinterpret_cast<unsigned*>(base);
    pb.buf[r] = reinterpret_cast<float4*>(base + kFlagBytes) + (size_t)parity * per_parity;
  }
}

int nb_mgpu_peer_push(nb_context* ctx, const nb_particles* d_local, float dt, float G, int do_drift, nb_stream s) {
  if (int rc = peer_prepare(ctx, d_local, "push")) return rc;
  PeerState* ps = ctx->peer;
  const unsigned step = ++ps->step;
  const int parity = (int)(step & 1u);
  PeerBufs pb{};
  peer_bufs(ctx, parity, pb);
  return launch_pack_push(ctx, d_local, pb, (size_t)ctx->mg_rank * ctx->mg_padded, ctx->mg_count, ctx->mg_padded, G, dt,
                          do_drift != 0, step, parity * NB_MAX_PEERS + ctx->mg_rank, ps->done_counter,
                          (cudaStream_t)s);
}

int nb_mgpu_peer_forces(nb_context* ctx, const nb_particles* d_local, float dt, float eps2, int do_kick, nb_stream s) {
  if (int rc = peer_prepare(ctx, d_local, "forces")) return rc;
  PeerState* ps = ctx->peer;
  if (ps->step == 0) {
    set_error("mgpu peer forces: nb_mgpu_peer_push has not run");
    return NB_ERR_INVALID;
  }
  cudaStream_t st = (cudaStream_t)s;
  const DirectCfg& cfg = direct_cfg(ctx->mg_n_total);
  const unsigned step = ps->step;
  const int parity = (int)(step & 1u);
  PeerBufs pb{};
  peer_bufs(ctx, parity, pb);
  const float4* src = pb.buf[ctx->mg_rank];
  const float4* tgt = src + (size_t)ctx->mg_rank * ctx->mg_padded;
  const int ptiles = (int)(ctx->mg_padded / cfg.JT);
  const int n_itiles = (int)((ctx->mg_count + cfg.IT() - 1) / cfg.IT());
  // local-source pass: needs nothing from the peers and overlaps their pushes
  ctx->n_pass = 1;
  ctx->pass[1].U = 0;
  if (int rc = launch_direct_pass(ctx, cfg, src, tgt, n_itiles, ctx->mg_rank * ptiles, ptiles, 0, 0, eps2, 0, st))
    return rc;
  if (ctx->mg_nranks > 1) {
    if (int rc = launch_peer_wait(ctx, pb.flags[ctx->mg_rank], parity * NB_MAX_PEERS, step, ctx->mg_nranks,
                                  ctx->mg_rank, ps->d_timeout, st))
      return rc;
    // the other R-1 slices in ring order: rank+1, ..., R-1, 0, ..., rank-1
    ctx->n_pass = 2;
    if (int rc = launch_direct_pass(ctx, cfg, src, tgt, n_itiles, (ctx->mg_rank + 1) * ptiles,
                                    (ctx->mg_nranks - 1 - ctx->mg_rank) * ptiles, 0, ctx->mg_rank * ptiles, eps2, 1, st))
      return rc;
  }
  return launch_reduce(ctx, cfg, d_local, ctx->mg_count, dt, do_kick != 0, st);
}

int nb_mgpu_peer_step(nb_context* ctx, const nb_particles* d_local, float dt, float G, float eps2, int do_step,
                      nb_stream s) {
  if (int rc = nb_mgpu_peer_push(ctx, d_local, dt, G, do_step, s)) return rc;
  return nb_mgpu_peer_forces(ctx, d_local, dt, eps2, do_step, s);
}

// ---- energy of a target-sharded run ---------------------------------------------------
// This rank's share of the pair sum; the caller all-reduces {KE_local, partial} and takes PE = -partial_sum / 2.
// gathered_src: the replicated packed source array of the NCCL path, or null for the peer exchange block of the
// step that ran last.  Synchronises.
int nb_mgpu_potential_partial(nb_context* ctx, const nb_particles* d_local, const float* gathered_src, float eps,
                              double* partial, nb_stream s) {
  if (!ctx || !d_local || !partial || ctx->mg_n_total == 0) {
    set_error("mgpu potential: nb_mgpu_layout not called or null argument");
    return NB_ERR_INVALID;
  }
  if (d_local->count != ctx->mg_count || (d_local->count && !d_local->mass)) {
    set_error("mgpu potential: local particle count %zu does not match the slice (%zu)", d_local->count, ctx->mg_count);
    return NB_ERR_INVALID;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  const float4* src = reinterpret_cast<const float4*>(gathered_src);
  if (!src) {
    if (int rc = peer_prepare(ctx, d_local, "potential")) return rc;
    if (ctx->peer->step == 0) {
      set_error("mgpu potential: no step or force evaluation has filled the exchange block yet");
      return NB_ERR_INVALID;
    }
    PeerBufs pb{};
    peer_bufs(ctx, (int)(ctx->peer->step & 1u), pb);
    src = pb.buf[ctx->mg_rank];
  }
  return potential_partial(ctx, src, (size_t)ctx->mg_nranks * ctx->mg_padded, (size_t)ctx->mg_rank * ctx->mg_padded,
                           ctx->mg_count, d_local->mass, eps, partial, (cudaStream_t)s);
}

// ---- target-sharded Barnes-Hut over replicated state --------------------------
// Every rank holds ALL bodies.  Per step every rank drifts all bodies and rebuilds the identical tree
// (deterministic kernels on identical inputs), traverses only the groups dealt to it, and stores the
// accelerations -- sorted order, float4 per body -- into every rank's exchange buffer from the
// traversal kernel's epilogue (NVLink peer stores, no collective).  After the per-peer flags are up,
// every rank scatters the complete acceleration array back to body order and kicks all bodies, so
// the replicas stay bit-identical without ever exchanging positions.  Buffers alternate by parity
// as in the direct path.
static int bhs_check(nb_context* ctx, const nb_particles* d, const char* what) {
  if (!ctx || !ctx->peer_bh || !ctx->peer_bh->block || !d) {
    set_error("mgpu bh %s: nb_mgpu_bh_alloc not called or null particles", what);
    return NB_ERR_INVALID;
  }
  if (d->count != ctx->bhs_n) {
    set_error("mgpu bh %s: %zu bodies, exchange block sized for %zu (every rank holds all bodies)", what, d->count,
              ctx->bhs_n);
    return NB_ERR_INVALID;
  }
  if (*ctx->peer_bh->h_timeout) {
    set_error("mgpu bh %s: a peer's accelerations did not arrive within 20 s (peer process gone?)", what);
    return NB_ERR_INTERNAL;
  }
  NB_CUDA(cudaSetDevice(ctx->device));
  for (int r = 0; r < ctx->bhs_nranks; r++)
    if (!ctx->peer_bh->peer_block[r]) {
      set_error("mgpu bh %s: rank %d's exchange block is not mapped (nb_mgpu_bh_open / _map)", what, r);
      return NB_ERR_INVALID;
    }
  return NB_OK;
}

int nb_mgpu_bh_alloc(nb_context* ctx, size_t n_total, int rank, int nranks, unsigned char handle_out[64]) {
  if (!ctx || n_total == 0 || nranks < 1 || nranks > NB_MAX_PEERS || rank < 0 || rank >= nranks || !handle_out) {
    set_error("mgpu bh alloc: bad arguments (at most %d ranks)", NB_MAX_PEERS);
    return NB_ERR_INVALID;
  }
  ctx->bhs_n = n_total;
  ctx->bhs_rank = rank;
  ctx->bhs_nranks = nranks;
  return peer_block_alloc(ctx, ctx->peer_bh, 2 * n_total * sizeof(float4), rank, handle_out);
}

int nb_mgpu_bh_open(nb_context* ctx, int peer_rank, const unsigned char handle[64]) {
  if (!ctx || !ctx->peer_bh || !handle || peer_rank < 0 || peer_rank >= ctx->bhs_nranks) {
    set_error("mgpu bh open: bad arguments or nb_mgpu_bh_alloc not called");
    return NB_ERR_INVALID;
  }
  return peer_block_open(ctx, ctx->peer_bh, ctx->bhs_rank, peer_rank, handle);
}

int nb_mgpu_bh_block(nb_context* ctx, void** block, size_t* bytes) {
  if (!ctx || !ctx->peer_bh || !ctx->peer_bh->block) {
    set_error("mgpu bh: nb_mgpu_bh_alloc not called");
    return NB_ERR_INVALID;
  }
  if (block) *block = ctx->peer_bh->block;
  if (bytes) *bytes = ctx->peer_bh->block_bytes;
  return NB_OK;
}

int nb_mgpu_bh_map(nb_context* ctx, int peer_rank, void* peer_block) {
  if (!ctx || !ctx->peer_bh || !peer_block || peer_rank < 0 || peer_rank >= ctx->bhs_nranks) {
    set_error("mgpu bh map: bad arguments or nb_mgpu_bh_alloc not called");
    return NB_ERR_INVALID;
  }
  if (peer_rank != ctx->bhs_rank) ctx->peer_bh->peer_block[peer_rank] = peer_block;
  return NB_OK;
}

int nb_mgpu_bh_close(nb_context* ctx) {
  if (!ctx) return NB_OK;
  return peer_block_close(ctx, ctx->peer_bh);
}

// first half: [store-acc + drift,] tree build, this rank's share of the traversal pushed to every rank
int nb_mgpu_bh_traverse(nb_context* ctx, const nb_particles* d_all, float dt, float theta, float G, float eps,
                        int do_drift, nb_stream s) {
  if (int rc = bhs_check(ctx, d_all, "traverse")) return rc;
  cudaStream_t st = (cudaStream_t)s;
  PeerState* ps = ctx->peer_bh;
  if (do_drift) {
    if (int rc = nb_store_accelerations(ctx, d_all, s)) return rc;
    if (int rc = nb_update_positions(ctx, d_all, dt, s)) return rc;
  }
  if (int rc = bh_build(ctx, d_all, st)) return rc;
  const unsigned step = ++ps->step;
  const int parity = (int)(step & 1u);
  BhShard sh{};
  sh.rank = ctx->bhs_rank;
  sh.nranks = ctx->bhs_nranks;
  for (int r = 0; r < ctx->bhs_nranks; r++) {
    unsigned char* base = static_cast<unsigned char*>(ps->peer_block[r]);
    sh.flags[r] = reinterpret_cast<unsigned*>(base);
    sh.out[r] = reinterpret_cast<float4*>(base + kFlagBytes) + (size_t)parity * ctx->bhs_n;
  }
  sh.flag_index = parity * NB_MAX_PEERS + ctx->bhs_rank;
  sh.step = step;
  sh.done_counter = ps->done_counter;
  if (int rc = bh_forces(ctx, d_all, theta, G, eps, st, &sh)) return rc;
  return bh_shard_inverse(ctx, st);
}

// second half: wait for every peer's share, scatter to body order [+ kick all bodies]
int nb_mgpu_bh_finish(nb_context* ctx, const nb_particles* d_all, float dt, int do_kick, nb_stream s) {
  if (int rc = bhs_check(ctx, d_all, "finish")) return rc;
  cudaStream_t st = (cudaStream_t)s;
  PeerState* ps = ctx->peer_bh;
  if (ps->step == 0) {
    set_error("mgpu bh finish: nb_mgpu_bh_traverse has not run");
    return NB_ERR_INVALID;
  }
  const int parity = (int)(ps->step & 1u);
  unsigned char* base = static_cast<unsigned char*>(ps->block);
  if (ctx->bhs_nranks > 1)
    if (int rc = bh_shard_wait(ctx, reinterpret_cast<const unsigned*>(base), parity * NB_MAX_PEERS, ps->step,
                               ctx->bhs_nranks, ctx->bhs_rank, ps->d_timeout, st))
      return rc;
  const float4* acc = reinterpret_cast<const float4*>(base + kFlagBytes) + (size_t)parity * ctx->bhs_n;
  return bh_shard_finish(ctx, acc, d_all, dt, do_kick != 0, st);
}

int nb_mgpu_bh_step(nb_context* ctx, const nb_particles* d_all, float dt, float theta, float G, float eps, int do_step,
                    nb_stream s) {
  if (int rc = nb_mgpu_bh_traverse(ctx, d_all, dt, theta, G, eps, do_step, s)) return rc;
  return nb_mgpu_bh_finish(ctx, d_all, dt, do_step, s);
}

}  // extern "C"
