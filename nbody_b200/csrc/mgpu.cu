// mgpu.cu -- target-sharded direct N^2 across the GPUs of one node (no counterpart in the
// reference, which is single-GPU: SURVEY.md sec. 2.2).  Rank r owns targets
// [r*N/R, (r+1)*N/R) and every rank holds a replicated packed source array
// float4[R][padded_slice].  The exchange itself (all-gather of each rank's slice) is done by the
// caller on its own stream -- NCCL through torch.distributed in nbody_b200/dist.py -- between
// nb_mgpu_pack_local and nb_mgpu_forces_remote; the local-source pass needs no remote data and
// overlaps the gather.
#include <cstring>

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
                            (cudaStream_t)s, nullptr);
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
                            (cudaStream_t)s, nullptr);
}

int nb_mgpu_finish(nb_context* ctx, const nb_particles* d_local, float dt, int do_kick, nb_stream s) {
  if (int rc = mg_check(ctx, d_local)) return rc;
  const DirectCfg& cfg = direct_cfg(ctx->mg_n_total);
  return launch_reduce(ctx, cfg, d_local, ctx->mg_count, dt, do_kick != 0, (cudaStream_t)s);
}

// ---- peer exchange: no collective library in the step ----------------------------------------------------
// Each rank owns ONE cudaMalloc'ed exchange block {flags[2][8] | replicated sources, parity 0 | parity 1}
// and maps every peer's block with CUDA IPC.  A step is three launches, as on one GPU:
//   pack_push_kernel   store-acc + drift + pack, stored into all R replicated buffers over NVLink, then
//                      this rank's step flag raised on every peer;
//   direct_force_kernel<GATED>  all N sources in ring order starting with the local slice; the TMA
//                      producer thread waits for a peer's flag when it first reaches that peer's tiles;
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

static PeerState* peer_get(nb_context* ctx) {
  if (!ctx->peer) ctx->peer = new PeerState();
  return ctx->peer;
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
  NB_CUDA(cudaSetDevice(ctx->device));
  nb_mgpu_peer_close(ctx);
  PeerState* ps = peer_get(ctx);
  ps->block_bytes = kFlagBytes + 2 * peer_buf_floats4(ctx) * sizeof(float4);
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
  ps->peer_block[ctx->mg_rank] = ps->block;
  return NB_OK;
}

int nb_mgpu_peer_open(nb_context* ctx, int peer_rank, const unsigned char handle[64]) {
  if (!ctx || !ctx->peer || !ctx->peer->block || !handle || peer_rank < 0 || peer_rank >= ctx->mg_nranks) {
    set_error("mgpu peer open: bad arguments or nb_mgpu_peer_alloc not called");
    return NB_ERR_INVALID;
  }
  if (peer_rank == ctx->mg_rank) return NB_OK;
  NB_CUDA(cudaSetDevice(ctx->device));
  PeerState* ps = ctx->peer;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  NB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  ps->peer_block[peer_rank] = p;
  ps->opened[peer_rank] = true;
  return NB_OK;
}

int nb_mgpu_peer_close(nb_context* ctx) {
  if (!ctx || !ctx->peer) return NB_OK;
  PeerState* ps = ctx->peer;
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
  ctx->peer = nullptr;
  return NB_OK;
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
  PeerGate gate{};
  gate.flags = pb.flags[ctx->mg_rank];
  gate.flag_base = parity * NB_MAX_PEERS;
  gate.step = step;
  gate.ptiles = ptiles;
  gate.self = ctx->mg_rank;
  gate.timeout_flag = ps->d_timeout;
  ctx->n_pass = 1;
  ctx->pass[1].U = 0;
  // ring order from the local slice: [rank, R) then [0, rank)
  if (int rc = launch_direct_pass(ctx, cfg, src, tgt, n_itiles, ctx->mg_rank * ptiles,
                                  (ctx->mg_nranks - ctx->mg_rank) * ptiles, 0, ctx->mg_rank * ptiles, eps2, 0, st,
                                  ctx->mg_nranks > 1 ? &gate : nullptr))
    return rc;
  return launch_reduce(ctx, cfg, d_local, ctx->mg_count, dt, do_kick != 0, st);
}

int nb_mgpu_peer_step(nb_context* ctx, const nb_particles* d_local, float dt, float G, float eps2, int do_step,
                      nb_stream s) {
  if (int rc = nb_mgpu_peer_push(ctx, d_local, dt, G, do_step, s)) return rc;
  return nb_mgpu_peer_forces(ctx, d_local, dt, eps2, do_step, s);
}

}  // extern "C"
