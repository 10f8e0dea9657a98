// mgpu.cu -- target-sharded direct N^2 across the GPUs of one node (no counterpart in the
// reference, which is single-GPU: SURVEY.md sec. 2.2).  Rank r owns targets
// [r*N/R, (r+1)*N/R) and every rank holds a replicated packed source array
// float4[R][padded_slice].  The exchange itself (all-gather of each rank's slice) is done by the
// caller on its own stream -- NCCL through torch.distributed in nbody_b200/dist.py -- between
// nb_mgpu_pack_local and nb_mgpu_forces_remote; the local-source pass needs no remote data and
// overlaps the gather.
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
  const int before = ctx->mg_rank * ptiles, after = (ctx->mg_nranks - 1 - ctx->mg_rank) * ptiles;
  ctx->n_pass = 2;
  return launch_direct_pass(ctx, cfg, src, tgt, n_itiles, 0, before, (ctx->mg_rank + 1) * ptiles, after, eps2, 1,
                            (cudaStream_t)s);
}

int nb_mgpu_finish(nb_context* ctx, const nb_particles* d_local, float dt, int do_kick, nb_stream s) {
  if (int rc = mg_check(ctx, d_local)) return rc;
  const DirectCfg& cfg = direct_cfg(ctx->mg_n_total);
  return launch_reduce(ctx, cfg, d_local, ctx->mg_count, dt, do_kick != 0, (cudaStream_t)s);
}

}  // extern "C"
