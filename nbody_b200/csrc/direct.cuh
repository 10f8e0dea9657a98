// direct.cuh -- host-side interface of direct.cu shared with mgpu.cu.
#pragma once
#include "common.cuh"

namespace nb {

struct DirectCfg {
  int T, I, JT;  // threads per CTA, targets per thread, sources per shared-memory tile
  int IT() const { return T * I; }
};

constexpr int NB_MAX_PEERS = 8;

// multi-GPU peer exchange (mgpu.cu): every rank's replicated source buffer and flag block, as mapped
// into this process (own allocation + cudaIpc mappings of the peers')
struct PeerBufs {
  float4* buf[NB_MAX_PEERS];
  unsigned* flags[NB_MAX_PEERS];
  int nranks, self;
};
const DirectCfg& direct_cfg(size_t n_total_sources);
int launch_pack_push(nb_context* ctx, const nb_particles* d, const PeerBufs& pb, size_t slot_off, size_t n,
                     size_t n_pad, float G, float dt, bool drift, unsigned step, int flag_index, unsigned* done_counter,
                     cudaStream_t st);
int launch_pack(nb_context* ctx, const nb_particles* d, float4* src, size_t n, size_t n_pad, float G, float dt,
                bool drift, cudaStream_t st);
int launch_direct_pass(nb_context* ctx, const DirectCfg& cfg, const float4* src, const float4* tgt, int n_itiles,
                       int ja0, int na, int jb0, int nbt, float eps2, int pass_index, cudaStream_t st);
int launch_peer_wait(nb_context* ctx, const unsigned* flags, int flag_base, unsigned step, int nranks, int self,
                     int* timeout_flag, cudaStream_t st);
// integrate.cu: this rank's share of the potential-energy pair sum of a target-sharded run
int potential_partial(nb_context* ctx, const float4* src, size_t n_src, size_t tgt_off, size_t n_tgt,
                      const float* mass_local, float eps, double* partial, cudaStream_t st);
int launch_reduce(nb_context* ctx, const DirectCfg& cfg, const nb_particles* d, size_t n, float dt, bool kick,
                  cudaStream_t st);

}  // namespace nb
