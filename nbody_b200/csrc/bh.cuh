// bh.cuh -- host-side interface of bh.cu shared with mgpu.cu (target-sharded Barnes-Hut).
#pragma once
#include "common.cuh"

namespace nb {

struct BhShard {
  int rank, nranks;
  float4* out[8];        // every rank's exchange buffer (parity already applied)
  unsigned* flags[8];    // every rank's flag block
  int flag_index;        // parity * 8 + rank
  unsigned step;
  unsigned* done_counter;
};

int bh_build(nb_context* ctx, const nb_particles* d, cudaStream_t st);
int bh_forces(nb_context* ctx, const nb_particles* d, float theta, float G, float eps, cudaStream_t st,
              const BhShard* shard);
int bh_shard_wait(nb_context* ctx, const unsigned* flags, int flag_base, unsigned step, int nranks, int self,
                  int* timeout_flag, cudaStream_t st);
int bh_shard_inverse(nb_context* ctx, cudaStream_t st);
int bh_shard_finish(nb_context* ctx, const float4* acc_sorted, const nb_particles* d, float dt, bool kick,
                    cudaStream_t st);

}  // namespace nb
