// comm.cuh -- NCCL communicator used by the data-parallel learner (src/network scope of the north star).
// NCCL is resolved with dlopen at run time so that the single-GPU library has no link-time dependency and, in a
// torch process, shares the libnccl.so.2 torch already loaded.
#pragma once
#include "common.cuh"

namespace gbl {

struct Comm {
  void* nccl_comm = nullptr;
  int num_ranks = 1;
  int rank = 0;
  int device = 0;
};

enum class RedOp { kSum, kMax };

void comm_get_unique_id(uint8_t out[128]);
Comm* comm_create(const uint8_t id[128], int num_ranks, int rank, int device);
void comm_free(Comm* c);
void comm_allreduce_i64(Comm* c, long long* buf, size_t count, RedOp op, cudaStream_t s);
void comm_allreduce_f32(Comm* c, float* buf, size_t count, RedOp op, cudaStream_t s);
void comm_allreduce_f64(Comm* c, double* buf, size_t count, RedOp op, cudaStream_t s);

}  // namespace gbl
