// Tensor-parallel plumbing: NCCL (loaded at run time) behind three calls.  One process per GPU;
// the host program exchanges the 128-byte unique id with whatever it has (torch.distributed,
// MPI, a file) and hands it to every rank.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

namespace rb {

struct TpInfo {
    int rank = 0, world = 1;
    void * comm = nullptr;           // ncclComm_t
};

bool tp_available();                                               // libnccl found
bool tp_unique_id(void * out128);                                  // ncclGetUniqueId
bool tp_comm_init(TpInfo & tp, const void * id128);                // ncclCommInitRank (device = current)
void tp_comm_destroy(TpInfo & tp);
// sum of `n` floats over the ranks, result on every rank (send may equal recv)
cudaError_t tp_allreduce(const TpInfo & tp, const float * send, float * recv, size_t n, cudaStream_t st);
// every rank contributes `n` floats; recv holds world * n floats in rank order
cudaError_t tp_allgather(const TpInfo & tp, const float * send, float * recv, size_t n, cudaStream_t st);

} // namespace rb
