// Tensor-parallel plumbing: NCCL (loaded at run time) behind three calls.  One process per GPU;
// the host program exchanges the 128-byte unique id with whatever it has (torch.distributed,
// MPI, a file) and hands it to every rank.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace rb {

constexpr int TP_MAX_WORLD = 8;
constexpr int TP_CTAS = 8;                   // CTAs of the one-shot collective kernel (each owns a slice of the vector)
constexpr size_t TP_NMAX = 65536;            // floats per rank slot of the peer-memory inbox

// Peer-memory region of one rank (cudaMalloc + CUDA IPC), written by the other ranks over NVLink:
//   flags [world][TP_CTAS] u32   epoch of the last vector rank r / CTA c has delivered
//   epoch [TP_CTAS] u32          local call counter per CTA
//   inbox [2][world][TP_NMAX] f32  two generations (epoch parity) of one slot per source rank
struct TpPeer {
    void *     base = nullptr;       // local region
    uint8_t *  peer[TP_MAX_WORLD] = {};    // every rank's region mapped into this process (peer[rank] == base)
    bool       connected = false;
};
inline size_t tp_region_flags_bytes() { return 4096; }
inline size_t tp_region_bytes(int world) { return tp_region_flags_bytes() + (size_t) 2 * world * TP_NMAX * sizeof(float); }

struct TpInfo {
    int rank = 0, world = 1;
    void * comm = nullptr;           // ncclComm_t
    TpPeer p2p;
};

bool tp_available();                                               // libnccl found
bool tp_unique_id(void * out128);                                  // ncclGetUniqueId
bool tp_comm_init(TpInfo & tp, const void * id128);                // ncclCommInitRank (device = current)
void tp_comm_destroy(TpInfo & tp);
// sum of `n` floats over the ranks, result on every rank (send may equal recv)
cudaError_t tp_allreduce(const TpInfo & tp, const float * send, float * recv, size_t n, cudaStream_t st);
// every rank contributes `n` floats; recv holds world * n floats in rank order
cudaError_t tp_allgather(const TpInfo & tp, const float * send, float * recv, size_t n, cudaStream_t st);

// ---- one-shot collectives over peer memory (tp_p2p.cu) -------------------------------------
// Allocates the local region and returns its 64-byte CUDA IPC handle; the host exchanges the
// handles of all ranks and calls tp_p2p_connect.  Afterwards tp_allreduce / tp_allgather of up
// to TP_NMAX floats run as ONE kernel that stores this rank's vector into every peer's inbox,
// raises a flag, waits for the peers' flags and reduces the slots in rank order (so every rank
// obtains the same bits); larger vectors keep using NCCL.
bool tp_p2p_alloc(TpInfo & tp, void * handle_out_64);
bool tp_p2p_connect(TpInfo & tp, const void * handles /* world x 64 bytes, rank order */);
void tp_p2p_destroy(TpInfo & tp);
cudaError_t tp_p2p_collective(const TpInfo & tp, const float * send, float * recv, size_t n, bool gather, cudaStream_t st);

} // namespace rb
