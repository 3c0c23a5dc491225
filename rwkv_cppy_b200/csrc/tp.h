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
    unsigned int * h_err = nullptr;  // host-mapped error word (cudaHostAlloc): a kernel that gave up waiting for a peer sets it,
    unsigned int * d_err = nullptr;  // the API reads and clears it after every synchronised call and fails the call
};
inline size_t tp_region_flags_bytes() { return 4096; }
// Low-latency inbox of the FUSED all-reduce (kernels.h TpFused; producer = epilogue of a row-split mat-vec,
// consumer = prologue of the next mat-vec): 2 generations (site parity) x world source slots x TP_LL_SLOT
// entries of {float value, u32 flag}, flag = pass counter << 8 | site.  A slot holds the D partial sums of the
// source rank followed by its D / world receptance slice (ffn sites).  Sites alternate generations, and a rank
// can only reach site s + 2 after every rank has produced site s + 1, i.e. consumed site s: no slot is
// overwritten while a peer still reads it.
constexpr size_t TP_LL_MAX_D = 8192;
constexpr size_t TP_LL_SLOT = TP_LL_MAX_D + TP_LL_MAX_D / 2;
inline size_t tp_region_ll_offset(int world) { return tp_region_flags_bytes() + (size_t) 2 * world * TP_NMAX * sizeof(float); }
// A second area of the same shape serves the stand-alone low-latency all-reduce kernel (tp_p2p.cu tp_ll_allreduce_kernel): entries
// {value, call counter} instead of a separate flag word, so a collective costs one NVLink store latency and no system-wide fence.
inline size_t tp_region_ll2_offset(int world) { return tp_region_ll_offset(world) + (size_t) 2 * world * TP_LL_SLOT * 8; }
inline size_t tp_region_bytes(int world) { return tp_region_ll2_offset(world) + (size_t) 2 * world * TP_LL_SLOT * 8; }
constexpr int TP_STEP_WORD = 1000;           // u32 index of the pass counter inside the flags page (local use only)
constexpr int TP_LL2_COUNTER_WORD = 512;     // u32 index of the per-CTA call counters of the low-latency all-reduce (TP_LL_SLOT / 256 of them)

struct TpInfo {
    int rank = 0, world = 1;
    void * comm = nullptr;           // ncclComm_t
    TpPeer p2p;
};

bool tp_available();                                               // libnccl found
bool tp_unique_id(void * out128);                                  // ncclGetUniqueId
bool tp_comm_init(TpInfo & tp, const void * id128);                // ncclCommInitRank (device = current)
void tp_comm_destroy(TpInfo & tp);
// sum of `n` floats over the ranks, result on every rank (send may equal recv).  low_latency: the vector is ONE row of the model
// (single-token decode): it may take the {value, call counter} kernel.  Longer vectors (several rows) must not: an inbox entry of that
// kernel is always validated against the call counter of the CTA that owns it, and only with one row per call do the plain form
// (entries [0, D)) and the channel-mix form (entries [0, 2D), two per thread) agree on the owner of every entry.
cudaError_t tp_allreduce(const TpInfo & tp, const float * send, float * recv, size_t n, cudaStream_t st, bool low_latency = false);
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
// Low-latency form for vectors of up to TP_LL_SLOT floats: recv[i] = sum over ranks (rank order) of send[i].  With combine_x the
// vector is [n / 2 partial sums | n / 2 gate values] and the kernel finishes the v4 channel mix in place of a separate launch:
// combine_x[i] += sum(send[n / 2 + i]) * sum(send[i])   (recv is not written).
bool tp_ll_allreduce_usable(const TpInfo & tp, size_t n);
cudaError_t tp_ll_allreduce(const TpInfo & tp, const float * send, float * recv, size_t n, float * combine_x, cudaStream_t st);
struct TpFused;
TpFused tp_fused_desc(const TpInfo & tp);                          // launch descriptor of the fused collectives (ll[0] == nullptr: peer memory not connected)
bool tp_error_pending(const TpInfo & tp);                          // true (and the word cleared) when a collective kernel timed out since the last call

} // namespace rb
