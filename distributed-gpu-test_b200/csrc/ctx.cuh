// The per-process context: what deepspeed.init_distributed() + set_device leave behind
// in the reference (train.py:59-61) -- a device binding and a communicator -- plus the
// side stream and events the bucket all-reduce is ordered with.
#pragma once
#include "common.cuh"
#include <nccl.h>

struct b200ds_ctx {
  int device = 0, rank = 0, world = 1;
  ncclComm_t comm = nullptr;
  cudaStream_t side = nullptr;       // all-reduce stream (non-blocking)
  cudaEvent_t ev_ready = nullptr;    // bucket complete on the caller's stream
  cudaEvent_t ev_done = nullptr;     // all-reduce complete on the side stream
  int64_t* iota = nullptr;           // device [32] = 0..31, identity order for the host-buffer step
  float* barrier_word = nullptr;     // device scratch for b200ds_barrier
  // host-batch prefetch (b200ds_train_step_host_prefetch): copy stream, "copy landed" event,
  // which host batch sits in which half of the staging buffers
  cudaStream_t copy = nullptr;
  cudaEvent_t ev_copied = nullptr;
  const float* pf_x = nullptr;
  const float* pf_y = nullptr;
  int pf_slot = 0;
  // the device part of that step as a CUDA graph per staging half (captured on `cap`, replayed on the
  // caller's stream); the Adam step number travels through step_word
  struct StepGraph {
    unsigned char key[192];
    size_t key_len = 0;
    cudaGraphExec_t exec = nullptr;
  } sgraph[2];
  cudaStream_t cap = nullptr;
  unsigned long long* step_word = nullptr;     // device
  unsigned long long* step_pinned = nullptr;   // pinned host
  bool step_word_known = false;                // the host knows what step_word holds (it is only written by this context's graphs and copies)
  int64_t step_word_value = 0;
  // peer-memory exchange (b200ds_ctx_enable_p2p): every rank's exchange buffer and flag row,
  // own entries cudaMalloc'ed here, the others opened from the peers' IPC handles
  float* p2p_xchg[B200DS_P2P_MAX_WORLD] = {};
  unsigned long long* p2p_flags[B200DS_P2P_MAX_WORLD] = {};
  int64_t p2p_cap = 0;               // floats per half of an exchange buffer; 0 = not enabled
  bool p2p_own = false;
  // a kernel that gives up waiting for a peer reports here (pinned host memory, mapped into the device: it stays
  // readable after the trap that follows): [0] = 1 + the rank waited for, [1] = the epoch
  unsigned long long* p2p_mailbox = nullptr;       // host pointer
  unsigned long long* p2p_mailbox_dev = nullptr;   // the same words as the device sees them
  unsigned long long p2p_timeout_ns = 0;           // bound of a flag wait (B200DS_P2P_TIMEOUT_MS, default 300 s)
  // host-batch prefetch: shape of the batch sitting in the staging half (a different shape is a miss)
  int64_t pf_batch = 0;
  int pf_F = 0;
};

namespace b200ds {
// NCCL is bound at run time (dlopen of the libnccl.so.2 torch already loaded, else the
// system one) so that the library loads on machines without NCCL when world == 1.
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  const char* (*GetErrorString)(ncclResult_t);
};
const NcclApi* nccl_api();   // nullptr (+ error string set) when NCCL cannot be loaded
}  // namespace b200ds
