// TEST INFRASTRUCTURE — placeholder for <nccl.h> in the CUDA-on-CPU emulation (single rank, no communication)
#pragma once
typedef struct emu_nccl_comm_* ncclComm_t;
typedef int ncclResult_t;
enum { ncclSuccess = 0 };
static inline const char* ncclGetErrorString(ncclResult_t) { return "emulated"; }
