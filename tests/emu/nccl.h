// TEST INFRASTRUCTURE — stands in for <nccl.h> in the CUDA-on-CPU emulation: the handful of NCCL calls comm.cu makes,
// carried over files in a temporary directory shared by the rank processes of one test (the "unique id" is its path).
// Sends write a file and return; receives inside a group are completed at ncclGroupEnd by polling for the peer's file;
// all-reduce sums the ranks' contributions in rank order.  Slow and simple on purpose.
#pragma once
#include <cstddef>
typedef struct emu_nccl_comm* ncclComm_t;
typedef int ncclResult_t;
enum { ncclSuccess = 0 };
enum ncclDataType_t { ncclInt32 = 4, ncclDouble = 8 };            // value = bytes per element
enum ncclRedOp_t { ncclSum = 0 };
struct ncclUniqueId { char internal[128]; };
const char* ncclGetErrorString(ncclResult_t);
ncclResult_t ncclGetUniqueId(ncclUniqueId* id);
ncclResult_t ncclCommInitRank(ncclComm_t* comm, int nranks, ncclUniqueId id, int rank);
ncclResult_t ncclCommDestroy(ncclComm_t comm);
ncclResult_t ncclGroupStart();
ncclResult_t ncclGroupEnd();
ncclResult_t ncclSend(const void* buf, size_t count, ncclDataType_t, int peer, ncclComm_t comm, void* stream);
ncclResult_t ncclRecv(void* buf, size_t count, ncclDataType_t, int peer, ncclComm_t comm, void* stream);
ncclResult_t ncclAllGather(const void* send, void* recv, size_t sendcount, ncclDataType_t, ncclComm_t comm, void* stream);
ncclResult_t ncclAllReduce(const void* send, void* recv, size_t count, ncclDataType_t, ncclRedOp_t, ncclComm_t comm, void* stream);
ncclResult_t ncclBroadcast(const void* send, void* recv, size_t count, ncclDataType_t, int root, ncclComm_t comm, void* stream);
