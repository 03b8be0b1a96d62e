// Inter-GPU communication of libfesom_b200.so: one process per GPU, NCCL over NVLink.
//   HaloPlan   = the reference's com_struct (src/gen_modules_partitioning.F90:17-29) as device index lists;
//   halo_exchange = exchange_nod / exchange_elem (src/gen_halo_exchange.F90): pack kernel -> grouped ncclSend/ncclRecv
//                   -> unpack kernel, all on the library stream;
//   comm_allreduce_sum = the solver's MPI_Allreduce / MPI_Iallreduce (lib/parms/src/bicgstab_ras.c:133,162,191,211).
#pragma once
#include "common.cuh"
#include <nccl.h>

struct HaloPlan {
    int nsp = 0, nrp = 0;                       // number of PEs to send to / receive from
    std::vector<int> sPE, soff, rPE, roff;      // PE ids and offsets into slist / rlist (size n+1)
    DevBuf<int> slist, rlist;                   // local 0-based entity ids
    DevBuf<double> sbuf, rbuf;                  // staging, sized on first use
    int nsend() const { return soff.empty() ? 0 : soff.back(); }
    int nrecv() const { return roff.empty() ? 0 : roff.back(); }
    void release() { slist.release(); rlist.release(); sbuf.release(); rbuf.release(); nsp = nrp = 0; sPE.clear(); soff.clear(); rPE.clear(); roff.clear(); }
};

int comm_nranks();
int comm_rank();
ncclComm_t comm_handle();
// build a plan from the reference's 1-based com_struct arrays
void halo_plan_build(HaloPlan& h, int rPEnum, const int* rPE, const int* rptr, const int* rlist, int sPEnum, const int* sPE,
                     const int* sptr, const int* slist);
// field: tiled [tile][nlp][32][ncomp] (entity = node or element); every level of the listed entities is exchanged
void halo_exchange_tiled(HaloPlan& h, double* field, int nlp, int ncomp, cudaStream_t st);
// flat vector (one value per entity): x[rlist] <- owner's x[slist]
void halo_exchange_flat(HaloPlan& h, double* x, cudaStream_t st);
void comm_allreduce_sum(double* dev_inout, int count, cudaStream_t st);

// ---- peer memory over NVLink (one NVSwitch node is the whole machine: every GPU can load/store every other GPU's HBM).
// A PeerBuf is one cudaMalloc'd buffer per rank whose CUDA IPC handles were all-gathered, so peer[r] is rank r's buffer
// mapped into this process (peer[me] = local).  The SSH solver uses it for its latency-bound traffic: the 16-byte
// all-reduces and the interface values of the SpMV are written straight into the peers' memory by the producing kernel
// and consumed after a sequence flag arrives — no NCCL call, no extra launch (ssh_solver.cu).  NCCL remains the
// transport of the bulk exchanges and the alternative path (FB200_SSH_COMM=nccl).
struct PeerBuf {
    void* local = nullptr; size_t bytes = 0; bool owned = false;
    std::vector<void*> peer;                    // [nranks]
    DevBuf<void*> d_peer;                       // the same table on the device
};
// collective over all ranks (NCCL all-gather of the handles on stream st).  base == nullptr: allocate (zeroed) `bytes`;
// else publish an existing cudaMalloc'd allocation (base must be the allocation's base address).  false = CUDA IPC is not
// available here (then the caller stays on NCCL).
bool peer_publish(PeerBuf& b, void* base, size_t bytes, cudaStream_t st);
void peer_release(PeerBuf& b);
bool peer_enabled();                            // more than one rank, not switched off by FB200_SSH_COMM=nccl, not the CPU emulation
#define NCCL_CHECK(x) do { ncclResult_t r_ = (x); if (r_ != ncclSuccess) fb_set_error(300 + (int)r_, std::string(#x) + ": " + ncclGetErrorString(r_)); } while (0)
