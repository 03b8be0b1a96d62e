// NCCL plumbing and halo exchange kernels (see comm.cuh).
#include "comm.cuh"
#include "fesom_b200.h"
#include <cstring>

static ncclComm_t g_comm = nullptr;
static int g_nranks = 1, g_rank = 0;
int comm_nranks() { return g_nranks; }
int comm_rank() { return g_rank; }
ncclComm_t comm_handle() { return g_comm; }

void halo_plan_build(HaloPlan& h, int rPEnum, const int* rPE, const int* rptr, const int* rlist, int sPEnum, const int* sPE,
                     const int* sptr, const int* slist) {
    h.release();
    h.nrp = rPEnum; h.nsp = sPEnum;
    h.rPE.assign(rPE, rPE + rPEnum); h.sPE.assign(sPE, sPE + sPEnum);
    h.roff.resize(rPEnum + 1); h.soff.resize(sPEnum + 1);
    for (int k = 0; k <= rPEnum; k++) h.roff[k] = rptr[k] - rptr[0];
    for (int k = 0; k <= sPEnum; k++) h.soff[k] = sptr[k] - sptr[0];
    std::vector<int> rl(h.nrecv()), sl(h.nsend());
    for (int k = 0; k < h.nrecv(); k++) rl[k] = rlist[k] - 1;
    for (int k = 0; k < h.nsend(); k++) sl[k] = slist[k] - 1;
    h.rlist.upload(rl); h.slist.upload(sl);
}

// one thread per (listed entity, level, component); buffer layout [entity][level][comp]
__global__ void __launch_bounds__(256) k_halo_pack(const double* __restrict__ field, double* __restrict__ buf, const int* __restrict__ list,
                                                   int count, int nlp, int ncomp) {
    const int per = nlp * ncomp;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)count * per) return;
    const int e = (int)(i / per), r = (int)(i - (long long)e * per), k = r / ncomp, cmp = r - k * ncomp;
    const int ent = list[e];
    buf[i] = field[((size_t)(ent >> 5) * nlp * TILE + (size_t)k * TILE + (ent & 31)) * ncomp + cmp];
}
__global__ void __launch_bounds__(256) k_halo_unpack(double* __restrict__ field, const double* __restrict__ buf, const int* __restrict__ list,
                                                     int count, int nlp, int ncomp) {
    const int per = nlp * ncomp;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)count * per) return;
    const int e = (int)(i / per), r = (int)(i - (long long)e * per), k = r / ncomp, cmp = r - k * ncomp;
    const int ent = list[e];
    field[((size_t)(ent >> 5) * nlp * TILE + (size_t)k * TILE + (ent & 31)) * ncomp + cmp] = buf[i];
}
__global__ void __launch_bounds__(256) k_flat_pack(const double* __restrict__ x, double* __restrict__ buf, const int* __restrict__ list, int count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) buf[i] = x[list[i]];
}
__global__ void __launch_bounds__(256) k_flat_unpack(double* __restrict__ x, const double* __restrict__ buf, const int* __restrict__ list, int count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) x[list[i]] = buf[i];
}

static void exchange_buffers(HaloPlan& h, size_t per, cudaStream_t st) {
    NCCL_CHECK(ncclGroupStart());
    for (int k = 0; k < h.nrp; k++)
        NCCL_CHECK(ncclRecv(h.rbuf.p + (size_t)h.roff[k] * per, (size_t)(h.roff[k + 1] - h.roff[k]) * per, ncclDouble, h.rPE[k], g_comm, st));
    for (int k = 0; k < h.nsp; k++)
        NCCL_CHECK(ncclSend(h.sbuf.p + (size_t)h.soff[k] * per, (size_t)(h.soff[k + 1] - h.soff[k]) * per, ncclDouble, h.sPE[k], g_comm, st));
    NCCL_CHECK(ncclGroupEnd());
}

void halo_exchange_tiled(HaloPlan& h, double* field, int nlp, int ncomp, cudaStream_t st) {
    if (g_nranks <= 1 || !fb_ok() || (h.nsend() == 0 && h.nrecv() == 0)) return;
    if (!g_comm) { fb_set_error(50, "halo exchange before fb200_comm_init"); return; }
    const size_t per = (size_t)nlp * ncomp;
    if (h.sbuf.n < (size_t)h.nsend() * per) h.sbuf.alloc((size_t)h.nsend() * per);
    if (h.rbuf.n < (size_t)h.nrecv() * per) h.rbuf.alloc((size_t)h.nrecv() * per);
    const long long ns = (long long)h.nsend() * per, nr = (long long)h.nrecv() * per;
    if (ns) { ProfScope ps("k_halo_pack", st); k_halo_pack<<<(unsigned)((ns + 255) / 256), 256, 0, st>>>(field, h.sbuf.p, h.slist.p, h.nsend(), nlp, ncomp); LAUNCH_CHECK(); }
    { ProfScope ps("nccl_sendrecv", st); exchange_buffers(h, per, st); }
    if (nr) { ProfScope ps("k_halo_unpack", st); k_halo_unpack<<<(unsigned)((nr + 255) / 256), 256, 0, st>>>(field, h.rbuf.p, h.rlist.p, h.nrecv(), nlp, ncomp); LAUNCH_CHECK(); }
}

void halo_exchange_flat(HaloPlan& h, double* x, cudaStream_t st) {
    if (g_nranks <= 1 || !fb_ok() || (h.nsend() == 0 && h.nrecv() == 0)) return;
    if (!g_comm) { fb_set_error(50, "halo exchange before fb200_comm_init"); return; }
    if (h.sbuf.n < (size_t)h.nsend()) h.sbuf.alloc(h.nsend());
    if (h.rbuf.n < (size_t)h.nrecv()) h.rbuf.alloc(h.nrecv());
    if (h.nsend()) { k_flat_pack<<<(h.nsend() + 255) / 256, 256, 0, st>>>(x, h.sbuf.p, h.slist.p, h.nsend()); LAUNCH_CHECK(); }
    exchange_buffers(h, 1, st);
    if (h.nrecv()) { k_flat_unpack<<<(h.nrecv() + 255) / 256, 256, 0, st>>>(x, h.rbuf.p, h.rlist.p, h.nrecv()); LAUNCH_CHECK(); }
}

void comm_allreduce_sum(double* dev_inout, int count, cudaStream_t st) {
    if (g_nranks <= 1 || !fb_ok()) return;
    NCCL_CHECK(ncclAllReduce(dev_inout, dev_inout, count, ncclDouble, ncclSum, g_comm, st));
}

extern "C" void ssh_drop_graphs();

extern "C" {

void fb200_comm_unique_id(char* id128) {
    ncclUniqueId id;
    NCCL_CHECK(ncclGetUniqueId(&id));
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id128, &id, sizeof(id));
}

void fb200_comm_init(int* nranks, int* rank, char* id128) {
    if (g_comm) { ssh_drop_graphs(); cudaDeviceSynchronize(); }
    if (g_comm) { ncclCommDestroy(g_comm); g_comm = nullptr; }
    g_nranks = *nranks; g_rank = *rank;
    if (g_nranks <= 1) return;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fb_set_error(12, "no CUDA device: libfesom_b200 has no CPU fallback"); return; }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    NCCL_CHECK(ncclCommInitRank(&g_comm, g_nranks, id, g_rank));
}

void fb200_comm_finalize(void) {
    ssh_drop_graphs();
    cudaDeviceSynchronize();
    if (g_comm) { ncclCommDestroy(g_comm); g_comm = nullptr; }
    g_nranks = 1; g_rank = 0;
}

}  // extern "C"
