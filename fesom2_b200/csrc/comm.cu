// NCCL plumbing and halo exchange kernels (see comm.cuh).
#include "comm.cuh"
#include "fesom_b200.h"
#include <cstring>
#include <cstdlib>

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

// ---- peer memory -------------------------------------------------------------------------------------
bool peer_enabled() {
#ifdef FB200_EMULATION
    return false;
#else
    if (g_nranks <= 1) return false;
    const char* e = getenv("FB200_SSH_COMM");
    return !(e && !strcmp(e, "nccl"));
#endif
}
void peer_release(PeerBuf& b) {
#ifndef FB200_EMULATION
    for (size_t r = 0; r < b.peer.size(); r++) if (b.peer[r] && (int)r != g_rank) cudaIpcCloseMemHandle(b.peer[r]);
#endif
    if (b.owned && b.local) cudaFree(b.local);
    b.peer.clear(); b.d_peer.release(); b.local = nullptr; b.bytes = 0; b.owned = false;
}
bool peer_publish(PeerBuf& b, void* base, size_t bytes, cudaStream_t st) {
#ifdef FB200_EMULATION
    (void)b; (void)base; (void)bytes; (void)st;
    return false;
#else
    peer_release(b);
    if (!peer_enabled() || !g_comm || !fb_ok()) return false;
    b.bytes = bytes;
    if (base) { b.local = base; b.owned = false; }
    else {
        if (cudaMalloc(&b.local, bytes) != cudaSuccess) { cudaGetLastError(); b.local = nullptr; return false; }
        b.owned = true;
        CUDA_CHECK(cudaMemsetAsync(b.local, 0, bytes, st));
    }
    cudaIpcMemHandle_t mine;
    int ok = cudaIpcGetMemHandle(&mine, b.local) == cudaSuccess ? 1 : 0;
    if (!ok) cudaGetLastError();
    // all-gather (ok flag + handle) so that every rank takes the same decision
    const size_t rec = sizeof(cudaIpcMemHandle_t) + 8;
    std::vector<char> hsend(rec, 0), hall(rec * g_nranks, 0);
    memcpy(hsend.data(), &ok, sizeof(int)); memcpy(hsend.data() + 8, &mine, sizeof(mine));
    DevBuf<char> d_s, d_a;
    d_s.upload(hsend); d_a.alloc(rec * g_nranks);
    NCCL_CHECK(ncclAllGather(d_s.p, d_a.p, rec, ncclChar, g_comm, st));
    CUDA_CHECK(cudaMemcpyAsync(hall.data(), d_a.p, hall.size(), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    d_s.release(); d_a.release();
    bool all_ok = fb_ok();
    for (int r = 0; r < g_nranks; r++) { int f = 0; memcpy(&f, hall.data() + rec * r, sizeof(int)); all_ok = all_ok && f == 1; }
    b.peer.assign(g_nranks, nullptr);
    if (all_ok) {
        for (int r = 0; r < g_nranks; r++) {
            if (r == g_rank) { b.peer[r] = b.local; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, hall.data() + rec * r + 8, sizeof(h));
            if (cudaIpcOpenMemHandle(&b.peer[r], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); b.peer[r] = nullptr; all_ok = false; }
        }
    }
    // a second agreement round: opening can fail on one rank only
    {
        int mine_ok = all_ok ? 1 : 0;
        DevBuf<int> d_f, d_g;
        std::vector<int> f1(1, mine_ok), fall(g_nranks, 0);
        d_f.upload(f1); d_g.alloc(g_nranks);
        NCCL_CHECK(ncclAllGather(d_f.p, d_g.p, 1, ncclInt32, g_comm, st));
        CUDA_CHECK(cudaMemcpyAsync(fall.data(), d_g.p, fall.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        d_f.release(); d_g.release();
        for (int v : fall) all_ok = all_ok && v == 1;
    }
    if (!all_ok) { peer_release(b); return false; }
    b.d_peer.upload(b.peer);
    return fb_ok();
#endif
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
