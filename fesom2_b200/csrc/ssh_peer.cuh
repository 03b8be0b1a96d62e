// Peer-memory transport of the SSH solver's small messages: structures and device code shared by the BiCGStab kernels
// (ssh_solver.cu) and the V-cycle (ssh_amg.cu, whose first kernel pushes / awaits the overlap rows' right-hand side).
#pragma once
#include "common.cuh"
#include <cstddef>

#define SBS 256                 // threads per block in the solver kernels (8 slices)

struct SolverScal {             // device-resident scalars
    double alpha, sigma, beta, rho, rho_alt, omega, tol2;
    double tmp0, tmp1;
    int done, its, maxits, comm_error;
    unsigned long long epoch;                      // number of this psolve call (host-written, the same on every rank): high half of the interface tags
    unsigned long long red_seq;                    // peer-memory transport: sequence number of the reductions (lives on across calls)
};

// Peer-memory transport of the solver's small messages (comm.cuh: PeerBuf).  Every rank owns one control buffer
//   RedSlot slots[4][nranks]      reduction k: rank r writes its two partial sums + k into slot [k & 3][r] of EVERY rank; a rank
//                                 that finds all nranks sequence numbers equal to k adds the partials in rank order, so all ranks
//                                 form bitwise the same sums and take the same decisions.  Slot k & 3 is reused by reduction
//                                 k + 4, which a rank can only reach after every peer has contributed to k + 1 .. k + 3, i.e.
//                                 after every peer has finished reading reduction k.
//   u64     flags[4][nranks]      interface values: the head of the consuming kernel (push_and_wait) writes this rank's send list straight
//                                 into the tails of the peers' vectors (x, pt, st, the V-cycle's b0), then the exchange's tag into
//                                 flags[vec][me] of each peer, and waits for the flags of its own source ranks.  The reductions between two pushes
//                                 of the same vector order the next push after the peers' reads (see psolve).
struct RedSlot { double a, b; unsigned long long seq, pad; };
struct PeerCtl {                // by value into the kernels.  mode 0: one rank; 1: NCCL all-reduce + k_apply_scalars; 2: peer memory
    int mode, nranks, me, nsrc;
    RedSlot* const* slots;      // [nranks] rank r's slot array
    unsigned long long* const* flags;   // [nranks] rank r's flag array
    const int* src;             // [nsrc] ranks this rank receives interface values from
};
// A peer that never arrives raises comm_error (psolve then reports error 48) instead of hanging the GPU.  The limit is ~60 s of SM
// clocks: the ranks of a job reach their first exchange seconds apart (host-side set-up, page-locking of large arrays), which is not a
// failure.  Once the flag is up no wait spins any more and the iteration loop ends (done = 1): the failing solve drains in milliseconds.
#define SPIN_LIMIT_CYCLES 120000000000ll
__device__ __forceinline__ bool spin_expired(SolverScal* sc, long long t0) {
    if (*(volatile int*)&sc->comm_error == 0 && clock64() - t0 <= SPIN_LIMIT_CYCLES) return false;
    sc->comm_error = 1; sc->done = 1;
    return true;
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) { return *(const volatile unsigned long long*)p; }
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) { *(volatile unsigned long long*)p = v; }
// Interface values over peer memory, fused into the CONSUMING kernel (no launch of its own).  The first `npb` blocks of the
// kernel write this rank's send list straight into the tails of the peers' vectors, the last of them to finish publishes the
// tag of this exchange in flags[vec][me] of every destination rank; then every block waits until the flags of its source ranks
// have reached the tag.  A tag is (psolve call number << 32) + a counter derived from the iteration number, which every rank
// knows without communication and which never decreases.  Pushing comes before waiting in every block, so two ranks waiting
// for each other always have their pushes under way; the pushing blocks are the lowest-numbered ones and are scheduled first.
struct PushArgs {
    int count, ndst, npb, nsrc;         // entries to send, destination ranks, pushing blocks (0: nothing to send), source ranks
    const int *slist, *s_peer, *s_dst, *dst_ranks, *src;
    double* const* peer_vec;            // [nranks] base of the peers' copies of the exchanged vector
    unsigned* counter;
};
// vec 0 x (initial guess), 1 pt, 2 st: once per iteration.  vec 3, the V-cycle's right-hand side on the overlap rows: first apply (mode 0),
// then per iteration mode 1 (its = k) and mode 2 (its = k + 1)  ->  1, 2, 3, 5, 6, 8, 9 ...
__device__ __forceinline__ unsigned long long halo_tag(const SolverScal* sc, int vec, int mode) {
    const unsigned long long its = (unsigned long long)sc->its;
    const unsigned long long k = vec != 3 ? its + 1 : (mode == 0 ? 1ull : mode == 1 ? 3 * its + 2 : 3 * its);
    return (sc->epoch << 32) + k;
}
__device__ __forceinline__ void push_and_wait(const PushArgs& pa, const double* x, const PeerCtl& pc, SolverScal* sc, int vec, int mode) {
    if (pc.mode != 2) return;
    const unsigned long long tag = halo_tag(sc, vec, mode);
    __shared__ bool last;
    if ((int)blockIdx.x < pa.npb) {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < pa.count; i += pa.npb * blockDim.x) pa.peer_vec[pa.s_peer[i]][pa.s_dst[i]] = x[pa.slist[i]];
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) last = atomicAdd(pa.counter, 1u) == (unsigned)pa.npb - 1;
        __syncthreads();
        if (last) {
            if (threadIdx.x == 0) *pa.counter = 0;
            __threadfence_system();
            if ((int)threadIdx.x < pa.ndst) st_volatile_u64(pc.flags[pa.dst_ranks[threadIdx.x]] + (size_t)vec * pc.nranks + pc.me, tag);
        }
    }
    if (threadIdx.x == 0) {
        const unsigned long long* f = pc.flags[pc.me] + (size_t)vec * pc.nranks;
        const long long t0 = clock64();
        for (int k = 0; k < pa.nsrc; k++)
            while (ld_volatile_u64(f + pa.src[k]) < tag) if (spin_expired(sc, t0)) break;
        __threadfence_system();
    }
    __syncthreads();
}

