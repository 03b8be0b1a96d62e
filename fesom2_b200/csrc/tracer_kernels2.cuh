// Interface between the step driver (tracer_kernels.cu) and the column kernels of tracer_kernels2.cu.
#pragma once
#include "common.cuh"
#include "tracer_dev.cuh"

#define MAXNT 2                               // tracers per pass
// the per-tracer fields of the tracers a pass works on (the tracer-independent fields stay in KP)
struct TrSet {
    double *T[MAXNT], *Told[MAXNT], *LO[MAXNT], *adf_h[MAXNT], *adf_v[MAXNT], *tstar[MAXNT], *eflx[MAXNT], *trz[MAXNT];
    double2 *txy[MAXNT], *fpm[MAXNT];
    const double* redi_v[MAXNT];
    int tr_num[MAXNT];
};

bool v2_usable(const TracerCtx& c);
void v2_ab2(TracerCtx& c, const TrSet& s, int nt, cudaStream_t st);
void v2_grad_elem(TracerCtx& c, const KP& p, const TrSet& s, int nt, int which, cudaStream_t st);
void v2_grad_z(TracerCtx& c, const KP& p, const TrSet& s, int nt, cudaStream_t st);
void v2_edge_adv(TracerCtx& c, const KP& p, const TrSet& s, int nt, bool fct, cudaStream_t st);
void v2_edge_diff(TracerCtx& c, const KP& p, const TrSet& s, int nt, cudaStream_t st);
void v2_lo(TracerCtx& c, const KP& p, const TrSet& s, int nt, bool fct, cudaStream_t st);
void v2_fct_bounds(TracerCtx& c, const KP& p, const TrSet& s, int nt, cudaStream_t st);
void v2_update(TracerCtx& c, const KP& p, const TrSet& s, int nt, bool fct, bool expl, cudaStream_t st);
