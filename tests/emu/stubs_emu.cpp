// TEST INFRASTRUCTURE — what the emulated tracer library needs from the files that are not emulated: comm.cu (single rank:
// no exchange) and the SSH solver's entry points (not emulated: they report an error).
#include "common.cuh"
#include "comm.cuh"

int comm_nranks() { return 1; }
int comm_rank() { return 0; }
ncclComm_t comm_handle() { return nullptr; }
void halo_plan_build(HaloPlan&, int, const int*, const int*, const int*, int, const int*, const int*, const int*) {}
void halo_exchange_tiled(HaloPlan&, double*, int, int, cudaStream_t) {}
void halo_exchange_flat(HaloPlan&, double*, cudaStream_t) {}
void comm_allreduce_sum(double*, int, cudaStream_t) {}
extern "C" {
void fb200_comm_unique_id(char*) {}
void fb200_comm_init(int*, int*, char*) {}
void fb200_comm_finalize(void) {}
void psolver_final(void) {}
}
double ssh_last_ms() { return 0.0; }
