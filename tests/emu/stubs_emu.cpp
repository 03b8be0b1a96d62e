// TEST INFRASTRUCTURE — what the emulated tracer library needs from the SSH solver files, which are not emulated
#include "common.cuh"
double ssh_last_ms() { return 0.0; }
extern "C" {
void ssh_drop_graphs() {}
void psolver_final(void) {}
}
