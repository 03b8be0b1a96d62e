// TEST INFRASTRUCTURE — CUDA-on-CPU emulation used by tests/test_emulated.py only.
// It lets the CPU test suite execute the SOURCE of the tracer kernels (fesom2_b200/csrc/*.cu, transformed by
// tests/emu/build_emu.py and compiled with g++ -ffp-contract=off) before any GPU time is spent: blocks run one after
// another, the threads of a block as ucontext fibers, so __syncthreads / __syncwarp / __shfl_down_sync behave as on the
// device.  It is NOT a fallback of the product: libfesom_b200.so never loads it and still fails loudly without a CUDA
// device; nothing under fesom2_b200/ refers to this directory.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __launch_bounds__(...)
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))
#define __CUDACC__ 1

struct dim3 { unsigned x, y, z; dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {} };
struct uint3 { unsigned x, y, z; };
extern uint3 threadIdx, blockIdx;
extern dim3 blockDim, gridDim;

struct double2 { double x, y; };
struct int2 { int x, y; };
struct int4 { int x, y, z, w; };
struct ulonglong2 { unsigned long long x, y; };
static inline double2 make_double2(double x, double y) { double2 r; r.x = x; r.y = y; return r; }
static inline int2 make_int2(int x, int y) { int2 r; r.x = x; r.y = y; return r; }
static inline int4 make_int4(int x, int y, int z, int w) { int4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
static inline ulonglong2 make_ulonglong2(unsigned long long x, unsigned long long y) { ulonglong2 r; r.x = x; r.y = y; return r; }

// ---- runtime API subset
typedef int cudaError_t;
typedef struct emu_stream_* cudaStream_t;
typedef struct emu_event_* cudaEvent_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyHostToHost, cudaMemcpyDefault };
enum { cudaEventDisableTiming = 2, cudaStreamNonBlocking = 1, cudaFuncAttributeMaxDynamicSharedMemorySize = 8, cudaHostRegisterDefault = 0 };
template <class T> static inline cudaError_t cudaMalloc(T** p, size_t n) { *p = (T*)malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaFree(void* p) { free(p); return 0; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemset(void* d, int v, size_t n) { memset(d, v, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = (cudaStream_t)malloc(1); return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t s) { free(s); return 0; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return 0; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = (cudaEvent_t)malloc(1); return 0; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { *e = (cudaEvent_t)malloc(1); return 0; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { free(e); return 0; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr) { return 0; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return 0; }
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return 0; }
static inline cudaError_t cudaGetLastError() { return 0; }
static inline cudaError_t cudaGetDevice(int* d) { *d = 0; return 0; }
template <class T> static inline cudaError_t cudaGetSymbolAddress(void** p, T& sym) { *p = (void*)&sym; return 0; }
static inline const char* cudaGetErrorString(cudaError_t) { return "emulated"; }
static inline cudaError_t cudaHostRegister(void*, size_t, unsigned) { return 0; }
static inline cudaError_t cudaHostUnregister(void*) { return 0; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return 0; }
template <class T> static inline cudaError_t cudaMallocHost(T** p, size_t n) { *p = (T*)malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaFreeHost(void* p) { free(p); return 0; }
// CUDA graphs are not emulated (tests/emu/__init__.py switches them off with fb200_psolver_use_graph(0)): capture reports an error
typedef struct emu_graph_* cudaGraph_t;
typedef struct emu_graph_exec_* cudaGraphExec_t;
enum { cudaStreamCaptureModeThreadLocal = 1 };
static inline cudaError_t cudaStreamBeginCapture(cudaStream_t, int) { return 900; }
static inline cudaError_t cudaStreamEndCapture(cudaStream_t, cudaGraph_t* g) { *g = nullptr; return 900; }
static inline cudaError_t cudaGraphInstantiate(cudaGraphExec_t*, cudaGraph_t, unsigned long long) { return 900; }
static inline cudaError_t cudaGraphDestroy(cudaGraph_t) { return 0; }
static inline cudaError_t cudaGraphExecDestroy(cudaGraphExec_t) { return 0; }
static inline cudaError_t cudaGraphLaunch(cudaGraphExec_t, cudaStream_t) { return 900; }
static inline void __threadfence() {}
static inline void __threadfence_system() {}
static inline long long clock64() { return 0; }
enum { cudaDevAttrMultiProcessorCount = 16 };
static inline cudaError_t cudaDeviceGetAttribute(int* v, int, int) { *v = 148; return 0; }

// ---- execution model
struct emu_cfg {
    dim3 grid, block; size_t smem;
    emu_cfg(dim3 g, dim3 b, size_t s = 0, cudaStream_t = nullptr) : grid(g), block(b), smem(s) {}
};
void emu_launch(const emu_cfg& cfg, const std::function<void()>& body);
void* emu_dyn_smem();
void emu_syncthreads();
void emu_syncwarp();
void emu_shfl_put(const void* v, size_t n);
void emu_shfl_get(void* v, size_t n, int lane);
int emu_lane();
#define __syncthreads() emu_syncthreads()
#define __syncwarp(...) emu_syncwarp()
template <class T> static inline T __shfl_down_sync(unsigned, T v, int delta) {
    emu_shfl_put(&v, sizeof(T)); emu_syncwarp();
    T r = v; const int src = emu_lane() + delta;
    if (src < 32) emu_shfl_get(&r, sizeof(T), src);
    emu_syncwarp();
    return r;
}
template <class T> static inline T __shfl_sync(unsigned, T v, int src) {
    emu_shfl_put(&v, sizeof(T)); emu_syncwarp();
    T r = v; emu_shfl_get(&r, sizeof(T), src & 31);
    emu_syncwarp();
    return r;
}
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int mask) {
    emu_shfl_put(&v, sizeof(T)); emu_syncwarp();
    T r = v; emu_shfl_get(&r, sizeof(T), (emu_lane() ^ mask) & 31);
    emu_syncwarp();
    return r;
}
static inline int __ffsll(long long v) { return __builtin_ffsll(v); }
static inline int __clzll(long long v) { return v ? __builtin_clzll((unsigned long long)v) : 64; }
template <class T> static inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
template <class T> static inline T __ldg(const T* p) { return *p; }
template <class T> static inline T __ldcg(const T* p) { return *p; }
using std::fma; using std::fabs; using std::floor; using std::sqrt;
static inline int max(int a, int b) { return a > b ? a : b; }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
