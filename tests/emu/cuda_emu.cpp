// TEST INFRASTRUCTURE — fiber scheduler of the CUDA-on-CPU emulation (see cuda_emu.h)
#include "cuda_emu.h"
#include <ucontext.h>
#include <vector>
#include <stdexcept>

uint3 threadIdx, blockIdx;
dim3 blockDim, gridDim;

namespace {
constexpr size_t STACK = 256 * 1024;
struct Fiber { ucontext_t ctx; char* stack = nullptr; int state = 0; /* 0 ready, 1 at block barrier, 2 at warp barrier, 3 done */ uint3 tid; int lin; };
std::vector<Fiber> fibers;
ucontext_t sched_ctx;
int cur = -1;
const std::function<void()>* cur_body = nullptr;
std::vector<char> dyn_smem;
std::vector<char> shfl_slots;            // [warp][32][16 bytes]
void trampoline() {
    (*cur_body)();
    fibers[cur].state = 3;
    swapcontext(&fibers[cur].ctx, &sched_ctx);
}
void yield_with(int st) {
    const int me = cur;
    fibers[me].state = st;
    swapcontext(&fibers[me].ctx, &sched_ctx);
    threadIdx = fibers[me].tid;              // restored by the scheduler as well; kept for clarity
}
}  // namespace

void* emu_dyn_smem() { return dyn_smem.data(); }
void emu_syncthreads() { yield_with(1); }
void emu_syncwarp() { yield_with(2); }
int emu_lane() { return fibers[cur].lin & 31; }
void emu_shfl_put(const void* v, size_t n) { if (n > 16) throw std::runtime_error("emu shfl: type too large"); memcpy(&shfl_slots[(size_t)fibers[cur].lin * 16], v, n); }
void emu_shfl_get(void* v, size_t n, int lane) { memcpy(v, &shfl_slots[(size_t)((fibers[cur].lin & ~31) + lane) * 16], n); }

void emu_launch(const emu_cfg& cfg, const std::function<void()>& body) {
    const int nt = (int)(cfg.block.x * cfg.block.y * cfg.block.z);
    if (nt <= 0 || nt > 1024) throw std::runtime_error("emu_launch: bad block size");
    if ((int)fibers.size() < nt) {
        const size_t old = fibers.size();
        fibers.resize(nt);
        for (size_t i = old; i < fibers.size(); i++) fibers[i].stack = (char*)malloc(STACK);
    }
    dyn_smem.assign(cfg.smem + 64, 0);
    shfl_slots.assign((size_t)((nt + 31) / 32) * 32 * 16, 0);
    blockDim = cfg.block; gridDim = cfg.grid;
    cur_body = &body;
    // FB200_EMU_ORDER=reverse: blocks and the threads of a block are scheduled back to front.  A kernel without atomics whose
    // result depends on the execution order has a race; running the parity tests in both orders is a cheap detector.
    static const bool reverse = getenv("FB200_EMU_ORDER") && !strcmp(getenv("FB200_EMU_ORDER"), "reverse");
    const unsigned long long nblocks = (unsigned long long)cfg.grid.x * cfg.grid.y * cfg.grid.z;
    for (unsigned long long bi = 0; bi < nblocks; bi++) {
        const unsigned long long b = reverse ? nblocks - 1 - bi : bi;
        const unsigned bx = (unsigned)(b % cfg.grid.x), by = (unsigned)((b / cfg.grid.x) % cfg.grid.y), bz = (unsigned)(b / ((unsigned long long)cfg.grid.x * cfg.grid.y));
        blockIdx.x = bx; blockIdx.y = by; blockIdx.z = bz;
        int lin = 0;
        for (unsigned tz = 0; tz < cfg.block.z; tz++) for (unsigned ty = 0; ty < cfg.block.y; ty++) for (unsigned tx = 0; tx < cfg.block.x; tx++, lin++) {
            Fiber& f = fibers[lin];
            f.tid.x = tx; f.tid.y = ty; f.tid.z = tz; f.lin = lin; f.state = 0;
            getcontext(&f.ctx);
            f.ctx.uc_stack.ss_sp = f.stack; f.ctx.uc_stack.ss_size = STACK; f.ctx.uc_link = &sched_ctx;
            makecontext(&f.ctx, trampoline, 0);
        }
        int done = 0;
        while (done < nt) {
            bool progressed = false;
            for (int ii = 0; ii < nt; ii++) {
                const int i = reverse ? nt - 1 - ii : ii;
                if (fibers[i].state != 0) continue;
                cur = i; threadIdx = fibers[i].tid;
                swapcontext(&sched_ctx, &fibers[i].ctx);
                progressed = true;
                if (fibers[i].state == 3) done++;
            }
            // release warp barriers: every live thread of the warp waits at one
            for (int w = 0; w * 32 < nt; w++) {
                int live = 0, at = 0;
                for (int i = w * 32; i < nt && i < w * 32 + 32; i++) { if (fibers[i].state != 3) live++; if (fibers[i].state == 2) at++; }
                if (live > 0 && at == live) { for (int i = w * 32; i < nt && i < w * 32 + 32; i++) if (fibers[i].state == 2) fibers[i].state = 0; progressed = true; }
            }
            // release the block barrier: every live thread waits at it
            {
                int live = 0, at = 0;
                for (int i = 0; i < nt; i++) { if (fibers[i].state != 3) live++; if (fibers[i].state == 1) at++; }
                if (live > 0 && at == live) { for (int i = 0; i < nt; i++) if (fibers[i].state == 1) fibers[i].state = 0; progressed = true; }
            }
            if (!progressed && done < nt) throw std::runtime_error("emu_launch: deadlock (divergent barrier)");
        }
    }
    cur_body = nullptr; cur = -1;
}
