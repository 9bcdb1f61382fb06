// TEST INFRASTRUCTURE: coroutine scheduler of the CUDA emulator (see cuda_emu.h).
#include <ucontext.h>

#include <vector>

#include "cuda_emu.h"

namespace emu {

uint3 g_threadIdx, g_blockIdx;
dim3 g_blockDim, g_gridDim;
unsigned char *g_dyn_smem = nullptr;

namespace {

enum State { READY, AT_BLOCK_BARRIER, AT_WARP_BARRIER, DONE };

struct Thread {
  ucontext_t ctx;
  std::vector<unsigned char> stack;
  State state;
  uint3 tid;
  uint32_t shfl_val;
};

constexpr size_t STACK = 256 * 1024;
std::vector<Thread> g_threads;
ucontext_t g_sched;
int g_cur = -1;
const std::function<void()> *g_body = nullptr;

void trampoline() {
  (*g_body)();
  g_threads[g_cur].state = DONE;
  swapcontext(&g_threads[g_cur].ctx, &g_sched);
}

void yield_as(State s) {
  g_threads[g_cur].state = s;
  swapcontext(&g_threads[g_cur].ctx, &g_sched);
}

void run_block(int nthreads) {
  for (int t = 0; t < nthreads; ++t) {
    Thread &th = g_threads[t];
    getcontext(&th.ctx);
    th.ctx.uc_stack.ss_sp = th.stack.data();
    th.ctx.uc_stack.ss_size = th.stack.size();
    th.ctx.uc_link = &g_sched;
    makecontext(&th.ctx, trampoline, 0);
    th.state = READY;
  }
  for (;;) {
    bool progressed = false;
    for (int t = 0; t < nthreads; ++t) {
      if (g_threads[t].state != READY) continue;
      g_cur = t;
      g_threadIdx = g_threads[t].tid;
      swapcontext(&g_sched, &g_threads[t].ctx);
      progressed = true;
    }
    // release warp barriers: all live lanes of the warp are waiting at it
    for (int w0 = 0; w0 < nthreads; w0 += 32) {
      int w1 = std::min(w0 + 32, nthreads), waiting = 0, live = 0;
      for (int t = w0; t < w1; ++t) {
        if (g_threads[t].state != DONE) ++live;
        if (g_threads[t].state == AT_WARP_BARRIER) ++waiting;
      }
      if (live && waiting == live) {
        for (int t = w0; t < w1; ++t)
          if (g_threads[t].state == AT_WARP_BARRIER) g_threads[t].state = READY;
        progressed = true;
      }
    }
    int live = 0, at_block = 0;
    for (int t = 0; t < nthreads; ++t) {
      if (g_threads[t].state != DONE) ++live;
      if (g_threads[t].state == AT_BLOCK_BARRIER) ++at_block;
    }
    if (live == 0) return;
    if (at_block == live) {
      for (int t = 0; t < nthreads; ++t)
        if (g_threads[t].state == AT_BLOCK_BARRIER) g_threads[t].state = READY;
      progressed = true;
    }
    if (!progressed) {
      fprintf(stderr, "cuda_emu: deadlock (divergent barrier) in block (%u,%u,%u)\n", g_blockIdx.x,
              g_blockIdx.y, g_blockIdx.z);
      abort();
    }
  }
}

}  // namespace

void sync_block() { yield_as(AT_BLOCK_BARRIER); }

uint32_t shfl_u32(uint32_t v, int src_lane) {
  int me = g_cur;
  g_threads[me].shfl_val = v;
  yield_as(AT_WARP_BARRIER);  // everyone has published
  int warp0 = me & ~31;
  int src = warp0 + (src_lane & 31);
  uint32_t out = (src < (int)g_threads.size() && g_threads[src].state != DONE) ? g_threads[src].shfl_val : v;
  yield_as(AT_WARP_BARRIER);  // everyone has read before the next publish
  return out;
}

void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body) {
  int nthreads = (int)(block.x * block.y * block.z);
  if ((int)g_threads.size() < nthreads) {
    size_t old = g_threads.size();
    g_threads.resize(nthreads);
    for (size_t t = old; t < g_threads.size(); ++t) g_threads[t].stack.resize(STACK);
  }
  std::vector<unsigned char> dyn(smem + 64);
  g_dyn_smem = dyn.data() + (64 - ((uintptr_t)dyn.data() % 64)) % 64;
  g_blockDim = block;
  g_gridDim = grid;
  g_body = &body;
  for (int t = 0; t < nthreads; ++t) {
    unsigned u = (unsigned)t;
    g_threads[t].tid = uint3{u % block.x, (u / block.x) % block.y, u / (block.x * block.y)};
  }
  for (unsigned bz = 0; bz < grid.z; ++bz)
    for (unsigned by = 0; by < grid.y; ++by)
      for (unsigned bx = 0; bx < grid.x; ++bx) {
        g_blockIdx = uint3{bx, by, bz};
        run_block(nthreads);
      }
  g_dyn_smem = nullptr;
  g_body = nullptr;
}

}  // namespace emu
