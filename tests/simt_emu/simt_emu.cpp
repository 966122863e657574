// Fiber scheduler of the SIMT emulator (see simt_emu.h). TEST INFRASTRUCTURE.
#include "simt_emu.h"
#include <stdio.h>
#include <stdlib.h>

emu_dim3 threadIdx, blockIdx, blockDim, gridDim;

namespace simt_emu {
static State g_state;
void set_site(int site) { State& s = st(); if (s.cur >= 0 && s.cur < (int)s.fibers.size()) s.fibers[s.cur].site = site; }
State& st() { return g_state; }
static const size_t kStack = 512 * 1024;

static void trampoline() {
  State& s = st();
  s.body();
  Fiber& f = s.fibers[s.cur];
  f.done = true;
  // a lane that leaves the kernel no longer takes part in barriers
  WarpState& w = s.warps[s.cur / 32];
  w.alive--;
  s.block_alive--;
  if (w.alive > 0 && w.arrived == w.alive) { w.arrived = 0; w.gen++; }
  if (s.block_alive > 0 && s.block_arrived == s.block_alive) { s.block_arrived = 0; s.block_gen++; }
  s.progressed = true;
  swapcontext(&f.ctx, &s.sched);
}

void yield() {
  State& s = st();
  swapcontext(&s.fibers[s.cur].ctx, &s.sched);
  threadIdx.x = (unsigned)s.cur;
}

unsigned char* dyn_smem() { return st().dyn; }

void warp_barrier() {
  State& s = st();
  WarpState& w = s.warps[s.cur / 32];
  unsigned gen = w.gen;
  const int g_site = s.fibers[s.cur].site;
  if (w.arrived == 0) {
    w.site = g_site;
  } else if (w.site != g_site) {
    fprintf(stderr, "simt_emu: lanes of one warp meet at different collectives: line %d vs line %d (lane %d, block %u)\n",
            w.site % 100000, g_site % 100000, s.cur % 32, blockIdx.x);
    abort();
  }
  if (++w.arrived == w.alive) {
    w.arrived = 0;
    w.gen++;
    s.progressed = true;
    return;
  }
  while (w.gen == gen) yield();
}

void block_barrier() {
  State& s = st();
  unsigned gen = s.block_gen;
  if (++s.block_arrived == s.block_alive) {
    s.block_arrived = 0;
    s.block_gen++;
    s.progressed = true;
    return;
  }
  while (s.block_gen == gen) yield();
}

void launch(dim3 grid3, int block, size_t smem, std::function<void()> body) {
  const int grid = (int)(grid3.x * grid3.y);
  State& s = st();
  s.body = body;
  if ((int)s.fibers.size() < block) {
    size_t old = s.fibers.size();
    s.fibers.resize(block);
    for (size_t i = old; i < (size_t)block; i++) s.fibers[i].stack = (char*)malloc(kStack);
  }
  s.dyn = (unsigned char*)aligned_alloc(256, ((smem + 255) & ~(size_t)255) + 256);
  gridDim.x = grid3.x; gridDim.y = grid3.y; gridDim.z = 1;
  blockDim.x = (unsigned)block; blockDim.y = blockDim.z = 1;
  threadIdx.y = threadIdx.z = blockIdx.z = 0;
  const int nwarps = (block + 31) / 32;
  for (int bid = 0; bid < grid; bid++) {
    blockIdx.x = (unsigned)bid % grid3.x;
    blockIdx.y = (unsigned)bid / grid3.x;
    s.warps.assign(nwarps, WarpState());
    for (int wi = 0; wi < nwarps; wi++) {
      int nl = block - wi * 32;
      s.warps[wi].alive = nl > 32 ? 32 : nl;
      s.warps[wi].arrived = 0;
      s.warps[wi].gen = 0;
    }
    s.block_alive = block;
    s.block_arrived = 0;
    s.block_gen = 0;
    for (int t = 0; t < block; t++) {
      Fiber& f = s.fibers[t];
      f.done = false;
      getcontext(&f.ctx);
      f.ctx.uc_stack.ss_sp = f.stack;
      f.ctx.uc_stack.ss_size = kStack;
      f.ctx.uc_link = &s.sched;
      makecontext(&f.ctx, (void (*)())trampoline, 0);
    }
    int remaining = block;
    while (remaining > 0) {
      s.progressed = false;
      remaining = 0;
      for (int t = 0; t < block; t++) {
        if (s.fibers[t].done) continue;
        s.cur = t;
        threadIdx.x = (unsigned)t;
        swapcontext(&s.sched, &s.fibers[t].ctx);
        if (!s.fibers[t].done) remaining++;
      }
      if (remaining > 0 && !s.progressed) {
        fprintf(stderr, "simt_emu: deadlock (divergent barrier) in block %d\n", bid);
        abort();
      }
    }
  }
  free(s.dyn);
  s.dyn = nullptr;
}
}  // namespace simt_emu
