// TEST INFRASTRUCTURE -- see cuda_runtime.h in this directory.  The fiber scheduler and the guarded allocator of
// the CUDA emulation shim.
#include "cuda_runtime.h"
#undef threadIdx
#undef blockIdx
#undef blockDim
#undef gridDim

#include <execinfo.h>
#include <signal.h>
#include <stdio.h>
#include <sys/mman.h>
#include <ucontext.h>
#include <unistd.h>

#include <map>
#include <mutex>
#include <vector>

namespace emu {

// ---- guarded "device" memory: [PROT_NONE page][ ... payload, end-aligned ][PROT_NONE page] -------------------
struct Alloc { char *base; size_t map_bytes; };
static std::map<void *, Alloc> g_allocs;   // shared by the emulated ranks (a context may be closed from another thread)
static std::mutex g_alloc_mutex;
static const size_t PAGE = 4096;
static bool g_handler = false;

static void on_segv(int, siginfo_t *si, void *)
{
  char msg[160];
  int n = snprintf(msg, sizeof msg, "\ncudaemu: invalid memory access at %p (out-of-bounds device access or a host bug)\n", si->si_addr);
  if (write(2, msg, n) < 0) {}
  void *bt[48];
  backtrace_symbols_fd(bt, backtrace(bt, 48), 2);
  _exit(139);
}

void *dev_alloc(size_t bytes)
{
  static std::mutex handler_mutex;
  std::unique_lock<std::mutex> hl(handler_mutex);
  if (!g_handler) {
    // the handler runs on a fiber's stack; give it its own so a fiber stack overflow is reported too
    static char altstack[1 << 16];
    stack_t ss;
    ss.ss_sp = altstack;
    ss.ss_size = sizeof altstack;
    ss.ss_flags = 0;
    sigaltstack(&ss, nullptr);
    struct sigaction sa;
    memset(&sa, 0, sizeof sa);
    sa.sa_sigaction = on_segv;
    sa.sa_flags = SA_SIGINFO | SA_ONSTACK;
    sigaction(SIGSEGV, &sa, nullptr);
    g_handler = true;
  }
  hl.unlock();
  if (bytes == 0) bytes = 1;
  const size_t payload = (bytes + 7) / 8 * 8;                 // keep 8-byte alignment of the start
  const size_t pages = (payload + PAGE - 1) / PAGE;
  const size_t map_bytes = (pages + 2) * PAGE;
  char *base = (char *) mmap(nullptr, map_bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
  if (base == MAP_FAILED) return nullptr;
  mprotect(base, PAGE, PROT_NONE);
  mprotect(base + (pages + 1) * PAGE, PAGE, PROT_NONE);
  char *p = base + (pages + 1) * PAGE - payload;
  memset(p, 0xCD, payload);                                   // uninitialised device memory is not zero
  std::lock_guard<std::mutex> lock(g_alloc_mutex);
  g_allocs[p] = {base, map_bytes};
  return p;
}

void dev_free(void *p)
{
  if (!p) return;
  std::lock_guard<std::mutex> lock(g_alloc_mutex);
  auto it = g_allocs.find(p);
  if (it == g_allocs.end()) {
    fprintf(stderr, "cudaemu: cudaFree of a pointer cudaMalloc did not return (%p)\n", p);
    abort();
  }
  munmap(it->second.base, it->second.map_bytes);
  g_allocs.erase(it);
}

// ---- fibers ------------------------------------------------------------------------------------------------
// x86-64: a six-register stack switch (System V callee-saved set); ucontext's swapcontext costs a signal-mask
// system call per switch, which dominated the run time.  Other hosts fall back to ucontext.
#if defined(__x86_64__)
#define EMU_ASM_SWITCH 1
extern "C" void emu_switch(void **save_sp, void *load_sp);
asm(R"(
.text
.globl emu_switch
.type emu_switch,@function
emu_switch:
  pushq %rbp
  pushq %rbx
  pushq %r12
  pushq %r13
  pushq %r14
  pushq %r15
  movq %rsp, (%rdi)
  movq %rsi, %rsp
  popq %r15
  popq %r14
  popq %r13
  popq %r12
  popq %rbx
  popq %rbp
  ret
.size emu_switch,.-emu_switch
)");
#else
#define EMU_ASM_SWITCH 0
#endif
struct Warp {
  int live = 0, count = 0;
  unsigned gen = 0, livemask = 0;
  unsigned long long scratch[32];
};
struct Fiber {
  ucontext_t ctx;
  void *sp = nullptr;
  char *stack = nullptr;
  bool done = true;
  ThreadState ts;
  const unsigned *wait_gen = nullptr;   // blocked until *wait_gen != wait_val
  unsigned wait_val = 0;
};
static const size_t STACK = 128 * 1024;
// fiber stacks are mappings: give them back when the emulated rank's thread ends (the map count of a process is limited)
struct FiberPool {
  std::vector<Fiber *> v;
  ~FiberPool();
};
static thread_local FiberPool g_pool;
#define g_fibers (g_pool.v)
static thread_local ucontext_t g_sched;
static thread_local void *g_sched_sp = nullptr;
static inline void to_sched(Fiber *f)
{
#if EMU_ASM_SWITCH
  emu_switch(&f->sp, g_sched_sp);
#else
  swapcontext(&f->ctx, &g_sched);
#endif
}
static inline void to_fiber(Fiber *f)
{
#if EMU_ASM_SWITCH
  emu_switch(&g_sched_sp, f->sp);
#else
  swapcontext(&g_sched, &f->ctx);
#endif
}
static thread_local Warp g_warps[32];
static thread_local int g_block_live = 0, g_block_count = 0;
static thread_local unsigned g_block_gen = 0;
static thread_local const std::function<void()> *g_body = nullptr;
static thread_local Fiber *g_cur = nullptr;
static thread_local ThreadState g_host_ts;
thread_local ThreadState *cur = nullptr;   // the host thread's own state until a kernel thread runs
thread_local void *dyn_smem = nullptr;
thread_local int launches = 0;
static thread_local std::vector<char> g_smem;

unsigned long long *warp_scratch() { return g_warps[cur->warp].scratch; }
unsigned warp_livemask() { return g_warps[cur->warp].livemask; }

static void yield_wait(const unsigned *gen, unsigned val)
{
  Fiber *f = g_cur;
  f->wait_gen = gen;
  f->wait_val = val;
  to_sched(f);
}

void warp_barrier()
{
  Warp &w = g_warps[cur->warp];
  const unsigned gen = w.gen;
  if (++w.count >= w.live) {
    w.count = 0;
    w.gen++;
    return;
  }
  yield_wait(&w.gen, gen);
}

void block_barrier()
{
  const unsigned gen = g_block_gen;
  if (++g_block_count >= g_block_live) {
    g_block_count = 0;
    g_block_gen++;
    return;
  }
  yield_wait(&g_block_gen, gen);
}

static void trampoline()
{
  Fiber *f = g_cur;
  (*g_body)();
  // an exited thread counts as arrived at every later rendezvous
  f->done = true;
  Warp &w = g_warps[f->ts.warp];
  w.live--;
  w.livemask &= ~(1u << f->ts.lane);
  if (w.count > 0 && w.count >= w.live) { w.count = 0; w.gen++; }
  g_block_live--;
  if (g_block_count > 0 && g_block_count >= g_block_live) { g_block_count = 0; g_block_gen++; }
  to_sched(f);
  abort();    // a finished fiber is never resumed
}

FiberPool::~FiberPool()
{
  for (Fiber *f : v) {
    if (f->stack) munmap(f->stack, STACK + PAGE);
    delete f;
  }
}

void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body)
{
  if (cur != nullptr && cur != &g_host_ts) {
    fprintf(stderr, "cudaemu: kernel launch from inside a kernel\n");
    abort();
  }
  const int nthreads = (int) (block.x * block.y * block.z);
  if (nthreads <= 0 || nthreads > 1024 || grid.x == 0 || grid.y == 0 || grid.z == 0) {
    fprintf(stderr, "cudaemu: invalid launch configuration grid (%u,%u,%u) block (%u,%u,%u)\n", grid.x, grid.y, grid.z, block.x, block.y, block.z);
    abort();
  }
  if (smem > 227 * 1024) {
    fprintf(stderr, "cudaemu: %zu bytes of dynamic shared memory exceed 227 KB\n", smem);
    abort();
  }
  launches++;
  while ((int) g_fibers.size() < nthreads) {
    Fiber *f = new Fiber();
    f->stack = (char *) mmap(nullptr, STACK + PAGE, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (f->stack == MAP_FAILED) {
      fprintf(stderr, "cudaemu: cannot map a fiber stack\n");
      abort();
    }
    mprotect(f->stack, PAGE, PROT_NONE);   // stack overflow faults
    g_fibers.push_back(f);
  }
  g_smem.assign(smem + 64, (char) 0xCD);
  dyn_smem = g_smem.data();
  g_body = &body;
  const int nwarps = (nthreads + 31) / 32;
  // CUDAEMU_ORDER=reverse runs the blocks of every launch last-to-first and resumes the threads of a block
  // last-to-first: CUDA promises no order, so results that change under it point at an ordering assumption
  // (a block reading what another block of the same launch writes, a thread relying on an earlier one)
  static const bool reverse = getenv("CUDAEMU_ORDER") && strcmp(getenv("CUDAEMU_ORDER"), "reverse") == 0;
  for (unsigned bz_ = 0; bz_ < grid.z; bz_++)
    for (unsigned by_ = 0; by_ < grid.y; by_++)
      for (unsigned bx_ = 0; bx_ < grid.x; bx_++) {
        const unsigned bx = reverse ? grid.x - 1 - bx_ : bx_, by = reverse ? grid.y - 1 - by_ : by_, bz = reverse ? grid.z - 1 - bz_ : bz_;
        for (int w = 0; w < nwarps; w++) {
          g_warps[w].live = std::min(32, nthreads - 32 * w);
          g_warps[w].count = 0;
          g_warps[w].livemask = g_warps[w].live == 32 ? 0xffffffffu : ((1u << g_warps[w].live) - 1u);
        }
        g_block_live = nthreads;
        g_block_count = 0;
        for (int t = 0; t < nthreads; t++) {
          Fiber *f = g_fibers[t];
          f->done = false;
          f->wait_gen = nullptr;
          f->ts.threadIdx = {(unsigned) t % block.x, ((unsigned) t / block.x) % block.y, (unsigned) t / (block.x * block.y)};
          f->ts.blockIdx = {bx, by, bz};
          f->ts.blockDim = block;
          f->ts.gridDim = grid;
          f->ts.lane = t & 31;
          f->ts.warp = t >> 5;
#if EMU_ASM_SWITCH
          {
            // first switch "returns" into trampoline with the stack as a call would leave it (rsp = 8 mod 16)
            uintptr_t top = ((uintptr_t) (f->stack + PAGE + STACK)) & ~(uintptr_t) 15;
            void **slot = (void **) (top - 16);
            slot[1] = nullptr;
            slot[0] = (void *) &trampoline;
            for (int r = 1; r <= 6; r++) slot[-r] = nullptr;
            f->sp = (void *) (slot - 6);
          }
#else
          getcontext(&f->ctx);
          f->ctx.uc_stack.ss_sp = f->stack + PAGE;
          f->ctx.uc_stack.ss_size = STACK;
          f->ctx.uc_link = nullptr;
          makecontext(&f->ctx, trampoline, 0);
#endif
        }
        int remaining = nthreads;
        while (remaining > 0) {
          bool progressed = false;
          for (int t_ = 0; t_ < nthreads; t_++) {
            const int t = reverse ? nthreads - 1 - t_ : t_;
            Fiber *f = g_fibers[t];
            if (f->done) continue;
            if (f->wait_gen) {
              if (*f->wait_gen == f->wait_val) continue;   // still blocked
              f->wait_gen = nullptr;
            }
            g_cur = f;
            cur = &f->ts;
            to_fiber(f);
            cur = &g_host_ts;
            progressed = true;
            if (f->done) remaining--;
          }
          if (!progressed) {
            fprintf(stderr, "cudaemu: deadlock in block (%u,%u,%u): every running thread waits at a barrier the others never reach\n", bx, by, bz);
            abort();
          }
        }
      }
  g_body = nullptr;
  g_cur = nullptr;
}

}    // namespace emu
