// emu_runtime.h -- TEST INFRASTRUCTURE ONLY.
//
// A minimal host-side SIMT emulator used to run the *same kernel source* that
// nvcc compiles for sm_100a inside the CPU-only build container, so that
// tests/ can check the kernels' indexing and barrier logic against the oracle
// before any GPU time is spent.  One ucontext fiber per CUDA thread; a
// scheduler pass resumes every fiber once, i.e. one pass == one
// __syncthreads() interval (all threads of a CTA must reach the same barriers,
// exactly CUDA's rule).  CTAs are distributed over host threads.
//
// Not part of the product: meta_pde_b200/_lib.py never loads the library built
// from this header, and nothing under meta_pde_b200/ includes it unless
// MPDE_EMU is defined by tests/emu/Makefile.
#pragma once
#include <ucontext.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>
#include <atomic>

namespace emu {

struct dim3_ { unsigned x = 1, y = 1, z = 1; dim3_() {} dim3_(unsigned a, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };

struct Tls {
  dim3_ tid, bid, bdim, gdim;
  unsigned char* smem = nullptr;
  ucontext_t* sched = nullptr;
  ucontext_t* self = nullptr;
};
inline Tls& tls() { static thread_local Tls t; return t; }

inline void syncthreads() {
  Tls& t = tls();
  swapcontext(t.self, t.sched);
}

struct FiberArg { const std::function<void()>* fn; bool* done; ucontext_t* self; ucontext_t* sched; };
inline void fiber_entry(unsigned lo, unsigned hi) {
  FiberArg* a = reinterpret_cast<FiberArg*>(((uintptr_t)hi << 32) | (uintptr_t)lo);
  (*a->fn)();
  *a->done = true;
  swapcontext(a->self, a->sched);
}

constexpr size_t kMaxSmem = 232448;   // 227 KB opt-in limit on sm_100
constexpr size_t kStack = 256 * 1024;

// run one CTA to completion
inline void run_cta(dim3_ grid, dim3_ block, dim3_ bid, size_t smem_bytes, const std::function<void()>& fn,
                    std::vector<char>& stacks, std::vector<unsigned char>& smem) {
  const unsigned nt = block.x * block.y * block.z;
  if (stacks.size() < (size_t)nt * kStack) stacks.resize((size_t)nt * kStack);
  if (smem.size() < smem_bytes + 64) smem.resize(smem_bytes + 64);
  // poison shared memory so that reads of unwritten locations show up as NaN
  memset(smem.data(), 0xFF, smem.size());
  unsigned char* sm = smem.data();
  sm += (16 - ((uintptr_t)sm & 15)) & 15;
  std::vector<ucontext_t> ctx(nt);
  std::vector<char> done_(nt, 0);
  std::vector<FiberArg> args(nt);
  ucontext_t sched;
  bool* done = reinterpret_cast<bool*>(done_.data());
  for (unsigned i = 0; i < nt; ++i) {
    getcontext(&ctx[i]);
    ctx[i].uc_stack.ss_sp = stacks.data() + (size_t)i * kStack;
    ctx[i].uc_stack.ss_size = kStack;
    ctx[i].uc_link = nullptr;
    args[i] = FiberArg{&fn, &done[i], &ctx[i], &sched};
    uintptr_t p = (uintptr_t)&args[i];
    makecontext(&ctx[i], (void (*)())fiber_entry, 2, (unsigned)(p & 0xffffffffu), (unsigned)(p >> 32));
  }
  Tls& t = tls();
  t.bdim = block; t.gdim = grid; t.bid = bid; t.smem = sm; t.sched = &sched;
  unsigned remaining = nt;
  while (remaining) {
    unsigned finished_this_pass = 0, ran = 0;
    for (unsigned i = 0; i < nt; ++i) {
      if (done[i]) continue;
      t.tid = dim3_(i % block.x, (i / block.x) % block.y, i / (block.x * block.y));
      t.self = &ctx[i];
      swapcontext(&sched, &ctx[i]);
      ++ran;
      if (done[i]) ++finished_this_pass;
    }
    if (finished_this_pass != 0 && finished_this_pass != ran) {
      fprintf(stderr, "[emu] barrier divergence: %u of %u threads exited while others wait\n", finished_this_pass, ran);
      abort();
    }
    remaining -= finished_this_pass;
  }
}

inline int host_threads() {
  const char* e = getenv("MPDE_EMU_THREADS");
  int n = e ? atoi(e) : (int)std::thread::hardware_concurrency();
  return n < 1 ? 1 : n;
}

inline void launch(dim3_ grid, dim3_ block, size_t smem_bytes, const std::function<void()>& fn) {
  if (smem_bytes > kMaxSmem) { fprintf(stderr, "[emu] %zu B dynamic shared memory exceeds the 227 KB limit\n", smem_bytes); abort(); }
  if (block.x * block.y * block.z > 1024) { fprintf(stderr, "[emu] block too large\n"); abort(); }
  const unsigned total = grid.x * grid.y * grid.z;
  std::atomic<unsigned> next{0};
  auto worker = [&]() {
    std::vector<char> stacks;
    std::vector<unsigned char> smem;
    for (;;) {
      unsigned b = next.fetch_add(1);
      if (b >= total) break;
      dim3_ bid(b % grid.x, (b / grid.x) % grid.y, b / (grid.x * grid.y));
      run_cta(grid, block, bid, smem_bytes, fn, stacks, smem);
    }
  };
  int nth = host_threads();
  if ((unsigned)nth > total) nth = (int)total;
  if (nth <= 1) { worker(); return; }
  std::vector<std::thread> th;
  for (int i = 0; i < nth; ++i) th.emplace_back(worker);
  for (auto& x : th) x.join();
}

}  // namespace emu

typedef emu::dim3_ dim3;
