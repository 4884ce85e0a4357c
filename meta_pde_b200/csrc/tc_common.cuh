// tc_common.cuh -- sm_100a building blocks for the tensor-core (tcgen05 / TMEM / bulk-copy) path:
// mbarriers with a watchdog, 1-D bulk copies global -> shared, shared-memory matrix descriptors for the
// 128-byte-swizzled canonical layouts (K-major and MN-major), instruction descriptors, tcgen05.mma issue,
// TMEM allocation and tcgen05.ld.  Inline PTX only; no CUTLASS.
//
// Layout conventions (bytes).  A "slab" is [rows][128 B] with rows grouped by 8 into 1024-byte swizzle atoms:
//   addr(row, b) = slab + (row / 8) * 1024 + (row % 8) * 128 + (((b / 16) ^ (row % 8)) * 16) + b % 16
// The same slab is a K-major operand (rows = M/N index, the 128 B run along K) and an MN-major operand
// (rows = K index, the 128 B run along M/N).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace mpde {
namespace tc {

#define TC_DEV __device__ __forceinline__

// The warp index as a value the compiler KNOWS to be warp-uniform (a shuffle result): role branches on it are uniform
// branches, so code under them may keep uniform registers live.  With `threadIdx.x >> 5` ptxas treated every role region
// as divergent and re-materialised the global-memory descriptor before each LDG / STG (2 R2UR per access: 9 % of the
// HVP kernel's instructions, round 2).
TC_DEV int warp_idx_uniform() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }

TC_DEV uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---- mbarrier ---------------------------------------------------------------------------------------
TC_DEV void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
TC_DEV void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
TC_DEV void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
TC_DEV void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// try_wait with a suspend-time hint: the hardware parks the thread until the phase completes or the hint (ns) expires.
// Without a hint the instruction returned after ~50 cycles (ncu r02c: 1.4e9 polls per HVP launch, 40 % of all executed
// warp instructions were spin-loop overhead of the three control warps, on a power-capped chip).
#ifndef TC_SUSPEND_HINT_NS
#define TC_SUSPEND_HINT_NS 1000000u
#endif
TC_DEV bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(TC_SUSPEND_HINT_NS)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must end in a trap (launch error), never in a hung GPU.
#ifndef TC_WATCHDOG_CYCLES
#define TC_WATCHDOG_CYCLES 4000000000LL
#endif
TC_DEV void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > TC_WATCHDOG_CYCLES) {
      printf("mbar_wait watchdog: block (%d,%d) thread %d bar %u parity %u\n", blockIdx.x, blockIdx.y, threadIdx.x,
             smem_u32(bar), parity);
      __trap();
    }
  }
}

// ---- register reallocation between warpgroups (all warps of a warpgroup execute the same one) ----------------
template <int N> TC_DEV void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> TC_DEV void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// ---- bulk copy global -> shared (1-D, completes on an mbarrier) --------------------------------------
TC_DEV void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// ask L2 for `bytes` (multiple of 16) of global memory starting at a 16-byte aligned address; no completion
TC_DEV void bulk_prefetch_l2(const void* src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (tcgen05.mma operand reads)
TC_DEV void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMEM ---------------------------------------------------------------------------------------------
// One full warp allocates `ncols` (power of two >= 32) columns; the base address lands in *slot (shared).
TC_DEV void tmem_alloc(uint32_t* slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
TC_DEV void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
TC_DEV void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
TC_DEV void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// all previously issued tcgen05.mma of this thread complete -> one arrive on `bar`
TC_DEV void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
TC_DEV void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
TC_DEV void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 consecutive 32-bit columns: thread i of the warp gets lane (taddr.lane + i), v[j] = column j.
// The warp may only touch the TMEM lane quarter 32 * (warp_id % 4).
TC_DEV void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
TC_DEV void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 16 consecutive columns, registers -> tensor memory (the epilogue parks its fp32 outputs in the accumulator
// columns it has just read: see tc_pde_kernel); tcgen05.wait::st before the same thread reads them back
TC_DEV void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
TC_DEV void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
      : "r"(taddr)
      : "memory");
}
TC_DEV uint32_t tmem_addr(uint32_t base, uint32_t lane, uint32_t col) { return base + (lane << 16) + col; }

// ---- descriptors ---------------------------------------------------------------------------------------
// Shared-memory matrix descriptor (64 bit): start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48),
// layout type [61,64) (2 = 128-byte swizzle).  Slabs must be 1024-byte aligned (base_offset = 0).
TC_DEV uint64_t smem_desc_sw128(uint32_t start_bytes, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((start_bytes >> 4) & 0x3FFFu);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// K-major operand inside one slab: 8-row groups are 1024 B apart (SBO); LBO is not used by swizzled K-major.
// `start` = slab + byte offset of the first K element of this MMA inside the 128-byte row (k0 * elem_bytes).
TC_DEV uint64_t desc_kmajor(uint32_t start) { return smem_desc_sw128(start, 16, 1024); }
// MN-major operand: 128-byte runs along M/N; the next 128 B of M/N are `slab_stride` bytes away (LBO), the next
// 8 K rows 1024 B away (SBO).  `start` = slab + (k0 / 8) * 1024.
TC_DEV uint64_t desc_mnmajor(uint32_t start, uint32_t slab_stride) { return smem_desc_sw128(start, slab_stride, 1024); }

// The same descriptors from precomputed halves: the upper word is constant (SBO 1024, version 1, 128-byte swizzle), the lower
// word is (start >> 4) | (LBO >> 4) << 16, so advancing an operand inside its slab is one 32-bit add on the lower word.
constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
TC_DEV uint32_t desc_lo_kmajor(uint32_t start) { return (start >> 4) | (1u << 16); }
TC_DEV uint32_t desc_lo_mnmajor(uint32_t start, uint32_t slab_stride) { return (start >> 4) | ((slab_stride >> 4) << 16); }
TC_DEV uint64_t desc_join(uint32_t lo) { return ((uint64_t)DESC_HI << 32) | (uint64_t)lo; }

// One lane of a CONVERGED warp (all 32 lanes must execute it).  The MMA issuer loops run warp-uniform and only the
// tcgen05.mma / tcgen05.commit instructions sit behind this predicate: issued from inside an `if (lane == 0)` region,
// ptxas cannot prove the descriptors uniform and wraps every UTCHMMA in a vote / elect / branch loop (~10 dependent
// instructions, ~90 cycles per MMA measured in round 2: slower than the 32-64 cycles an N = 64 / 128 MMA executes).
TC_DEV bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}

enum : uint32_t { FMT_F16 = 0, FMT_BF16 = 1, FMT_TF32 = 2 };
// Instruction descriptor (32 bit): D format F32 [4,6)=1, A format [7,10), B format [10,13), A MN-major bit 15,
// B MN-major bit 16, N>>3 [17,23), M>>4 [24,29).
__host__ __device__ constexpr uint32_t make_idesc(uint32_t fmt, uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]; one thread issues.
TC_DEV void mma_f16_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
TC_DEV void mma_tf32_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// byte offset of (row, byte-in-row b < 128) inside a swizzled slab
TC_DEV uint32_t sw128_off(uint32_t row, uint32_t b) {
  return (row >> 3) * 1024u + (row & 7u) * 128u + ((((b >> 4) ^ (row & 7u)) << 4) | (b & 15u));
}

// x = hi + lo with hi, lo in fp16 (x pre-scaled into the fp16 range by the caller); packed (hi | lo << 16)
TC_DEV void split_f16(float x, __half& hi, __half& lo) {
  hi = __float2half_rn(x);
  lo = __float2half_rn(x - __half2float(hi));
}

}  // namespace tc
}  // namespace mpde
