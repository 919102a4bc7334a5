// sm_100a primitives used by the tensor-core convolution kernels: mbarrier,
// cp.async (LDGSTS) with zero-fill, cp.async.bulk (UBLKCP), tcgen05 alloc /
// mma / commit / ld, UMMA shared-memory and instruction descriptors.
// Bit layouts follow the PTX ISA "tcgen05" chapter (smem matrix descriptor,
// instruction descriptor for .kind::f16).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mk {
namespace sm100 {

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// one lane of a converged warp (the compiler keeps the guarded code on the uniform datapath)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// ---- mbarrier ---------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
// Suspend-time hint of the sleeping waits.  A sleeper occasionally misses its wake-up (seen
// as one launch in ~15 000 taking 2 ms longer with the former 10 ms hint: the last waiter of a
// CTA, with no further barrier event to wake it, slept until the hardware's limit); with
// 20 us the same event costs at most a few microseconds and ordinary waits (well under the
// hint) are unchanged.
constexpr uint32_t kSuspendHintNs = 20000u;
// Potentially-blocking test of a phase: the suspend-time hint lets the hardware
// park the thread until the phase completes instead of busy-polling.  (A bare
// try_wait/test_wait loop in two dozen warps eats the issue slots of the warps
// that do the work: measured 2x on the whole kernel.)
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(kSuspendHintNs)
      : "memory");
  return ok != 0;
}
// try_wait without a suspend-time hint (the hardware's own short time limit): for
// barriers that several warps wait on at once.  The hinted form sleeps on the
// barrier (NANOSLEEP.SYNCS); with two warps asleep on one barrier a phase
// completion was seen to wake only one, the other slept out its 10 ms hint
// (measured on the dy-tile barrier of the weight-gradient kernel).
__device__ __forceinline__ bool mbar_try_wait_short(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must surface as an error (trap -> sticky CUDA
// error, reported by the next API call), never as a hung GPU.  The report goes
// to a word of mapped host memory (it survives the dead context); everything is
// inline - a call in this path would force spills into the hot loops.
static __device__ unsigned int *mk_hang_slot = nullptr;   // set once per translation unit by the launcher
// `sleep_ns` > 0 backs off between polls: the hinted try_wait (NANOSLEEP.SYNCS) comes
// back on every barrier event of the CTA, so a warp that waits long by construction
// (the epilogue: a whole group) re-polls thousands of times - 20 % of all
// instructions of the forward kernel (ncu source view) in a kernel bound by issue.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity, int budget = 1, uint32_t sleep_ns = 0) {
  if (mbar_try_wait(bar, parity)) return;
  uint32_t spins = 0;
  long long t0 = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (sleep_ns) __nanosleep(sleep_ns);
    if ((++spins & 0xFFFFu) == 0) {                   // clock read off the common path
      const long long now = clock64();
      if (t0 == 0) {
        t0 = now;
      } else if (now - t0 > 10000000000LL * budget) {  // ~5 s
        unsigned int *h = mk_hang_slot;
        if (h) {
          h[1] = blockIdx.x | (blockIdx.y << 16);
          h[2] = threadIdx.x >> 5;
          h[3] = (smem_u32(bar) << 1) | parity;
          h[0] = 0xDEADBA44u;
          __threadfence_system();
        }
        asm volatile("trap;");
      }
    }
  }
}

// multi-waiter wait: short polls, same watchdog
__device__ __forceinline__ void mbar_wait_shared(uint64_t *bar, uint32_t parity, int budget = 1) {
  uint32_t spins = 0;
  long long t0 = 0;
  while (!mbar_try_wait_short(bar, parity)) {
    if ((++spins & 0xFFFFu) == 0) {
      const long long now = clock64();
      if (t0 == 0) {
        t0 = now;
      } else if (now - t0 > 10000000000LL * budget) {
        unsigned int *h = mk_hang_slot;
        if (h) {
          h[1] = blockIdx.x | (blockIdx.y << 16);
          h[2] = threadIdx.x >> 5;
          h[3] = (smem_u32(bar) << 1) | parity;
          h[0] = 0xDEADBA44u;
          __threadfence_system();
        }
        asm volatile("trap;");
      }
    }
  }
}

// Warp-level wait: one lane polls, the rest park on the warp barrier.  Polling
// with all 32 lanes of two dozen warps saturates the barrier unit and slows the
// arrivals everybody is waiting for (measured: 0.82 ms of pure spin overhead
// per 0.2 M stages).
__device__ __forceinline__ void mbar_wait_warp(uint64_t *bar, uint32_t parity, int budget = 1,
                                               uint32_t sleep_ns = 0) {
  mbar_wait(bar, parity, budget, sleep_ns);
  __syncwarp();
}
// ---- shared-space forms (32-bit shared address) -----------------------------------
// The hot producer loops keep shared addresses as 32-bit integers: through a
// generic pointer the compiler emits generic LD/ST and a cvta per barrier
// operation, and those loops are bound by instruction issue.
__device__ __forceinline__ uint32_t lds_u8(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
  return v;
}
__device__ __forceinline__ int4 lds_v4(uint32_t saddr) {
  int4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(saddr)
               : "memory");
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t saddr, uint32_t v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
  return v;
}
__device__ __forceinline__ void mbar_arrive_expect_tx_a(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc_a(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_a(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity), "r"(kSuspendHintNs)
      : "memory");
  return ok != 0;
}
// mbar_wait_warp on a shared address; the (rare) slow path is the pointer form with its watchdog
__device__ __forceinline__ void mbar_wait_warp_a(uint32_t bar, uint32_t parity) {
  if (!mbar_try_wait_a(bar, parity)) mbar_wait((uint64_t *)__cvta_shared_to_generic(bar), parity);
  __syncwarp();
}

// ---- async copies -----------------------------------------------------------
// 16-byte LDGSTS; src_bytes == 0 writes 16 zero bytes without touching memory
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void *src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(src_bytes)
               : "memory");
}
// arrive on `bar` once all cp.async of this thread issued so far have landed
// (counts as one of the barrier's expected arrivals)
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t *bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// wait until at most N of this thread's committed cp.async groups are still in flight
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
// bulk global -> shared copy (UBLKCP), completion as tx bytes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// make generic-proxy / cp.async writes to shared memory visible to the async
// proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- TMA (cp.async.bulk.tensor) -------------------------------------------------
// tile::gather4: four rows (row indices r.x..r.w of dimension 1) x one box of
// columns starting at `col`, written as 4 consecutive box rows at dst with the
// tensor map's swizzle; out-of-range rows / columns are zero-filled.  Completion
// is counted on `bar` as the full box bytes.
__device__ __forceinline__ void tma_gather4(uint32_t dst_smem, const void *tmap, int32_t col, int4 r,
                                            uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes.cta_group::1 "
      "[%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(dst_smem),
      "l"(tmap), "r"(col), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w), "r"(smem_u32(bar))
      : "memory");
}
// plain 2-D tiled load: box at (col, row)
__device__ __forceinline__ void tma_gather4_a(uint32_t dst_smem, const void *tmap, int32_t col, int4 r, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes.cta_group::1 "
      "[%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(dst_smem),
      "l"(tmap), "r"(col), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst_smem, const void *tmap, int32_t col, int32_t row,
                                            uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3}], [%4];" ::"r"(dst_smem),
      "l"(tmap), "r"(col), "r"(row), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void *tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

// ---- tcgen05 ------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// whole warp; writes the TMEM base address to *slot (shared memory)
__device__ __forceinline__ void tmem_alloc_512(uint32_t *slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot))
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_512(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], bf16/f16 inputs, fp32 accumulate; one thread issues
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on `bar` when every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// TMEM -> registers: this warp's 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- descriptors ------------------------------------------------------------------
// Shared-memory matrix descriptor, 128-byte swizzle, canonical layout made of
// 8-row x 128-byte atoms (1024 B, base 1024-aligned), atoms `sbo_bytes` apart.
//   bits [0,14)  start address >> 4      bits [16,30) leading byte offset >> 4
//   bits [32,46) stride byte offset >> 4 bits [46,48) version = 1 (sm_100)
//   bits [61,64) layout type: 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor for .kind::f16: D fp32, A/B both `fmt` (0 = f16, 1 = bf16),
// a_major / b_major: 0 = K-major, 1 = MN-major.
__host__ __device__ __forceinline__ uint32_t idesc_f16(int fmt, int M, int N, int a_major, int b_major) {
  uint32_t d = 0;
  d |= 1u << 4;                       // c_format = F32
  d |= (uint32_t)fmt << 7;            // a_format
  d |= (uint32_t)fmt << 10;           // b_format
  d |= (uint32_t)a_major << 15;
  d |= (uint32_t)b_major << 16;
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}

}  // namespace sm100
}  // namespace mk
