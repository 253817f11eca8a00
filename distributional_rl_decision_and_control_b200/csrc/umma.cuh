// umma.cuh — tcgen05 / TMA / mbarrier helpers shared by the tensor-core kernels (K8 linear_umma.cu,
// K9 policy_umma.cu, K10 critic_umma.cu).  sm_100a inline PTX only.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace asvrl {

// device status codes a kernel leaves in its status word when a bounded wait expires (include/asvrl.h:
// surfaced as ASVRL_E_DEVICE by asvrl_umma_status; nothing traps, nothing hangs)
enum : int { UM_OK = 0, UM_TIMEOUT_RING = 1, UM_TIMEOUT_ACC = 2, UM_TIMEOUT_INPUT = 3, UM_TIMEOUT_K8 = 4, UM_BAD_SMEM = 5 };

__device__ __forceinline__ uint32_t um_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t um_globaltimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}

// one lane of the (converged) warp; the others skip.  Keeping the MMA warp converged and electing only around the
// tcgen05 instructions lets the compiler hold descriptors and addresses in uniform registers (no R2UR per operand).
__device__ __forceinline__ bool um_elect() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}\n" : "=r"(pred));
  return pred != 0u;
}

// ---- mbarrier ----
__device__ __forceinline__ void um_mbar_init(uint32_t mbar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void um_mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void um_mbar_arrive(uint32_t mbar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}\n" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void um_mbar_arrive_expect_tx(uint32_t mbar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}\n" ::"r"(mbar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ uint32_t um_mbar_try(uint32_t mbar, uint32_t parity) {
  uint32_t done;
  asm volatile("{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
               : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
  return done;
}

// non-blocking: has the phase of this parity completed?
__device__ __forceinline__ uint32_t um_mbar_test(uint32_t mbar, uint32_t parity) {
  uint32_t done;
  asm volatile("{\n\t.reg .pred P;\n\tmbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
               : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
  return done;
}

// Bounded wait.  `abort_flag` (shared memory, one per CTA) makes every later wait of the CTA give up at
// once after the first expiry, so a broken pipeline drains in microseconds; the first code wins the
// status word.  Returns false when the wait was abandoned.
constexpr uint64_t UM_TIMEOUT_NS = 2000000000ull;  // 2 s: three orders above any legitimate wait
__device__ __forceinline__ bool um_mbar_wait(uint32_t mbar, uint32_t parity, volatile uint32_t* abort_flag, int* status,
                                             int code) {
  if (um_mbar_try(mbar, parity)) return true;
  const uint64_t t0 = um_globaltimer();
  for (uint32_t it = 1;; ++it) {
    if (um_mbar_try(mbar, parity)) return true;
    if ((it & 63u) == 0u) {
      if (*abort_flag) return false;
      if (um_globaltimer() - t0 > UM_TIMEOUT_NS) {
        *abort_flag = 1u;
        if (status) atomicCAS(status, 0, code);
        return false;
      }
    }
  }
}

// ---- TMA: 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP) ----
__device__ __forceinline__ void um_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}

// ---- TMA: 1-D bulk copy shared -> global (bulk async-group completion) ----
__device__ __forceinline__ void um_bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void um_bulk_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
// at most N of this thread's bulk groups still READING their shared-memory source
template <int N>
__device__ __forceinline__ void um_bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;\n" ::"n"(N) : "memory"); }
template <int N>
__device__ __forceinline__ void um_bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;\n" ::"n"(N) : "memory"); }

// named barrier among `threads` threads (whole warps) of the CTA; id 1..15 (0 is __syncthreads)
__device__ __forceinline__ void um_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(threads) : "memory");
}

// ---- proxies and tcgen05 fences ----
__device__ __forceinline__ void um_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void um_tc_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void um_tc_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

// ---- tensor memory ----
template <int COLS>
__device__ __forceinline__ void um_tmem_alloc(uint32_t slot_smem) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(slot_smem), "n"(COLS) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void um_tmem_free(uint32_t tmem) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "n"(COLS) : "memory");
}

// 32 consecutive accumulator columns of this thread's TMEM lane (warp w reads lanes 32 (w % 4) ...)
__device__ __forceinline__ void um_tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
__device__ __forceinline__ void um_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void um_tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

// ---- UMMA descriptors ----
// shared-memory matrix descriptor: K-major operand tile of 128-byte rows, SWIZZLE_128B, 8-row groups
// 1024 B apart (SBO), descriptor version 1 (sm_100); the tile base must be 1024-byte aligned and the
// K step inside the 128-byte row is taken by advancing the start address (32 B per MMA)
__device__ __forceinline__ uint64_t um_desc_sw128(uint32_t saddr) {
  const uint32_t lo = (saddr >> 4) & 0x3FFFu;
  const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}
// the same tile read MN-major (the 64 elements of a 128-byte row run along M / N, the rows along K): the MMA's K step
// of 16 is two 8-row groups SBO = 1024 B apart; `lbo` = distance to the tile holding the next 64 M / N elements
__device__ __forceinline__ uint64_t um_desc_sw128_mn(uint32_t saddr, uint32_t lbo) {
  const uint32_t lo = ((saddr >> 4) & 0x3FFFu) | (((lbo >> 4) & 0x3FFFu) << 16);
  const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}
// byte offset of 16-byte chunk c (0..7) of row r in such a tile
__host__ __device__ __forceinline__ uint32_t um_sw128_off(int r, int c) {
  return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}

// instruction descriptors: D fp32, A / B both K-major, dense, no negate
__host__ __device__ constexpr uint32_t um_idesc_f16(int M, int N) {   // kind::f16, A and B fp16
  return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__host__ __device__ constexpr uint32_t um_idesc_f16_mn(int M, int N) {   // kind::f16, A and B both MN-major
  return um_idesc_f16(M, N) | (1u << 15) | (1u << 16);
}
__host__ __device__ constexpr uint32_t um_idesc_tf32(int M, int N) {  // kind::tf32
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void um_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
// arrive on `mbar` once every tcgen05.mma this thread issued so far has completed
__device__ __forceinline__ void um_commit(uint32_t mbar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(mbar) : "memory");
}

// ---- fp16 two-piece split: x = hi + lo with hi = fp16(x), lo = fp16(x - hi): 22 mantissa bits ----
__device__ __forceinline__ void um_split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 f = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}
__device__ __forceinline__ void um_sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 um_lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
// split 8 consecutive K elements of one row and store them as chunk `off` of the hi tile and of the
// lo tile `lo_delta` bytes after it
__device__ __forceinline__ void um_put8(uint32_t hi_addr, uint32_t lo_delta, const float (&v)[8]) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) um_split2(v[2 * i], v[2 * i + 1], h[i], l[i]);
  um_sts128(hi_addr, h[0], h[1], h[2], h[3]);
  um_sts128(hi_addr + lo_delta, l[0], l[1], l[2], l[3]);
}

// Sums of 32 columns (one per register) over groups of 2^LOGG consecutive lanes.  Afterwards registers
// v[0 .. (32 >> LOGG) - 1] of a lane hold the group sums of columns (lane & (2^LOGG - 1)) * (32 >> LOGG) + i.
template <int LOGG>
__device__ __forceinline__ void um_seg_reduce(float (&v)[32], int lane) {
  int n = 32;
#pragma unroll
  for (int t = 0; t < LOGG; ++t) {
    const int off = 1 << (LOGG - 1 - t);
    n >>= 1;
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (i < n) {
        const float keep = up ? v[i + n] : v[i];
        const float send = up ? v[i] : v[i + n];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
  }
}


}  // namespace asvrl
