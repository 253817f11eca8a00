// Stand-alone probe for DESIGN.md section 8 items 1 / 2 (NOT part of libasvrl.so): one 128 x 128 x 64 tile
// D = A * B^T on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM) from
// fp32 operands that the CTA's own threads stage in shared memory in the 128-byte-swizzled K-major
// layout (no TMA: the operands are split into tf32 hi + lo parts on the way in, which TMA cannot do).
// Two results are produced: the plain one-pass TF32 product and the three-pass split product
// (Alo*Bhi + Ahi*Blo + Ahi*Bhi) that recovers fp32-level accuracy.  Both are compared with a float64
// host reference.  Every wait is bounded: a wrong descriptor makes the probe report, not hang.
//
// build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_tf32_probe umma_tf32_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

constexpr int M = 128, N = 128, K = 64;
constexpr int ATOM_K = 32;                 // tf32 elements in one 128-byte swizzle row
constexpr int ATOM_BYTES = 128 * 128;      // 128 rows x 128 B
constexpr int UMMA_K = 8;                  // tf32 elements per tcgen05.mma (32 B)
constexpr int TMEM_COLS = 256;             // [0,128): one pass, [128,256): three passes

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// shared-memory matrix descriptor, K-major, SWIZZLE_128B: 8-row groups 1024 B apart (SBO), LBO unused
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  const uint32_t lo = (saddr >> 4) & 0x3FFFu;
  const uint32_t hi = (1024u >> 4) | (1u << 14) /* descriptor version 1 (sm_100) */ | (2u << 29) /* SWIZZLE_128B */;
  return ((uint64_t)hi << 32) | lo;
}

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// stage one fp32 operand (rows x K, K contiguous) as tf32 hi / lo tiles in the swizzled layout
__device__ void stage(const float* __restrict__ g, char* hi, char* lo, int tid) {
  for (int i = tid; i < 128 * (K / 4); i += 128) {
    const int r = i / (K / 4), q = i % (K / 4);       // q: 16-byte chunk along K
    const int kb = q / 8, c = q % 8;
    const float4 v = *reinterpret_cast<const float4*>(g + (size_t)r * K + q * 4);
    float4 h, l;
    h.x = to_tf32(v.x); h.y = to_tf32(v.y); h.z = to_tf32(v.z); h.w = to_tf32(v.w);
    l.x = to_tf32(v.x - h.x); l.y = to_tf32(v.y - h.y); l.z = to_tf32(v.z - h.z); l.w = to_tf32(v.w - h.w);
    const int off = kb * ATOM_BYTES + (r / 8) * 1024 + (r % 8) * 128 + ((c ^ (r % 8)) * 16);
    *reinterpret_cast<float4*>(hi + off) = h;
    *reinterpret_cast<float4*>(lo + off) = l;
  }
}

__global__ void __launch_bounds__(128, 1)
probe_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D1,
             float* __restrict__ D3, int* __restrict__ status) {
  extern __shared__ __align__(1024) char smem[];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(8) uint64_t mbar;
  const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
  char* base = (char*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);
  char *a_hi = base, *a_lo = base + 2 * ATOM_BYTES, *b_hi = base + 4 * ATOM_BYTES, *b_lo = base + 6 * ATOM_BYTES;

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tmem_base_slot)), "n"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;\n");
  }
  stage(A, a_hi, a_lo, tid);
  stage(B, b_hi, b_lo, tid);
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");   // generic-proxy stores -> visible to the MMA's async proxy
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tmem_base_slot;

  if (tid == 0) {
    // instruction descriptor: D = f32, A = B = tf32, both K-major, N = 128, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
    uint32_t acc1 = 0, acc3 = 0;
    for (int kb = 0; kb < K / ATOM_K; ++kb)
      for (int ks = 0; ks < ATOM_K / UMMA_K; ++ks) {
        const uint32_t o = kb * ATOM_BYTES + ks * UMMA_K * 4;
        umma_tf32(tmem, make_desc(ah + o), make_desc(bh + o), idesc, acc1);
        acc1 = 1;
        umma_tf32(tmem + 128, make_desc(al + o), make_desc(bh + o), idesc, acc3);   // small terms first
        acc3 = 1;
        umma_tf32(tmem + 128, make_desc(ah + o), make_desc(bl + o), idesc, 1);
      }
    for (int kb = 0; kb < K / ATOM_K; ++kb)
      for (int ks = 0; ks < ATOM_K / UMMA_K; ++ks) {
        const uint32_t o = kb * ATOM_BYTES + ks * UMMA_K * 4;
        umma_tf32(tmem + 128, make_desc(ah + o), make_desc(bh + o), idesc, 1);
      }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(&mbar)) : "memory");
  }
  // bounded wait for the MMAs (phase 0)
  uint32_t done = 0;
  for (int it = 0; it < (1 << 22) && !done; ++it)
    asm volatile("{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
                 : "=r"(done) : "r"(smem_u32(&mbar)), "r"(0u) : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  if (!done && tid == 0) *status = 1;
  if (__syncthreads_and(done)) {
    // warp w owns TMEM lanes [32 w, 32 w + 32) = rows of D; 8 columns per load
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < TMEM_COLS; c0 += 8) {
      uint32_t v[8];
      const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                   : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
      float* dst = (c0 < 128 ? D1 : D3) + (size_t)row * N + (c0 & 127);
      for (int i = 0; i < 8; ++i) dst[i] = __uint_as_float(v[i]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "n"(TMEM_COLS));
}

int main() {
  std::vector<float> A(M * K), B(N * K);
  uint32_t s = 12345u;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) * (1.0f / 16777216.0f)) * 2.0f - 1.0f; };
  for (auto& x : A) x = rnd();
  for (auto& x : B) x = rnd();
  float *dA, *dB, *d1, *d3;
  int* dst;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4);
  cudaMalloc(&d1, M * N * 4); cudaMalloc(&d3, M * N * 4); cudaMalloc(&dst, 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
  cudaMemset(d1, 0xFF, M * N * 4); cudaMemset(d3, 0xFF, M * N * 4); cudaMemset(dst, 0, 4);
  const int smem = 8 * ATOM_BYTES + 1024;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe_kernel<<<1, 128, smem>>>(dA, dB, d1, d3, dst);
  cudaError_t e = cudaDeviceSynchronize();
  int st = 0;
  cudaMemcpy(&st, dst, 4, cudaMemcpyDeviceToHost);
  printf("launch: %s, mbarrier %s\n", cudaGetErrorString(e), st ? "TIMED OUT" : "completed");
  if (e != cudaSuccess || st) return 1;
  std::vector<float> D1(M * N), D3(M * N);
  cudaMemcpy(D1.data(), d1, M * N * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(D3.data(), d3, M * N * 4, cudaMemcpyDeviceToHost);
  double e1 = 0, e3 = 0, ref_max = 0;
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      double r = 0;
      for (int k = 0; k < K; ++k) r += (double)A[i * K + k] * (double)B[j * K + k];
      ref_max = fmax(ref_max, fabs(r));
      e1 = fmax(e1, fabs(D1[i * N + j] - r));
      e3 = fmax(e3, fabs(D3[i * N + j] - r));
    }
  printf("128x128x64 tile: max|ref| %.3f, max abs error one-pass tf32 %.3e, three-pass split %.3e\n", ref_max, e1, e3);
  printf("D1[0][0..3] = %f %f %f %f\n", D1[0], D1[1], D1[2], D1[3]);
  const bool ok = e1 < 2e-2 && e3 < 2e-5;
  printf("%s\n", ok ? "PASS" : "FAIL");
  return ok ? 0 : 2;
}
