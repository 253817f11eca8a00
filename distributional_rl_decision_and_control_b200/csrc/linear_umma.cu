// linear_umma.cu — K8: y = [relu](x * W^T [+ bias]) for the policy MLPs' dense layers, fp32 in / fp32 out,
// on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM) as a THREE-PASS
// TF32 SPLIT: x = x_hi + x_lo, W = W_hi + W_lo (each part tf32-representable),
//     x * W^T ~= x_lo * W_hi^T + x_hi * W_lo^T + x_hi * W_hi^T          (fp32 accumulation)
// whose error is ~2^-21 |x| |W| per product — the level of an fp32 SIMT GEMM (measured on a 128 x 128 x 64
// tile: 3e-7 relative, against 3e-4 for one pass; profiles/probes/umma_tf32_probe.cu).  The layer
// replaces, in the collection path, a cuBLAS sgemm plus a bias / ReLU pass (reference lines
// rfarl/rfarl/policy/AC_IQN_model.py:315-316: relu(hidden_layer(f)), relu(hidden_layer_2(f))).
//
// One CTA (128 threads) owns 128 rows of x and all out_dim <= 256 columns.  Per 32-wide K atom the
// threads stage x and W from global memory through registers into shared memory, already split
// into hi / lo parts and in the 128-byte-swizzled K-major layout the MMA reads (TMA cannot split
// on the fly, so the staging is done by the threads; W is re-read from L2 by every CTA).  One
// thread issues the 4 x 3 MMAs of the atom and commits them to an mbarrier; while they run the
// CTA's threads already have the next atom's global loads in flight.  Epilogue: tcgen05.ld of the
// thread's row, bias, ReLU, 64-byte stores.  At out_dim = 128 three CTAs per SM (64 KB shared memory and
// 128 TMEM columns each) overlap one another's staging, MMA and epilogue phases.
// Bound: HBM — 4 (in_dim + out_dim) bytes per row; the MMA work is ~3 % of the tensor pipe.
#include "common.cuh"

namespace asvrl {

constexpr int LU_TPB = 128;
constexpr int LU_ROWS = 128;          // rows of x per CTA = UMMA M
constexpr int LU_ATOM_K = 32;         // fp32 / tf32 elements in one 128-byte swizzle row
constexpr int LU_UMMA_K = 8;          // elements per tcgen05.mma kind::tf32
constexpr int LU_MAX_N = 256;

__device__ __forceinline__ uint32_t lu_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// round to tf32's 10-bit mantissa, nearest with ties away from zero: what cvt.rna.tf32.f32 returns for
// finite values, in two integer instructions (ptxas expands the cvt into ~8 with its NaN / inf cases;
// it was a third of the kernel's instructions, profiles/r01_k8_v21_source.txt)
__device__ __forceinline__ float lu_tf32(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}

__device__ __forceinline__ void lu_sts(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row groups 1024 B apart, version 1
__device__ __forceinline__ uint64_t lu_desc(uint32_t saddr) {
  const uint32_t lo = (saddr >> 4) & 0x3FFFu;
  const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}

__device__ __forceinline__ void lu_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

// device status word of K8 (0 = ok), read back by asvrl_linear_tf32x3_status: a wait that expires is reported
// there — the launch finishes with unspecified output instead of hanging or killing the caller's context
static __device__ int lu_status_word = 0;

__device__ __forceinline__ void lu_wait(uint32_t mbar, uint32_t parity) {
  uint32_t done = 0;
  for (int it = 0; it < (1 << 26) && !done; ++it)
    asm volatile("{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
                 : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
  if (!done) atomicCAS(&lu_status_word, 0, 4 /* UM_TIMEOUT_K8 */);  // an MMA that never completes: reported, not trapped
}

// split + swizzled store of one 16-byte chunk (row r, chunk c of the atom)
__device__ __forceinline__ void lu_put(uint32_t hi, uint32_t lo, int r, int c, float4 v) {
  float4 h, l;
  h.x = lu_tf32(v.x); h.y = lu_tf32(v.y); h.z = lu_tf32(v.z); h.w = lu_tf32(v.w);
  l.x = lu_tf32(v.x - h.x); l.y = lu_tf32(v.y - h.y); l.z = lu_tf32(v.z - h.z); l.w = lu_tf32(v.w - h.w);
  const int off = (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4);
  lu_sts(hi + off, h);
  lu_sts(lo + off, l);
}

// NB = out_dim / 16 chunks-of-rows of W each thread stages: W rows = out_dim, 8 chunks per row,
// 128 threads -> out_dim * 8 / 128 = out_dim / 16 float4 per thread
template <int TMEM_COLS>
__global__ void __launch_bounds__(LU_TPB, 3)
linear_tf32x3_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                     int relu, int rows, int K, int N, float* __restrict__ out) {
  extern __shared__ __align__(1024) char lu_raw[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t mbar_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  char* base = (char*)(((uintptr_t)lu_raw + 1023) & ~(uintptr_t)1023);
  const int a_bytes = LU_ROWS * 128, b_bytes = N * 128;
  const uint32_t a_hi = lu_smem(base), a_lo = a_hi + a_bytes, b_hi = a_hi + 2 * a_bytes, b_lo = b_hi + b_bytes;
  const uint32_t mbar = lu_smem(&mbar_s);
  const long row0 = (long)blockIdx.x * LU_ROWS;

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(lu_smem(&tmem_slot)), "n"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(mbar));
    asm volatile("fence.mbarrier_init.release.cluster;\n");
  }

  // register staging of one K atom: 8 float4 of x (128 rows x 8 chunks / 128 threads) and up to
  // 16 float4 of W (N rows x 8 chunks / 128 threads)
  constexpr int XA = LU_ROWS * 8 / LU_TPB, WA_MAX = LU_MAX_N * 8 / LU_TPB;
  float4 xr[XA], wr[WA_MAX];
  const int wa = N * 8 / LU_TPB;
  // thread t stages chunk (t & 7) of rows (t >> 3) + 16 i
  const float* xp = x + (row0 + (tid >> 3)) * K + (tid & 7) * 4;
  const float* wp = w + (long)(tid >> 3) * K + (tid & 7) * 4;
  const long step16 = 16L * K;
  const int my_rows = (int)(rows - row0 - (tid >> 3) < 0 ? 0 : (rows - row0 - (tid >> 3) + 15) / 16);  // valid i
  auto load_atom = [&](int kb) {
    const float* xa = xp + kb * LU_ATOM_K;
    const float* wa_p = wp + kb * LU_ATOM_K;
#pragma unroll
    for (int i = 0; i < XA; ++i)
      xr[i] = i < my_rows ? __ldg(reinterpret_cast<const float4*>(xa + i * step16)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < WA_MAX; ++i)
      if (i < wa) wr[i] = __ldg(reinterpret_cast<const float4*>(wa_p + i * step16));
  };
  auto store_atom = [&]() {
#pragma unroll
    for (int i = 0; i < XA; ++i) {
      const int q = i * LU_TPB + tid;
      lu_put(a_hi, a_lo, q >> 3, q & 7, xr[i]);
    }
#pragma unroll
    for (int i = 0; i < WA_MAX; ++i)
      if (i < wa) {
        const int q = i * LU_TPB + tid;
        lu_put(b_hi, b_lo, q >> 3, q & 7, wr[i]);
      }
  };

  const int nkb = K / LU_ATOM_K;
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(LU_ROWS >> 4) << 24);
  load_atom(0);
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tmem = tmem_slot;

  for (int kb = 0; kb < nkb; ++kb) {
    if (kb > 0) lu_wait(mbar, (uint32_t)((kb - 1) & 1));   // the MMAs still reading the tiles are done
    store_atom();
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;\n");
    __syncthreads();
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;\n");
      const uint32_t ah = a_hi, al = a_lo, bh = b_hi, bl = b_lo;
#pragma unroll
      for (int ks = 0; ks < LU_ATOM_K / LU_UMMA_K; ++ks) {
        const uint32_t o = ks * LU_UMMA_K * 4;
        lu_mma(tmem, lu_desc(al + o), lu_desc(bh + o), idesc, (kb | ks) ? 1u : 0u);   // small terms first
        lu_mma(tmem, lu_desc(ah + o), lu_desc(bl + o), idesc, 1u);
        lu_mma(tmem, lu_desc(ah + o), lu_desc(bh + o), idesc, 1u);
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(mbar) : "memory");
    }
    if (kb + 1 < nkb) load_atom(kb + 1);   // global loads in flight while the tensor pipe works
  }
  lu_wait(mbar, (uint32_t)((nkb - 1) & 1));
  asm volatile("tcgen05.fence::after_thread_sync;\n");

  // epilogue: warp w owns TMEM lanes [32 w, 32 w + 32) = rows; 16 columns per load
  const long row = row0 + warp * 32 + lane;
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t v[16];
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
    if (row < rows) {
      float4* dst = reinterpret_cast<float4*>(out + row * N + c0);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float4 o = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                               __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
        if (bias) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(bias + c0) + i);
          o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
        }
        if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
        dst[i] = o;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "n"(TMEM_COLS));
}

}  // namespace asvrl

using namespace asvrl;

extern "C" int asvrl_linear_tf32x3(const float* x, const float* w, const float* bias, int relu, long rows,
                                   int in_dim, int out_dim, float* out, void* stream) {
  if (!x || !w || !out) return fail(ASVRL_E_BADARG, "asvrl_linear_tf32x3: null pointer");
  if (rows < 1 || in_dim < LU_ATOM_K || in_dim % LU_ATOM_K || out_dim < 16 || out_dim % 16 || out_dim > LU_MAX_N)
    return fail(ASVRL_E_BADARG, "asvrl_linear_tf32x3: rows %ld, in_dim %d (multiple of 32), out_dim %d (multiple of 16, <= 256)",
                rows, in_dim, out_dim);
  if (((uintptr_t)x | (uintptr_t)w | (uintptr_t)out | (uintptr_t)bias) & 15)
    return fail(ASVRL_E_BADARG, "asvrl_linear_tf32x3: pointers must be 16-byte aligned");
  const long grid = (rows + LU_ROWS - 1) / LU_ROWS;
  if (grid > 0x7FFFFFFFL) return fail(ASVRL_E_BADARG, "asvrl_linear_tf32x3: too many rows");
  const size_t smem = 2 * (size_t)LU_ROWS * 128 + 2 * (size_t)out_dim * 128 + 1024;
  auto launch = [&](auto kern) -> int {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "asvrl_linear_tf32x3: %zu B shared memory: %s", smem, cudaGetErrorString(e));
    kern<<<(int)grid, LU_TPB, smem, (cudaStream_t)stream>>>(x, w, bias, relu, (int)rows, in_dim, out_dim, out);
    return check_launch("asvrl_linear_tf32x3");
  };
  if (rows > 0x7FFFFFFFL) return fail(ASVRL_E_BADARG, "asvrl_linear_tf32x3: too many rows");
  if (out_dim <= 32) return launch(linear_tf32x3_kernel<32>);
  if (out_dim <= 64) return launch(linear_tf32x3_kernel<64>);
  if (out_dim <= 128) return launch(linear_tf32x3_kernel<128>);
  return launch(linear_tf32x3_kernel<256>);
}

// synchronous read of K8's device status word (0 = every bounded wait completed)
extern "C" int asvrl_linear_tf32x3_status(int* status_out) {
  if (!status_out) return fail(ASVRL_E_BADARG, "asvrl_linear_tf32x3_status: null pointer");
  cudaError_t e = cudaMemcpyFromSymbol(status_out, lu_status_word, sizeof(int));
  if (e != cudaSuccess) return fail((int)e, "asvrl_linear_tf32x3_status: %s", cudaGetErrorString(e));
  return 0;
}
