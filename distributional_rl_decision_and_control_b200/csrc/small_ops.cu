// small_ops.cu — the two narrow layers at the ends of the AC-IQN update's dense chain, forward and backward in
// one launch each (sm_100a).  Both have an input or output of width 2, which cuBLAS serves badly (a 128 x 2 weight
// gradient over 1024 rows took 27 us) and autograd turns into a string of element-wise launches:
//
//   action encoder  rfarl/rfarl/policy/AC_IQN_model.py:399, 472   aenc = relu(Wa a + ba),  a in R^2 -> R^D
//   actor head      rfarl/rfarl/policy/AC_IQN_model.py:318-321    act = s atan(W3 h + b3),  h in R^D -> R^2
//
// The batch is the learner's B sampled states (1024), so every kernel is latency bound: the forward kernels are
// one thread per output element / one warp per row; the backward kernels give every CTA >= 16 rows, keep per-CTA
// partial column sums and let the last CTA add them in CTA order (deterministic), replacing per launch:
//   action encoder forward   sgemm + K5                       (2 launches)
//   action encoder backward  K5 backward + bmm + sum (+ gemm) (3 - 4)
//   actor head forward       addmm + atan + mul               (3)
//   actor head backward      4 element-wise + gemm + bmm + 2 sums (8)
#include "common.cuh"

namespace asvrl {

constexpr int SO_T = 1024;

// aenc[b][j] = relu(wa[j][0] a[b][0] + wa[j][1] a[b][1] + ba[j])
__global__ void __launch_bounds__(256)
action_encoder_fwd_kernel(const float2* __restrict__ act, const float2* __restrict__ wa, const float* __restrict__ ba,
                          int B, int D, float* __restrict__ aenc) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= B * D) return;
  const int b = i / D, j = i - b * D;
  const float2 a = act[b], w = wa[j];
  aenc[i] = fmaxf(fmaf(w.y, a.y, fmaf(w.x, a.x, ba[j])), 0.0f);
}

// The last CTA to arrive (ticket in `counter`, left at zero) returns true with the others' partials visible.
__device__ __forceinline__ bool so_last_cta(uint32_t* counter) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

// dpre = daenc . 1[aenc > 0];  dwa[j][q] = sum_b dpre[b][j] a[b][q];  dba[j] = sum_b dpre[b][j];
// dact[b][q] = sum_j dpre[b][j] wa[j][q]  (nullable).  A CTA owns `rpc` consecutive rows; per-CTA partial sums
// part[cta][3][D] are added by the last CTA in CTA order (deterministic).  D <= 256, SO_T % D == 0.
__global__ void __launch_bounds__(SO_T)
action_encoder_bwd_kernel(const float* __restrict__ daenc, const float* __restrict__ aenc, const float2* __restrict__ act,
                          const float2* __restrict__ wa, int B, int D, int rpc, float2* __restrict__ dwa,
                          float* __restrict__ dba, float2* __restrict__ dact, float* __restrict__ part,
                          uint32_t* __restrict__ counter) {
  __shared__ float sm[3 * SO_T];
  const int tid = threadIdx.x, j = tid % D, p0 = tid / D, parts = SO_T / D;
  const int b0 = blockIdx.x * rpc, b1 = min(B, b0 + rpc);
  float s0 = 0.0f, s1 = 0.0f, sb = 0.0f;
  for (int b = b0 + p0; b < b1; b += parts) {
    const size_t i = (size_t)b * D + j;
    const float g = aenc[i] > 0.0f ? daenc[i] : 0.0f;
    const float2 a = act[b];
    s0 = fmaf(g, a.x, s0);
    s1 = fmaf(g, a.y, s1);
    sb += g;
  }
  sm[tid] = s0; sm[SO_T + tid] = s1; sm[2 * SO_T + tid] = sb;
  __syncthreads();
  if (p0 == 0) {
    for (int p = 1; p < parts; ++p) { s0 += sm[p * D + j]; s1 += sm[SO_T + p * D + j]; sb += sm[2 * SO_T + p * D + j]; }
    float* pr = part + (size_t)blockIdx.x * 3 * D;
    pr[j] = s0; pr[D + j] = s1; pr[2 * D + j] = sb;
  }
  if (dact) {
    const int lane = tid & 31, warp = tid >> 5;
    for (int b = b0 + warp; b < b1; b += SO_T / 32) {
      float d0 = 0.0f, d1 = 0.0f;
      for (int c = lane; c < D; c += 32) {
        const size_t i = (size_t)b * D + c;
        const float g = aenc[i] > 0.0f ? daenc[i] : 0.0f;
        const float2 w = wa[c];
        d0 = fmaf(g, w.x, d0);
        d1 = fmaf(g, w.y, d1);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        d0 += __shfl_xor_sync(0xffffffffu, d0, o);
        d1 += __shfl_xor_sync(0xffffffffu, d1, o);
      }
      if (lane == 0) dact[b] = make_float2(d0, d1);
    }
  }
  if (!so_last_cta(counter)) return;
  if (tid < D) {
    float t0 = 0.0f, t1 = 0.0f, tb = 0.0f;
#pragma unroll 8
    for (int g = 0; g < (int)gridDim.x; ++g) {
      const float* pr = part + (size_t)g * 3 * D;
      t0 += __ldcg(pr + tid); t1 += __ldcg(pr + D + tid); tb += __ldcg(pr + 2 * D + tid);
    }
    dwa[tid] = make_float2(t0, t1);
    dba[tid] = tb;
  }
  if (tid == 0) *counter = 0u;
}

// y[b][q] = w3[q] . h[b] + b3[q];  act[b][q] = s atan(y[b][q]).  One warp per row.
__global__ void __launch_bounds__(256)
atan_head_fwd_kernel(const float* __restrict__ h, const float* __restrict__ w3, const float* __restrict__ b3, float s,
                     int B, int D, float2* __restrict__ y, float2* __restrict__ act) {
  const int lane = threadIdx.x & 31;
  const int b = (blockIdx.x * 256 + threadIdx.x) >> 5;
  if (b >= B) return;
  float a0 = 0.0f, a1 = 0.0f;
  for (int c = lane; c < D; c += 32) {
    const float v = h[(size_t)b * D + c];
    a0 = fmaf(v, w3[c], a0);
    a1 = fmaf(v, w3[D + c], a1);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
  }
  if (lane == 0) {
    const float y0 = a0 + b3[0], y1 = a1 + b3[1];
    y[b] = make_float2(y0, y1);
    act[b] = make_float2(s * atanf(y0), s * atanf(y1));
  }
}

// dy = dact . s / (1 + y^2);  dh[b] = dy[b][0] w3[0] + dy[b][1] w3[1];  dw3[q][c] = sum_b dy[b][q] h[b][c];
// db3[q] = sum_b dy[b][q].  CTAs of `rpc` rows, partials part[cta][2 D + 2], last CTA sums in a fixed order.
__global__ void __launch_bounds__(SO_T)
atan_head_bwd_kernel(const float2* __restrict__ dact, const float2* __restrict__ y, const float* __restrict__ h,
                     const float* __restrict__ w3, float s, int B, int D, int rpc, float* __restrict__ dh,
                     float* __restrict__ dw3, float* __restrict__ db3, float* __restrict__ part,
                     uint32_t* __restrict__ counter) {
  __shared__ float sm[2 * SO_T];
  const int tid = threadIdx.x, c = tid % D, p0 = tid / D, parts = SO_T / D;
  const int b0 = blockIdx.x * rpc, b1 = min(B, b0 + rpc);
  const float w0 = w3[c], w1 = w3[D + c];
  float s0 = 0.0f, s1 = 0.0f, t0 = 0.0f, t1 = 0.0f;
  for (int b = b0 + p0; b < b1; b += parts) {
    const float2 g = dact[b], yy = y[b];
    const float d0 = g.x * s / fmaf(yy.x, yy.x, 1.0f), d1 = g.y * s / fmaf(yy.y, yy.y, 1.0f);
    const size_t i = (size_t)b * D + c;
    const float hv = h[i];
    if (dh) dh[i] = fmaf(d1, w1, d0 * w0);
    s0 = fmaf(d0, hv, s0);
    s1 = fmaf(d1, hv, s1);
    t0 += d0;   // the same in every column thread; column 0's copy is the one kept
    t1 += d1;
  }
  __shared__ float st[2 * 32];   // parts <= 32
  sm[tid] = s0; sm[SO_T + tid] = s1;
  if (c == 0) { st[p0] = t0; st[32 + p0] = t1; }
  __syncthreads();
  float* pr = part + (size_t)blockIdx.x * (2 * D + 2);
  if (p0 == 0) {
    for (int p = 1; p < parts; ++p) { s0 += sm[p * D + c]; s1 += sm[SO_T + p * D + c]; }
    pr[c] = s0; pr[D + c] = s1;
  }
  if (tid < 2) {
    float t = 0.0f;
    for (int p = 0; p < parts; ++p) t += st[32 * tid + p];
    pr[2 * D + tid] = t;
  }
  if (!so_last_cta(counter)) return;
  if (tid < D) {
    float a0 = 0.0f, a1 = 0.0f;
#pragma unroll 8
    for (int g = 0; g < (int)gridDim.x; ++g) {
      const float* q = part + (size_t)g * (2 * D + 2);
      a0 += __ldcg(q + tid); a1 += __ldcg(q + D + tid);
    }
    dw3[tid] = a0;
    dw3[D + tid] = a1;
  }
  if (tid >= SO_T - 64) {   // the last two warps: db3[0], db3[1] over the <= 64 CTAs, fixed shuffle tree
    const int q2 = (tid - (SO_T - 64)) >> 5, lane = tid & 31;
    float t = 0.0f;
    for (int g = lane; g < (int)gridDim.x; g += 32) t += __ldcg(part + (size_t)g * (2 * D + 2) + 2 * D + q2);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) db3[q2] = t;
  }
  if (tid == 0) *counter = 0u;
}

// out[0] = -mean(x): the actor loss (AGENT:423, "-Q.mean()") by ONE CTA in a fixed order (torch: mean + neg = two
// launches, and their autograd nodes cost four more; the loss' gradient is the constant -1 / n, which the caller keeps).
__global__ void __launch_bounds__(SO_T)
neg_mean_kernel(const float* __restrict__ x, long n, float* __restrict__ out) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float acc = 0.0f;
  if ((reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const long n4 = n >> 2;
    for (long i0 = 0; i0 < n4; i0 += 8L * SO_T) {
      float4 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const long i = i0 + (long)k * SO_T + tid;
        v[k] = i < n4 ? __ldg(x4 + i) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) acc += (v[k].x + v[k].y) + (v[k].z + v[k].w);
    }
    for (long i = (n4 << 2) + tid; i < n; i += SO_T) acc += __ldg(x + i);
  } else {
    for (long i = tid; i < n; i += SO_T) acc += __ldg(x + i);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float wsum[SO_T / 32];
  if (lane == 0) wsum[warp] = acc;
  __syncthreads();
  if (warp == 0) {
    float t = lane < SO_T / 32 ? wsum[lane] : 0.0f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) out[0] = -t / (float)n;
  }
}

}  // namespace asvrl

using namespace asvrl;

static int so_dims(const char* what, int batch, int dim) {
  if (batch < 1 || dim < 32 || dim > 256 || SO_T % dim != 0)
    return fail(ASVRL_E_BADARG, "%s: batch >= 1 and dim a divisor of %d in [32, 256] (B=%d D=%d)", what, SO_T, batch, dim);
  return 0;
}

extern "C" int asvrl_action_encoder_fwd(const float* actions, const float* wa, const float* ba, int batch, int dim,
                                        float* aenc, void* stream) {
  if (!actions || !wa || !ba || !aenc) return fail(ASVRL_E_BADARG, "asvrl_action_encoder_fwd: null pointer");
  if (int rc = so_dims("asvrl_action_encoder_fwd", batch, dim)) return rc;
  const long n = (long)batch * dim;
  action_encoder_fwd_kernel<<<(int)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const float2*)actions, (const float2*)wa, ba, batch, dim, aenc);
  return check_launch("asvrl_action_encoder_fwd");
}

static int so_grid(int batch, int* rpc) {   // CTAs of >= 16 rows, at most 64 of them
  int g = (batch + 15) / 16;
  g = g > 64 ? 64 : (g < 1 ? 1 : g);
  *rpc = (batch + g - 1) / g;
  return (batch + *rpc - 1) / *rpc;
}

extern "C" long asvrl_narrow_bwd_partials(int dim) { return 64L * (3 * (long)dim + 64); }   /* floats */

extern "C" int asvrl_action_encoder_bwd(const float* daenc, const float* aenc, const float* actions, const float* wa,
                                        int batch, int dim, float* dwa, float* dba, float* dact, float* partials,
                                        uint32_t* counter, void* stream) {
  if (!daenc || !aenc || !actions || !wa || !dwa || !dba || !partials || !counter)
    return fail(ASVRL_E_BADARG, "asvrl_action_encoder_bwd: null pointer");
  if (int rc = so_dims("asvrl_action_encoder_bwd", batch, dim)) return rc;
  int rpc;
  const int grid = so_grid(batch, &rpc);
  action_encoder_bwd_kernel<<<grid, SO_T, 0, (cudaStream_t)stream>>>(daenc, aenc, (const float2*)actions, (const float2*)wa,
                                                                    batch, dim, rpc, (float2*)dwa, dba, (float2*)dact,
                                                                    partials, counter);
  return check_launch("asvrl_action_encoder_bwd");
}

extern "C" int asvrl_atan_head_fwd(const float* h, const float* w3, const float* b3, float atan_scale, int batch, int dim,
                                   float* y, float* actions, void* stream) {
  if (!h || !w3 || !b3 || !y || !actions) return fail(ASVRL_E_BADARG, "asvrl_atan_head_fwd: null pointer");
  if (int rc = so_dims("asvrl_atan_head_fwd", batch, dim)) return rc;
  atan_head_fwd_kernel<<<(batch * 32 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(h, w3, b3, atan_scale, batch, dim,
                                                                                   (float2*)y, (float2*)actions);
  return check_launch("asvrl_atan_head_fwd");
}

extern "C" int asvrl_atan_head_bwd(const float* dact, const float* y, const float* h, const float* w3, float atan_scale,
                                   int batch, int dim, float* dh, float* dw3, float* db3, float* partials,
                                   uint32_t* counter, void* stream) {
  if (!dact || !y || !h || !w3 || !dw3 || !db3 || !partials || !counter)
    return fail(ASVRL_E_BADARG, "asvrl_atan_head_bwd: null pointer");
  if (int rc = so_dims("asvrl_atan_head_bwd", batch, dim)) return rc;
  int rpc;
  const int grid = so_grid(batch, &rpc);
  atan_head_bwd_kernel<<<grid, SO_T, 0, (cudaStream_t)stream>>>((const float2*)dact, (const float2*)y, h, w3, atan_scale,
                                                               batch, dim, rpc, dh, dw3, db3, partials, counter);
  return check_launch("asvrl_atan_head_bwd");
}

extern "C" int asvrl_neg_mean(const float* x, long n, float* out, void* stream) {
  if (!x || !out) return fail(ASVRL_E_BADARG, "asvrl_neg_mean: null pointer");
  if (n < 1 || n > (1L << 24)) return fail(ASVRL_E_BADARG, "asvrl_neg_mean: n = %ld outside [1, 2^24]", n);
  neg_mean_kernel<<<1, SO_T, 0, (cudaStream_t)stream>>>(x, n, out);
  return check_launch("asvrl_neg_mean");
}
