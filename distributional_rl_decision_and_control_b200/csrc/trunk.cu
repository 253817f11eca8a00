// trunk.cu — K5: the element-wise / reduction glue of the IQN critic's quantile trunk, forward and
// backward (sm_100a).  The three dense layers stay cuBLAS GEMMs WITHOUT epilogue; everything
// between them is one HBM pass per layer here instead of the ~12 launches autograd makes of it.
//
// Replaces (reference paths under rfarl/rfarl/policy/AC_IQN_model.py):
//   :453-460  x = (features.unsqueeze(1) * relu(cos_embedding(cos))).view(B*N, 256)
//   :470-474  x = relu(hidden_layer(x)); x = (action_encoder(a).unsqueeze(1) * x).view(B*N, 128)
//   :476-480  x = relu(hidden_layer_2(x)); quantiles = output_layer(x).view(B, N)
// and the autograd backward of those lines (bias gradients, ReLU masks, the per-sample reduction
// of the broadcast factor's gradient over the N quantile rows).
//
//   scale_relu : out[m][d] = scale[m / N][d] * relu(pre[m][d] + bias[d])            m < B*N
//   head       : z[m]      = sum_d relu(pre[m][d] + bias[d]) * w[d] + b_out
// Bound: HBM.  scale_relu fwd moves 8 B per element (read pre, write out); bwd 12-16 B (read dout,
// pre; write dpre); head fwd 4 B, bwd 8 B.  All cross-CTA reductions (bias / head-weight
// gradients) are per-CTA partials summed by the last CTA in a fixed order: deterministic.
#include "common.cuh"

namespace asvrl {

constexpr int TR_TPB = 256;
constexpr int TR_RPC = 64;    // rows per CTA in head_bwd

__device__ __forceinline__ float4 f4add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 f4mul(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 f4relu(float4 a) { return make_float4(fmaxf(a.x, 0.f), fmaxf(a.y, 0.f), fmaxf(a.z, 0.f), fmaxf(a.w, 0.f)); }
__device__ __forceinline__ float4 f4gate(float4 a, float4 v) {  // v where a > 0 (ReLU'), else 0
  return make_float4(a.x > 0.f ? v.x : 0.f, a.y > 0.f ? v.y : 0.f, a.z > 0.f ? v.z : 0.f, a.w > 0.f ? v.w : 0.f);
}
__device__ __forceinline__ float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }

// ---------------------------------------------------------------------------------------
// forward: grid-stride over the M * D/4 float4 elements
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TR_TPB)
scale_relu_fwd_kernel(const float4* __restrict__ pre, const float4* __restrict__ bias,
                      const float4* __restrict__ scale, int N, int M, int D4, float4* __restrict__ out) {
  const int total = M * D4;
  for (int i = blockIdx.x * TR_TPB + threadIdx.x; i < total; i += gridDim.x * TR_TPB) {
    const int m = i / D4, c = i - m * D4, b = m / N;
    const float4 a = f4add(pre[i], bias[c]);
    out[i] = f4mul(scale[b * D4 + c], f4relu(a));
  }
}

// Sum over the CTA's row lanes (threadIdx.x / D4) of one float4 per thread; the result is valid
// in the threads of row lane 0.  sm: [RL][D4] float4.
__device__ __forceinline__ float4 lane_sum(float4 v, float4* sm, int c, int rl, int RL, int D4) {
  sm[rl * D4 + c] = v;
  __syncthreads();
  float4 s = f4zero();
  if (rl == 0)
    for (int q = 0; q < RL; ++q) s = f4add(s, sm[q * D4 + c]);
  __syncthreads();
  return s;
}

// The last CTA to arrive sums the per-CTA partial rows part[g][D4] (g < n_part) in a fixed order.
__device__ __forceinline__ bool last_cta(uint32_t* counter) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}
__device__ __forceinline__ void reduce_partials(const float4* part, int n_part, float4* dst, float4* sm, int c,
                                                int rl, int RL, int D4) {
  float4 s = f4zero();
#pragma unroll 8
  for (int g = rl; g < n_part; g += RL) s = f4add(s, __ldcg(part + (size_t)g * D4 + c));
  s = lane_sum(s, sm, c, rl, RL, D4);
  if (rl == 0) dst[c] = s;
}

// ---------------------------------------------------------------------------------------
// backward of scale_relu.  A CTA owns `spb` consecutive samples (2); D/4 threads span a row, the
// 256 / (D/4) row lanes stride over the N quantile rows of a sample.
//   dpre[m][d]   = dout[m][d] * scale[b][d] * 1[pre + bias > 0]          (nullable)
//   dscale[b][d] = sum_n dout[m][d] * relu(pre[m][d] + bias[d])          (nullable)
//   dbias[d]     = sum_m dpre[m][d]                                      (nullable; needs part)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TR_TPB)
scale_relu_bwd_kernel(const float4* __restrict__ dout, const float4* __restrict__ pre,
                      const float4* __restrict__ bias, const float4* __restrict__ scale, int N, int B,
                      int D4, int spb, float4* __restrict__ dpre, float4* __restrict__ dscale,
                      float4* __restrict__ dbias, float4* __restrict__ part, uint32_t* counter) {
  extern __shared__ float4 tr_sm[];
  const int c = threadIdx.x % D4, rl = threadIdx.x / D4, RL = TR_TPB / D4;
  const float4 bi = bias[c];
  float4 db = f4zero();
  for (int s = 0; s < spb; ++s) {
    const int b = blockIdx.x * spb + s;
    if (b >= B) break;  // CTA-uniform
    const float4 sc = scale[b * D4 + c];
    float4 ds = f4zero();
#pragma unroll 4
    for (int n = rl; n < N; n += RL) {
      const size_t i = ((size_t)b * N + n) * D4 + c;
      const float4 g = dout[i];
      const float4 a = f4add(pre[i], bi);
      ds = f4add(ds, f4mul(g, f4relu(a)));
      const float4 dp = f4gate(a, f4mul(g, sc));
      db = f4add(db, dp);
      if (dpre) dpre[i] = dp;
    }
    if (dscale) {
      ds = lane_sum(ds, tr_sm, c, rl, RL, D4);
      if (rl == 0) dscale[b * D4 + c] = ds;
    }
  }
  if (!dbias) return;
  db = lane_sum(db, tr_sm, c, rl, RL, D4);
  if (rl == 0) part[(size_t)blockIdx.x * D4 + c] = db;
  if (last_cta(counter)) {
    reduce_partials(part, gridDim.x, dbias, tr_sm, c, rl, RL, D4);
    if (threadIdx.x == 0) *counter = 0u;  // left zero for the next call
  }
}

// ---------------------------------------------------------------------------------------
// head forward: one warp per row
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TR_TPB)
head_fwd_kernel(const float4* __restrict__ pre, const float4* __restrict__ bias, const float4* __restrict__ w,
                const float* __restrict__ b_out, int M, int D4, float* __restrict__ z) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * TR_TPB) >> 5;
  const float bo = b_out ? b_out[0] : 0.0f;
  for (int m = (blockIdx.x * TR_TPB + threadIdx.x) >> 5; m < M; m += warps) {
    float acc = 0.0f;
    for (int c = lane; c < D4; c += 32) {
      const float4 r = f4relu(f4add(pre[(size_t)m * D4 + c], bias[c]));
      const float4 wv = w[c];
      acc += (r.x * wv.x + r.y * wv.y) + (r.z * wv.z + r.w * wv.w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) z[m] = acc + bo;
  }
}

// ---------------------------------------------------------------------------------------
// head backward.  A CTA owns TR_RPC consecutive rows.
//   dpre[m][d] = dz[m] * w[d] * 1[pre + bias > 0]
//   dw[d]      = sum_m dz[m] * relu(pre[m][d] + bias[d])      (nullable, with dbias / db_out / part)
//   dbias[d]   = sum_m dpre[m][d] ;  db_out = sum_m dz[m]
// part: [grid][2 * D4 + 1] float4 (dw row, dbias row, then db_out in .x of the last one)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TR_TPB)
head_bwd_kernel(const float* __restrict__ dz, const float4* __restrict__ pre, const float4* __restrict__ bias,
                const float4* __restrict__ w, int M, int D4, float4* __restrict__ dpre,
                float4* __restrict__ dw, float4* __restrict__ dbias, float* __restrict__ db_out,
                float4* __restrict__ part, uint32_t* counter) {
  extern __shared__ float4 tr_sm[];
  const int c = threadIdx.x % D4, rl = threadIdx.x / D4, RL = TR_TPB / D4;
  const float4 bi = bias[c], wv = w[c];
  float4 aw = f4zero(), ab = f4zero();
  float ao = 0.0f;
  const int m0 = blockIdx.x * TR_RPC, m1 = min(M, m0 + TR_RPC);
#pragma unroll 4
  for (int m = m0 + rl; m < m1; m += RL) {
    const float g = dz[m];
    const float4 g4 = make_float4(g, g, g, g);
    const size_t i = (size_t)m * D4 + c;
    const float4 a = f4add(pre[i], bi);
    const float4 dp = f4gate(a, f4mul(g4, wv));
    dpre[i] = dp;
    aw = f4add(aw, f4mul(g4, f4relu(a)));
    ab = f4add(ab, dp);
    if (c == 0) ao += g;
  }
  if (!dw) return;
  const int stride = 2 * D4 + 1;
  aw = lane_sum(aw, tr_sm, c, rl, RL, D4);
  ab = lane_sum(ab, tr_sm, c, rl, RL, D4);
  const float4 o4 = lane_sum(make_float4(ao, 0.f, 0.f, 0.f), tr_sm, c, rl, RL, D4);
  if (rl == 0) {
    part[(size_t)blockIdx.x * stride + c] = aw;
    part[(size_t)blockIdx.x * stride + D4 + c] = ab;
    if (c == 0) part[(size_t)blockIdx.x * stride + 2 * D4] = o4;
  }
  if (last_cta(counter)) {
    float4 sw = f4zero(), sb = f4zero();
    float so = 0.0f;
#pragma unroll 4
    for (int g = rl; g < (int)gridDim.x; g += RL) {
      sw = f4add(sw, __ldcg(part + (size_t)g * stride + c));
      sb = f4add(sb, __ldcg(part + (size_t)g * stride + D4 + c));
      if (c == 0) so += __ldcg(part + (size_t)g * stride + 2 * D4).x;
    }
    sw = lane_sum(sw, tr_sm, c, rl, RL, D4);
    sb = lane_sum(sb, tr_sm, c, rl, RL, D4);
    const float4 so4 = lane_sum(make_float4(so, 0.f, 0.f, 0.f), tr_sm, c, rl, RL, D4);
    if (rl == 0) {
      dw[c] = sw;
      dbias[c] = sb;
      if (c == 0) db_out[0] = so4.x;
    }
    if (threadIdx.x == 0) *counter = 0u;
  }
}

}  // namespace asvrl

using namespace asvrl;

static int scale_relu_spb(int n_tau) {  // samples per CTA.  More rows per CTA (fewer partials) measured
  (void)n_tau;                           // slower inside the update's graph: per-sample barriers serialise
  return 2;
}

static int trunk_dims_ok(const char* what, long rows, int dim) {
  if (rows < 1 || dim < 4 || dim % 4 != 0 || dim / 4 > TR_TPB || TR_TPB % (dim / 4) != 0)
    return fail(ASVRL_E_BADARG, "%s: rows >= 1 and dim a multiple of 4 with (dim/4) dividing %d (rows=%ld dim=%d)",
                what, TR_TPB, rows, dim);
  if (rows * (long)(dim / 4) >= (1L << 31)) return fail(ASVRL_E_LIMIT, "%s: rows * dim / 4 must stay below 2^31", what);
  return 0;
}

extern "C" int asvrl_iqn_scale_relu_fwd(const float* pre, const float* bias, const float* scale, int batch,
                                        int n_tau, int dim, float* out, void* stream) {
  if (!pre || !bias || !scale || !out) return fail(ASVRL_E_BADARG, "asvrl_iqn_scale_relu_fwd: null pointer");
  if (batch < 1 || n_tau < 1) return fail(ASVRL_E_BADARG, "asvrl_iqn_scale_relu_fwd: bad sizes");
  const long M = (long)batch * n_tau;
  if (int rc = trunk_dims_ok("asvrl_iqn_scale_relu_fwd", M, dim)) return rc;
  const long total = M * (dim / 4);
  long grid = (total + TR_TPB - 1) / TR_TPB;
  if (grid > 148 * 8) grid = 148 * 8;
  scale_relu_fwd_kernel<<<(int)grid, TR_TPB, 0, (cudaStream_t)stream>>>(
      (const float4*)pre, (const float4*)bias, (const float4*)scale, n_tau, (int)M, dim / 4, (float4*)out);
  return check_launch("asvrl_iqn_scale_relu_fwd");
}

extern "C" int asvrl_iqn_scale_relu_bwd(const float* dout, const float* pre, const float* bias,
                                        const float* scale, int batch, int n_tau, int dim, float* dpre,
                                        float* dscale, float* dbias, float* partials, uint32_t* counter,
                                        void* stream) {
  if (!dout || !pre || !bias || !scale) return fail(ASVRL_E_BADARG, "asvrl_iqn_scale_relu_bwd: null input");
  if (!dpre && !dscale && !dbias) return fail(ASVRL_E_BADARG, "asvrl_iqn_scale_relu_bwd: no output requested");
  if (dbias && (!partials || !counter))
    return fail(ASVRL_E_BADARG, "asvrl_iqn_scale_relu_bwd: dbias needs partials and counter");
  if (batch < 1 || n_tau < 1) return fail(ASVRL_E_BADARG, "asvrl_iqn_scale_relu_bwd: bad sizes");
  if (int rc = trunk_dims_ok("asvrl_iqn_scale_relu_bwd", (long)batch * n_tau, dim)) return rc;
  const int spb = scale_relu_spb(n_tau);
  const int grid = (batch + spb - 1) / spb;
  const size_t smem = (size_t)TR_TPB * sizeof(float4);
  scale_relu_bwd_kernel<<<grid, TR_TPB, smem, (cudaStream_t)stream>>>(
      (const float4*)dout, (const float4*)pre, (const float4*)bias, (const float4*)scale, n_tau, batch, dim / 4,
      spb, (float4*)dpre, (float4*)dscale, (float4*)dbias, (float4*)partials, counter);
  return check_launch("asvrl_iqn_scale_relu_bwd");
}

extern "C" long asvrl_iqn_scale_relu_bwd_partials(int batch, int dim) {  /* floats: the bound over every n_tau */
  return (long)((batch + 1) / 2) * dim;
}

extern "C" int asvrl_iqn_head_fwd(const float* pre, const float* bias, const float* w, const float* b_out,
                                  int rows, int dim, float* z, void* stream) {
  if (!pre || !bias || !w || !z) return fail(ASVRL_E_BADARG, "asvrl_iqn_head_fwd: null pointer");
  if (int rc = trunk_dims_ok("asvrl_iqn_head_fwd", rows, dim)) return rc;
  long grid = ((long)rows * 32 + TR_TPB - 1) / TR_TPB;
  if (grid > 148 * 8) grid = 148 * 8;
  head_fwd_kernel<<<(int)grid, TR_TPB, 0, (cudaStream_t)stream>>>((const float4*)pre, (const float4*)bias,
                                                                  (const float4*)w, b_out, rows, dim / 4, z);
  return check_launch("asvrl_iqn_head_fwd");
}

extern "C" int asvrl_iqn_head_bwd(const float* dz, const float* pre, const float* bias, const float* w,
                                  int rows, int dim, float* dpre, float* dw, float* dbias, float* db_out,
                                  float* partials, uint32_t* counter, void* stream) {
  if (!dz || !pre || !bias || !w || !dpre) return fail(ASVRL_E_BADARG, "asvrl_iqn_head_bwd: null pointer");
  if ((dw || dbias || db_out) && !(dw && dbias && db_out && partials && counter))
    return fail(ASVRL_E_BADARG, "asvrl_iqn_head_bwd: dw, dbias, db_out, partials and counter go together");
  if (int rc = trunk_dims_ok("asvrl_iqn_head_bwd", rows, dim)) return rc;
  const int grid = (rows + TR_RPC - 1) / TR_RPC;
  const size_t smem = (size_t)TR_TPB * sizeof(float4);
  head_bwd_kernel<<<grid, TR_TPB, smem, (cudaStream_t)stream>>>(
      dz, (const float4*)pre, (const float4*)bias, (const float4*)w, rows, dim / 4, (float4*)dpre, (float4*)dw,
      (float4*)dbias, db_out, (float4*)partials, counter);
  return check_launch("asvrl_iqn_head_bwd");
}

extern "C" long asvrl_iqn_head_bwd_partials(int rows, int dim) {  /* floats */
  return (long)((rows + TR_RPC - 1) / TR_RPC) * (2 * (dim / 4) + 1) * 4;
}

// ---------------------------------------------------------------------------------------
// Actor head of the collection path (no gradient): from the pre-activation of the last hidden
// layer to the epsilon-greedy action in one pass (ACIQN:318-321 + AGENT:207-225 for every ASV):
//   a_greedy[m][j] = atan_scale * atan(sum_d relu(pre[m][d] + bias[d]) * w[j][d] + b_out[j]),  j < 2
//   with probability eps (Philox stream 5 keyed on seed, counter, row): a = lo + (hi - lo) * u
// One warp per row; replaces the ReLU pass, the 128 -> 2 GEMM, atan, scale, two rand, compare,
// where and the range arithmetic (about 10 launches and two 33 MB round trips at 65 536 ASVs).
// ---------------------------------------------------------------------------------------
namespace asvrl {

__global__ void __launch_bounds__(TR_TPB)
actor_head_kernel(const float4* __restrict__ pre, const float4* __restrict__ bias, const float4* __restrict__ w,
                  const float* __restrict__ b_out, float atan_scale, const float* __restrict__ eps_dev, float eps,
                  float lo0, float hi0, float lo1, float hi1, uint64_t seed, uint64_t counter,
                  const uint64_t* __restrict__ counter_dev, int M, int D4, float2* __restrict__ actions) {
  if (counter_dev) counter += *counter_dev;
  if (eps_dev) eps = *eps_dev;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * TR_TPB) >> 5;
  const float bo0 = b_out[0], bo1 = b_out[1];
  for (int m = (blockIdx.x * TR_TPB + threadIdx.x) >> 5; m < M; m += warps) {
    float a0 = 0.0f, a1 = 0.0f;
    for (int c = lane; c < D4; c += 32) {
      const float4 r = f4relu(f4add(pre[(size_t)m * D4 + c], bias[c]));
      const float4 w0 = w[c], w1 = w[D4 + c];
      a0 += (r.x * w0.x + r.y * w0.y) + (r.z * w0.z + r.w * w0.w);
      a1 += (r.x * w1.x + r.y * w1.y) + (r.z * w1.z + r.w * w1.w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a0 += __shfl_xor_sync(0xffffffffu, a0, o);
      a1 += __shfl_xor_sync(0xffffffffu, a1, o);
    }
    if (lane == 0) {
      float2 act = make_float2(atan_scale * atanf(a0 + bo0), atan_scale * atanf(a1 + bo1));
      if (eps > 0.0f) {
        const uint4 u = philox_at(seed, STREAM_EXPLORE, counter, (int64_t)m, 0u, 0u, 0u);
        if (u01(u.x) < eps) act = make_float2(lo0 + (hi0 - lo0) * u01(u.y), lo1 + (hi1 - lo1) * u01(u.z));
      }
      actions[m] = act;
    }
  }
}

}  // namespace asvrl

extern "C" int asvrl_actor_head(const float* pre, const float* bias, const float* w, const float* b_out,
                                float atan_scale, const float* eps_dev, float eps, const float* action_lo,
                                const float* action_hi, uint64_t seed, uint64_t counter,
                                const uint64_t* counter_dev, int rows, int dim, float* actions, void* stream) {
  if (!pre || !bias || !w || !b_out || !action_lo || !action_hi || !actions)
    return fail(ASVRL_E_BADARG, "asvrl_actor_head: null pointer");
  if (int rc = trunk_dims_ok("asvrl_actor_head", rows, dim)) return rc;
  long grid = ((long)rows * 32 + TR_TPB - 1) / TR_TPB;
  if (grid > 148 * 8) grid = 148 * 8;
  asvrl::actor_head_kernel<<<(int)grid, TR_TPB, 0, (cudaStream_t)stream>>>(
      (const float4*)pre, (const float4*)bias, (const float4*)w, b_out, atan_scale, eps_dev, eps, action_lo[0],
      action_hi[0], action_lo[1], action_hi[1], seed, counter, counter_dev, rows, dim / 4, (float2*)actions);
  return check_launch("asvrl_actor_head");
}
