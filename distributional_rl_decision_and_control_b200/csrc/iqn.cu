// iqn.cu — K3 (tau sampling + cosine embedding) and K4 (fused Bellman target, pairwise N x N'
// TD-error, quantile-Huber loss and its backward) for AC-IQN / IQN (sm_100a).
//
// Replaces (reference paths under rfarl/rfarl/):
//   policy/AC_IQN_model.py:410-426, policy/IQN_model.py:56-72   calc_cos
//   agent.py:401-412, 455-467, 701-707                          loss core + autograd backward
// The B x N x N' tensors the reference materialises (td_error, |td|, huber, mask, quantile
// loss, plus autograd's saved copies) never exist here: a warp owns one sample, holds Z_i and
// tau_i in registers, broadcasts the N' targets by shuffle and reduces in registers.
#include <math.h>

#include "common.cuh"

namespace asvrl {

// ---------------------------------------------------------------------------------------
// K4.  One warp per sample b.  loss_b = (1/N') sum_j sum_i |tau_i - 1[d<0]| huber(d)/kappa,
// grad_bi = -(1/(B N')) sum_j |tau_i - 1[d<0]| clamp(d,-kappa,kappa)/kappa   
// The scalar loss is reduced by the last CTA to finish, in a fixed order (deterministic).
// ---------------------------------------------------------------------------------------
constexpr int QH_WARPS = 4;

__global__ void __launch_bounds__(QH_WARPS * 32)
qhuber_kernel(const float* __restrict__ z, const float* __restrict__ tau,
              const float* __restrict__ z_next, const float* __restrict__ rew,
              const float* __restrict__ done, float gamma, float kappa, int B, int N, int Np,
              float* __restrict__ loss_per_sample, float* __restrict__ grad_z,
              float* __restrict__ loss, uint32_t* workspace) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.x * QH_WARPS + warp;
  const float inv_k = 1.0f / kappa;
  if (b < B) {
    const float r = rew[b], keep = 1.0f - done[b];
    const float gscale = -1.0f / ((float)B * (float)Np);
    float lsum = 0.0f;
    for (int i0 = 0; i0 < N; i0 += 32) {
      const int i = i0 + lane;
      const bool vi = i < N;
      const float zi = vi ? z[(size_t)b * N + i] : 0.0f;
      const float ti = vi ? tau[(size_t)b * N + i] : 0.0f;
      float li = 0.0f, gi = 0.0f;
      for (int j0 = 0; j0 < Np; j0 += 32) {
        const int j = j0 + lane;
        // AGENT:401  T = r + (gamma * Z') * (1 - done)
        const float Tl = j < Np ? r + (gamma * z_next[(size_t)b * Np + j]) * keep : 0.0f;
        const int nj = min(32, Np - j0);
        for (int jj = 0; jj < nj; ++jj) {
          const float T = __shfl_sync(0xffffffffu, Tl, jj);
          const float d = T - zi;                                   // AGENT:406
          const float ad = fabsf(d);
          const float h = ad <= kappa ? 0.5f * d * d : kappa * (ad - 0.5f * kappa);  // AGENT:705
          const float wgt = fabsf(ti - (d < 0.0f ? 1.0f : 0.0f));   // AGENT:409
          li += wgt * h * inv_k;
          gi += wgt * fminf(fmaxf(d, -kappa), kappa) * inv_k;
        }
      }
      if (vi) {
        grad_z[(size_t)b * N + i] = gi * gscale;
        lsum += li;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
    if (lane == 0) loss_per_sample[b] = lsum / (float)Np;  // AGENT:411
  }
  // ---- deterministic mean over the batch by the last CTA (AGENT:412) ----
  __shared__ bool is_last;
  __shared__ float part[QH_WARPS * 32];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(workspace, 1u) == gridDim.x - 1;
  __syncthreads();
  if (is_last) {
    __threadfence();
    float acc = 0.0f;
    for (int i = threadIdx.x; i < B; i += blockDim.x) acc += __ldcg(loss_per_sample + i);
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int sft = (QH_WARPS * 32) / 2; sft > 0; sft >>= 1) {
      if (threadIdx.x < sft) part[threadIdx.x] += part[threadIdx.x + sft];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      loss[0] = part[0] / (float)B;
      *workspace = 0u;  // left zero for the next call
    }
  }
}

// ---------------------------------------------------------------------------------------
// K3.  A warp owns 32 consecutive (b, n) rows: every lane draws / loads its tau, then the warp
// walks the 32 rows and writes each row's n_cos cosines with 128-byte coalesced stores.
// cos is evaluated exactly as torch does at ACIQN:424: float32 product tau * float32(pi*i),
// then cosf (no fast-math, no fused multiply).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tau_cos_kernel(const float* __restrict__ tau_in, uint64_t seed, uint64_t counter, float cvar,
               int rows, int n_cos, float* __restrict__ tau_out, float* __restrict__ cos_out,
               const uint64_t* __restrict__ counter_dev) {
  if (counter_dev) counter += *counter_dev;
  const int lane = threadIdx.x & 31;
  const int warps_per_grid = (gridDim.x * blockDim.x) >> 5;
  const int n_groups = (rows + 31) / 32;
  for (int grp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; grp < n_groups; grp += warps_per_grid) {
    const int row = grp * 32 + lane;
    float t = 0.0f;
    if (row < rows) {
      if (tau_in) {
        t = tau_in[row];
      } else {  // torch.rand(B, N) * cvar, ACIQN:419-422
        uint4 w = philox_at(seed, STREAM_TAU, counter, (int64_t)(row >> 2), 0u, 0u, 0u);
        uint32_t word = (row & 3) == 0 ? w.x : (row & 3) == 1 ? w.y : (row & 3) == 2 ? w.z : w.w;
        t = __fmul_rn(u01(word), cvar);
      }
      if (tau_out) tau_out[row] = t;
    }
    const int nrow = min(32, rows - grp * 32);
    for (int rr = 0; rr < nrow; ++rr) {
      const float tr = __shfl_sync(0xffffffffu, t, rr);
      float* dst = cos_out + (size_t)(grp * 32 + rr) * n_cos;
      for (int i = lane; i < n_cos; i += 32) {
        const float pi_i = (float)(3.14159265358979323846 * (double)i);  // ACIQN:397
        dst[i] = cosf(__fmul_rn(tr, pi_i));
      }
    }
  }
}

// n_cos % 4 == 0 (the shipped 64): one thread per float4 of cosines, n_cos/4 threads per row, so
// a launch at C3 is 524 288 independent threads instead of 1 024 warps walking 32 rows each.
// Every thread of a row re-derives the row's tau (one Philox block, shared by 4 rows).
__global__ void __launch_bounds__(256)
tau_cos4_kernel(const float* __restrict__ tau_in, uint64_t seed, uint64_t counter, float cvar,
                int rows, int nc4, float* __restrict__ tau_out, float4* __restrict__ cos_out,
                const uint64_t* __restrict__ counter_dev) {
  if (counter_dev) counter += *counter_dev;
  const int total = rows * nc4;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    const int row = i / nc4, c = i - row * nc4;
    float t;
    if (tau_in) {
      t = tau_in[row];
    } else {  // torch.rand(B, N) * cvar, ACIQN:419-422
      const uint4 w = philox_at(seed, STREAM_TAU, counter, (int64_t)(row >> 2), 0u, 0u, 0u);
      const uint32_t word = (row & 3) == 0 ? w.x : (row & 3) == 1 ? w.y : (row & 3) == 2 ? w.z : w.w;
      t = __fmul_rn(u01(word), cvar);
    }
    if (c == 0 && tau_out) tau_out[row] = t;
    const double PI = 3.14159265358979323846;
    const int i0 = 4 * c;
    float4 v;
    v.x = cosf(__fmul_rn(t, (float)(PI * (double)(i0 + 0))));  // ACIQN:397, 424
    v.y = cosf(__fmul_rn(t, (float)(PI * (double)(i0 + 1))));
    v.z = cosf(__fmul_rn(t, (float)(PI * (double)(i0 + 2))));
    v.w = cosf(__fmul_rn(t, (float)(PI * (double)(i0 + 3))));
    cos_out[i] = v;
  }
}

}  // namespace asvrl

using namespace asvrl;

extern "C" int asvrl_qhuber_fwd_bwd(const float* z, const float* tau, const float* z_next,
                                    const float* r, const float* done, float gamma, float kappa,
                                    int batch, int n, int n_next, float* loss_per_sample,
                                    float* grad_z, float* loss, uint32_t* workspace, void* stream) {
  if (!z || !tau || !z_next || !r || !done || !loss_per_sample || !grad_z || !loss || !workspace)
    return fail(ASVRL_E_BADARG, "asvrl_qhuber_fwd_bwd: null pointer");
  if (batch < 1 || n < 1 || n_next < 1 || !(kappa > 0.0f))
    return fail(ASVRL_E_BADARG, "asvrl_qhuber_fwd_bwd: bad sizes (B=%d N=%d N'=%d kappa=%g)", batch, n, n_next, kappa);
  const int grid = (batch + QH_WARPS - 1) / QH_WARPS;
  qhuber_kernel<<<grid, QH_WARPS * 32, 0, (cudaStream_t)stream>>>(
      z, tau, z_next, r, done, gamma, kappa, batch, n, n_next, loss_per_sample, grad_z, loss, workspace);
  return check_launch("asvrl_qhuber_fwd_bwd");
}

extern "C" int asvrl_iqn_tau_cos(const float* tau_in, uint64_t seed, uint64_t counter, float cvar,
                                 int batch, int n_tau, int n_cos, float* tau_out, float* cos_out,
                                 const uint64_t* counter_dev, void* stream) {
  if (!cos_out) return fail(ASVRL_E_BADARG, "asvrl_iqn_tau_cos: null cos_out");
  if (batch < 1 || n_tau < 1 || n_cos < 1) return fail(ASVRL_E_BADARG, "asvrl_iqn_tau_cos: bad sizes");
  const int rows = batch * n_tau;
  if (n_cos % 4 == 0 && (long)rows * (n_cos / 4) < (1L << 31)) {
    const long total = (long)rows * (n_cos / 4);
    long grid = (total + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    tau_cos4_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>(tau_in, seed, counter, cvar, rows, n_cos / 4,
                                                                 tau_out, (float4*)cos_out, counter_dev);
    return check_launch("asvrl_iqn_tau_cos");
  }
  const int groups = (rows + 31) / 32;
  int grid = (groups + 7) / 8;  // 8 warps per CTA
  if (grid > 148 * 8) grid = 148 * 8;
  tau_cos_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(tau_in, seed, counter, cvar, rows, n_cos,
                                                         tau_out, cos_out, counter_dev);
  return check_launch("asvrl_iqn_tau_cos");
}
