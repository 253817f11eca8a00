// optim.cu — K7: gradient-norm clipping and the Adam update of one optimiser in two launches
// (sm_100a), over a list of parameter tensors passed by value (no flattening copies).
//
// Replaces (reference paths under rfarl/rfarl/agent.py):
//   :414-416, :424-427 and twins   torch.nn.utils.clip_grad_norm_(params, 0.5); optimizer.step()
// i.e. torch's foreach norm / stack / norm / clamp / mul chain and the foreach Adam chain
// (about 30 launches per optimiser step) with the same arithmetic:
//   total = || g ||_2 * grad_scale ;  c = min(1, max_norm / (total + 1e-6))
//   g' = g * grad_scale * c
//   m <- m + (1 - b1)(g' - m) ;  v <- b2 v + (1 - b2) g'^2 ;  t <- t + 1   (t per tensor, as torch)
//   p <- p - (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// grad_scale = 1 / world_size when the gradients were SUMMED by the all-reduce.  Bound: HBM, 16 B
// read + 12 B written per parameter; at the 67 249 + 50 354 parameters of AC-IQN it is launch
// latency.  The sum of squares is reduced per CTA and then by the last CTA in a fixed order.
#include "common.cuh"

namespace asvrl {

constexpr int OPT_TPB = 256;
constexpr int OPT_CHUNK = 4 * OPT_TPB;  // elements per (tensor, chunk) work item
constexpr int OPT_GRID = 128;           // CTAs (and partial sums) at most

struct AdamList {
  int n;
  float* param[ASVRL_OPT_MAX_TENSORS];
  const float* grad[ASVRL_OPT_MAX_TENSORS];
  float* exp_avg[ASVRL_OPT_MAX_TENSORS];
  float* exp_avg_sq[ASVRL_OPT_MAX_TENSORS];
  float* step[ASVRL_OPT_MAX_TENSORS];  // per-tensor step count t (torch keeps one per parameter)
  long long numel[ASVRL_OPT_MAX_TENSORS];
  int chunk_end[ASVRL_OPT_MAX_TENSORS];  // running count of OPT_CHUNK-sized work items up to and including tensor t
  int n_chunks;
};

// work item c -> (tensor t, first element): every CTA owns whole chunks, so its loads are independent
// of each other (one DRAM round trip) instead of a serial walk over the tensor list
__device__ __forceinline__ int chunk_tensor(const AdamList& L, int c, long long& first) {
  int t = 0;
  while (c >= L.chunk_end[t]) ++t;
  first = (long long)(c - (t ? L.chunk_end[t - 1] : 0)) * OPT_CHUNK;
  return t;
}

// scal: [0] clip * grad_scale, then per tensor t: [1 + 2t] lr / (1 - b1^t_t), [2 + 2t] 1 / sqrt(1 - b2^t_t)
__global__ void __launch_bounds__(OPT_TPB)
grad_norm_kernel(const __grid_constant__ AdamList L, float max_norm, float lr, float beta1, float beta2,
                 float grad_scale, float* __restrict__ norm_out,
                 float* __restrict__ scal, float* __restrict__ part, uint32_t* counter) {
  float acc = 0.0f;
  for (int c = blockIdx.x; c < L.n_chunks; c += gridDim.x) {
    long long first;
    const int t = chunk_tensor(L, c, first);
    const float* g = L.grad[t];
    const long long n = L.numel[t];
#pragma unroll
    for (int q = 0; q < OPT_CHUNK / OPT_TPB; ++q) {
      const long long i = first + q * OPT_TPB + threadIdx.x;
      const float v = i < n ? g[i] : 0.0f;
      acc = fmaf(v, v, acc);
    }
  }
  __shared__ float sm[OPT_TPB];
  __shared__ bool is_last;
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int s = OPT_TPB / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    part[blockIdx.x] = sm[0];
    __threadfence();
    is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  for (int t = threadIdx.x; t < L.n; t += OPT_TPB) {  // a tensor without gradient is not in the list: its t stays
    const float tt = *L.step[t] + 1.0f;
    *L.step[t] = tt;
    scal[1 + 2 * t] = lr / (1.0f - powf(beta1, tt));
    scal[2 + 2 * t] = 1.0f / sqrtf(1.0f - powf(beta2, tt));
  }
  if (threadIdx.x != 0) return;
  float total = 0.0f;
  for (int g = 0; g < (int)gridDim.x; ++g) total += __ldcg(part + g);
  const float norm = sqrtf(total) * grad_scale;
  const float clip = fminf(max_norm / (norm + 1e-6f), 1.0f);  // clip_grad_norm_: clamp(max=1.0)
  if (norm_out) norm_out[0] = norm;
  scal[0] = clip * grad_scale;
  *counter = 0u;
}

__global__ void __launch_bounds__(OPT_TPB)
adam_kernel(const __grid_constant__ AdamList L, float beta1, float beta2, float eps,
            const float* __restrict__ scal) {
  const float gs = scal[0];
  for (int c = blockIdx.x; c < L.n_chunks; c += gridDim.x) {
    long long first;
    const int t = chunk_tensor(L, c, first);
    const float step_size = scal[1 + 2 * t], inv_bc2_sqrt = scal[2 + 2 * t];
    const float* g = L.grad[t];
    float *p = L.param[t], *m = L.exp_avg[t], *v = L.exp_avg_sq[t];
    const long long n = L.numel[t];
    float gq[OPT_CHUNK / OPT_TPB], mq[OPT_CHUNK / OPT_TPB], vq[OPT_CHUNK / OPT_TPB], pq[OPT_CHUNK / OPT_TPB];
#pragma unroll
    for (int q = 0; q < OPT_CHUNK / OPT_TPB; ++q) {  // every load of the chunk before the first store
      const long long i = first + q * OPT_TPB + threadIdx.x;
      const bool ok = i < n;
      gq[q] = ok ? g[i] : 0.0f; mq[q] = ok ? m[i] : 0.0f; vq[q] = ok ? v[i] : 0.0f; pq[q] = ok ? p[i] : 0.0f;
    }
#pragma unroll
    for (int q = 0; q < OPT_CHUNK / OPT_TPB; ++q) {
      const long long i = first + q * OPT_TPB + threadIdx.x;
      if (i < n) {
        const float gi = gq[q] * gs;
        const float mi = mq[q] + (1.0f - beta1) * (gi - mq[q]);
        const float vi = beta2 * vq[q] + (1.0f - beta2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        p[i] = pq[q] - step_size * (mi / (sqrtf(vi) * inv_bc2_sqrt + eps));
      }
    }
  }
}

}  // namespace asvrl

using namespace asvrl;

extern "C" int asvrl_clip_adam_step(const asvrl_adam_tensors* tensors, float max_norm, float lr, float beta1,
                                    float beta2, float eps, float grad_scale, float* norm_out,
                                    float* scratch, uint32_t* counter, void* stream) {
  if (!tensors || !scratch || !counter) return fail(ASVRL_E_BADARG, "asvrl_clip_adam_step: null pointer");
  if (tensors->n < 0 || tensors->n > ASVRL_OPT_MAX_TENSORS)
    return fail(ASVRL_E_LIMIT, "asvrl_clip_adam_step: at most %d tensors per call", ASVRL_OPT_MAX_TENSORS);
  if (!(max_norm > 0.0f) || !(lr >= 0.0f) || !(beta1 >= 0.0f && beta1 < 1.0f) || !(beta2 >= 0.0f && beta2 < 1.0f) ||
      !(eps >= 0.0f) || !(grad_scale > 0.0f))
    return fail(ASVRL_E_BADARG, "asvrl_clip_adam_step: bad hyper-parameter");
  if (tensors->n == 0) return 0;
  AdamList L;
  L.n = tensors->n;
  long long total = 0;  // work items so far
  for (int t = 0; t < L.n; ++t) {
    if (!tensors->param[t] || !tensors->grad[t] || !tensors->exp_avg[t] || !tensors->exp_avg_sq[t] ||
        !tensors->step[t] || tensors->numel[t] < 0)
      return fail(ASVRL_E_BADARG, "asvrl_clip_adam_step: tensor %d has a null pointer or negative size", t);
    L.param[t] = tensors->param[t]; L.grad[t] = tensors->grad[t];
    L.exp_avg[t] = tensors->exp_avg[t]; L.exp_avg_sq[t] = tensors->exp_avg_sq[t]; L.step[t] = tensors->step[t];
    L.numel[t] = tensors->numel[t];
    total += (tensors->numel[t] + OPT_CHUNK - 1) / OPT_CHUNK;
    if (total > 0x7fffffffLL) return fail(ASVRL_E_LIMIT, "asvrl_clip_adam_step: too many elements");
    L.chunk_end[t] = (int)total;
  }
  L.n_chunks = (int)total;
  if (L.n_chunks == 0) return 0;
  const int grid = L.n_chunks > OPT_GRID ? OPT_GRID : L.n_chunks;
  cudaStream_t s = (cudaStream_t)stream;
  float* scal = scratch;                                     // [1 + 2 * ASVRL_OPT_MAX_TENSORS]
  float* part = scratch + 1 + 2 * ASVRL_OPT_MAX_TENSORS;     // [OPT_GRID]
  grad_norm_kernel<<<grid, OPT_TPB, 0, s>>>(L, max_norm, lr, beta1, beta2, grad_scale, norm_out, scal,
                                            part, counter);
  adam_kernel<<<grid, OPT_TPB, 0, s>>>(L, beta1, beta2, eps, scal);
  return check_launch("asvrl_clip_adam_step");
}
