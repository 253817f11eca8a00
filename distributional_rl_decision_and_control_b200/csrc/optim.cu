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
#include <stdlib.h>

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
  long long off[ASVRL_OPT_MAX_TENSORS];  // flat offset of tensor t in a packed bucket (multi-GPU path)
  int n_chunks;
};

constexpr int P2P_FAULT_WORD = ASVRL_MAX_PEERS;  // flag word after the per-rank epochs: sticky "a rank timed out"

struct PeerList {
  int world, rank;
  float* bucket[ASVRL_MAX_PEERS];   // every rank's symmetric buffer: [half 0 | half 1 | flag words]
  long long half;                   // floats per half
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
                 float* __restrict__ scal, float* __restrict__ part, float* __restrict__ skipped, uint32_t* counter) {
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
  __shared__ bool good;
  if (threadIdx.x == 0) {
    float total = 0.0f;
    for (int g = 0; g < (int)gridDim.x; ++g) total += __ldcg(part + g);
    const float norm = sqrtf(total) * grad_scale;
    const float clip = fminf(max_norm / (norm + 1e-6f), 1.0f);  // clip_grad_norm_: clamp(max=1.0)
    if (norm_out) norm_out[0] = norm;
    good = norm <= 3.0e38f;
    scal[0] = good ? clip * grad_scale : __int_as_float(0x7fc00000);  // fminf drops a NaN: keep it
    if (!good) skipped[0] += 1.0f;  // sticky evidence for ClipAdam.check(): this step changed nothing
    *counter = 0u;
  }
  __syncthreads();
  if (!good) return;  // a skipped step advances neither the moments nor the step counts
  for (int t = threadIdx.x; t < L.n; t += OPT_TPB) {  // a tensor without gradient is not in the list: its t stays
    const float tt = *L.step[t] + 1.0f;
    *L.step[t] = tt;
    scal[1 + 2 * t] = lr / (1.0f - powf(beta1, tt));
    scal[2 + 2 * t] = 1.0f / sqrtf(1.0f - powf(beta2, tt));
  }
}

__global__ void __launch_bounds__(OPT_TPB)
adam_kernel(const __grid_constant__ AdamList L, float beta1, float beta2, float eps,
            const float* __restrict__ scal) {
  const float gs = scal[0];
  if (!(gs == gs)) return;  // NaN norm (a peer never arrived, or a non-finite gradient): leave the parameters alone
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


// ---------------------------------------------------------------------------------------
// Multi-GPU variant: the all-reduce of the flat gradient bucket is done by the optimiser's own
// kernels over NVLink peer memory instead of an NCCL call (one-shot: every rank reads all
// ranks' buckets and adds them in rank order, so all replicas see bit-identical sums).
//   launch 1  pack: gradients -> this rank's bucket half (epoch & 1); the last CTA publishes
//             flag[my rank] = epoch in every peer's buffer (release, system scope)
//   launch 2  wait until all `world` flags in the local buffer reach the epoch (acquire), then
//             sum_r bucket_r[i] -> grad_sum[i] and the sum of squares; last CTA: clip factor,
//             bias corrections, step counts, epoch += 1
//   launch 3  adam_kernel on grad_sum with grad_scale = 1 / world folded into the clip factor
// Double buffering by epoch parity makes the reuse safe: a rank can write half (e & 1) again only
// after its step e+1 saw every peer's flag e+1, which each peer sets after finishing step e.
// A wait that exceeds ~2^26 polls gives up, poisons the norm with NaN instead of hanging, and raises a sticky
// fault word in every rank's buffer so that all replicas skip from then on (never a silent divergence).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(OPT_TPB)
pack_signal_kernel(const __grid_constant__ AdamList L, const __grid_constant__ PeerList P,
                   const uint32_t* __restrict__ epoch_dev, uint32_t* counter) {
  const uint32_t e = *epoch_dev + 1u;
  float* dst = P.bucket[P.rank] + (e & 1u) * P.half;
  for (int c = blockIdx.x; c < L.n_chunks; c += gridDim.x) {
    long long first;
    const int t = chunk_tensor(L, c, first);
    const float* g = L.grad[t];
    const long long n = L.numel[t];
#pragma unroll
    for (int q = 0; q < OPT_CHUNK / OPT_TPB; ++q) {
      const long long i = first + q * OPT_TPB + threadIdx.x;
      if (i < n) dst[L.off[t] + i] = g[i];
    }
  }
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence_system();
  if ((int)threadIdx.x < P.world) {
    uint32_t* flags = (uint32_t*)(P.bucket[threadIdx.x] + 2 * P.half);
    st_release_sys(flags + P.rank, e);
  }
  if (threadIdx.x == 0) *counter = 0u;
}

__global__ void __launch_bounds__(OPT_TPB)
grad_norm_p2p_kernel(const __grid_constant__ AdamList L, const __grid_constant__ PeerList P, float max_norm,
                     float lr, float beta1, float beta2, float grad_scale, uint32_t* epoch_dev,
                     float* __restrict__ grad_sum, float* __restrict__ norm_out, float* __restrict__ scal,
                     float* __restrict__ part, float* __restrict__ skipped, uint32_t* counter, uint32_t max_polls) {
  const uint32_t e = *epoch_dev + 1u;
  __shared__ int timed_out;
  if (threadIdx.x == 0) timed_out = 0;
  __syncthreads();
  const uint32_t* my_flags = (const uint32_t*)(P.bucket[P.rank] + 2 * P.half);
  if ((int)threadIdx.x < P.world) {  // one thread per peer flag
    const uint32_t* flag = my_flags + threadIdx.x;
    unsigned polls = 0;
    while ((int)(ld_acquire_sys(flag) - e) < 0) {
      if (++polls > max_polls) { timed_out = 1; break; }
      __nanosleep(64);
    }
  }
  __syncthreads();
  // A rank that gave up waiting raises the STICKY fault word in every rank's buffer (its own included): this
  // step and every later one is skipped on every rank that sees it, so the replicas cannot drift apart
  // silently — ClipAdam.check() turns the word into an error.  (A late peer reads the word after its own
  // pack + wait, long after the timeout that raised it.)
  if (timed_out && (int)threadIdx.x < P.world)
    st_release_sys((uint32_t*)(P.bucket[threadIdx.x] + 2 * P.half) + P2P_FAULT_WORD, 1u);
  __shared__ int faulted;
  if (threadIdx.x == 0) faulted = timed_out || ld_acquire_sys(my_flags + P2P_FAULT_WORD) != 0u;
  __syncthreads();
  const long long base = (long long)(e & 1u) * P.half;
  float acc = 0.0f;
  for (int c = blockIdx.x; c < L.n_chunks; c += gridDim.x) {
    long long first;
    const int t = chunk_tensor(L, c, first);
    const long long n = L.numel[t];
#pragma unroll
    for (int q = 0; q < OPT_CHUNK / OPT_TPB; ++q) {
      const long long i = first + q * OPT_TPB + threadIdx.x;
      if (i < n) {
        float s = 0.0f;
        for (int r = 0; r < P.world; ++r) s += __ldcg(P.bucket[r] + base + L.off[t] + i);  // rank order: same bits everywhere
        grad_sum[L.off[t] + i] = s;
        acc = fmaf(s, s, acc);
      }
    }
  }
  if (faulted) acc = __int_as_float(0x7fc00000);  // NaN: visible in norm_out, no hang
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
  __shared__ bool good;
  if (threadIdx.x == 0) {
    float total = 0.0f;
    for (int g = 0; g < (int)gridDim.x; ++g) total += __ldcg(part + g);
    const float norm = sqrtf(total) * grad_scale;
    const float clip = fminf(max_norm / (norm + 1e-6f), 1.0f);
    if (norm_out) norm_out[0] = norm;
    good = norm <= 3.0e38f;
    scal[0] = good ? clip * grad_scale : __int_as_float(0x7fc00000);  // fminf drops a NaN: keep it
    if (!good) skipped[0] += 1.0f;
    *epoch_dev = e;  // every CTA has read the old value (they all arrived at the counter); the exchange stays in lock step
    *counter = 0u;
  }
  __syncthreads();
  if (!good) return;  // a skipped step advances neither the moments nor the step counts
  for (int t = threadIdx.x; t < L.n; t += OPT_TPB) {
    const float tt = *L.step[t] + 1.0f;
    *L.step[t] = tt;
    scal[1 + 2 * t] = lr / (1.0f - powf(beta1, tt));
    scal[2 + 2 * t] = 1.0f / sqrtf(1.0f - powf(beta2, tt));
  }
}

}  // namespace asvrl

using namespace asvrl;

static int build_list(const char* what, const asvrl_adam_tensors* tensors, AdamList* Lp) {
  AdamList& L = *Lp;
  if (tensors->n < 0 || tensors->n > ASVRL_OPT_MAX_TENSORS)
    return fail(ASVRL_E_LIMIT, "%s: at most %d tensors per call", what, ASVRL_OPT_MAX_TENSORS);
  L.n = tensors->n;
  long long total = 0, off = 0;  // work items / elements so far
  for (int t = 0; t < L.n; ++t) {
    if (!tensors->param[t] || !tensors->grad[t] || !tensors->exp_avg[t] || !tensors->exp_avg_sq[t] ||
        !tensors->step[t] || tensors->numel[t] < 0)
      return fail(ASVRL_E_BADARG, "%s: tensor %d has a null pointer or negative size", what, t);
    L.param[t] = tensors->param[t]; L.grad[t] = tensors->grad[t];
    L.exp_avg[t] = tensors->exp_avg[t]; L.exp_avg_sq[t] = tensors->exp_avg_sq[t]; L.step[t] = tensors->step[t];
    L.numel[t] = tensors->numel[t];
    L.off[t] = off;
    off += tensors->numel[t];
    total += (tensors->numel[t] + OPT_CHUNK - 1) / OPT_CHUNK;
    if (total > 0x7fffffffLL) return fail(ASVRL_E_LIMIT, "%s: too many elements", what);
    L.chunk_end[t] = (int)total;
  }
  L.n_chunks = (int)total;
  return 0;
}

static int check_hyper(const char* what, float max_norm, float lr, float beta1, float beta2, float eps,
                       float grad_scale) {
  if (!(max_norm > 0.0f) || !(lr >= 0.0f) || !(beta1 >= 0.0f && beta1 < 1.0f) || !(beta2 >= 0.0f && beta2 < 1.0f) ||
      !(eps >= 0.0f) || !(grad_scale > 0.0f))
    return fail(ASVRL_E_BADARG, "%s: bad hyper-parameter", what);
  return 0;
}

extern "C" int asvrl_clip_adam_step(const asvrl_adam_tensors* tensors, float max_norm, float lr, float beta1,
                                    float beta2, float eps, float grad_scale, float* norm_out,
                                    float* scratch, uint32_t* counter, void* stream) {
  if (!tensors || !scratch || !counter) return fail(ASVRL_E_BADARG, "asvrl_clip_adam_step: null pointer");
  if (int rc = check_hyper("asvrl_clip_adam_step", max_norm, lr, beta1, beta2, eps, grad_scale)) return rc;
  AdamList L;
  if (int rc = build_list("asvrl_clip_adam_step", tensors, &L)) return rc;
  if (L.n_chunks == 0) return 0;
  const int grid = L.n_chunks > OPT_GRID ? OPT_GRID : L.n_chunks;
  cudaStream_t s = (cudaStream_t)stream;
  float* scal = scratch;                                     // [1 + 2 * ASVRL_OPT_MAX_TENSORS]
  float* part = scratch + 1 + 2 * ASVRL_OPT_MAX_TENSORS;     // [OPT_GRID]
  grad_norm_kernel<<<grid, OPT_TPB, 0, s>>>(L, max_norm, lr, beta1, beta2, grad_scale, norm_out, scal,
                                            part, scratch + ASVRL_OPT_SCRATCH_FLOATS - 1, counter);
  adam_kernel<<<grid, OPT_TPB, 0, s>>>(L, beta1, beta2, eps, scal);
  return check_launch("asvrl_clip_adam_step");
}

extern "C" int asvrl_clip_adam_step_p2p(const asvrl_adam_tensors* tensors, const asvrl_peer_buckets* peers,
                                        float* grad_sum, uint32_t* epoch, float max_norm, float lr, float beta1,
                                        float beta2, float eps, float* norm_out, float* scratch,
                                        uint32_t* counter, void* stream) {
  const char* what = "asvrl_clip_adam_step_p2p";
  if (!tensors || !peers || !grad_sum || !epoch || !scratch || !counter) return fail(ASVRL_E_BADARG, "%s: null pointer", what);
  if (peers->world < 1 || peers->world > ASVRL_MAX_PEERS || peers->rank < 0 || peers->rank >= peers->world)
    return fail(ASVRL_E_BADARG, "%s: world in [1, %d] and rank in [0, world)", what, ASVRL_MAX_PEERS);
  const float grad_scale = 1.0f / (float)peers->world;
  if (int rc = check_hyper(what, max_norm, lr, beta1, beta2, eps, grad_scale)) return rc;
  AdamList L;
  if (int rc = build_list(what, tensors, &L)) return rc;
  if (L.n_chunks == 0) return 0;
  const long long total = L.off[L.n - 1] + L.numel[L.n - 1];
  if (total > peers->half_floats) return fail(ASVRL_E_BADARG, "%s: %lld gradient floats exceed the bucket half (%lld)", what, total, peers->half_floats);
  PeerList P;
  P.world = peers->world; P.rank = peers->rank; P.half = peers->half_floats;
  for (int r = 0; r < P.world; ++r) {
    if (!peers->bucket[r]) return fail(ASVRL_E_BADARG, "%s: null peer bucket %d", what, r);
    P.bucket[r] = peers->bucket[r];
  }
  const int grid = L.n_chunks > OPT_GRID ? OPT_GRID : L.n_chunks;
  cudaStream_t s = (cudaStream_t)stream;
  float* scal = scratch;
  float* part = scratch + 1 + 2 * ASVRL_OPT_MAX_TENSORS;
  pack_signal_kernel<<<grid, OPT_TPB, 0, s>>>(L, P, epoch, counter);
  uint32_t max_polls = 1u << 26;  // ~4 s at 64 ns per poll; ASVRL_P2P_TIMEOUT_POLLS overrides (tests)
  if (const char* env = getenv("ASVRL_P2P_TIMEOUT_POLLS")) {
    const long v = atol(env);
    if (v > 0 && v < 0x7fffffffL) max_polls = (uint32_t)v;
  }
  grad_norm_p2p_kernel<<<grid, OPT_TPB, 0, s>>>(L, P, max_norm, lr, beta1, beta2, grad_scale, epoch, grad_sum,
                                                norm_out, scal, part, scratch + ASVRL_OPT_SCRATCH_FLOATS - 1,
                                                counter, max_polls);
  AdamList LA = L;  // Adam reads the all-reduced sum
  for (int t = 0; t < LA.n; ++t) LA.grad[t] = grad_sum + LA.off[t];
  adam_kernel<<<grid, OPT_TPB, 0, s>>>(LA, beta1, beta2, eps, scal);
  return check_launch(what);
}
