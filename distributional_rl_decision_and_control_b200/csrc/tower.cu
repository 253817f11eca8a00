// tower.cu — K6: the observation tower shared by the actor and the critic, forward and backward
// in one launch each (sm_100a).
//
// Replaces (reference paths under rfarl/rfarl/policy/AC_IQN_model.py):
//   :284-308 / :428-451  self_encoder (7 -> 56, ReLU), object_encoder on each of the 5 padded
//                        object rows (5 -> 40, ReLU), the mask product and the concat to 256
// and the autograd backward of those lines with respect to the two encoders' parameters (the
// observations are data: no input gradient exists).
//
//   feats[b][j]              = relu(Ws[j] . self[b] + bs[j])                       j < SF
//   feats[b][SF + k*OF + f]  = mask[b][k] >= 0.5 ? relu(Wo[f] . obj[b][k] + bo[f]) : 0
// The contractions have K = 7 and K = 5: this is HBM / launch bound work, not a GEMM.  Forward
// moves 4*(SD + K*OD + K) B in and 4*W B out per sample; backward reads feats, dfeats and the
// observations once.  The parameter gradients are per-CTA partials summed by the last CTA in a
// fixed order (deterministic).
#include <stdlib.h>

#include "common.cuh"

namespace asvrl {

constexpr int TW_MAXIN = 8;  // self_dimension, object_dimension <= 8

struct TowerDims {
  int B, SD, OD, K, SF, OF, W;  // W = SF + K * OF
};

// Thread j owns feature column j: its weight row (<= 8 values) and bias live in registers, the CTA
// walks batch rows b = blockIdx.x, += gridDim.x.  No index division, no shared memory; the
// observation loads of a row are warp broadcasts and the 4*W-byte feature row is one coalesced store.
__global__ void __launch_bounds__(1024)
tower_fwd_kernel(const float* __restrict__ self_obs, const float* __restrict__ objects,
                 const float* __restrict__ mask, const float* __restrict__ Ws, const float* __restrict__ bs,
                 const float* __restrict__ Wo, const float* __restrict__ bo, TowerDims d,
                 float* __restrict__ feats) {
  const int j = threadIdx.x;
  if (j >= d.W) return;
  const bool own = j < d.SF;
  const int k = own ? 0 : (j - d.SF) / d.OF;
  const int f = own ? j : (j - d.SF) - k * d.OF;
  const int nin = own ? d.SD : d.OD;
  float w[TW_MAXIN];
#pragma unroll
  for (int q = 0; q < TW_MAXIN; ++q) w[q] = q < nin ? (own ? Ws[f * d.SD + q] : Wo[f * d.OD + q]) : 0.0f;
  const float bias = own ? bs[f] : bo[f];
#pragma unroll 4
  for (int b = blockIdx.x; b < d.B; b += gridDim.x) {  // unrolled: four rows' loads in flight per thread
    const float* x = own ? self_obs + (size_t)b * d.SD : objects + ((size_t)b * d.K + k) * d.OD;
    float acc = bias;
#pragma unroll
    for (int q = 0; q < TW_MAXIN; ++q)
      if (q < nin) acc = fmaf(w[q], x[q], acc);
    acc = fmaxf(acc, 0.0f);
    if (!own && !(mask[(size_t)b * d.K + k] >= 0.5f)) acc = 0.0f;
    feats[(size_t)b * d.W + j] = acc;
  }
}

// Thread (rl, j): feature column j, row lane rl; the CTA's lanes stride over the batch rows.  The
// lanes and the K object slots (which share one encoder) are summed in shared memory, so a CTA
// publishes one partial per parameter: part[cta][SF + OF][TW_MAXIN + 1] (weight row, then bias).
// The last CTA sums the partials in a fixed order, one thread per parameter.
__global__ void __launch_bounds__(1024)
tower_bwd_kernel(const float* __restrict__ dfeats, const float* __restrict__ feats,
                 const float* __restrict__ self_obs, const float* __restrict__ objects, TowerDims d,
                 int W32, float* __restrict__ dWs, float* __restrict__ dbs, float* __restrict__ dWo,
                 float* __restrict__ dbo, float* __restrict__ part, uint32_t* counter) {
  extern __shared__ float tw_sm[];  // [RL][W][PW]
  constexpr int PW = TW_MAXIN + 1;
  const int j = threadIdx.x % W32, rl = threadIdx.x / W32, RL = blockDim.x / W32;
  if (j < d.W) {
    float acc[TW_MAXIN], accb = 0.0f;
#pragma unroll
    for (int q = 0; q < TW_MAXIN; ++q) acc[q] = 0.0f;
    const bool own = j < d.SF;
    const int k = own ? 0 : (j - d.SF) / d.OF;
    const int nin = own ? d.SD : d.OD;
#pragma unroll 4   // four rows' loads in flight per thread
    for (int b = blockIdx.x * RL + rl; b < d.B; b += gridDim.x * RL) {
      const size_t i = (size_t)b * d.W + j;
      const float g = feats[i] > 0.0f ? dfeats[i] : 0.0f;  // ReLU' and the mask: a masked row is 0
      const float* x = own ? self_obs + (size_t)b * d.SD : objects + ((size_t)b * d.K + k) * d.OD;
      accb += g;
#pragma unroll
      for (int q = 0; q < TW_MAXIN; ++q)
        if (q < nin) acc[q] = fmaf(g, x[q], acc[q]);
    }
    float* p = tw_sm + ((size_t)rl * d.W + j) * PW;
#pragma unroll
    for (int q = 0; q < TW_MAXIN; ++q) p[q] = acc[q];
    p[TW_MAXIN] = accb;
  }
  __syncthreads();
  const int C = d.SF + d.OF, n_val = C * PW;
  for (int o = threadIdx.x; o < n_val; o += blockDim.x) {
    const int c = o / PW, q = o - c * PW;
    const int reps = c < d.SF ? 1 : d.K;
    const int col0 = c < d.SF ? c : d.SF + (c - d.SF);
    float s = 0.0f;
    for (int kk = 0; kk < reps; ++kk)
      for (int r = 0; r < RL; ++r) s += tw_sm[((size_t)r * d.W + col0 + kk * d.OF) * PW + q];
    part[(size_t)blockIdx.x * n_val + o] = s;
  }
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  for (int o = threadIdx.x; o < n_val; o += blockDim.x) {
    const int c = o / PW, q = o - c * PW;
    const int nin = c < d.SF ? d.SD : d.OD;
    if (q < TW_MAXIN && q >= nin) continue;
    float s = 0.0f;
#pragma unroll 16
    for (int g = 0; g < (int)gridDim.x; ++g) s += __ldcg(part + (size_t)g * n_val + o);
    if (c < d.SF) { if (q == TW_MAXIN) dbs[c] = s; else dWs[c * d.SD + q] = s; }
    else { const int f = c - d.SF; if (q == TW_MAXIN) dbo[f] = s; else dWo[f * d.OD + q] = s; }
  }
  if (threadIdx.x == 0) *counter = 0u;
}

}  // namespace asvrl

using namespace asvrl;

static int tower_dims(const char* what, int batch, int self_dim, int obj_dim, int max_obj, int self_feat,
                      int obj_feat, TowerDims* d) {
  if (batch < 1 || self_dim < 1 || obj_dim < 1 || max_obj < 1 || self_feat < 1 || obj_feat < 1)
    return fail(ASVRL_E_BADARG, "%s: bad sizes", what);
  if (self_dim > TW_MAXIN || obj_dim > TW_MAXIN)
    return fail(ASVRL_E_LIMIT, "%s: self_dim, obj_dim <= %d", what, TW_MAXIN);
  const int W = self_feat + max_obj * obj_feat;
  if (W > 1024) return fail(ASVRL_E_LIMIT, "%s: feature width %d > 1024", what, W);
  *d = TowerDims{batch, self_dim, obj_dim, max_obj, self_feat, obj_feat, W};
  return 0;
}

static int tower_bwd_grid(int batch) {
  static int cap = 0;
  if (cap == 0) {
    const char* e = getenv("ASVRL_TOWER_BWD_GRID");   // tuning knob; the partial buffer is sized for the default cap
    cap = e ? atoi(e) : 64;   // 64 CTAs: 8.7 us at B = 1024 (32: 9.4; profiles/probes/small_kernels_probe.py)
    if (cap < 1 || cap > 64) cap = 64;
  }
  int g = (batch + 15) / 16;
  return g > cap ? cap : (g < 1 ? 1 : g);
}

extern "C" int asvrl_tower_fwd(const float* self_obs, const float* objects, const float* mask,
                               const float* w_self, const float* b_self, const float* w_obj,
                               const float* b_obj, int batch, int self_dim, int obj_dim, int max_obj,
                               int self_feat, int obj_feat, float* feats, void* stream) {
  if (!self_obs || !objects || !mask || !w_self || !b_self || !w_obj || !b_obj || !feats)
    return fail(ASVRL_E_BADARG, "asvrl_tower_fwd: null pointer");
  TowerDims d;
  if (int rc = tower_dims("asvrl_tower_fwd", batch, self_dim, obj_dim, max_obj, self_feat, obj_feat, &d)) return rc;
  const int threads = (d.W + 31) / 32 * 32;
  const int per_sm = 2048 / threads > 0 ? 2048 / threads : 1;
  int grid = 148 * per_sm;
  if (grid > d.B) grid = d.B;
  tower_fwd_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>(self_obs, objects, mask, w_self, b_self, w_obj,
                                                                b_obj, d, feats);
  return check_launch("asvrl_tower_fwd");
}

extern "C" int asvrl_tower_bwd(const float* dfeats, const float* feats, const float* self_obs,
                               const float* objects, int batch, int self_dim, int obj_dim, int max_obj,
                               int self_feat, int obj_feat, float* dw_self, float* db_self, float* dw_obj,
                               float* db_obj, float* partials, uint32_t* counter, void* stream) {
  if (!dfeats || !feats || !self_obs || !objects || !dw_self || !db_self || !dw_obj || !db_obj || !partials ||
      !counter)
    return fail(ASVRL_E_BADARG, "asvrl_tower_bwd: null pointer");
  TowerDims d;
  if (int rc = tower_dims("asvrl_tower_bwd", batch, self_dim, obj_dim, max_obj, self_feat, obj_feat, &d)) return rc;
  const int W32 = (d.W + 31) / 32 * 32, RL = 1024 / W32;
  const size_t smem = (size_t)RL * d.W * (TW_MAXIN + 1) * sizeof(float);  // <= 36 KB
  tower_bwd_kernel<<<tower_bwd_grid(batch), RL * W32, smem, (cudaStream_t)stream>>>(
      dfeats, feats, self_obs, objects, d, W32, dw_self, db_self, dw_obj, db_obj, partials, counter);
  return check_launch("asvrl_tower_bwd");
}

extern "C" long asvrl_tower_bwd_partials(int batch, int max_obj, int self_feat, int obj_feat) {  /* floats */
  (void)max_obj;
  int g = (batch + 15) / 16;   // the bound over every grid the launch may choose (<= 64 CTAs)
  g = g > 64 ? 64 : (g < 1 ? 1 : g);
  return (long)g * (self_feat + obj_feat) * (TW_MAXIN + 1);
}

// ---------------------------------------------------------------------------------------------------------------------
// asvrl_tower_bwd_static: asvrl_tower_bwd for callers that can neither allocate nor bring scratch memory (CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  The per-CTA partial sums and the arrival ticket of
// tower_bwd_kernel live in a file-scope device array instead of the caller's `partials` / `counter`:
//   tw_static_scratch = [0] the ticket (a 256-byte header), then 64 CTAs x TW_STATIC_MAXC columns x 9 floats
// = 295 168 bytes, held by every GPU that loads the library until the process exits.  Host code must not pass the
// array's NAME to a kernel (on the host it is a stand-in symbol, not a device address): the device address is asked for
// once per device with cudaGetSymbolAddress, with that device current, and cached.  No cudaMalloc / cudaFree / library
// allocation on this path.  Every call shares the array: calls that overlap on the GPU (two streams, two threads) race.
// ---------------------------------------------------------------------------------------------------------------------
namespace asvrl {
constexpr int TW_STATIC_MAXC = ASVRL_TOWER_BWD_STATIC_MAX_FEATURES;   // self_feat + obj_feat <= 128 (the reference's tower: 56 + 40)
constexpr int TW_STATIC_HDR = 64;     // floats in front of the partials (the ticket word, 256-byte aligned rest)
constexpr long TW_STATIC_FLOATS = TW_STATIC_HDR + 64L * TW_STATIC_MAXC * (TW_MAXIN + 1);
static __device__ __align__(256) float tw_static_scratch[TW_STATIC_FLOATS];
}  // namespace asvrl

static float* tw_static_address(const char* what, int* rc_out) {
  constexpr int MAX_DEV = 64;
  static float* ptrs[MAX_DEV] = {nullptr};   // per device ordinal; a racing first call resolves to the same address
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
    *rc_out = fail(e != cudaSuccess ? (int)e : ASVRL_E_LIMIT, "%s: cudaGetDevice: %s (device %d)", what,
                   cudaGetErrorString(e), dev);
    return nullptr;
  }
  float* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    e = cudaGetSymbolAddress((void**)&p, tw_static_scratch);   // the current device's copy of the array
    if (e != cudaSuccess || !p) {
      *rc_out = fail(e != cudaSuccess ? (int)e : ASVRL_E_NODEV, "%s: cudaGetSymbolAddress: %s", what, cudaGetErrorString(e));
      return nullptr;
    }
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);
  }
  *rc_out = 0;
  return p;
}

extern "C" int asvrl_tower_bwd_static(const float* dfeats, const float* feats, const float* self_obs,
                                      const float* objects, int batch, int self_dim, int obj_dim, int max_obj,
                                      int self_feat, int obj_feat, float* dw_self, float* db_self, float* dw_obj,
                                      float* db_obj, void* stream) {
  const char* what = "asvrl_tower_bwd_static";
  if (!dfeats || !feats || !self_obs || !objects || !dw_self || !db_self || !dw_obj || !db_obj)
    return fail(ASVRL_E_BADARG, "%s: null pointer", what);
  TowerDims d;
  if (int rc = tower_dims(what, batch, self_dim, obj_dim, max_obj, self_feat, obj_feat, &d)) return rc;
  if (self_feat + obj_feat > TW_STATIC_MAXC)
    return fail(ASVRL_E_LIMIT, "%s: self_feat + obj_feat = %d exceeds the static scratch (%d columns); use asvrl_tower_bwd "
                "with caller scratch", what, self_feat + obj_feat, TW_STATIC_MAXC);
  int rc = 0;
  float* scratch = tw_static_address(what, &rc);
  if (!scratch) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // the ticket must start at zero; a finished launch leaves it so, an aborted one may not: clear it in stream order
  cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(uint32_t), st);
  if (e != cudaSuccess) return fail((int)e, "%s: cudaMemsetAsync: %s", what, cudaGetErrorString(e));
  const int W32 = (d.W + 31) / 32 * 32, RL = 1024 / W32;
  const size_t smem = (size_t)RL * d.W * (TW_MAXIN + 1) * sizeof(float);
  tower_bwd_kernel<<<tower_bwd_grid(batch), RL * W32, smem, st>>>(
      dfeats, feats, self_obs, objects, d, W32, dw_self, db_self, dw_obj, db_obj, scratch + TW_STATIC_HDR,
      reinterpret_cast<uint32_t*>(scratch));
  return check_launch(what);
}
