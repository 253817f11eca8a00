// c51.cu — the greedy action of Rainbow's dueling C51 head for every ASV of a batched step, in one pass over
// the two output layers' products (sm_100a).
//
// Replaces, per collection step (reference: one batch-1 forward per ASV),
//   rfarl/rfarl/policy/Rainbow_model.py:127-135  q = v + a - a.mean(1); softmax over the atoms
//   rfarl/rfarl/agent.py:235-246                 act_rainbow: (p * support).sum(2).argmax(1)
// which on (M, 25, 51) tensors is a dozen element-wise / softmax / reduction launches moving ~0.3 GB each at
// M = 65 536 (and a separate bias pass over the (M, 1275) advantage product).  Here one warp owns one row:
// lanes run over the atoms (two per lane, atoms <= 64), the mean over the actions is a register sum, and
// each action's softmax expectation is three warp reductions.  Biases are added here, so the two products
// come straight from a bias-free GEMM.  Bound: HBM, 4 (atoms + A atoms) B per row read once.
#include "common.cuh"

namespace asvrl {

constexpr int C51_MAX_A = 32;
constexpr int C51_WARPS = 8;

// One warp per row.  The row's A x atoms advantages (+ bias) are staged in shared memory with coalesced loads (the
// atom count is not a multiple of 32), then lanes run over the atoms: conflict-free column sums for the mean over
// the actions, and per action a softmax expectation = three interleaved warp reductions.
__global__ void __launch_bounds__(C51_WARPS * 32)
c51_greedy_kernel(const float* __restrict__ zv, const float* __restrict__ za, const float* __restrict__ bv,
                  const float* __restrict__ ba, const float* __restrict__ support, int rows, int A, int atoms,
                  int* __restrict__ action, float* __restrict__ q_out) {
  extern __shared__ float c51_sm[];   // [warp][A * atoms]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int warps = gridDim.x * C51_WARPS;
  const int n = A * atoms;
  float* sm = c51_sm + (size_t)w * n;
  const int j0 = lane, j1 = lane + 32;
  const bool h0 = j0 < atoms, h1 = j1 < atoms;
  const float s0 = h0 ? support[j0] : 0.0f, s1 = h1 ? support[j1] : 0.0f;
  const float bv0 = h0 ? bv[j0] : 0.0f, bv1 = h1 ? bv[j1] : 0.0f;
  const float inv = 1.0f / (float)A;
  for (int m = blockIdx.x * C51_WARPS + w; m < rows; m += warps) {
    const float* ar = za + (size_t)m * n;
    for (int e = lane; e < n; e += 32) sm[e] = ar[e] + __ldg(ba + e);
    __syncwarp();
    float m0 = 0.0f, m1 = 0.0f;
    for (int a = 0; a < A; ++a) {
      if (h0) m0 += sm[a * atoms + j0];
      if (h1) m1 += sm[a * atoms + j1];
    }
    const float v0 = (h0 ? zv[(size_t)m * atoms + j0] + bv0 : 0.0f) - m0 * inv;   // v - mean_a(a), per atom
    const float v1 = (h1 ? zv[(size_t)m * atoms + j1] + bv1 : 0.0f) - m1 * inv;
    float best = -INFINITY;
    int best_a = 0;
    for (int a = 0; a < A; ++a) {
      const float q0 = h0 ? v0 + sm[a * atoms + j0] : -INFINITY, q1 = h1 ? v1 + sm[a * atoms + j1] : -INFINITY;
      float mx = fmaxf(q0, q1);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      const float e0 = h0 ? __expf(q0 - mx) : 0.0f, e1 = h1 ? __expf(q1 - mx) : 0.0f;
      float den = e0 + e1, num = e0 * s0 + e1 * s1;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        den += __shfl_xor_sync(0xffffffffu, den, o);
        num += __shfl_xor_sync(0xffffffffu, num, o);
      }
      const float ev = num / den;   // sum_j softmax(q)_j support_j
      if (q_out && lane == 0) q_out[(size_t)m * A + a] = ev;
      if (ev > best) { best = ev; best_a = a; }   // first maximum, as torch.argmax
    }
    if (lane == 0) action[m] = best_a;
    __syncwarp();
  }
}

}  // namespace asvrl

using namespace asvrl;

extern "C" int asvrl_c51_greedy(const float* zv, const float* za, const float* bv, const float* ba, const float* support,
                                long rows, int n_actions, int atoms, int* action, float* q_out, void* stream) {
  if (!zv || !za || !bv || !ba || !support || !action) return fail(ASVRL_E_BADARG, "asvrl_c51_greedy: null pointer");
  if (rows < 1 || rows > 0x7FFFFFFFL || n_actions < 1 || n_actions > C51_MAX_A || atoms < 1 || atoms > 64)
    return fail(ASVRL_E_BADARG, "asvrl_c51_greedy: rows >= 1, 1 <= actions <= %d, 1 <= atoms <= 64 (rows=%ld A=%d atoms=%d)",
                C51_MAX_A, rows, n_actions, atoms);
  long grid = (rows + C51_WARPS - 1) / C51_WARPS;
  if (grid > 148 * 4) grid = 148 * 4;
  const size_t smem = (size_t)C51_WARPS * n_actions * atoms * sizeof(float);   // <= 64 KB
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(c51_greedy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 32 * 64 * 4);
    if (e != cudaSuccess) return fail((int)e, "asvrl_c51_greedy: %s", cudaGetErrorString(e));
    attr = true;
  }
  c51_greedy_kernel<<<(int)grid, C51_WARPS * 32, smem, (cudaStream_t)stream>>>(zv, za, bv, ba, support, (int)rows,
                                                                               n_actions, atoms, action, q_out);
  return check_launch("asvrl_c51_greedy");
}
