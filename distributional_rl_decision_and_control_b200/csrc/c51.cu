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

// One warp per row, rows double-buffered in shared memory.  The row's A x atoms advantages (5.1 KB at 25 x 51, not a
// multiple of 16 bytes, so 4-byte cp.async) are fetched for the NEXT row while the current one is reduced: with
// register-staged loads a warp had 8 requests in flight and then waited (ncu: 84 % of the stall samples on the staging
// loop, 13 % of the DRAM rate, 320 us for 65 536 rows); now every warp keeps a whole row in flight.
// Per row: lanes first run over the ATOMS (conflict-free column sums: v - mean_a(a) per atom, left in shared memory),
// then over the ACTIONS: lane a walks its own action's atoms sequentially (bank (a * atoms + j) mod 32 is conflict-free
// for odd atom counts) — a softmax expectation without cross-lane reductions; the argmax is the only warp reduction.
__device__ __forceinline__ void c51_cp4(float* dst_smem, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src));
}

__global__ void __launch_bounds__(C51_WARPS * 32)
c51_greedy_kernel(const float* __restrict__ zv, const float* __restrict__ za, const float* __restrict__ bv,
                  const float* __restrict__ ba, const float* __restrict__ support, int rows, int A, int atoms,
                  int zv_ld, int apb, int za_ld, long za_bs, int* __restrict__ action, float* __restrict__ q_out) {
  extern __shared__ float c51_sm[];   // [warp][2][A * atoms] rows | [warp][64] v - mean | [A * atoms] ba | [64] support
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int warps = gridDim.x * C51_WARPS;
  const int n = A * atoms;
  float* buf = c51_sm + (size_t)w * 2 * n;
  float* vm = c51_sm + (size_t)C51_WARPS * 2 * n + w * 64;
  float* bas = c51_sm + (size_t)C51_WARPS * (2 * n + 64);
  float* sup = bas + n;
  for (int e = threadIdx.x; e < n; e += blockDim.x) bas[e] = ba[e];
  if (threadIdx.x < 64) sup[threadIdx.x] = threadIdx.x < atoms ? support[threadIdx.x] : 0.0f;
  __syncthreads();
  const int j0 = lane, j1 = lane + 32;
  const bool h0 = j0 < atoms, h1 = j1 < atoms;
  const float bv0 = h0 ? bv[j0] : 0.0f, bv1 = h1 ? bv[j1] : 0.0f;
  float bs0 = 0.0f, bs1 = 0.0f;   // column sums of the advantage bias
  for (int a = 0; a < A; ++a) {
    if (h0) bs0 += bas[a * atoms + j0];
    if (h1) bs1 += bas[a * atoms + j1];
  }
  const float inv = 1.0f / (float)A;
  // za is stored in blocks of `apb` actions: block k holds [rows][za_ld] with the actions' atoms packed at the front
  // (one block of all A actions with za_ld = A * atoms is the dense layout)
  auto fetch = [&](int m, float* dst) {
    for (int a0 = 0, k = 0; a0 < A; a0 += apb, ++k) {
      const int cnt = min(apb, A - a0) * atoms;
      const float* ar = za + (size_t)k * za_bs + (size_t)m * za_ld;
      float* d = dst + a0 * atoms;
      for (int e = lane; e < cnt; e += 32) c51_cp4(d + e, ar + e);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int m = blockIdx.x * C51_WARPS + w, x = 0;
  if (m < rows) fetch(m, buf);
  for (; m < rows; m += warps, x ^= 1) {
    float* sm = buf + (size_t)x * n;
    const float z0 = h0 ? __ldg(zv + (size_t)m * zv_ld + j0) : 0.0f, z1 = h1 ? __ldg(zv + (size_t)m * zv_ld + j1) : 0.0f;
    if (m + warps < rows) {
      fetch(m + warps, buf + (size_t)(x ^ 1) * n);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncwarp();
    float m0 = 0.0f, m1 = 0.0f;
#pragma unroll 5
    for (int a = 0; a < A; ++a) {
      if (h0) m0 += sm[a * atoms + j0];
      if (h1) m1 += sm[a * atoms + j1];
    }
    if (h0) vm[j0] = (z0 + bv0) - (m0 + bs0) * inv;   // v - mean_a(a + ba), per atom
    if (h1) vm[j1] = (z1 + bv1) - (m1 + bs1) * inv;
    __syncwarp();
    float ev = -INFINITY;
    if (lane < A) {
      const float* mine = sm + lane * atoms;
      const float* mb = bas + lane * atoms;
      float mx = -INFINITY;
#pragma unroll 4
      for (int j = 0; j < atoms; ++j) mx = fmaxf(mx, vm[j] + (mine[j] + mb[j]));
      float den = 0.0f, num = 0.0f;
#pragma unroll 4
      for (int j = 0; j < atoms; ++j) {
        const float e = __expf(vm[j] + (mine[j] + mb[j]) - mx);
        den += e;
        num = fmaf(e, sup[j], num);
      }
      ev = num / den;   // sum_j softmax(q)_j support_j
      if (q_out) q_out[(size_t)m * A + lane] = ev;
    }
    // first maximum over the actions, as torch.argmax
    float best = ev;
    int best_a = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oa = __shfl_xor_sync(0xffffffffu, best_a, o);
      if (ob > best || (ob == best && oa < best_a)) { best = ob; best_a = oa; }
    }
    if (lane == 0) action[m] = best_a;
    __syncwarp();   // every lane is done with this buffer before the next iteration's fetch overwrites it
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Rainbow's prioritised n-step memory, batched append (policy/replay_memory_rainbow.py:125-131 of the reference, one
// row per ASV of a batched env step; utils/replay_memory_rainbow.py: append_step).  A warp owns a lane (= one ASV):
// the 36-float row [self 7 | objects 25 | n_obj | action | reward | nonterminal] goes to time slot `head`, the lane's
// episode clock, priority and the global counters follow.  The last CTA (ticket) advances the ring words, so the
// whole append is ONE launch that replays inside a CUDA graph; it was ~30 torch launches (cat, index_copy x 3, where,
// sums, clamps: ~150 us per batched step of 65 536 ASVs).
//   words: [0] head, [1] filled, [2] count (int64 each; the host tensors of the Python class)
//   scratch: [0] ticket, [1] active lanes of this step (zero between calls)
// ---------------------------------------------------------------------------------------------------------------
constexpr int RB_W = 36, RB_WARPS = 8;
__global__ void __launch_bounds__(RB_WARPS * 32)
rainbow_append_kernel(const float* __restrict__ o_self, const float* __restrict__ o_obj, const float* __restrict__ o_mask,
                      const int* __restrict__ actions, const float* __restrict__ reward,
                      const uint32_t* __restrict__ status, int lanes, long slots, long capacity,
                      float* __restrict__ rows, int* __restrict__ tstep, float* __restrict__ prio,
                      int* __restrict__ lane_t, const float* __restrict__ max_priority, long long* __restrict__ head,
                      long long* __restrict__ filled, long long* __restrict__ count, unsigned int* __restrict__ scratch) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long t = (long)head[0];
  const float maxp = __ldg(max_priority);
  int n_active = 0;
  for (int l = blockIdx.x * RB_WARPS + w; l < lanes; l += gridDim.x * RB_WARPS) {
    const uint32_t st = __ldg(status + l);
    const bool active = (st & ASVRL_S_WAS_DEACTIVATED) == 0u && ((st >> ASVRL_S_INFO_SHIFT) & 15u) != (uint32_t)ASVRL_INFO_ABSENT;
    const bool done = (st & ASVRL_S_DONE) != 0u;
    float* row = rows + ((size_t)t * lanes + l) * RB_W;
    row[lane] = lane < 7 ? __ldg(o_self + (size_t)l * 7 + lane) : __ldg(o_obj + (size_t)l * 25 + (lane - 7));
    if (lane < 4) {
      float v;
      if (lane == 0) {
        const float* mk = o_mask + (size_t)l * 5;
        v = __ldg(mk) + __ldg(mk + 1) + __ldg(mk + 2) + __ldg(mk + 3) + __ldg(mk + 4);
      } else if (lane == 1) v = (float)__ldg(actions + l);
      else if (lane == 2) v = __ldg(reward + l);
      else v = done ? 0.0f : 1.0f;
      row[32 + lane] = v;
    }
    if (lane == 0) {
      const int lt = lane_t[l];
      tstep[(size_t)t * lanes + l] = active ? lt : 0;
      prio[(size_t)t * lanes + l] = active ? maxp : 0.0f;
      lane_t[l] = (active && !done) ? lt + 1 : 0;
      n_active += active ? 1 : 0;
    }
  }
  if (lane == 0 && n_active) atomicAdd(&scratch[1], (unsigned int)n_active);
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&scratch[0], 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last || threadIdx.x != 0) return;
  __threadfence();
  const long long c = count[0] + (long long)atomicExch(&scratch[1], 0u);
  count[0] = c < (long long)capacity ? c : (long long)capacity;
  head[0] = (t + 1) % slots;
  const long long f = filled[0] + 1;
  filled[0] = f < (long long)slots ? f : (long long)slots;
  scratch[0] = 0u;
}

}  // namespace asvrl

using namespace asvrl;

extern "C" int asvrl_c51_greedy(const float* zv, const float* za, const float* bv, const float* ba, const float* support,
                                long rows, int n_actions, int atoms, int zv_ld, int za_actions_per_block, int za_ld,
                                long za_block_stride, int* action, float* q_out, void* stream) {
  if (!zv || !za || !bv || !ba || !support || !action) return fail(ASVRL_E_BADARG, "asvrl_c51_greedy: null pointer");
  if (rows < 1 || rows > 0x7FFFFFFFL || n_actions < 1 || n_actions > C51_MAX_A || atoms < 1 || atoms > 64)
    return fail(ASVRL_E_BADARG, "asvrl_c51_greedy: rows >= 1, 1 <= actions <= %d, 1 <= atoms <= 64 (rows=%ld A=%d atoms=%d)",
                C51_MAX_A, rows, n_actions, atoms);
  if (zv_ld == 0) zv_ld = atoms;
  if (za_actions_per_block == 0) { za_actions_per_block = n_actions; za_ld = n_actions * atoms; za_block_stride = 0; }
  if (zv_ld < atoms || za_actions_per_block < 1 || za_actions_per_block > n_actions ||
      za_ld < za_actions_per_block * atoms || za_block_stride < 0)
    return fail(ASVRL_E_BADARG, "asvrl_c51_greedy: zv_ld %d, %d actions per block, za_ld %d, block stride %ld", zv_ld,
                za_actions_per_block, za_ld, za_block_stride);
  long grid = (rows + C51_WARPS - 1) / C51_WARPS;
  if (grid > 148 * 2) grid = 148 * 2;   // two 88 KB CTAs per SM
  const size_t n = (size_t)n_actions * atoms;
  const size_t smem = ((size_t)C51_WARPS * (2 * n + 64) + n + 64) * sizeof(float);   // 88 KB at 25 x 51
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(c51_greedy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (C51_WARPS * (2 * C51_MAX_A * 64 + 64) + C51_MAX_A * 64 + 64) * 4);
    if (e != cudaSuccess) return fail((int)e, "asvrl_c51_greedy: %s", cudaGetErrorString(e));
    attr = true;
  }
  c51_greedy_kernel<<<(int)grid, C51_WARPS * 32, smem, (cudaStream_t)stream>>>(zv, za, bv, ba, support, (int)rows,
                                                                               n_actions, atoms, zv_ld, za_actions_per_block,
                                                                               za_ld, za_block_stride, action, q_out);
  return check_launch("asvrl_c51_greedy");
}

extern "C" int asvrl_rainbow_append(const float* obs_self, const float* obs_objects, const float* obs_mask,
                                    const int* actions, const float* reward, const uint32_t* status, int lanes, long slots,
                                    long capacity, float* rows, int* tstep, float* prio, int* lane_t,
                                    const float* max_priority, long long* head, long long* filled, long long* count,
                                    uint32_t* scratch, void* stream) {
  if (!obs_self || !obs_objects || !obs_mask || !actions || !reward || !status || !rows || !tstep || !prio || !lane_t ||
      !max_priority || !head || !filled || !count || !scratch)
    return fail(ASVRL_E_BADARG, "asvrl_rainbow_append: null pointer");
  if (lanes < 1 || slots < 1 || capacity < 1) return fail(ASVRL_E_BADARG, "asvrl_rainbow_append: bad sizes");
  int grid = (lanes + RB_WARPS - 1) / RB_WARPS;
  if (grid > 148 * 8) grid = 148 * 8;
  rainbow_append_kernel<<<grid, RB_WARPS * 32, 0, (cudaStream_t)stream>>>(obs_self, obs_objects, obs_mask, actions, reward,
                                                                          status, lanes, slots, capacity, rows, tstep, prio,
                                                                          lane_t, max_priority, head, filled, count, scratch);
  return check_launch("asvrl_rainbow_append");
}
