// replay.cu — K2: GPU-resident uniform replay buffer with on-device sampling (sm_100a).
//
// Replaces (reference paths under rfarl/rfarl/):
//   utils/replay_buffer.py:22-24   ReplayBuffer.add        (deque of python tuples)
//   utils/replay_buffer.py:26-69   sample + state_batch    (random.sample, pad to 5 + mask)
//   agent.py:360-368, 387-392      state_to_tensor + the five host->device copies per train()
//   policy/trainer.py:157-166      which transitions are stored (robots active at step start)
//
// Layout: a ring of dense float32[70] rows (280 B) in HBM.  push flattens (row, element) items
// over a CTA, sample gives one warp per row; every store is a contiguous, coalesced run and the
// work is pure HBM traffic (560 B per pushed transition, ~600 B per sampled one).
#include "common.cuh"

namespace asvrl {

constexpr int PUSH_G = 64;    // ASVs per push group (one count slot, one scatter CTA)
constexpr int SCAT_T = 256;   // threads of a scatter CTA
constexpr int SCAN_T = 256;   // threads of the last-CTA scan

// valid transition <=> the ASV was on the map when the step started (TRAINER:157-159)
__device__ __forceinline__ bool row_valid(uint32_t status) {
  return !(status & ASVRL_S_WAS_DEACTIVATED) &&
         ((status >> ASVRL_S_INFO_SHIFT) & 15u) != (uint32_t)ASVRL_INFO_ABSENT;
}

// pass 1: valid count of every 64-ASV group (4 groups per CTA); the last CTA to finish turns them
// into exclusive offsets (deterministic (env, robot) order) and advances the ring counters.
// workspace: [0] ticket, [1] total, [2..3] base ring position (u64), [4..] per-group offsets
__global__ void __launch_bounds__(SCAN_T)
replay_count_kernel(const uint32_t* __restrict__ status, int n_asv, int n_groups, int64_t capacity,
                    unsigned long long* head_size, uint32_t* ws) {
  const int i = blockIdx.x * SCAN_T + threadIdx.x;
  const bool v = i < n_asv && row_valid(status[i]);
  const unsigned bal = __ballot_sync(0xffffffffu, v);
  __shared__ uint32_t wcnt[SCAN_T / 32];
  if ((threadIdx.x & 31) == 0) wcnt[threadIdx.x >> 5] = __popc(bal);
  __syncthreads();
  __shared__ bool is_last;
  if (threadIdx.x < SCAN_T / PUSH_G) {
    const int gidx = blockIdx.x * (SCAN_T / PUSH_G) + threadIdx.x;
    if (gidx < n_groups) ws[4 + gidx] = wcnt[2 * threadIdx.x] + wcnt[2 * threadIdx.x + 1];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    is_last = atomicAdd(&ws[0], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // exclusive scan of n_groups counts by one CTA: chunked serial + shared scan
  __shared__ uint32_t warp_tot[SCAN_T / 32];
  const int per = (n_groups + SCAN_T - 1) / SCAN_T;
  const int b0 = min(threadIdx.x * per, n_groups), b1 = min(b0 + per, n_groups);
  uint32_t s = 0;
#pragma unroll 4
  for (int b = b0; b < b1; ++b) s += __ldcg(&ws[4 + b]);
  // block-wide exclusive scan of the per-thread sums: shuffles inside a warp, then over the 8 warp totals
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[w] = inc;
  __syncthreads();
  uint32_t run = inc - s, total_new = 0;
#pragma unroll
  for (int t = 0; t < SCAN_T / 32; ++t) {
    const uint32_t c = warp_tot[t];
    if (t < w) run += c;
    total_new += c;
  }
  if (threadIdx.x == 0) {
    const unsigned long long base = head_size[0];
    ws[1] = total_new;
    ws[2] = (uint32_t)base; ws[3] = (uint32_t)(base >> 32);
    const unsigned long long total = base + total_new;
    head_size[0] = total;
    head_size[1] = total < (unsigned long long)capacity ? total : (unsigned long long)capacity;
    ws[0] = 0u;  // ticket left clean for the next call
  }
  for (int b = b0; b < b1; ++b) { uint32_t c = __ldcg(&ws[4 + b]); ws[4 + b] = run; run += c; }
}

// pass 2: a CTA owns one 64-ASV group; the (row, element) items are flattened over its threads, so
// every load is independent of the others and every store lands in a contiguous 280-byte row
__global__ void __launch_bounds__(SCAT_T)
replay_scatter_kernel(float* __restrict__ rows, int64_t capacity, const float* __restrict__ obs_self,
                      const float* __restrict__ obs_obj, const float* __restrict__ obs_mask,
                      const void* __restrict__ actions, int discrete, const float* __restrict__ reward,
                      const float* __restrict__ nself, const float* __restrict__ nobj,
                      const float* __restrict__ nmask, const uint32_t* __restrict__ status, int n_asv,
                      const uint32_t* __restrict__ ws, const float* __restrict__ tself,
                      const float* __restrict__ tobj, const float* __restrict__ tmask,
                      const uint8_t* __restrict__ ended, int rpe) {
  __shared__ long long dst_row[PUSH_G];
  __shared__ uint8_t use_term[PUSH_G];
  __shared__ uint32_t cnt0;
  const int i0 = blockIdx.x * PUSH_G;
  const int tid = threadIdx.x;
  if (tid < PUSH_G) {
    const int i = i0 + tid;
    const bool v = i < n_asv && row_valid(status[i]);
    const unsigned bal = __ballot_sync(0xffffffffu, v);
    if (tid == 0) cnt0 = __popc(bal);
    __syncwarp();
    // two warps: ranks of the second one start after the first one's count
    asm volatile("bar.sync 1, 64;");
    const uint32_t before = __popc(bal & ((1u << (tid & 31)) - 1u)) + (tid >= 32 ? cnt0 : 0u);
    const unsigned long long base = (unsigned long long)ws[2] | ((unsigned long long)ws[3] << 32);
    const unsigned long long pos = (base + ws[4 + blockIdx.x] + before) % (unsigned long long)capacity;
    dst_row[tid] = v ? (long long)pos : -1;
    use_term[tid] = (ended && i < n_asv) ? ended[i / rpe] : 0;
  }
  __syncthreads();
  // One warp per row, eight rows per warp.  A row is 70 floats = three passes of a warp (q = lane, lane + 32,
  // lane + 64), so what a lane copies is fixed for the whole launch: its three source arrays, offsets and per-ASV
  // strides are computed ONCE, and a row costs each lane three loads and three stores at `base + i * stride`.  (A flat
  // (row, element) loop spent ~150 instructions per element on index arithmetic and source selection: 21.5 M warp
  // instructions for 4.6 M elements, issue bound at 35 us; profiles/r02_scatter_*.)  The two object counts of a row
  // (RB:60-67: the mask is "first n ones") are summed from five lanes' mask loads of the third pass by shuffles.
  static_assert(ASVRL_ROW_OBJ == 7 && ASVRL_ROW_NOBJ == 32 && ASVRL_ROW_ACT == 33 && ASVRL_ROW_REW == 35 &&
                ASVRL_ROW_NSELF == 36 && ASVRL_ROW_NOBJ_F == 43 && ASVRL_ROW_NNOBJ == 68 && ASVRL_ROW_DONE == 69 &&
                ASVRL_ROW == 70 && PUSH_G == 64 && SCAT_T == 256, "lane roles of the scatter kernel");
  const int lane = tid & 31, w = tid >> 5;
  const bool have_term = ended != nullptr;
  // pass 0, q = lane: own state (7) | objects 0 .. 24
  const float* s0 = lane < 7 ? obs_self + lane : obs_obj + (lane - 7);
  const int st0 = lane < 7 ? 7 : 25;
  // pass 1, q = 32 + lane: [count] | action (2) | reward | next own state (7) | next objects 0 .. 20
  const float *s1a, *s1b;
  int st1, kind1 = 0;   // 0 float, 1 int32 action id, 3 constant zero, 4 the object count
  if (lane == 0) { s1a = s1b = reward; st1 = 1; kind1 = 4; }
  else if (lane < 3) {
    s1a = s1b = discrete ? (const float*)actions : (const float*)actions + (lane - 1);
    st1 = discrete ? 1 : 2;
    kind1 = discrete ? (lane == 1 ? 1 : 3) : 0;
  } else if (lane == 3) { s1a = s1b = reward; st1 = 1; }
  else if (lane < 11) { s1a = nself + (lane - 4); s1b = have_term ? tself + (lane - 4) : s1a; st1 = 7; }
  else { s1a = nobj + (lane - 11); s1b = have_term ? tobj + (lane - 11) : s1a; st1 = 25; }
  // pass 2, q = 64 + lane: next objects 21 .. 24 | [next count] | done; lanes 8 .. 12 / 16 .. 20 fetch the two masks
  const float *s2a = reward, *s2b = reward;
  int st2 = 1, kind2 = 5;   // 0 float, 2 status word -> done, 4 next object count, 5 nothing to store
  if (lane < 4) { s2a = nobj + (21 + lane); s2b = have_term ? tobj + (21 + lane) : s2a; st2 = 25; kind2 = 0; }
  else if (lane == 4) { kind2 = 4; }
  else if (lane == 5) { s2a = s2b = (const float*)status; kind2 = 2; }
  else if (lane >= 8 && lane < 13) { s2a = s2b = obs_mask + (lane - 8); st2 = 5; }
  else if (lane >= 16 && lane < 21) { s2a = nmask + (lane - 16); s2b = have_term ? tmask + (lane - 16) : s2a; st2 = 5; }
  constexpr int RPW = PUSH_G / (SCAT_T / 32), RB = 4;   // rows per warp, rows whose loads are in flight together
#pragma unroll
  for (int r0 = 0; r0 < RPW; r0 += RB) {
    uint32_t v0[RB], v1[RB], v2[RB];
    long long d[RB];
#pragma unroll
    for (int j = 0; j < RB; ++j) {
      const int r = w * RPW + r0 + j;
      d[j] = dst_row[r];
      v0[j] = v1[j] = v2[j] = 0u;
      if (d[j] >= 0) {   // warp-uniform
        const size_t i = (size_t)(i0 + r);
        const bool t = use_term[r];  // the env ended: its next state is the kept terminal observation
        v0[j] = __ldg(reinterpret_cast<const uint32_t*>(s0 + i * st0));
        v1[j] = __ldg(reinterpret_cast<const uint32_t*>((t ? s1b : s1a) + i * st1));
        v2[j] = __ldg(reinterpret_cast<const uint32_t*>((t ? s2b : s2a) + i * st2));
      }
    }
#pragma unroll
    for (int j = 0; j < RB; ++j) {
      if (d[j] < 0) continue;
      float* row = rows + d[j] * ASVRL_ROW;
      const float mk = __uint_as_float(v2[j]);
      float n_obj = __shfl_sync(0xffffffffu, mk, 8), n_nobj = __shfl_sync(0xffffffffu, mk, 16);
#pragma unroll
      for (int q = 1; q < 5; ++q) {
        n_obj += __shfl_sync(0xffffffffu, mk, 8 + q);
        n_nobj += __shfl_sync(0xffffffffu, mk, 16 + q);
      }
      row[lane] = __uint_as_float(v0[j]);
      float f1 = __uint_as_float(v1[j]);
      if (kind1 == 1) f1 = (float)(int32_t)v1[j];
      else if (kind1 == 3) f1 = 0.0f;
      else if (kind1 == 4) f1 = n_obj;
      row[32 + lane] = f1;
      if (kind2 != 5) {
        float f2 = mk;
        if (kind2 == 2) f2 = (v2[j] & ASVRL_S_DONE) ? 1.0f : 0.0f;
        else if (kind2 == 4) f2 = n_nobj;
        row[64 + lane] = f2;
      }
    }
  }
}

// keyed Feistel permutation of [0, 2^bits) with cycle walking down to [0, n): a bijection, so the
// first `batch` images of 0..batch-1 are DISTINCT uniform indices (random.sample, RB:28)
__device__ __forceinline__ uint64_t feistel_perm(uint64_t x, uint64_t n, int half_bits, uint2 key) {
  const uint32_t mask = half_bits >= 32 ? 0xffffffffu : ((1u << half_bits) - 1u);
  do {
    uint32_t l = (uint32_t)(x >> half_bits) & mask, r = (uint32_t)x & mask;
#pragma unroll
    for (int round = 0; round < 4; ++round) {
      uint4 f = philox4x32_10(make_uint4(r, (uint32_t)round, 0x52504c59u, 0u), key);
      uint32_t nl = r, nr = l ^ (f.x & mask);
      l = nl; r = nr;
    }
    x = ((uint64_t)l << half_bits) | r;
  } while (x >= n);
  return x;
}

__global__ void __launch_bounds__(256)
replay_sample_kernel(const float* __restrict__ rows, const unsigned long long* __restrict__ head_size,
                     int64_t capacity, const int64_t* __restrict__ indices, uint64_t seed,
                     uint64_t counter, int batch, float* __restrict__ self_b, float* __restrict__ obj_b,
                     float* __restrict__ mask_b, float* __restrict__ act_b, float* __restrict__ rew_b,
                     float* __restrict__ nself_b, float* __restrict__ nobj_b, float* __restrict__ nmask_b,
                     float* __restrict__ done_b, int64_t* __restrict__ indices_out,
                     const uint64_t* __restrict__ counter_dev) {
  const int lane = threadIdx.x & 31;
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (b >= batch) return;
  const uint64_t size = head_size[1];
  if (size == 0) return;
  if (counter_dev) counter += *counter_dev;
  uint64_t idx;
  if (indices) {
    idx = (uint64_t)indices[b] % size;
  } else {
    int bits = 2;
    while (bits < 64 && (1ull << bits) < size) ++bits;
    if (bits & 1) ++bits;
    // the key mixes the draw counter so successive batches use unrelated permutations
    uint4 kk = philox_at(seed, STREAM_REPLAY, counter, 0, 0u, 0u, 0u);
    idx = feistel_perm((uint64_t)b % size, size, bits / 2, make_uint2(kk.x, kk.y));
  }
  if (lane == 0 && indices_out) indices_out[b] = (int64_t)idx;
  const float* row = rows + (size_t)idx * ASVRL_ROW;
  // 70 floats: lanes take q, q+32, q+64
  for (int q = lane; q < ASVRL_ROW; q += 32) {
    const float v = row[q];
    if (q < ASVRL_ROW_OBJ) self_b[(size_t)b * 7 + q] = v;
    else if (q < ASVRL_ROW_NOBJ) obj_b[(size_t)b * 25 + (q - ASVRL_ROW_OBJ)] = v;
    else if (q == ASVRL_ROW_NOBJ) {
#pragma unroll
      for (int t = 0; t < 5; ++t) mask_b[(size_t)b * 5 + t] = (float)t < v ? 1.0f : 0.0f;  // RB:65
    } else if (q < ASVRL_ROW_REW) act_b[(size_t)b * 2 + (q - ASVRL_ROW_ACT)] = v;
    else if (q == ASVRL_ROW_REW) rew_b[b] = v;
    else if (q < ASVRL_ROW_NOBJ_F) nself_b[(size_t)b * 7 + (q - ASVRL_ROW_NSELF)] = v;
    else if (q < ASVRL_ROW_NNOBJ) nobj_b[(size_t)b * 25 + (q - ASVRL_ROW_NOBJ_F)] = v;
    else if (q == ASVRL_ROW_NNOBJ) {
#pragma unroll
      for (int t = 0; t < 5; ++t) nmask_b[(size_t)b * 5 + t] = (float)t < v ? 1.0f : 0.0f;
    } else done_b[b] = v;
  }
}

}  // namespace asvrl

using namespace asvrl;

extern "C" int asvrl_replay_push(float* rows, int64_t capacity, unsigned long long* head_size,
                                 const float* obs_self, const float* obs_objects,
                                 const float* obs_mask, const void* actions, int discrete,
                                 const float* reward, const float* next_self,
                                 const float* next_objects, const float* next_mask,
                                 const uint32_t* status, int n_asv, uint32_t* workspace,
                                 const float* term_self, const float* term_objects,
                                 const float* term_mask, const uint8_t* ended, int robots_per_env,
                                 void* stream) {
  if (!rows || !head_size || !obs_self || !obs_objects || !obs_mask || !actions || !reward ||
      !next_self || !next_objects || !next_mask || !status || !workspace)
    return fail(ASVRL_E_BADARG, "asvrl_replay_push: null pointer");
  if (capacity < 1 || n_asv < 0) return fail(ASVRL_E_BADARG, "asvrl_replay_push: bad sizes");
  if ((int64_t)n_asv > capacity) return fail(ASVRL_E_LIMIT, "asvrl_replay_push: n_asv %d exceeds capacity", n_asv);
  if (ended && (!term_self || !term_objects || !term_mask || robots_per_env < 1))
    return fail(ASVRL_E_BADARG, "asvrl_replay_push: ended needs term_* and robots_per_env >= 1");
  if (n_asv == 0) return 0;
  const int n_groups = (n_asv + PUSH_G - 1) / PUSH_G;
  cudaStream_t s = (cudaStream_t)stream;
  replay_count_kernel<<<(n_asv + SCAN_T - 1) / SCAN_T, SCAN_T, 0, s>>>(status, n_asv, n_groups, capacity,
                                                                      head_size, workspace);
  int rc = check_launch("asvrl_replay_push(count)");
  if (rc) return rc;
  replay_scatter_kernel<<<n_groups, SCAT_T, 0, s>>>(rows, capacity, obs_self, obs_objects, obs_mask,
                                                actions, discrete, reward, next_self, next_objects,
                                                next_mask, status, n_asv, workspace, term_self,
                                                term_objects, term_mask, ended,
                                                robots_per_env < 1 ? 1 : robots_per_env);
  return check_launch("asvrl_replay_push(scatter)");
}

extern "C" int asvrl_replay_sample(const float* rows, const unsigned long long* head_size,
                                   int64_t capacity, const int64_t* indices, uint64_t seed,
                                   uint64_t counter, int batch, float* self_b, float* obj_b,
                                   float* mask_b, float* act_b, float* rew_b, float* nself_b,
                                   float* nobj_b, float* nmask_b, float* done_b,
                                   int64_t* indices_out, const uint64_t* counter_dev, void* stream) {
  if (!rows || !head_size || !self_b || !obj_b || !mask_b || !act_b || !rew_b || !nself_b ||
      !nobj_b || !nmask_b || !done_b)
    return fail(ASVRL_E_BADARG, "asvrl_replay_sample: null pointer");
  if (batch < 1 || capacity < 1) return fail(ASVRL_E_BADARG, "asvrl_replay_sample: bad sizes");
  const int grid = (batch * 32 + 255) / 256;
  replay_sample_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      rows, head_size, capacity, indices, seed, counter, batch, self_b, obj_b, mask_b, act_b, rew_b,
      nself_b, nobj_b, nmask_b, done_b, indices_out, counter_dev);
  return check_launch("asvrl_replay_sample");
}
