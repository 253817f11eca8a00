// policy_umma.cu — K9: the collection policy forward of the AC-IQN / DDPG actor for every ASV in ONE
// persistent, warp-specialised launch (sm_100a: tcgen05.mma kind::f16, accumulators in TMEM, weights
// by TMA bulk copy, mbarrier pipelines).
//
// Replaces, for all E*R ASVs of a batched step (reference: one batch-1 forward per ASV),
//   rfarl/rfarl/policy/AC_IQN_model.py:284-323  Actor.forward: self / object encoders, mask, concat,
//                                               256 -> 128 -> 128 (ReLU), 128 -> 2, (2/pi) atan
//   rfarl/rfarl/agent.py:207-225                act_ac_iqn's epsilon-greedy choice
// i.e. K6 + 2 x K8 + asvrl_actor_head of round 1 (4 launches and two 33 MB activation round trips at
// 65 536 ASVs).  HBM traffic is now the algorithmic minimum: 148 B in + 8 B out per ASV.
//
// Arithmetic.  The dense layers are fp32 in the reference and the parity bar is 1e-5, so every
// operand is split in two fp16 pieces, x = x_hi + x_lo (x_hi = fp16(x), x_lo = fp16(x - x_hi); 22
// mantissa bits) and three products are accumulated in fp32 in TMEM, small terms first:
//     x W^T ~= x_lo W_hi^T + x_hi W_lo^T + x_hi W_hi^T.
// W is scaled by 2^6 before the split (its low piece would otherwise be an fp16 subnormal) and the
// accumulator by 2^-6 afterwards, both exact.  Measured against the float64 module path: 8e-7 on
// the actions (an fp32 SIMT GEMM: 1.4e-6).  fp16 pieces halve shared memory and tensor time
// against the tf32 split of K8.
//
// Mapping.  One CTA per SM walks 128-row tiles (tile = blockIdx.x + i * gridDim.x).  16 warps:
//   warp 0      W producer: one thread streams the pre-split, pre-swizzled weight chunks (32 KB: hi and
//               lo tile of 128 out x 64 k) into the ring with cp.async.bulk
//   warp 1      MMA issuer: one thread, 12 tcgen05.mma per chunk, tcgen05.commit to the ring's empty
//               barrier and to the accumulator-full barriers
//   warp 2      input producer: the tile's raw observations (18.9 KB) by three bulk copies
//   warp 3      TMEM allocation
//   warps 4-7   epilogue: TMEM -> bias, ReLU, split -> layer 2's A chunks in the ring; then TMEM ->
//               bias, ReLU, 128 -> 2, atan, epsilon-greedy (Philox stream 5) -> 8 B per ASV
//   warps 8-15  tower: lane = row, the encoders' contractions of length 7 / 5 straight into the
//               swizzled fp16 A chunks of layer 1 (weights broadcast from shared memory)
// The ring has 3 stages of 64 KB (A_hi, A_lo, W_hi, W_lo tiles of 128 rows x 128 B, SWIZZLE_128B,
// K-major); a tile is 4 chunks of layer 1 and 2 of layer 2.  The MMA order is software pipelined
// across tiles — L1(0), L1(1), L2(0), L1(2), L2(1), ... — with both accumulators double buffered in
// TMEM (4 x 128 columns), so the tensor pipe works on the next tile's first layer while the epilogue
// warps turn this tile's first-layer accumulator into the second layer's operand.
#include "umma.cuh"

namespace asvrl {

constexpr int PF_THREADS = 512;
constexpr int PF_ROWS = 128;                       // rows per tile = UMMA M
constexpr int PF_SD = 7, PF_OD = 5, PF_NOBJ = 5, PF_SF = 56, PF_OF = 40, PF_W = 256, PF_H = 128;
constexpr int PF_TILE_B = PF_ROWS * 128;           // one operand tile: 128 rows x 128 B = 16 KB
constexpr int PF_CHUNK_B = 2 * PF_TILE_B;          // hi | lo pieces of a 128 x 64 operand chunk
constexpr int PF_STAGE_B = 4 * PF_TILE_B;          // A_hi | A_lo | W_hi | W_lo
constexpr int PF_STAGES = 3;
constexpr int PF_L1_CHUNKS = PF_W / 64, PF_L2_CHUNKS = PF_H / 64;
constexpr int PF_IMAGE_B = (PF_L1_CHUNKS + PF_L2_CHUNKS) * 2 * PF_TILE_B;   // 192 KB packed weights
constexpr float PF_WSCALE = 64.0f, PF_WUNSCALE = 1.0f / 64.0f;
constexpr float PF_F16_MAX = 60000.0f;

// shared-memory map (bytes from the 1024-aligned base)
constexpr int PF_OFF_RING = 0;
constexpr int PF_OFF_SELF = PF_STAGES * PF_STAGE_B;                   // 196608
constexpr int PF_IN_SELF_B = PF_ROWS * PF_SD * 4;                     // 3584
constexpr int PF_IN_OBJ_B = PF_ROWS * PF_NOBJ * PF_OD * 4;            // 12800
constexpr int PF_IN_MASK_B = PF_ROWS * PF_NOBJ * 4;                   // 2560
constexpr int PF_OFF_OBJ = PF_OFF_SELF + PF_IN_SELF_B;
constexpr int PF_OFF_MASK = PF_OFF_OBJ + PF_IN_OBJ_B;
constexpr int PF_OFF_TABS = PF_OFF_MASK + PF_IN_MASK_B;               // self table [56][8] floats
constexpr int PF_OFF_TABO = PF_OFF_TABS + PF_SF * 32;                 // object table [40][8] floats
constexpr int PF_OFF_B1 = PF_OFF_TABO + PF_OF * 32;
constexpr int PF_OFF_B2 = PF_OFF_B1 + PF_H * 4;
constexpr int PF_OFF_WOUT = PF_OFF_B2 + PF_H * 4;                     // [2][128]
constexpr int PF_OFF_BAR = PF_OFF_WOUT + 2 * PF_H * 4;
constexpr int PF_NBAR = 5 * PF_STAGES + 2 + 8;   // full, empty x {W, tower team 0, tower team 1, epilogue}, input, accumulators
constexpr int PF_OFF_MISC = PF_OFF_BAR + PF_NBAR * 8;                 // tmem slot, abort flag
constexpr int PF_SMEM_B = PF_OFF_MISC + 16 + 1024;                    // + alignment slack
static_assert(PF_SMEM_B <= 232448, "K9 shared memory");
static_assert(PF_OFF_BAR % 8 == 0 && PF_OFF_TABS % 16 == 0 && PF_OFF_OBJ % 16 == 0 && PF_OFF_MASK % 16 == 0, "alignment");

// position of a chunk in the ring's consumption order: L1(0), then for every tile i: L1(i+1) (when it
// exists) followed by L2(i)
__device__ __forceinline__ int pf_pos_l1(int i) { return i == 0 ? 0 : 6 * i - 2; }
__device__ __forceinline__ int pf_pos_l2(int i, int n) { return i + 1 < n ? 6 * i + 8 : 6 * i + 4; }
// who fills the A tiles of position `pos` (n tiles in this CTA): 0 / 1 = tower team (chunk parity), 2 = epilogue.
// An mbarrier wait tells phases apart by parity only, so a waiter may be at most one phase away from the
// barrier: every producer role therefore has its OWN "stage free" barrier per stage, and the MMA thread
// commits the release of position p to the barrier of the role that owns position p + PF_STAGES.
__device__ __forceinline__ int pf_role(int pos, int n) {
  if (pos < 4) return pos & 1;
  const int q = pos - 4, blk = q / 6, r = q - 6 * blk;
  return (blk + 1 < n && r < 4) ? (r & 1) : 2;
}

struct PolicyArgs {
  const float *self_obs, *objects, *mask;                       // [rows][7], [rows][5][5], [rows][5]
  const float *w_self, *b_self, *w_obj, *b_obj;                 // encoders (nn.Linear layout)
  const float *b1, *b2, *w_out, *b_out;                         // hidden biases, head [2][128], [2]
  const unsigned char* image;                                   // packed W1 | W2 chunks (policy_pack_kernel)
  float* hidden;                                                  // non-null: store the second hidden layer [rows][128], no head
  unsigned char* hidden_img;                                      // non-null: ... as a tcgen05 operand image [tile][2][32 KB] (hi | lo fp16, swizzled)
  float atan_scale, eps;
  const float* eps_dev;
  float lo0, hi0, lo1, hi1;
  uint64_t seed, counter;
  const uint64_t* counter_dev;
  int rows, n_tiles, bulk_ok;
  int* status;
  float* actions;                                               // [rows][2]
};

// ---- weight image: chunk c of layer L = hi tile (n = 0..127 rows of 128 B, swizzled) then lo tile ----
__global__ void __launch_bounds__(256)
policy_pack_kernel(const float* __restrict__ w1, const float* __restrict__ w2, unsigned char* __restrict__ image) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;   // one 16-byte chunk (8 k) of one row
  const int n1 = PF_L1_CHUNKS * PF_H * 8;
  if (t >= n1 + PF_L2_CHUNKS * PF_H * 8) return;
  const bool l1 = t < n1;
  const int u = l1 ? t : t - n1;
  const int c = u / (PF_H * 8), n = (u / 8) % PF_H, oc = u & 7;
  const int K = l1 ? PF_W : PF_H;
  const float* src = (l1 ? w1 : w2) + (size_t)n * K + c * 64 + oc * 8;
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = fminf(fmaxf(src[i] * PF_WSCALE, -PF_F16_MAX), PF_F16_MAX);
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) um_split2(v[2 * i], v[2 * i + 1], h[i], l[i]);
  unsigned char* tile = image + (size_t)((l1 ? 0 : PF_L1_CHUNKS) + c) * (2 * PF_TILE_B) + um_sw128_off(n, oc);
  *reinterpret_cast<uint4*>(tile) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(tile + PF_TILE_B) = make_uint4(l[0], l[1], l[2], l[3]);
}

// ---- tower octets: 8 consecutive features of this thread's row -> one 16-byte chunk of A_hi / A_lo ----
// tab: [feature][8] floats; self rows {w0..w6, b}, object rows {w0..w4, b, 0, 0}.  FMA order = K6's.
template <bool SELF>
__device__ __forceinline__ void pf_octets(const float4* __restrict__ tab, int oct0, int n_oct, int col0, uint32_t a_row,
                                          int r7, float x0, float x1, float x2, float x3, float x4, float x5, float x6,
                                          bool on) {
#pragma unroll 1
  for (int t = 0; t < n_oct; ++t) {
    const float4* w = tab + (oct0 + t) * 16;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 wa = w[2 * j], wb = w[2 * j + 1];
      float acc;
      if (SELF) {
        acc = wb.w;
        acc = fmaf(wa.x, x0, acc); acc = fmaf(wa.y, x1, acc); acc = fmaf(wa.z, x2, acc); acc = fmaf(wa.w, x3, acc);
        acc = fmaf(wb.x, x4, acc); acc = fmaf(wb.y, x5, acc); acc = fmaf(wb.z, x6, acc);
      } else {
        acc = wb.y;
        acc = fmaf(wa.x, x0, acc); acc = fmaf(wa.y, x1, acc); acc = fmaf(wa.z, x2, acc); acc = fmaf(wa.w, x3, acc);
        acc = fmaf(wb.x, x4, acc);
      }
      v[j] = on ? fminf(fmaxf(acc, 0.0f), PF_F16_MAX) : 0.0f;
    }
    um_put8(a_row + (uint32_t)(((col0 + t) ^ r7) << 4), PF_TILE_B, v);
  }
}

__global__ void __launch_bounds__(PF_THREADS, 1)
policy_forward_kernel(const PolicyArgs a) {
  extern __shared__ __align__(1024) unsigned char pf_raw[];
  unsigned char* base = pf_raw + ((1024u - (um_smem(pf_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space
  const uint32_t sb = um_smem(base);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  const uint32_t bar0 = sb + PF_OFF_BAR;
  auto bar_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_empty_w = [&](int s) { return bar0 + 8u * (PF_STAGES + s); };                 // W tiles of stage s free
  auto bar_empty_a = [&](int role, int s) { return bar0 + 8u * ((2 + role) * PF_STAGES + s); };  // A tiles free, per role
  const uint32_t bar_in_full = bar0 + 8u * (5 * PF_STAGES), bar_in_empty = bar_in_full + 8u;
  auto bar_d1_full = [&](int b) { return bar_in_empty + 8u + 8u * b; };
  auto bar_d1_empty = [&](int b) { return bar_in_empty + 24u + 8u * b; };
  auto bar_d2_full = [&](int b) { return bar_in_empty + 40u + 8u * b; };
  auto bar_d2_empty = [&](int b) { return bar_in_empty + 56u + 8u * b; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base + PF_OFF_MISC);
  volatile uint32_t* abort_flag = reinterpret_cast<volatile uint32_t*>(base + PF_OFF_MISC + 4);

  // ---- one-time setup ----
  if (tid == 0) {
    for (int s = 0; s < PF_STAGES; ++s) {
      um_mbar_init(bar_full(s), 1 + 4);   // W producer's expect_tx arrive + 4 producing warps
      um_mbar_init(bar_empty_w(s), 1);    // tcgen05.commit
      for (int role = 0; role < 3; ++role) um_mbar_init(bar_empty_a(role, s), 1);
    }
    um_mbar_init(bar_in_full, 1);
    um_mbar_init(bar_in_empty, 8);        // the 8 tower warps
    for (int b = 0; b < 2; ++b) {
      um_mbar_init(bar_d1_full(b), 1);
      um_mbar_init(bar_d1_empty(b), 4);   // the 4 epilogue warps
      um_mbar_init(bar_d2_full(b), 1);
      um_mbar_init(bar_d2_empty(b), 4);
    }
    *abort_flag = 0u;
    um_mbar_init_fence();
  }
  if (warp == 3) um_tmem_alloc<512>(um_smem(tmem_slot));
  {
    // encoder tables, biases and the head into shared memory
    float* tabs = reinterpret_cast<float*>(base + PF_OFF_TABS);
    for (int i = tid; i < PF_SF * 8; i += PF_THREADS) {
      const int f = i >> 3, q = i & 7;
      tabs[i] = q < PF_SD ? a.w_self[f * PF_SD + q] : a.b_self[f];
    }
    float* tabo = reinterpret_cast<float*>(base + PF_OFF_TABO);
    for (int i = tid; i < PF_OF * 8; i += PF_THREADS) {
      const int f = i >> 3, q = i & 7;
      tabo[i] = q < PF_OD ? a.w_obj[f * PF_OD + q] : (q == PF_OD ? a.b_obj[f] : 0.0f);
    }
    float* b1 = reinterpret_cast<float*>(base + PF_OFF_B1);
    float* b2 = reinterpret_cast<float*>(base + PF_OFF_B2);
    float* wo = reinterpret_cast<float*>(base + PF_OFF_WOUT);
    for (int i = tid; i < PF_H; i += PF_THREADS) { b1[i] = a.b1[i]; b2[i] = a.b2[i]; }
    for (int i = tid; i < 2 * PF_H; i += PF_THREADS) wo[i] = a.w_out ? a.w_out[i] : 0.0f;
  }
  um_tc_before();
  __syncthreads();
  um_tc_after();
  const uint32_t tmem = *tmem_slot;

  const int n_my = ((int)blockIdx.x < a.n_tiles) ? (a.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == 0) {
    // =============== W producer ===============
    if (lane == 0) {
      int pos = 0;
      uint32_t ph = 0;   // bit s: parity of the next wait on stage s
      bool ok = true;
      auto emit = [&](int first_chunk, int n_chunks) {
        for (int c = 0; c < n_chunks && ok; ++c, ++pos) {
          const int s = pos % PF_STAGES;
          if (pos >= PF_STAGES) {
            ok = um_mbar_wait(bar_empty_w(s), (ph >> s) & 1u, abort_flag, a.status, UM_TIMEOUT_RING);
            ph ^= 1u << s;
          }
          if (!ok) break;
          um_mbar_arrive_expect_tx(bar_full(s), 2 * PF_TILE_B);
          um_bulk_g2s(sb + PF_OFF_RING + s * PF_STAGE_B + 2 * PF_TILE_B,
                      a.image + (size_t)(first_chunk + c) * (2 * PF_TILE_B), 2 * PF_TILE_B, bar_full(s));
        }
      };
      if (n_my > 0) emit(0, PF_L1_CHUNKS);
      for (int i = 0; i < n_my && ok; ++i) {
        if (i + 1 < n_my) emit(0, PF_L1_CHUNKS);
        emit(PF_L1_CHUNKS, PF_L2_CHUNKS);
      }
    }
  } else if (warp == 1) {
    // =============== MMA issuer ===============
    if (lane == 0) {
      constexpr uint32_t idesc = um_idesc_f16(PF_ROWS, PF_H);
      int pos = 0;
      bool ok = true;
      auto layer = [&](uint32_t d_tmem, int n_chunks) {
        for (int c = 0; c < n_chunks && ok; ++c, ++pos) {
          const int s = pos % PF_STAGES;
          ok = um_mbar_wait(bar_full(s), (pos / PF_STAGES) & 1, abort_flag, a.status, UM_TIMEOUT_RING);
          if (!ok) break;
          um_tc_after();
          const uint32_t st = sb + PF_OFF_RING + s * PF_STAGE_B;
          const uint32_t ah = st, al = st + PF_TILE_B, wh = st + 2 * PF_TILE_B, wl = st + 3 * PF_TILE_B;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {   // 4 x (K = 16 fp16 = 32 B) per 128-byte row
            const uint32_t o = ks * 32;
            um_mma_f16(d_tmem, um_desc_sw128(al + o), um_desc_sw128(wh + o), idesc, (c | ks) ? 1u : 0u);  // small terms first
            um_mma_f16(d_tmem, um_desc_sw128(ah + o), um_desc_sw128(wl + o), idesc, 1u);
            um_mma_f16(d_tmem, um_desc_sw128(ah + o), um_desc_sw128(wh + o), idesc, 1u);
          }
          um_commit(bar_empty_w(s));
          if (pos + PF_STAGES < 6 * n_my) um_commit(bar_empty_a(pf_role(pos + PF_STAGES, n_my), s));
        }
      };
      auto l1 = [&](int i) {
        const int b = i & 1;
        ok = ok && um_mbar_wait(bar_d1_empty(b), ((i >> 1) & 1) ^ 1, abort_flag, a.status, UM_TIMEOUT_ACC);
        um_tc_after();
        layer(tmem + (uint32_t)(b * PF_H), PF_L1_CHUNKS);
        um_commit(bar_d1_full(b));
      };
      auto l2 = [&](int i) {
        const int b = i & 1;
        ok = ok && um_mbar_wait(bar_d2_empty(b), ((i >> 1) & 1) ^ 1, abort_flag, a.status, UM_TIMEOUT_ACC);
        um_tc_after();
        layer(tmem + (uint32_t)(2 * PF_H + b * PF_H), PF_L2_CHUNKS);
        um_commit(bar_d2_full(b));
      };
      if (n_my > 0) l1(0);
      for (int i = 0; i < n_my && ok; ++i) {
        if (i + 1 < n_my) l1(i + 1);
        l2(i);
      }
    }
  } else if (warp == 2) {
    // =============== input producer ===============
    bool ok = true;
    for (int i = 0; i < n_my && ok; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int row0 = tile * PF_ROWS;
      const int nrows = min(PF_ROWS, a.rows - row0);
      ok = um_mbar_wait(bar_in_empty, (i & 1) ^ 1, abort_flag, a.status, UM_TIMEOUT_INPUT);
      if (!ok) break;
      if (nrows == PF_ROWS && a.bulk_ok) {
        if (lane == 0) {
          um_mbar_arrive_expect_tx(bar_in_full, PF_IN_SELF_B + PF_IN_OBJ_B + PF_IN_MASK_B);
          um_bulk_g2s(sb + PF_OFF_SELF, a.self_obs + (size_t)row0 * PF_SD, PF_IN_SELF_B, bar_in_full);
          um_bulk_g2s(sb + PF_OFF_OBJ, a.objects + (size_t)row0 * PF_NOBJ * PF_OD, PF_IN_OBJ_B, bar_in_full);
          um_bulk_g2s(sb + PF_OFF_MASK, a.mask + (size_t)row0 * PF_NOBJ, PF_IN_MASK_B, bar_in_full);
        }
      } else {
        // ragged last tile (or pointers that are not 16-byte aligned): plain loads, zero fill
        float* ds = reinterpret_cast<float*>(base + PF_OFF_SELF);
        float* dob = reinterpret_cast<float*>(base + PF_OFF_OBJ);
        float* dm = reinterpret_cast<float*>(base + PF_OFF_MASK);
        const float* ss = a.self_obs + (size_t)row0 * PF_SD;
        const float* so = a.objects + (size_t)row0 * PF_NOBJ * PF_OD;
        const float* sm = a.mask + (size_t)row0 * PF_NOBJ;
        for (int j = lane; j < PF_ROWS * PF_SD; j += 32) ds[j] = j < nrows * PF_SD ? ss[j] : 0.0f;
        for (int j = lane; j < PF_ROWS * PF_NOBJ * PF_OD; j += 32) dob[j] = j < nrows * PF_NOBJ * PF_OD ? so[j] : 0.0f;
        for (int j = lane; j < PF_ROWS * PF_NOBJ; j += 32) dm[j] = j < nrows * PF_NOBJ ? sm[j] : 0.0f;
        __syncwarp();
        if (lane == 0) um_mbar_arrive(bar_in_full);
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // =============== epilogue warps: TMEM lanes [32 q, 32 q + 32) ===============
    const int q = warp - 4;
    const int r = 32 * q + lane, r7 = r & 7;
    const uint32_t row_off = (uint32_t)((r >> 3) * 1024 + r7 * 128);
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16);
    const float4* b1 = reinterpret_cast<const float4*>(base + PF_OFF_B1);
    const float4* b2 = reinterpret_cast<const float4*>(base + PF_OFF_B2);
    const float4* w0 = reinterpret_cast<const float4*>(base + PF_OFF_WOUT);
    const float4* w1 = w0 + PF_H / 4;
    const float bo0 = a.b_out[0], bo1 = a.b_out[1];
    const uint64_t counter = a.counter + (a.counter_dev ? *a.counter_dev : 0ull);
    const float eps = a.eps_dev ? *a.eps_dev : a.eps;
    bool ok = true;
    uint32_t ph = 0;
    for (int i = 0; i < n_my && ok; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int b = i & 1;
      // ---- epilogue 1: D1 -> h1 = relu(D1 / 64 + b1) -> the two A chunks of layer 2 ----
      ok = um_mbar_wait(bar_d1_full(b), (i >> 1) & 1, abort_flag, a.status, UM_TIMEOUT_ACC);
      if (!ok) break;
      um_tc_after();
      const int p2 = pf_pos_l2(i, n_my);
      for (int j = 0; j < PF_L2_CHUNKS && ok; ++j) {
        const int pos = p2 + j, s = pos % PF_STAGES;
        if (pos >= PF_STAGES) {
          ok = um_mbar_wait(bar_empty_a(2, s), (ph >> s) & 1u, abort_flag, a.status, UM_TIMEOUT_RING);
          ph ^= 1u << s;
        }
        if (!ok) break;
        const uint32_t a_row = sb + PF_OFF_RING + s * PF_STAGE_B + row_off;
#pragma unroll
        for (int hlf = 0; hlf < 2; ++hlf) {
          uint32_t v[32];
          um_tmem_ld32(lane_addr + (uint32_t)(b * PF_H + 64 * j + 32 * hlf), v);
          um_tmem_ld_wait();
#pragma unroll
          for (int oc = 0; oc < 4; ++oc) {
            float h[8];
            const float4 ba = b1[(64 * j + 32 * hlf + 8 * oc) / 4], bb = b1[(64 * j + 32 * hlf + 8 * oc) / 4 + 1];
            const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
#pragma unroll
            for (int e = 0; e < 8; ++e)
              h[e] = fminf(fmaxf(fmaf(__uint_as_float(v[8 * oc + e]), PF_WUNSCALE, bias[e]), 0.0f), PF_F16_MAX);
            um_put8(a_row + (uint32_t)(((4 * hlf + oc) ^ r7) << 4), PF_TILE_B, h);
          }
        }
        um_fence_async_smem();
        __syncwarp();
        if (lane == 0) um_mbar_arrive(bar_full(s));
      }
      if (!ok) break;
      um_tc_before();
      __syncwarp();
      if (lane == 0) um_mbar_arrive(bar_d1_empty(b));
      // ---- epilogue 2: D2 -> relu(D2 / 64 + b2) . w_out -> atan -> epsilon-greedy ----
      ok = um_mbar_wait(bar_d2_full(b), (i >> 1) & 1, abort_flag, a.status, UM_TIMEOUT_ACC);
      if (!ok) break;
      um_tc_after();
      float a0 = 0.0f, a1 = 0.0f;
      const long m = (long)tile * PF_ROWS + r;
      float4* hrow = (a.hidden && m < a.rows) ? reinterpret_cast<float4*>(a.hidden + (size_t)m * PF_H) : nullptr;
      unsigned char* irow = a.hidden_img ? a.hidden_img + (size_t)tile * 2 * PF_CHUNK_B : nullptr;
      const bool live_row = m < a.rows;
#pragma unroll
      for (int c0 = 0; c0 < PF_H; c0 += 32) {
        uint32_t v[32];
        um_tmem_ld32(lane_addr + (uint32_t)(2 * PF_H + b * PF_H + c0), v);
        um_tmem_ld_wait();
#pragma unroll
        for (int gp = 0; gp < 4; ++gp) {
          float h[8];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int g = 2 * gp + half;
            const float4 bb = b2[c0 / 4 + g], u0 = w0[c0 / 4 + g], u1 = w1[c0 / 4 + g];
            const float h0 = fmaxf(fmaf(__uint_as_float(v[4 * g + 0]), PF_WUNSCALE, bb.x), 0.0f);
            const float h1 = fmaxf(fmaf(__uint_as_float(v[4 * g + 1]), PF_WUNSCALE, bb.y), 0.0f);
            const float h2 = fmaxf(fmaf(__uint_as_float(v[4 * g + 2]), PF_WUNSCALE, bb.z), 0.0f);
            const float h3 = fmaxf(fmaf(__uint_as_float(v[4 * g + 3]), PF_WUNSCALE, bb.w), 0.0f);
            if (hrow) hrow[c0 / 4 + g] = make_float4(h0, h1, h2, h3);   // the policy family's shared trunk output
            h[4 * half + 0] = h0; h[4 * half + 1] = h1; h[4 * half + 2] = h2; h[4 * half + 3] = h3;
            a0 = fmaf(h0, u0.x, a0); a0 = fmaf(h1, u0.y, a0); a0 = fmaf(h2, u0.z, a0); a0 = fmaf(h3, u0.w, a0);
            a1 = fmaf(h0, u1.x, a1); a1 = fmaf(h1, u1.y, a1); a1 = fmaf(h2, u1.z, a1); a1 = fmaf(h3, u1.w, a1);
          }
          if (irow) {   // octet (c0 + 8 gp) / 8 of the row, two fp16 pieces, in the K-major 128-byte-swizzled tile layout
            const int k0 = c0 + 8 * gp;
            unsigned char* dst = irow + (size_t)(k0 >> 6) * PF_CHUNK_B + um_sw128_off(r, (k0 & 63) >> 3);
            uint32_t hh[4], ll[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) um_split2(live_row ? h[2 * i] : 0.0f, live_row ? h[2 * i + 1] : 0.0f, hh[i], ll[i]);
            *reinterpret_cast<uint4*>(dst) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
            *reinterpret_cast<uint4*>(dst + PF_TILE_B) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
          }
        }
      }
      um_tc_before();
      __syncwarp();
      if (lane == 0) um_mbar_arrive(bar_d2_empty(b));
      if (m < a.rows && !a.hidden && !a.hidden_img) {
        float2 act = make_float2(a.atan_scale * atanf(a0 + bo0), a.atan_scale * atanf(a1 + bo1));
        if (eps > 0.0f) {
          const uint4 u = philox_at(a.seed, STREAM_EXPLORE, counter, (int64_t)m, 0u, 0u, 0u);
          if (u01(u.x) < eps) act = make_float2(a.lo0 + (a.hi0 - a.lo0) * u01(u.y), a.lo1 + (a.hi1 - a.lo1) * u01(u.z));
        }
        reinterpret_cast<float2*>(a.actions)[m] = act;
      }
    }
  } else if (warp >= 8) {
    // =============== tower warps: rows [32 g, 32 g + 32), chunks h and h + 2 of layer 1 ===============
    const int tw = warp - 8, g = tw & 3, h = tw >> 2;
    const int r = 32 * g + lane, r7 = r & 7;
    const uint32_t row_off = (uint32_t)((r >> 3) * 1024 + r7 * 128);
    const float4* tabs = reinterpret_cast<const float4*>(base + PF_OFF_TABS);
    const float4* tabo = reinterpret_cast<const float4*>(base + PF_OFF_TABO);
    const float* in_s = reinterpret_cast<const float*>(base + PF_OFF_SELF) + r * PF_SD;
    const float* in_o = reinterpret_cast<const float*>(base + PF_OFF_OBJ) + r * PF_NOBJ * PF_OD;
    const float* in_m = reinterpret_cast<const float*>(base + PF_OFF_MASK) + r * PF_NOBJ;
    bool ok = true;
    uint32_t ph = 0;
    for (int i = 0; i < n_my && ok; ++i) {
      ok = um_mbar_wait(bar_in_full, i & 1, abort_flag, a.status, UM_TIMEOUT_INPUT);
      if (!ok) break;
      float s[PF_SD], o[PF_NOBJ][PF_OD];
      bool on[PF_NOBJ];
#pragma unroll
      for (int k = 0; k < PF_SD; ++k) s[k] = in_s[k];
#pragma unroll
      for (int k = 0; k < PF_NOBJ; ++k) {
#pragma unroll
        for (int d = 0; d < PF_OD; ++d) o[k][d] = in_o[k * PF_OD + d];
        on[k] = in_m[k] >= 0.5f;
      }
      __syncwarp();
      if (lane == 0) um_mbar_arrive(bar_in_empty);
      const int p1 = pf_pos_l1(i);
#pragma unroll
      for (int rep = 0; rep < 2 && ok; ++rep) {
        const int cc = h + 2 * rep;
        const int pos = p1 + cc, st = pos % PF_STAGES;
        if (pos >= PF_STAGES) {
          ok = um_mbar_wait(bar_empty_a(h, st), (ph >> st) & 1u, abort_flag, a.status, UM_TIMEOUT_RING);
          ph ^= 1u << st;
        }
        if (!ok) break;
        const uint32_t a_row = sb + PF_OFF_RING + st * PF_STAGE_B + row_off;
        // feature octets 8 cc .. 8 cc + 7: octets 0-6 are the self encoder, then 5 per object
#define PF_OBJ(k, o0, n, c0) pf_octets<false>(tabo, o0, n, c0, a_row, r7, o[k][0], o[k][1], o[k][2], o[k][3], o[k][4], 0.f, 0.f, on[k])
        if (cc == 0) {
          pf_octets<true>(tabs, 0, 7, 0, a_row, r7, s[0], s[1], s[2], s[3], s[4], s[5], s[6], true);
          PF_OBJ(0, 0, 1, 7);
        } else if (cc == 1) {
          PF_OBJ(0, 1, 4, 0);
          PF_OBJ(1, 0, 4, 4);
        } else if (cc == 2) {
          PF_OBJ(1, 4, 1, 0);
          PF_OBJ(2, 0, 5, 1);
          PF_OBJ(3, 0, 2, 6);
        } else {
          PF_OBJ(3, 2, 3, 0);
          PF_OBJ(4, 0, 5, 3);
        }
#undef PF_OBJ
        um_fence_async_smem();
        __syncwarp();
        if (lane == 0) um_mbar_arrive(bar_full(st));
      }
    }
  }

  um_tc_before();
  __syncthreads();
  if (warp == 3) {
    um_tc_after();
    um_tmem_free<512>(tmem);
  }
}

static __device__ int pf_status_word = 0;

}  // namespace asvrl

using namespace asvrl;

extern "C" long asvrl_policy_forward_workspace(void) { return (long)PF_IMAGE_B; }

static int pf_run(const char* what, const float* self_obs, const float* objects, const float* mask,
                  const float* w_self, const float* b_self, const float* w_obj, const float* b_obj,
                  const float* w1, const float* b1, const float* w2, const float* b2,
                  const float* w_out, const float* b_out, float atan_scale, const float* eps_dev,
                  float eps, const float* action_lo, const float* action_hi, uint64_t seed,
                  uint64_t counter, const uint64_t* counter_dev, long rows, int repack,
                  void* workspace, float* actions, float* hidden, void* hidden_img, void* stream) {
  if (!self_obs || !objects || !mask || !w_self || !b_self || !w_obj || !b_obj || !w1 || !b1 || !w2 || !b2 || !workspace)
    return fail(ASVRL_E_BADARG, "%s: null pointer", what);
  if (rows < 1 || rows > 0x7FFFFFFFL - PF_ROWS) return fail(ASVRL_E_BADARG, "asvrl_policy_forward: rows %ld", rows);
  if ((uintptr_t)workspace & 127) return fail(ASVRL_E_BADARG, "asvrl_policy_forward: workspace must be 128-byte aligned");
  static int n_sm = 0;
  cudaError_t e;
  if (n_sm == 0) {
    int dev = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess || (e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess)
      return fail((int)e, "asvrl_policy_forward: %s", cudaGetErrorString(e));
    e = cudaFuncSetAttribute(policy_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PF_SMEM_B);
    if (e != cudaSuccess) { n_sm = 0; return fail((int)e, "asvrl_policy_forward: %d B shared memory: %s", PF_SMEM_B, cudaGetErrorString(e)); }
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (repack) {
    const int n = (PF_L1_CHUNKS + PF_L2_CHUNKS) * PF_H * 8;
    policy_pack_kernel<<<(n + 255) / 256, 256, 0, st>>>(w1, w2, (unsigned char*)workspace);
    if (int rc = check_launch("asvrl_policy_forward (pack)")) return rc;
  }
  PolicyArgs a;
  a.self_obs = self_obs; a.objects = objects; a.mask = mask;
  a.w_self = w_self; a.b_self = b_self; a.w_obj = w_obj; a.b_obj = b_obj;
  a.b1 = b1; a.b2 = b2; a.w_out = w_out; a.b_out = b_out;
  a.image = (const unsigned char*)workspace;
  a.atan_scale = atan_scale; a.eps = eps; a.eps_dev = eps_dev;
  a.lo0 = action_lo[0]; a.hi0 = action_hi[0]; a.lo1 = action_lo[1]; a.hi1 = action_hi[1];
  a.seed = seed; a.counter = counter; a.counter_dev = counter_dev;
  a.rows = (int)rows; a.n_tiles = (int)((rows + PF_ROWS - 1) / PF_ROWS);
  a.bulk_ok = (((uintptr_t)self_obs | (uintptr_t)objects | (uintptr_t)mask) & 15) == 0;
  int* status = nullptr;
  if (cudaGetSymbolAddress((void**)&status, pf_status_word) != cudaSuccess) status = nullptr;
  a.status = status;
  a.actions = actions;
  a.hidden = hidden;
  a.hidden_img = (unsigned char*)hidden_img;
  const int grid = a.n_tiles < n_sm ? a.n_tiles : n_sm;
  policy_forward_kernel<<<grid, PF_THREADS, PF_SMEM_B, st>>>(a);
  return check_launch(what);
}

extern "C" int asvrl_policy_forward(const float* self_obs, const float* objects, const float* mask,
                                    const float* w_self, const float* b_self, const float* w_obj, const float* b_obj,
                                    const float* w1, const float* b1, const float* w2, const float* b2,
                                    const float* w_out, const float* b_out, float atan_scale, const float* eps_dev,
                                    float eps, const float* action_lo, const float* action_hi, uint64_t seed,
                                    uint64_t counter, const uint64_t* counter_dev, long rows, int repack,
                                    void* workspace, float* actions, void* stream) {
  if (!w_out || !b_out || !action_lo || !action_hi || !actions)
    return fail(ASVRL_E_BADARG, "asvrl_policy_forward: null pointer");
  return pf_run("asvrl_policy_forward", self_obs, objects, mask, w_self, b_self, w_obj, b_obj, w1, b1, w2, b2, w_out, b_out,
                atan_scale, eps_dev, eps, action_lo, action_hi, seed, counter, counter_dev, rows, repack, workspace, actions,
                nullptr, nullptr, stream);
}

// The trunk alone: tower -> 256 -> 128 -> 128 (ReLU) of the same launch, the second hidden layer written out.  Every
// policy of the family (DQN:50-77, SAC's actor, Rainbow's two noisy streams) shares this front end and differs in the
// layers after it.
extern "C" int asvrl_policy_hidden(const float* self_obs, const float* objects, const float* mask, const float* w_self,
                                   const float* b_self, const float* w_obj, const float* b_obj, const float* w1,
                                   const float* b1, const float* w2, const float* b2, long rows, int repack,
                                   void* workspace, float* hidden, void* hidden_img, void* stream) {
  if ((!hidden && !hidden_img) || ((uintptr_t)hidden & 15) || ((uintptr_t)hidden_img & 127))
    return fail(ASVRL_E_BADARG, "asvrl_policy_hidden: hidden (16-byte aligned) and / or hidden_img (128-byte aligned)");
  static const float zero2[2] = {0.0f, 0.0f};
  // the head's operands are read into shared memory by every CTA: point them at the (valid) second-layer bias
  return pf_run("asvrl_policy_hidden", self_obs, objects, mask, w_self, b_self, w_obj, b_obj, w1, b1, w2, b2, nullptr, b2,
                1.0f, nullptr, 0.0f, zero2, zero2, 0, 0, nullptr, rows, repack, workspace, nullptr, hidden, hidden_img, stream);
}

// synchronous read of K9's device status word (0 = every bounded wait completed)
extern "C" int asvrl_policy_forward_status(int* status_out) {
  if (!status_out) return fail(ASVRL_E_BADARG, "asvrl_policy_forward_status: null pointer");
  cudaError_t e = cudaMemcpyFromSymbol(status_out, pf_status_word, sizeof(int));
  if (e != cudaSuccess) return fail((int)e, "asvrl_policy_forward_status: %s", cudaGetErrorString(e));
  return 0;
}
