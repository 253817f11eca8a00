// c51_head_umma.cu — K13: Rainbow's dueling C51 head for a whole env batch in ONE persistent tcgen05 launch (sm_100a).
//
// Replaces (reference paths under rfarl/rfarl/):
//   policy/Rainbow_model.py:127-135   a = output_layer_a(h_a).view(-1, A, atoms); q = v + a - a.mean(1); softmax(q, 2)
//   agent.py:235-246                  act_rainbow: (dist * support).sum(2).argmax(1), per ASV
// and, in this repository, five K8 launches for the 128 -> A * atoms advantage layer (0.27 ms at 65 536 rows, 347 MB
// written and read back) plus asvrl_c51_greedy (0.17 ms).  The (rows, A, atoms) product never exists in HBM: it is
// accumulated in TMEM, 128 columns (two actions) at a time, and reduced to one expected value per action by the thread
// that owns the row.
//
// Algebra.  mean_a(a_a) = h_a . mean_a(W_a) + mean_a(b_a) is itself a linear layer, so with
//   T[j]   = (z_v[j] + b_v[j]) - (h_a . Wbar[j] + bbar[j])                       (per row, per atom)
// every action is independent: logits_a[j] = h_a . W_a[j] + b_a[j] + T[j]; ev_a = sum_j softmax(logits_a)_j support_j.
// The weight image holds `units` of 128 output columns: unit 0 = [Wbar | W_0], unit u = [W_(2u-1) | W_(2u)]; atoms are
// padded to 64 columns per action with zero rows.  z_v (the value stream's product, 51 of 64 columns) comes from K8.
//
// Arithmetic: the fp16 two-piece split of K9 - K12 (x = hi + lo, three products, fp32 accumulation), weights pre-scaled
// by 2^6.  The A operand (h_a) arrives as the operand image asvrl_policy_hidden writes: two 32 KB chunks per 128-row
// tile, moved by bulk copy.  Structure: tiles are taken two at a time (slots) so that one pass over the weight image
// (2 x 32 KB ring; the stream out of L2 is what bounds the kernel) serves both; warp 0 = producer, warps 1 and 10 = the
// slots' MMA issuers, warps 2 - 5 / 6 - 9 = the slots' epilogue warps (one per TMEM lane quadrant, lane = row).
#include "umma.cuh"

namespace asvrl {

constexpr int CH_ROWS = 128, CH_K = 128, CH_UNIT_N = 128, CH_COLS = 64;   // rows per tile, hidden width, columns per unit / action
constexpr int CH_TILE_B = CH_ROWS * 128;          // 16 KB: 128 rows x 64 fp16
constexpr int CH_CHUNK_B = 2 * CH_TILE_B;         // hi | lo
constexpr int CH_MAX_A = 32, CH_MAX_UNITS = (CH_MAX_A + 2) / 2;
constexpr int CH_SLOTS = 2;                       // tiles that share one pass over the weight image
constexpr int CH_WSTAGES = 2, CH_DBUF = 2;        // weight ring stages; TMEM buffers of CH_SLOTS x 128 columns
constexpr int CH_THREADS = 352;   // producer, MMA issuer of slot 0, 8 epilogue warps, MMA issuer of slot 1
constexpr float CH_WSCALE = 64.0f, CH_WUNSCALE = 1.0f / 64.0f;
// workspace: weight image [units][2 k-chunks][32 KB] | tb [64] (b_v - bbar) | ba [A][64] | support [64]
constexpr int CH_TAB_FLOATS = 64 + CH_MAX_A * 64 + 64;
constexpr long CH_WS_B = (long)CH_MAX_UNITS * 2 * CH_CHUNK_B + CH_TAB_FLOATS * 4;
// shared memory
constexpr int CH_OFF_A = 0;                                   // per slot: 2 chunks
constexpr int CH_OFF_WR = CH_SLOTS * 2 * CH_CHUNK_B;          // weight ring
constexpr int CH_OFF_BAR = CH_OFF_WR + CH_WSTAGES * CH_CHUNK_B;
constexpr int CH_NBAR = 2 * CH_SLOTS + 2 * CH_WSTAGES + 2 * CH_DBUF * CH_SLOTS;
constexpr int CH_OFF_MISC = CH_OFF_BAR + CH_NBAR * 8;         // tmem slot, abort flag
constexpr int CH_OFF_TAB = CH_OFF_MISC + 16;                  // tb | ba | support (CH_TAB_FLOATS floats, 16-byte aligned)
constexpr int CH_SMEM_B = CH_OFF_TAB + CH_TAB_FLOATS * 4 + 1024;
static_assert(CH_OFF_TAB % 16 == 0, "table alignment");
static_assert(CH_SMEM_B <= 232448, "K13 shared memory");

struct C51HeadArgs {
  const unsigned char* h_img;      // [tiles][2][32 KB]
  const unsigned char* image;      // packed weights + tables (c51_head_pack_kernel)
  const float* zv;                 // [rows][zv_ld]
  int zv_ld, rows, n_tiles, A, atoms, units;
  int* action;                     // [rows]
  float* q_out;                    // [rows][A] nullable
  int* status;
};

// weight image: unit u, k-chunk c: 128 output rows n (n < 64: first action of the unit, n >= 64: second) x 64 k.
// action index of (u, half): u == 0 ? (half ? 0 : MEAN) : 2 u - 1 + half; rows beyond `atoms` and actions beyond A are 0.
__global__ void __launch_bounds__(256)
c51_head_pack_kernel(const float* __restrict__ wa, const float* __restrict__ ba, const float* __restrict__ bv,
                     const float* __restrict__ support, int A, int atoms, int units, unsigned char* __restrict__ image) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;   // one 16-byte octet (8 k) of one row of one chunk
  const int n_oct = units * 2 * CH_UNIT_N * 8;
  float* tab = reinterpret_cast<float*>(image + (size_t)units * 2 * CH_CHUNK_B);
  if (t < CH_TAB_FLOATS) {
    float v = 0.0f;
    if (t < 64) {                                   // tb[j] = b_v[j] - mean_a b_a[j]
      if (t < atoms) {
        float s = 0.0f;
        for (int a = 0; a < A; ++a) s += ba[a * atoms + t];
        v = bv[t] - s / (float)A;
      }
    } else if (t < 64 + CH_MAX_A * 64) {            // ba padded to 64 columns per action
      const int a = (t - 64) >> 6, j = (t - 64) & 63;
      if (a < A && j < atoms) v = ba[a * atoms + j];
      else if (j >= atoms) v = -INFINITY;   // padded atoms: logit -inf, weight exp(-inf) = 0 without a predicate
    } else {
      const int j = t - 64 - CH_MAX_A * 64;
      if (j < atoms) v = support[j];
    }
    tab[t] = v;
  }
  if (t >= n_oct) return;
  const int oc = t & 7, n = (t >> 3) & (CH_UNIT_N - 1), c = (t >> 10) & 1, u = t >> 11;
  const int half = n >> 6, j = n & 63;
  const int act = u == 0 ? (half ? 0 : -1) : 2 * u - 1 + half;   // -1: the mean over the actions
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = 64 * c + 8 * oc + e;
    float x = 0.0f;
    if (j < atoms) {
      if (act >= 0) {
        if (act < A) x = wa[((size_t)act * atoms + j) * CH_K + k];
      } else {
        float s = 0.0f;
        for (int a = 0; a < A; ++a) s += wa[((size_t)a * atoms + j) * CH_K + k];
        x = s / (float)A;
      }
    }
    v[e] = x * CH_WSCALE;
  }
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) um_split2(v[2 * i], v[2 * i + 1], h[i], l[i]);
  unsigned char* tile = image + (size_t)(u * 2 + c) * CH_CHUNK_B + um_sw128_off(n, oc);
  *reinterpret_cast<uint4*>(tile) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(tile + CH_TILE_B) = make_uint4(l[0], l[1], l[2], l[3]);
}

static __device__ int ch_status_word = 0;

// -DCH_TRACE (debug builds, profiles/probes/k13_trace.py): CTA 0's timeline, [writer][event] = (clock << 8) | tag
#ifdef CH_TRACE
__device__ unsigned long long ch_trace_buf[4][512];
__device__ int ch_trace_n[4];
__device__ unsigned long long ch_cta_ns[256][2];   // globaltimer at the start / end of every CTA
#define CH_TR(writer, tag) do { if (blockIdx.x == 0) { const int i_ = ch_trace_n[writer]; if (i_ < 512) { \
    ch_trace_buf[writer][i_] = ((unsigned long long)clock64() << 8) | (unsigned)(tag); ch_trace_n[writer] = i_ + 1; } } } while (0)
#else
#define CH_TR(writer, tag) do { } while (0)
#endif

// step -> unit of the image: 0 first, then 1 .. units - 1 rotated by `rot`
__device__ __forceinline__ int ch_unit(int step, int units, int rot) {
  if (step == 0) return 0;
  int u = step + rot;
  if (u >= units) u -= units - 1;
  return u;
}

// One half (32 atoms) of an action for the thread's row: logits x[j] = acc[j] / 64 + ba[j] + T[j] (padded atoms carry
// ba = -inf), their maximum, and sum exp(x - max), sum exp(x - max) support[j] with four independent accumulators.
// The bias and support reads (shared memory, broadcast) are issued before the TMEM load is waited for.
__device__ __forceinline__ void ch_half(uint32_t taddr, const float* T32, uint32_t ba_s, uint32_t sup_s, float& mx,
                                        float& den, float& num) {
  uint32_t v[32];
  um_tmem_ld32(taddr, v);
  uint4 b[8];
#pragma unroll
  for (int g = 0; g < 8; ++g) b[g] = um_lds128(ba_s + 16u * g);
  um_tmem_ld_wait();
  float x[32];
#pragma unroll
  for (int g = 0; g < 8; ++g) {
    const uint32_t bw[4] = {b[g].x, b[g].y, b[g].z, b[g].w};
#pragma unroll
    for (int i = 0; i < 4; ++i) x[4 * g + i] = fmaf(__uint_as_float(v[4 * g + i]), CH_WUNSCALE, __uint_as_float(bw[i])) + T32[4 * g + i];
  }
  float m4[4] = {x[0], x[1], x[2], x[3]};
#pragma unroll
  for (int j = 4; j < 32; ++j) m4[j & 3] = fmaxf(m4[j & 3], x[j]);
  mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
  float d4[4] = {0.0f, 0.0f, 0.0f, 0.0f}, n4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
  const float sh = -mx * 1.4426950408889634f;
#pragma unroll
  for (int g = 0; g < 8; ++g) {
    const uint4 sv = um_lds128(sup_s + 16u * g);
    const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float e;   // exp(x - mx) <= 1: the raw MUFU form (exp2f adds five instructions of range handling per element)
      asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fmaf(x[4 * g + i], 1.4426950408889634f, sh)));
      d4[i] += e;
      n4[i] = fmaf(e, __uint_as_float(sw[i]), n4[i]);
    }
  }
  den = (d4[0] + d4[1]) + (d4[2] + d4[3]);
  num = (n4[0] + n4[1]) + (n4[2] + n4[3]);
}

// expected value of one action: softmax over the atoms, expectation over the support (two halves, merged)
__device__ __forceinline__ float ch_action_value(uint32_t taddr, const float (&T)[64], uint32_t ba_s, uint32_t sup_s, int atoms) {
  float m0, d0, n0;
  ch_half(taddr, T, ba_s, sup_s, m0, d0, n0);
  if (atoms <= 32) return n0 / d0;
  float m1, d1, n1;
  ch_half(taddr + 32u, T + 32, ba_s + 128u, sup_s + 128u, m1, d1, n1);
  const float mx = fmaxf(m0, m1);
  const float f0 = __expf(m0 - mx), f1 = __expf(m1 - mx);
  return fmaf(n0, f0, n1 * f1) / fmaf(d0, f0, d1 * f1);
}

__global__ void __launch_bounds__(CH_THREADS, 1)
c51_head_kernel(const C51HeadArgs a) {
  extern __shared__ __align__(1024) unsigned char ch_raw[];
  unsigned char* base = ch_raw + ((1024u - (um_smem(ch_raw) & 1023u)) & 1023u);
  const uint32_t sb = um_smem(base);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t bar0 = sb + CH_OFF_BAR;
  auto bar_a_full = [&](int sl) { return bar0 + 8u * sl; };
  auto bar_a_free = [&](int sl) { return bar0 + 8u * (CH_SLOTS + sl); };
  auto bar_w_full = [&](int st) { return bar0 + 8u * (2 * CH_SLOTS + st); };
  auto bar_w_empty = [&](int st) { return bar0 + 8u * (2 * CH_SLOTS + CH_WSTAGES + st); };
  auto bar_d_full = [&](int b, int sl) { return bar0 + 8u * (2 * CH_SLOTS + 2 * CH_WSTAGES + CH_SLOTS * b + sl); };
  auto bar_d_empty = [&](int b, int sl) { return bar0 + 8u * (2 * CH_SLOTS + 2 * CH_WSTAGES + CH_SLOTS * (CH_DBUF + b) + sl); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base + CH_OFF_MISC);
  volatile uint32_t* abort_flag = reinterpret_cast<volatile uint32_t*>(base + CH_OFF_MISC + 4);

#ifdef CH_TRACE
  if (tid == 0 && blockIdx.x < 256) ch_cta_ns[blockIdx.x][0] = um_globaltimer();
#endif
  if (tid == 0) {
    for (int sl = 0; sl < CH_SLOTS; ++sl) { um_mbar_init(bar_a_full(sl), 1); um_mbar_init(bar_a_free(sl), 1); }
    for (int st = 0; st < CH_WSTAGES; ++st) { um_mbar_init(bar_w_full(st), 1); um_mbar_init(bar_w_empty(st), CH_SLOTS); }
    for (int b = 0; b < CH_DBUF; ++b)
      for (int sl = 0; sl < CH_SLOTS; ++sl) { um_mbar_init(bar_d_full(b, sl), 1); um_mbar_init(bar_d_empty(b, sl), 4); }
    *abort_flag = 0u;
    um_mbar_init_fence();
  }
  if (warp == 1) um_tmem_alloc<512>(um_smem(tmem_slot));
  {
    const float* tab = reinterpret_cast<const float*>(a.image + (size_t)a.units * 2 * CH_CHUNK_B);
    float* tab_s = reinterpret_cast<float*>(base + CH_OFF_TAB);
    for (int i2 = tid; i2 < CH_TAB_FLOATS; i2 += CH_THREADS) tab_s[i2] = __ldg(tab + i2);
  }
  um_tc_before();
  __syncthreads();
  um_tc_after();
  const uint32_t tmem = *tmem_slot;
  // Every CTA streams the whole weight image per tile.  After unit 0 (the mean, needed first) CTA b walks the other
  // units from a different start, so that at any instant the 148 SMs read different lines of the image instead of
  // all hammering the same L2 slices.
  const int rot = a.units > 1 ? (int)(blockIdx.x % (unsigned)(a.units - 1)) : 0;
  const int n_my = ((int)blockIdx.x < a.n_tiles) ? (a.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto wait = [&](uint32_t bar, uint32_t parity, int code) { (void)um_mbar_wait(bar, parity, abort_flag, a.status, code); };

  // Tiles are taken two at a time ("slots"): every weight stage serves both tiles' MMAs, which halves the weight stream
  // out of L2 — the stream, at ~41 GB/s per SM with all SMs pulling the same image, is what bounds this kernel
  // (profiles/probes/k13_trace.py), not the tensor pipe and not the epilogue.
  const int n_pairs = (n_my + CH_SLOTS - 1) / CH_SLOTS;
  if (warp == 0) {
    // =============== producer: the pair's A images, then every weight chunk through the ring ===============
    if (lane == 0) {
      int pos = 0;
      for (int p = 0; p < n_pairs; ++p) {
        for (int sl = 0; sl < CH_SLOTS; ++sl) {
          const int i = CH_SLOTS * p + sl;
          if (i >= n_my) break;
          const int tile = (int)blockIdx.x + i * (int)gridDim.x;
          if (p > 0) wait(bar_a_free(sl), (uint32_t)(p - 1) & 1u, UM_TIMEOUT_INPUT);
          const uint32_t dst = sb + CH_OFF_A + (uint32_t)sl * 2 * CH_CHUNK_B;
          um_mbar_arrive_expect_tx(bar_a_full(sl), 2 * CH_CHUNK_B);
          um_bulk_g2s(dst, a.h_img + (size_t)tile * 2 * CH_CHUNK_B, CH_CHUNK_B, bar_a_full(sl));
          um_bulk_g2s(dst + CH_CHUNK_B, a.h_img + ((size_t)tile * 2 + 1) * CH_CHUNK_B, CH_CHUNK_B, bar_a_full(sl));
        }
        for (int c = 0; c < 2 * a.units; ++c, ++pos) {
          const int st = pos % CH_WSTAGES;
          if (pos >= CH_WSTAGES) wait(bar_w_empty(st), (uint32_t)((pos / CH_WSTAGES) - 1) & 1u, UM_TIMEOUT_RING);
          um_mbar_arrive_expect_tx(bar_w_full(st), CH_CHUNK_B);
          const int uu = ch_unit(c >> 1, a.units, rot);
          um_bulk_g2s(sb + CH_OFF_WR + st * CH_CHUNK_B, a.image + (size_t)(2 * uu + (c & 1)) * CH_CHUNK_B, CH_CHUNK_B,
                      bar_w_full(st));
          CH_TR(0, c);
        }
      }
    }
  } else if (warp == 1 || warp == 10) {
    // =============== MMA issuers, one per tile slot ===============
    // A tcgen05.mma costs its issuing thread ~100 cycles whatever its shape (68 cycles of tensor work at N = 128), so
    // each slot has its own issuer; the warp stays converged and one elected lane issues (uniform registers).  Both
    // issuers walk every weight stage; the stage is free when both have committed (or, for an idle slot, arrived).
    {
      const int sl = warp == 1 ? 0 : 1;
      constexpr uint32_t idesc = um_idesc_f16(CH_ROWS, CH_UNIT_N);
      int pos = 0, gu = 0;
      for (int p = 0; p < n_pairs; ++p) {
        const bool active = CH_SLOTS * p + sl < n_my;
        if (active) wait(bar_a_full(sl), (uint32_t)p & 1u, UM_TIMEOUT_INPUT);
        for (int u = 0; u < a.units; ++u) {
          const int buf = gu % CH_DBUF;
          if (active && gu >= CH_DBUF) wait(bar_d_empty(buf, sl), (uint32_t)((gu / CH_DBUF) - 1) & 1u, UM_TIMEOUT_ACC);
          if (lane == 0 && sl == 0) CH_TR(1, 2 * u);
          const uint32_t d_tmem = tmem + (uint32_t)(buf * CH_SLOTS * CH_UNIT_N + sl * CH_UNIT_N);
#pragma unroll 1
          for (int c = 0; c < 2; ++c, ++pos) {
            const int st = pos % CH_WSTAGES;
            wait(bar_w_full(st), (uint32_t)(pos / CH_WSTAGES) & 1u, UM_TIMEOUT_RING);
            um_tc_after();
            const uint32_t wh = sb + CH_OFF_WR + (uint32_t)st * CH_CHUNK_B, wl = wh + CH_TILE_B;
            const uint32_t ah = sb + CH_OFF_A + (uint32_t)(sl * 2 + c) * CH_CHUNK_B, al = ah + CH_TILE_B;
            if (um_elect()) {
              if (active) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                  const uint32_t o = ks * 32;
                  um_mma_f16(d_tmem, um_desc_sw128(al + o), um_desc_sw128(wh + o), idesc, (c == 0 && ks == 0) ? 0u : 1u);
                  um_mma_f16(d_tmem, um_desc_sw128(ah + o), um_desc_sw128(wl + o), idesc, 1u);
                  um_mma_f16(d_tmem, um_desc_sw128(ah + o), um_desc_sw128(wh + o), idesc, 1u);
                }
                um_commit(bar_w_empty(st));
                if (c == 1) um_commit(bar_d_full(buf, sl));
              } else {
                um_mbar_arrive(bar_w_empty(st));
              }
            }
            __syncwarp();
          }
          if (lane == 0 && sl == 0) CH_TR(1, 2 * u + 1);
          if (active) ++gu;
        }
        if (active && um_elect()) um_commit(bar_a_free(sl));
        __syncwarp();
      }
    }
  } else {
    // =============== epilogue: lane = row of TMEM quadrant q of tile slot sl; every unit of its tile ===============
    const int q = warp & 3, sl = (warp - 2) >> 2, r = 32 * q + lane;
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(sl * CH_UNIT_N);
    const float4* tb4 = reinterpret_cast<const float4*>(base + CH_OFF_TAB);
    const uint32_t ba_s = sb + CH_OFF_TAB + 64u * 4u, sup_s = ba_s + (uint32_t)(CH_MAX_A * 64 * 4);
    int gu = 0;
    for (int p = 0; p < n_pairs; ++p) {
      const int i = CH_SLOTS * p + sl;
      const bool have = i < n_my;   // the last pair may hold one tile only
      if (!have) break;
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const long m = (long)tile * CH_ROWS + r;
      const bool live = have && m < (long)a.rows;
      const float4* z4 = reinterpret_cast<const float4*>(a.zv + (size_t)(live ? m : 0) * a.zv_ld);
      float T[64];
      float best = -INFINITY;
      int best_a = 0x7fffffff;
      for (int u = 0; u < a.units; ++u, ++gu) {
        const int buf = gu % CH_DBUF;
        wait(bar_d_full(buf, sl), (uint32_t)(gu / CH_DBUF) & 1u, UM_TIMEOUT_ACC);
        um_tc_after();
        if (lane == 0 && q == 2) CH_TR(2 + sl, 2 * u);
        const uint32_t t0 = lane_addr + (uint32_t)(buf * CH_SLOTS * CH_UNIT_N);
        const int uu = ch_unit(u, a.units, rot);
        if (have) {
          if (u == 0) {
            // T[j] = z_v[j] + (b_v[j] - bbar[j]) - h_a . Wbar[j]
            uint32_t v0[32], v1[32];
            um_tmem_ld32(t0, v0);
            um_tmem_ld32(t0 + 32u, v1);
            um_tmem_ld_wait();
#pragma unroll
            for (int g = 0; g < 16; ++g) {
              const float4 z = 4 * g < a.atoms ? __ldg(z4 + g) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
              const float4 b = tb4[g];
              const float zz[4] = {z.x, z.y, z.z, z.w}, bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const int j = 4 * g + k;
                const float acc = __uint_as_float(j < 32 ? v0[j & 31] : v1[j & 31]);
                T[j] = (zz[k] + bb[k]) - acc * CH_WUNSCALE;
              }
            }
          }
#pragma unroll 1
          for (int half = 0; half < 2; ++half) {
            const int act = 2 * uu - 1 + half;   // unit 0: [mean | action 0]
            if (act >= 0 && act < a.A) {
              const float ev = ch_action_value(t0 + (uint32_t)(64 * half), T, ba_s + (uint32_t)(256 * act), sup_s, a.atoms);
              if (a.q_out && live) a.q_out[(size_t)m * a.A + act] = ev;
              if (ev > best || (ev == best && act < best_a)) { best = ev; best_a = act; }   // the first maximum, as torch.argmax
            }
          }
        }
        if (lane == 0 && q == 2) CH_TR(2 + sl, 2 * u + 1);
        um_tc_before();
        __syncwarp();
        if (lane == 0) um_mbar_arrive(bar_d_empty(buf, sl));
      }
      if (live) a.action[m] = best_a == 0x7fffffff ? 0 : best_a;
    }
  }
  um_tc_before();
  __syncthreads();
#ifdef CH_TRACE
  if (tid == 0 && blockIdx.x < 256) ch_cta_ns[blockIdx.x][1] = um_globaltimer();
#endif
  if (warp == 1) {
    um_tc_after();
    um_tmem_free<512>(tmem);
  }
}

}  // namespace asvrl

using namespace asvrl;

extern "C" long asvrl_c51_head_workspace(void) { return CH_WS_B; }

extern "C" int asvrl_c51_head(const void* h_img, const float* zv, int zv_ld, const float* wa, const float* ba,
                              const float* bv, const float* support, long rows, int n_actions, int atoms, void* workspace,
                              int* action, float* q_out, void* stream) {
  if (!h_img || !zv || !wa || !ba || !bv || !support || !workspace || !action)
    return fail(ASVRL_E_BADARG, "asvrl_c51_head: null pointer");
  if (rows < 1 || rows > 0x7FFFFFFFL - CH_ROWS || n_actions < 1 || n_actions > CH_MAX_A || atoms < 1 || atoms > CH_COLS)
    return fail(ASVRL_E_BADARG, "asvrl_c51_head: rows >= 1, 1 <= actions <= %d, 1 <= atoms <= %d (rows=%ld A=%d atoms=%d)",
                CH_MAX_A, CH_COLS, rows, n_actions, atoms);
  if (zv_ld < 64 || (zv_ld & 3) || ((uintptr_t)zv & 15))
    return fail(ASVRL_E_BADARG, "asvrl_c51_head: zv rows must be 16-byte aligned and >= 64 floats apart (zv_ld %d)", zv_ld);
  if (((uintptr_t)h_img | (uintptr_t)workspace) & 127)
    return fail(ASVRL_E_BADARG, "asvrl_c51_head: h_img and workspace must be 128-byte aligned");
  static int n_sm = 0;
  cudaError_t e;
  if (n_sm == 0) {
    int dev = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess || (e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess)
      return fail((int)e, "asvrl_c51_head: %s", cudaGetErrorString(e));
    e = cudaFuncSetAttribute(c51_head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CH_SMEM_B);
    if (e != cudaSuccess) { n_sm = 0; return fail((int)e, "asvrl_c51_head: %d B shared memory: %s", CH_SMEM_B, cudaGetErrorString(e)); }
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int units = (n_actions + 2) / 2;   // [mean | a0], [a1 | a2], ...
  {
    int n = units * 2 * CH_UNIT_N * 8;
    if (n < CH_TAB_FLOATS) n = CH_TAB_FLOATS;
    c51_head_pack_kernel<<<(n + 255) / 256, 256, 0, st>>>(wa, ba, bv, support, n_actions, atoms, units, (unsigned char*)workspace);
    if (int rc = check_launch("asvrl_c51_head (pack)")) return rc;
  }
  C51HeadArgs a;
  a.h_img = (const unsigned char*)h_img; a.image = (const unsigned char*)workspace;
  a.zv = zv; a.zv_ld = zv_ld; a.rows = (int)rows; a.n_tiles = (int)((rows + CH_ROWS - 1) / CH_ROWS);
  a.A = n_actions; a.atoms = atoms; a.units = units;
  a.action = action; a.q_out = q_out;
  int* status = nullptr;
  if (cudaGetSymbolAddress((void**)&status, ch_status_word) != cudaSuccess) status = nullptr;
  a.status = status;
  const int grid = a.n_tiles < n_sm ? a.n_tiles : n_sm;
  c51_head_kernel<<<grid, CH_THREADS, CH_SMEM_B, st>>>(a);
  return check_launch("asvrl_c51_head");
}

extern "C" int asvrl_c51_head_status(int* status_out) {
  if (!status_out) return fail(ASVRL_E_BADARG, "asvrl_c51_head_status: null pointer");
  cudaError_t e = cudaMemcpyFromSymbol(status_out, ch_status_word, sizeof(int));
  if (e != cudaSuccess) return fail((int)e, "asvrl_c51_head_status: %s", cudaGetErrorString(e));
  return 0;
}

#ifdef CH_TRACE
extern "C" __attribute__((visibility("default"))) int asvrl_c51_head_cta_times(unsigned long long* ns) {
  cudaDeviceSynchronize();
  return (int)cudaMemcpyFromSymbol(ns, ch_cta_ns, sizeof(unsigned long long) * 256 * 2);
}
extern "C" __attribute__((visibility("default"))) int asvrl_c51_head_trace(unsigned long long* events, int* counts) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(events, ch_trace_buf, sizeof(unsigned long long) * 4 * 512);
  cudaMemcpyFromSymbol(counts, ch_trace_n, sizeof(int) * 4);
  int zero[4] = {0, 0, 0, 0};
  cudaMemcpyToSymbol(ch_trace_n, zero, sizeof zero);
  return 0;
}
#endif
