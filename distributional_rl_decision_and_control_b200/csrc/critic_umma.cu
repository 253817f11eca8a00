// critic_umma.cu — K10: the IQN critic's quantile trunk, forward, for all B*N quantile rows in ONE persistent,
// warp-specialised launch (sm_100a: tcgen05.mma kind::f16, accumulators in TMEM, weights by TMA bulk copy,
// mbarrier pipelines).  The cosine embedding input is generated in the operand loader: the (B, N, 64) table
// of K3 never exists in HBM, and no activation leaves the SM.
//
// Replaces, per critic evaluation (reference: rfarl/rfarl/policy/AC_IQN_model.py),
//   :410-426  calc_cos: tau ~ U(0,1) * cvar, cos(tau * pi * i), i < 64                      (K3)
//   :453-460  x = features.unsqueeze(1) * relu(cos_embedding(cos))                           (GEMM + K5)
//   :470-474  x = relu(hidden_layer(x)); x = action_encoder(a).unsqueeze(1) * x              (GEMM + K5)
//   :476-480  x = relu(hidden_layer_2(x)); quantiles = output_layer(x)                       (GEMM + K5 head)
// and, with aenc == nullptr, the same trunk of rfarl/rfarl/policy/IQN_model.py:74-86 up to the last hidden
// layer.  7 launches and ~150 us at B*N = 32 768 become one.
//
// Arithmetic: as K9 — every operand is split in two fp16 pieces (22 mantissa bits), three products per K step
// accumulate in fp32 in TMEM (x_lo W_hi + x_hi W_lo + x_hi W_hi), W pre-scaled by 2^6.  No clamping: an
// activation beyond fp16's range gives a NaN quantile, which the caller's finiteness check reports.
//
// Mapping.  One CTA per SM walks 128-row tiles (tile = blockIdx.x + i * gridDim.x); even and odd tiles of a CTA
// run in two SLOTS, each with its own 8 epilogue warps, two 32 KB operand buffers and 256 TMEM columns, so one
// slot's CUDA-core work overlaps the other's tensor work.  20 warps:
//   warps 0, 3  W producers, one per slot: one thread streams the packed weights as half chunks (16 KB = hi | lo tile
//               of 64 out x 64 k) into the slot's own 3-stage ring with cp.async.bulk — a slot consumes its weights in
//               a fixed order, whatever the other slot does
//   warps 1, 2  MMA issuers, one per slot (a tcgen05.mma costs its issuing thread ~90 cycles whatever its shape: two
//               issuers, and N = 64 halves of a chunk as two independent accumulation chains).  The warp stays converged
//               and elects a lane around the tcgen05 instructions, so descriptors live in uniform registers.  Each slot
//               runs at its own pace: one slot's tensor work lies under the other's CUDA-core work.
//               Warp 2 also allocates / frees TMEM
//   warps 4-11  slot 0, warps 12-19 slot 1: warp e of a slot owns TMEM lanes 32 (e % 4) .. and the column half
//               e / 4.  Per tile: E0 tau -> cos -> A chunk; E1 D1 -> feats * relu(. + bc) -> 4 A chunks of layer 2;
//               E2 D2 -> aenc * relu(. + bh) -> 2 A chunks of layer 3; E3 D3 -> relu(. + bh2) . wo + bo -> z.
// A slot's operand buffers alternate with a per-slot use counter u (buffer u & 1): the epilogue waits for the
// buffer's previous reader (x_free, tcgen05.commit), writes, arrives on a_full; the MMA thread waits a_full,
// issues, commits x_free.  TMEM per slot: D1 = columns [0, 256) (two N = 128 halves), D2 reuses [0, 128) once
// the first two layer-2 chunks are built, D3 reuses [128, 256).
//
// SAVE mode (the learner's differentiable pass): every operand image the backward kernels need — cos, H1, H2
// and R3 = relu(P3 + bh2) as swizzled fp16 hi | lo chunks, exactly the bytes the MMAs read — goes to HBM
// by cp.async.bulk from the operand buffer itself (one elected thread per slot, bulk async-groups).
#include "umma.cuh"

namespace asvrl {

constexpr int CF_THREADS = 640;
constexpr int CF_ROWS = 128;                       // rows per tile = UMMA M
constexpr int CF_NCOS = 64, CF_W = 256, CF_H = 128;
constexpr int CF_TILE_B = CF_ROWS * 128;           // one fp16 piece of a [128 x 64] chunk: 16 KB
constexpr int CF_CHUNK_B = 2 * CF_TILE_B;          // hi | lo
constexpr int CF_WSTAGES = 3;                      // per slot, of half chunks
constexpr int CF_HALF_B = CF_CHUNK_B / 2;          // hi | lo of 64 out x 64 k
constexpr int CF_PIECE_B = CF_TILE_B / 2;          // 64 rows of one fp16 piece
constexpr int CF_W_CHUNKS = 8;                     // Wc: 2 (n halves), Wh: 4 (k chunks), Wh2: 2
constexpr int CF_IMAGE_B = CF_W_CHUNKS * CF_CHUNK_B;   // 256 KB packed weights
constexpr float CF_WSCALE = 64.0f, CF_WUNSCALE = 1.0f / 64.0f;

constexpr int CF_OFF_X = 0;                                        // [slot][buffer] operand chunks
constexpr int CF_OFF_WR = 4 * CF_CHUNK_B;                          // weight ring
constexpr int CF_OFF_BAR = CF_OFF_WR + 2 * CF_WSTAGES * CF_HALF_B; // 229376
constexpr int CF_NBAR = 2 * (8 + 2 * CF_WSTAGES);                  // per slot a_full[2] x_free[2] d_full[4] w_full[3] w_empty[3]
constexpr int CF_OFF_MISC = CF_OFF_BAR + CF_NBAR * 8;              // tmem slot, abort flag
constexpr int CF_OFF_ZPART = CF_OFF_MISC + 16;                     // [slot][128] floats
constexpr int CF_SMEM_B = CF_OFF_ZPART + 2 * CF_ROWS * 4 + 1024;   // + alignment slack
static_assert(CF_SMEM_B <= 232448, "K10 shared memory");

#ifdef CF_TRACE
// timeline of CTA 0 for profiles/probes/k10_trace.py: [writer][event] = (globaltimer << 8) | tag
__device__ unsigned long long cf_trace_buf[8][2048];
__device__ int cf_trace_n[8];
#define CF_TR(writer, tag)                                                                              \
  do {                                                                                                   \
    if (blockIdx.x == 0) {                                                                               \
      const int n_ = cf_trace_n[writer];                                                                 \
      if (n_ < 2048) { cf_trace_buf[writer][n_] = (um_globaltimer() << 8) | (unsigned)(tag); cf_trace_n[writer] = n_ + 1; } \
    }                                                                                                    \
  } while (0)
#else
#define CF_TR(writer, tag) do {} while (0)
#endif

struct CriticArgs {
  const float* tau_in;                        // [rows] or null (Philox stream 3)
  uint64_t seed, counter;
  const uint64_t* counter_dev;
  float cvar;
  const float *feats, *aenc;                  // [B][256], [B][128] (null: unit scale)
  const float *bc, *bh, *bh2, *wo, *bo;       // biases, head row [128], head bias [1]
  const unsigned char* image;                 // packed Wc | Wh | Wh2 chunks (critic_pack_kernel)
  int rows, n_tau, batch, n_tiles;
  int* status;
  float *tau_out, *z;                         // [rows] (tau_out nullable)
  unsigned char *cos_img, *h1_img, *h2_img, *r3_img;   // SAVE mode: [tile][1 | 4 | 2 | 2][32 KB]
};

// ---- weight image: chunk c = hi tile (128 rows of 128 B, swizzled) then lo tile ----
// Chunks 0 .. 7 are K10's (Wc as two N halves, Wh as four K chunks, Wh2 as two).  With `bwd_image` the same launch
// also writes K11's six chunks (the transposed operands of the backward chain, critic_bwd_umma.cu): chunk 0, 1:
// B[n = i][k = j - 64 c] = Wh2[j][i]; chunk 2 + 2 nh + c: B[n][k] = Wh[64 c + k][128 nh + n].
__global__ void __launch_bounds__(256)
critic_pack_kernel(const float* __restrict__ wc, const float* __restrict__ wh, const float* __restrict__ wh2,
                   unsigned char* __restrict__ image, unsigned char* __restrict__ bwd_image,
                   const float4* __restrict__ feats4, int n_feats4, float* __restrict__ fpart) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;   // one 16-byte chunk (8 k) of one row of one chunk
  const int n_fwd = CF_W_CHUNKS * CF_H * 8;
  if (fpart) {
    // K11 scales its dP1 image by a power of two chosen from max feats (critic_bwd_umma.cu): every CTA of this launch
    // leaves the maximum of its grid-stride share in fpart[blockIdx.x] (64 slots; the unused ones are zeroed), K11's
    // prologue takes the maximum of the 64.  No extra launch, no atomics, order independent.
    __shared__ float wmax[8];
    float m = 0.0f;
#pragma unroll 4
    for (int i = t; i < n_feats4; i += (int)(gridDim.x * blockDim.x)) {
      const float4 f = __ldg(feats4 + i);
      m = fmaxf(fmaxf(m, fmaxf(f.x, f.y)), fmaxf(f.z, f.w));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < 8; ++w) m = fmaxf(m, wmax[w]);
      if (blockIdx.x < 64) fpart[blockIdx.x] = m;
    }
    if (blockIdx.x == 0 && threadIdx.x >= gridDim.x && threadIdx.x < 64) fpart[threadIdx.x] = 0.0f;
  }
  if (t >= n_fwd + (bwd_image ? 6 * CF_H * 8 : 0)) return;
  const bool bwd = t >= n_fwd;
  const int u = bwd ? t - n_fwd : t;
  const int c = u / (CF_H * 8), n = (u / 8) % CF_H, oc = u & 7;
  float v[8];
  if (!bwd) {
    const float* src;
    if (c < 2) src = wc + (size_t)(128 * c + n) * CF_NCOS + oc * 8;
    else if (c < 6) src = wh + (size_t)n * CF_W + (c - 2) * 64 + oc * 8;
    else src = wh2 + (size_t)n * CF_H + (c - 6) * 64 + oc * 8;
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = src[e] * CF_WSCALE;
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = oc * 8 + e;
      v[e] = (c < 2 ? wh2[(size_t)(64 * c + k) * CF_H + n]
                    : wh[(size_t)(64 * ((c - 2) & 1) + k) * CF_W + 128 * ((c - 2) >> 1) + n]) * CF_WSCALE;
    }
  }
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) um_split2(v[2 * i], v[2 * i + 1], h[i], l[i]);
  unsigned char* tile = (bwd ? bwd_image : image) + (size_t)c * CF_CHUNK_B + um_sw128_off(n, oc);
  *reinterpret_cast<uint4*>(tile) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(tile + CF_TILE_B) = make_uint4(l[0], l[1], l[2], l[3]);
}

// cos(tau * fl32(pi i)) for i = I0 .. I0 + 31 of this thread's row -> octets OC0 .. OC0 + 3 of the A chunk.
// x = fl32(tau * fl32(pi i)) is the reference's rounded argument (ACIQN:397, 424; up to 64 pi) and the cosine is cosf,
// as torch's.  A faster form was measured and REJECTED (-DCF_FAST_COS keeps it for the record): MUFU.COS of the argument
// reduced to [-pi, pi] by a two-constant Cody-Waite step is a third of cosf's 21 instructions (K10 29.5 -> 24.8 us at
// C3, 1.27 -> 1.06 ms at 2.1 M rows) and keeps the quantiles within 1e-5, but its ~6e-7 absolute error per cosine
// (cosf: 6e-8) flips ten times more layer-1 ReLU gates: on the reference's observation ranges the parameter gradients
// then differ from float64 autograd by 5e-4 .. 7e-3 of their scale instead of 1e-6 .. 4e-6 (profiles/probes/critic_grad_err.py).
__device__ __forceinline__ float cf_cos(float x) {
#ifdef CF_FAST_COS
  const float k = rintf(x * 0.15915494309189535f);
  float r = fmaf(k, -6.28125f, x);           // k <= 33: k * 6.28125 is exact
  r = fmaf(k, -1.9353071795864769e-3f, r);   // 2 pi - 6.28125
  return __cosf(r);
#else
  return cosf(x);
#endif
}
template <int I0>
__device__ __forceinline__ void cf_cos_octets(float tau, uint32_t a_row, int r7) {
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float pi_i = (float)(3.14159265358979323846 * (double)(I0 + 8 * t + e));   // ACIQN:397
      v[e] = cf_cos(__fmul_rn(tau, pi_i));                                              // ACIQN:424
    }
    um_put8(a_row + (uint32_t)(((I0 / 8 + t) ^ r7) << 4), CF_TILE_B, v);
  }
}

template <bool SAVE, bool AENC, bool MEAN>
__global__ void __launch_bounds__(CF_THREADS, 1)
critic_forward_kernel(const CriticArgs a) {
  extern __shared__ __align__(1024) unsigned char cf_raw[];
  unsigned char* base = cf_raw + ((1024u - (um_smem(cf_raw) & 1023u)) & 1023u);
  const uint32_t sb = um_smem(base);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform: role code stays in the uniform datapath

  const uint32_t bar0 = sb + CF_OFF_BAR;
  auto bar_a_full = [&](int s, int x) { return bar0 + 8u * (14 * s + x); };
  auto bar_x_free = [&](int s, int x) { return bar0 + 8u * (14 * s + 2 + x); };
  auto bar_d_full = [&](int s, int d) { return bar0 + 8u * (14 * s + 4 + d); };
  auto bar_w_full = [&](int s, int st) { return bar0 + 8u * (14 * s + 8 + st); };
  auto bar_w_empty = [&](int s, int st) { return bar0 + 8u * (14 * s + 8 + CF_WSTAGES + st); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base + CF_OFF_MISC);
  volatile uint32_t* abort_flag = reinterpret_cast<volatile uint32_t*>(base + CF_OFF_MISC + 4);
  float* zpart = reinterpret_cast<float*>(base + CF_OFF_ZPART);

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      for (int x = 0; x < 2; ++x) {
        um_mbar_init(bar_a_full(s, x), SAVE ? 1 : 8);   // SAVE: the slot's elected thread, after the slot barrier
        um_mbar_init(bar_x_free(s, x), 1);
      }
      for (int d = 0; d < 4; ++d) um_mbar_init(bar_d_full(s, d), 1);
      for (int st = 0; st < CF_WSTAGES; ++st) {
        um_mbar_init(bar_w_full(s, st), 1);
        um_mbar_init(bar_w_empty(s, st), 1);
      }
    }
    *abort_flag = 0u;
    um_mbar_init_fence();
  }
  if (warp == 2) um_tmem_alloc<512>(um_smem(tmem_slot));
  um_tc_before();
  __syncthreads();
  um_tc_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  const int n_my = ((int)blockIdx.x < a.n_tiles) ? (a.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  // A timed-out wait sets the CTA's abort flag and the status word; every later wait then returns at once and the
  // roles run their loops to the end on garbage, so no named barrier is left short of a participant.
  auto wait = [&](uint32_t bar, uint32_t parity, int code) { (void)um_mbar_wait(bar, parity, abort_flag, a.status, code); };

  if (warp == 0 || warp == 3) {
    // =============== W producers: slot 0 (warp 0), slot 1 (warp 3) ===============
    if (lane == 0) {
      const int s = warp == 0 ? 0 : 1;
      int wpos = 0;
      for (int i = s; i < n_my; i += 2)
        for (int h = 0; h < 2 * CF_W_CHUNKS; ++h, ++wpos) {   // half chunk h: chunk h / 2, output rows [64 (h & 1), +64)
          const int st = wpos % CF_WSTAGES;
          if (wpos >= CF_WSTAGES) wait(bar_w_empty(s, st), (uint32_t)((wpos / CF_WSTAGES) - 1) & 1u, UM_TIMEOUT_RING);
          um_mbar_arrive_expect_tx(bar_w_full(s, st), CF_HALF_B);
          CF_TR(3 + s, h);
          const unsigned char* src = a.image + (size_t)(h >> 1) * CF_CHUNK_B + (h & 1) * CF_PIECE_B;
          const uint32_t dst = sb + CF_OFF_WR + (uint32_t)(s * CF_WSTAGES + st) * CF_HALF_B;
          um_bulk_g2s(dst, src, CF_PIECE_B, bar_w_full(s, st));                              // hi
          um_bulk_g2s(dst + CF_PIECE_B, src + CF_TILE_B, CF_PIECE_B, bar_w_full(s, st));     // lo
        }
    }
  } else if (warp == 1 || warp == 2) {
    // =============== MMA issuers: slot 0 (warp 1), slot 1 (warp 2) ===============
    // The whole warp runs the (uniform) loop; one elected lane issues the tcgen05 instructions.  One issuing warp per slot:
    // a tcgen05.mma costs its issuing thread ~90 cycles whatever its shape, so two warps issue twice as fast as one, and
    // each follows its own slot's order with plain blocking waits.
    {
      const int s = warp - 1;
      constexpr uint32_t idesc = um_idesc_f16(CF_ROWS, 64);
      int wpos = 0;
      uint32_t u0 = 0u;
      for (int it = s; it < n_my; it += 2) {
        if (SAVE && it >= 2)
          for (uint32_t j = 0; j < 2u; ++j) {   // the previous tile's R3 chunks passed through the buffers without an MMA
            const uint32_t use = u0 - 2u + j;
            wait(bar_a_full(s, (int)(use & 1u)), (use >> 1) & 1u, UM_TIMEOUT_INPUT);
            if (um_elect()) um_mbar_arrive(bar_x_free(s, (int)(use & 1u)));
            __syncwarp();
          }
#pragma unroll 1
        for (int c = 0; c < CF_W_CHUNKS; ++c) {
          // operand use of chunk c: the cos chunk serves chunks 0 and 1, then one use per chunk
          const uint32_t use = u0 + (uint32_t)(c < 2 ? 0 : c - 1);
          if (c != 1) {
            wait(bar_a_full(s, (int)(use & 1u)), (use >> 1) & 1u, UM_TIMEOUT_INPUT);
            // layer 2 accumulates into columns [0,128): both chunks built from them must exist first
            if (c == 2) wait(bar_a_full(s, (int)((use + 1u) & 1u)), ((use + 1u) >> 1) & 1u, UM_TIMEOUT_INPUT);
          }
          const int st0 = wpos % CF_WSTAGES, st1 = (wpos + 1) % CF_WSTAGES;
          wait(bar_w_full(s, st0), (uint32_t)(wpos / CF_WSTAGES) & 1u, UM_TIMEOUT_RING);
          wait(bar_w_full(s, st1), (uint32_t)((wpos + 1) / CF_WSTAGES) & 1u, UM_TIMEOUT_RING);
          um_tc_after();
          if (lane == 0) CF_TR(0, s * 32 + c);
          const uint32_t d0 = tmem + (uint32_t)(256 * s + ((c == 1 || c >= 6) ? 128 : 0)), d1 = d0 + 64u;
          const bool first = c < 3 || c == 6;
          const uint32_t ah = sb + CF_OFF_X + (uint32_t)(2 * s + (int)(use & 1u)) * CF_CHUNK_B, al = ah + CF_TILE_B;
          const uint32_t wh0 = sb + CF_OFF_WR + (uint32_t)(s * CF_WSTAGES + st0) * CF_HALF_B, wl0 = wh0 + CF_PIECE_B;
          const uint32_t wh1 = sb + CF_OFF_WR + (uint32_t)(s * CF_WSTAGES + st1) * CF_HALF_B, wl1 = wh1 + CF_PIECE_B;
          if (um_elect()) {
            // the two 64-column halves are independent accumulation chains: alternate between them
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {   // 4 x (K = 16 fp16 = 32 B) per 128-byte row
              const uint32_t o = ks * 32;
              const uint32_t acc0 = (first && ks == 0) ? 0u : 1u;
              um_mma_f16(d0, um_desc_sw128(al + o), um_desc_sw128(wh0 + o), idesc, acc0);  // small terms first
              um_mma_f16(d1, um_desc_sw128(al + o), um_desc_sw128(wh1 + o), idesc, acc0);
              um_mma_f16(d0, um_desc_sw128(ah + o), um_desc_sw128(wl0 + o), idesc, 1u);
              um_mma_f16(d1, um_desc_sw128(ah + o), um_desc_sw128(wl1 + o), idesc, 1u);
              um_mma_f16(d0, um_desc_sw128(ah + o), um_desc_sw128(wh0 + o), idesc, 1u);
              um_mma_f16(d1, um_desc_sw128(ah + o), um_desc_sw128(wh1 + o), idesc, 1u);
            }
            um_commit(bar_w_empty(s, st0));
            um_commit(bar_w_empty(s, st1));
            if (c >= 1) um_commit(bar_x_free(s, (int)(use & 1u)));
            if (c == 0) um_commit(bar_d_full(s, 0));
            else if (c == 1) um_commit(bar_d_full(s, 1));
            else if (c == 5) um_commit(bar_d_full(s, 2));
            else if (c == 7) um_commit(bar_d_full(s, 3));
          }
          __syncwarp();
          wpos += 2;
        }
        u0 += SAVE ? 9u : 7u;
      }
    }
  } else if (warp >= 4) {
    // =============== epilogue warps ===============
    const int ew = warp - 4, s = ew >> 3, e = ew & 7, q = e & 3, hh = e >> 2;
    const int r = 32 * q + lane, r7 = r & 7;
    const uint32_t row_off = (uint32_t)((r >> 3) * 1024 + r7 * 128);
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(256 * s);
    const uint32_t xbase = sb + CF_OFF_X + (uint32_t)(2 * s) * CF_CHUNK_B;
    const bool elected = (e == 0 && lane == 0);
    const int grp_bar = 1 + s, pair_bar = 3 + 4 * s + q;
    const float4* bc4 = reinterpret_cast<const float4*>(a.bc);
    const float4* bh4 = reinterpret_cast<const float4*>(a.bh);
    const float4* bh24 = reinterpret_cast<const float4*>(a.bh2);
    const float4* wo4 = reinterpret_cast<const float4*>(a.wo);
    const float bo = MEAN ? 0.0f : __ldg(a.bo);
    const uint64_t counter = a.counter + (a.counter_dev ? *a.counter_dev : 0ull);
    uint32_t u = 0;

    // buffer of use u, once its previous reader (use u - 2) is done with it
    // (`newer_stores`: bulk stores committed after the one that read this buffer last — 1 in the running order)
    auto acquire = [&](bool newer_store = true) -> uint32_t {
      const uint32_t x = u & 1u;
      if (u >= 2u) wait(bar_x_free(s, (int)x), ((u >> 1) - 1u) & 1u, UM_TIMEOUT_RING);
      if (SAVE) {
        if (elected) {   // the bulk store of use u - 2 has read the buffer
          if (newer_store) um_bulk_wait_read<1>(); else um_bulk_wait_read<0>();
        }
        um_bar_sync(grp_bar, 256);
      }
      return xbase + x * CF_CHUNK_B;
    };
    // hand the chunk to the MMA thread (and, SAVE, to HBM at `img`)
    auto publish = [&](uint32_t xa, unsigned char* img) {
      um_fence_async_smem();
      um_tc_before();
      if (SAVE) {
        um_bar_sync(grp_bar, 256);
        if (elected) {
          um_bulk_s2g(img, xa, CF_CHUNK_B);
          um_bulk_commit();
          um_mbar_arrive(bar_a_full(s, (int)(u & 1u)));
        }
      } else {
        __syncwarp();
        if (lane == 0) um_mbar_arrive(bar_a_full(s, (int)(u & 1u)));
      }
      ++u;
    };

    for (int i = s; i < n_my; i += 2) {
      const uint32_t kpar = (uint32_t)(i >> 1) & 1u;
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const long m = (long)tile * CF_ROWS + r;
      const bool live = m < (long)a.rows;
      const int b = (int)((live ? m : (long)a.rows - 1) / a.n_tau);
      const float4* frow = reinterpret_cast<const float4*>(a.feats + (size_t)b * CF_W);
      const float4* arow = AENC ? reinterpret_cast<const float4*>(a.aenc + (size_t)b * CF_H) : nullptr;

      // ---- E0: tau, cos(tau pi i) -> the A chunk of layer 1 ----
      float tau = 0.0f;
      if (live) {
        if (a.tau_in) {
          tau = a.tau_in[m];
        } else {   // torch.rand(B, N) * cvar, ACIQN:419-422; the same draw as K3
          const uint4 w = philox_at(a.seed, STREAM_TAU, counter, (int64_t)(m >> 2), 0u, 0u, 0u);
          const uint32_t word = (m & 3) == 0 ? w.x : (m & 3) == 1 ? w.y : (m & 3) == 2 ? w.z : w.w;
          tau = __fmul_rn(u01(word), a.cvar);
        }
        if (hh == 0 && a.tau_out) a.tau_out[m] = tau;
      }
      {
        if (elected) CF_TR(1 + s, 0);
        const uint32_t xa = acquire();
        if (elected) CF_TR(1 + s, 1);
        if (hh == 0) cf_cos_octets<0>(tau, xa + row_off, r7); else cf_cos_octets<32>(tau, xa + row_off, r7);
        publish(xa, SAVE ? a.cos_img + (size_t)tile * CF_CHUNK_B : nullptr);
        if (elected) CF_TR(1 + s, 2);
      }

      // ---- E1: D1 -> H1 = feats[b] * relu(D1 / 64 + bc) -> the four A chunks of layer 2 ----
      wait(bar_d_full(s, 0), kpar, UM_TIMEOUT_ACC);
      if (elected) CF_TR(1 + s, 3);
      for (int j = 0; j < 4; ++j) {
        if (j == 1) wait(bar_d_full(s, 1), kpar, UM_TIMEOUT_ACC);   // also: layer 1 has read the cos chunk's buffer
        um_tc_after();
        const uint32_t xa = acquire();
        if (elected) CF_TR(1 + s, 4 + 2 * j);
        uint32_t v[32];
        um_tmem_ld32(lane_addr + (uint32_t)(64 * j + 32 * hh), v);
        um_tmem_ld_wait();
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int c4 = (64 * j + 32 * hh + 8 * t) / 4;
          const float4 fa = __ldg(frow + c4), fb = __ldg(frow + c4 + 1), ba = __ldg(bc4 + c4), bb = __ldg(bc4 + c4 + 1);
          const float f[8] = {fa.x, fa.y, fa.z, fa.w, fb.x, fb.y, fb.z, fb.w};
          const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
          float h[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) h[k] = f[k] * fmaxf(fmaf(__uint_as_float(v[8 * t + k]), CF_WUNSCALE, bias[k]), 0.0f);
          um_put8(xa + row_off + (uint32_t)(((4 * hh + t) ^ r7) << 4), CF_TILE_B, h);
        }
        publish(xa, SAVE ? a.h1_img + ((size_t)tile * 4 + j) * CF_CHUNK_B : nullptr);
        if (elected) CF_TR(1 + s, 5 + 2 * j);
      }

      // ---- E2: D2 -> H2 = aenc[b] * relu(D2 / 64 + bh) -> the two A chunks of layer 3 ----
      wait(bar_d_full(s, 2), kpar, UM_TIMEOUT_ACC);
      if (elected) CF_TR(1 + s, 12);
      for (int j = 0; j < 2; ++j) {
        um_tc_after();
        const uint32_t xa = acquire();
        uint32_t v[32];
        um_tmem_ld32(lane_addr + (uint32_t)(64 * j + 32 * hh), v);
        um_tmem_ld_wait();
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int c4 = (64 * j + 32 * hh + 8 * t) / 4;
          const float4 ba = __ldg(bh4 + c4), bb = __ldg(bh4 + c4 + 1);
          float4 fa = make_float4(1.f, 1.f, 1.f, 1.f), fb = fa;
          if (AENC) { fa = __ldg(arow + c4); fb = __ldg(arow + c4 + 1); }
          const float f[8] = {fa.x, fa.y, fa.z, fa.w, fb.x, fb.y, fb.z, fb.w};
          const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
          float h[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) h[k] = f[k] * fmaxf(fmaf(__uint_as_float(v[8 * t + k]), CF_WUNSCALE, bias[k]), 0.0f);
          um_put8(xa + row_off + (uint32_t)(((4 * hh + t) ^ r7) << 4), CF_TILE_B, h);
        }
        publish(xa, SAVE ? a.h2_img + ((size_t)tile * 2 + j) * CF_CHUNK_B : nullptr);
        if (elected) CF_TR(1 + s, 13 + j);
      }

      // ---- E3: D3 -> R3 = relu(D3 / 64 + bh2); z = R3 . wo + bo.  This warp: columns [64 hh, 64 hh + 64) ----
      wait(bar_d_full(s, 3), kpar, UM_TIMEOUT_ACC);
      if (elected) CF_TR(1 + s, 15);
      um_tc_after();
      float acc = 0.0f;
      uint32_t xr[2] = {0u, 0u};
      if (SAVE) {   // both R3 chunks are acquired up front (use u and u + 1): this warp fills chunk hh only
        xr[0] = acquire();
        ++u;
        xr[1] = acquire(false);
        --u;
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t v[32];
        um_tmem_ld32(lane_addr + (uint32_t)(128 + 64 * hh + 32 * half), v);
        um_tmem_ld_wait();
        if (MEAN) {
          // IQN's act (AGENT:227-250): the action values are the mean over the 32 quantile rows of a LINEAR output layer, so
          // the mean is taken here, of the last hidden layer; the warp's 32 lanes are exactly one sample's rows
          float p[32];
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int c4 = (64 * hh + 32 * half + 8 * t) / 4;
            const float4 ba = __ldg(bh24 + c4), bb = __ldg(bh24 + c4 + 1);
            const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
#pragma unroll
            for (int k = 0; k < 8; ++k) p[8 * t + k] = fmaxf(fmaf(__uint_as_float(v[8 * t + k]), CF_WUNSCALE, bias[k]), 0.0f);
          }
          um_seg_reduce<5>(p, lane);
          if (live) a.z[(size_t)b * CF_H + 64 * hh + 32 * half + lane] = p[0] * (1.0f / 32.0f);
          continue;
        }
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int c4 = (64 * hh + 32 * half + 8 * t) / 4;
          const float4 ba = __ldg(bh24 + c4), bb = __ldg(bh24 + c4 + 1), wa = __ldg(wo4 + c4), wb = __ldg(wo4 + c4 + 1);
          const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
          const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
          float h[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            h[k] = fmaxf(fmaf(__uint_as_float(v[8 * t + k]), CF_WUNSCALE, bias[k]), 0.0f);
            acc = fmaf(h[k], w[k], acc);
          }
          if (SAVE) um_put8(xr[hh] + row_off + (uint32_t)(((4 * half + t) ^ r7) << 4), CF_TILE_B, h);
        }
      }
      if (SAVE) {
        publish(xr[0], a.r3_img + ((size_t)tile * 2 + 0) * CF_CHUNK_B);
        publish(xr[1], a.r3_img + ((size_t)tile * 2 + 1) * CF_CHUNK_B);
      }
      um_tc_before();
      if (!MEAN) {
        if (hh == 1) zpart[s * CF_ROWS + r] = acc;
        um_bar_sync(pair_bar, 64);
        if (hh == 0 && live) a.z[m] = acc + zpart[s * CF_ROWS + r] + bo;
      }
      if (elected) CF_TR(1 + s, 16);
    }
    if (SAVE && elected) um_bulk_wait<0>();
  }

  um_tc_before();
  __syncthreads();
  if (warp == 2) {
    um_tc_after();
    um_tmem_free<512>(tmem);
  }
}

static __device__ int cf_status_word = 0;

}  // namespace asvrl

using namespace asvrl;

extern "C" long asvrl_critic_forward_workspace(void) { return (long)CF_IMAGE_B; }
extern "C" long asvrl_critic_forward_chunk_bytes(void) { return (long)CF_CHUNK_B; }

extern "C" int asvrl_critic_forward(const float* tau_in, uint64_t seed, uint64_t counter, const uint64_t* counter_dev,
                                    float cvar, const float* feats, const float* aenc, const float* wc, const float* bc,
                                    const float* wh, const float* bh, const float* wh2, const float* bh2, const float* wo,
                                    const float* bo, int batch, int n_tau, int head, int repack, void* workspace,
                                    void* bwd_workspace, float* tau_out, float* z, void* cos_img, void* h1_img,
                                    void* h2_img, void* r3_img, void* stream) {
  if (!feats || !wc || !bc || !wh || !bh || !wh2 || !bh2 || !workspace || !z || (head == 0 && (!wo || !bo)))
    return fail(ASVRL_E_BADARG, "asvrl_critic_forward: null pointer");
  if (head != 0 && head != 1) return fail(ASVRL_E_BADARG, "asvrl_critic_forward: head must be 0 or 1");
  if (head == 1 && (n_tau != 32 || aenc || cos_img || h1_img || h2_img || r3_img))
    return fail(ASVRL_E_BADARG, "asvrl_critic_forward: head 1 (mean of the last hidden layer) takes n_tau = 32, no aenc, no images");
  if (batch < 1 || n_tau < 1 || (long)batch * n_tau > 0x7FFFFFFFL - CF_ROWS)
    return fail(ASVRL_E_BADARG, "asvrl_critic_forward: bad sizes (B=%d N=%d)", batch, n_tau);
  const bool save = cos_img || h1_img || h2_img || r3_img;
  if (save && !(cos_img && h1_img && h2_img && r3_img))
    return fail(ASVRL_E_BADARG, "asvrl_critic_forward: the four operand images go together");
  if (bwd_workspace && !repack) return fail(ASVRL_E_BADARG, "asvrl_critic_forward: bwd_workspace is written by the pack launch (repack != 0)");
  if (((uintptr_t)workspace | (uintptr_t)bwd_workspace | (uintptr_t)cos_img | (uintptr_t)h1_img | (uintptr_t)h2_img | (uintptr_t)r3_img) & 127)
    return fail(ASVRL_E_BADARG, "asvrl_critic_forward: workspace and images must be 128-byte aligned");
  if (((uintptr_t)feats | (uintptr_t)aenc | (uintptr_t)bc | (uintptr_t)bh | (uintptr_t)bh2 | (uintptr_t)wo | (uintptr_t)z) & 15)
    return fail(ASVRL_E_BADARG, "asvrl_critic_forward: feats, aenc, biases and wo must be 16-byte aligned");
  static int n_sm = 0;
  cudaError_t e;
  if (n_sm == 0) {
    int dev = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess || (e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess)
      return fail((int)e, "asvrl_critic_forward: %s", cudaGetErrorString(e));
    if ((e = cudaFuncSetAttribute(critic_forward_kernel<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF_SMEM_B)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(critic_forward_kernel<false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF_SMEM_B)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(critic_forward_kernel<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF_SMEM_B)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(critic_forward_kernel<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF_SMEM_B)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(critic_forward_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF_SMEM_B)) != cudaSuccess) {
      n_sm = 0;
      return fail((int)e, "asvrl_critic_forward: %d B shared memory: %s", CF_SMEM_B, cudaGetErrorString(e));
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (repack) {
    const int n = (CF_W_CHUNKS + (bwd_workspace ? 6 : 0)) * CF_H * 8;
    // with a backward workspace the launch also leaves max feats for K11 behind the packed image (64 partial maxima)
    float* fpart = bwd_workspace ? reinterpret_cast<float*>((unsigned char*)bwd_workspace + 6 * CF_CHUNK_B) : nullptr;
    critic_pack_kernel<<<(n + 255) / 256, 256, 0, st>>>(wc, wh, wh2, (unsigned char*)workspace, (unsigned char*)bwd_workspace,
                                                        reinterpret_cast<const float4*>(feats), batch * (CF_W / 4), fpart);
    if (int rc = check_launch("asvrl_critic_forward (pack)")) return rc;
  }
  CriticArgs a;
  a.tau_in = tau_in; a.seed = seed; a.counter = counter; a.counter_dev = counter_dev; a.cvar = cvar;
  a.feats = feats; a.aenc = aenc; a.bc = bc; a.bh = bh; a.bh2 = bh2; a.wo = wo; a.bo = bo;
  a.image = (const unsigned char*)workspace;
  a.rows = batch * n_tau; a.n_tau = n_tau; a.batch = batch; a.n_tiles = (a.rows + CF_ROWS - 1) / CF_ROWS;
  int* status = nullptr;
  if (cudaGetSymbolAddress((void**)&status, cf_status_word) != cudaSuccess) status = nullptr;
  a.status = status;
  a.tau_out = tau_out; a.z = z;
  a.cos_img = (unsigned char*)cos_img; a.h1_img = (unsigned char*)h1_img;
  a.h2_img = (unsigned char*)h2_img; a.r3_img = (unsigned char*)r3_img;
  const int grid = a.n_tiles < n_sm ? a.n_tiles : n_sm;
  if (head == 1) {
    critic_forward_kernel<false, false, true><<<grid, CF_THREADS, CF_SMEM_B, st>>>(a);
  } else if (save) {
    if (aenc) critic_forward_kernel<true, true, false><<<grid, CF_THREADS, CF_SMEM_B, st>>>(a);
    else critic_forward_kernel<true, false, false><<<grid, CF_THREADS, CF_SMEM_B, st>>>(a);
  } else {
    if (aenc) critic_forward_kernel<false, true, false><<<grid, CF_THREADS, CF_SMEM_B, st>>>(a);
    else critic_forward_kernel<false, false, false><<<grid, CF_THREADS, CF_SMEM_B, st>>>(a);
  }
  return check_launch("asvrl_critic_forward");
}

#ifdef CF_TRACE
// debug build only: copies CTA 0's timeline out ([8][2048] events, [8] counts) and resets it
extern "C" __attribute__((visibility("default"))) int asvrl_critic_forward_trace(unsigned long long* events, int* counts) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(events, cf_trace_buf, sizeof(unsigned long long) * 8 * 2048);
  cudaMemcpyFromSymbol(counts, cf_trace_n, sizeof(int) * 8);
  int zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  cudaMemcpyToSymbol(cf_trace_n, zero, sizeof zero);
  return 0;
}
#endif

// synchronous read of K10's device status word (0 = every bounded wait completed)
extern "C" int asvrl_critic_forward_status(int* status_out) {
  if (!status_out) return fail(ASVRL_E_BADARG, "asvrl_critic_forward_status: null pointer");
  cudaError_t e = cudaMemcpyFromSymbol(status_out, cf_status_word, sizeof(int));
  if (e != cudaSuccess) return fail((int)e, "asvrl_critic_forward_status: %s", cudaGetErrorString(e));
  return 0;
}
