// critic_bwd_umma.cu — K11: the backward chain of the IQN critic's quantile trunk (the activation half of the autograd
// backward of rfarl/rfarl/policy/AC_IQN_model.py:453-480; reference: loss.backward() at rfarl/rfarl/agent.py:413, 425)
// for all B*N quantile rows in ONE persistent, warp-specialised tcgen05 launch.  Per 128-row tile, with the operand
// images K10 saved (R3, H2, H1) brought by bulk copies into the very buffers the step's output overwrites:
//   B0  dP3 = S dz . wo . 1[R3 > 0]                              -> A chunks of   dH2 = dP3 Wh2          (tensor)
//   B1  dP2 = dH2 . aenc[b] . 1[H2 > 0];  daenc[b] = sum_n dH2 . H2 / aenc[b]
//                                                                 -> A chunks of   dH1 = dP2 Wh           (tensor)
//   B2  dP1 = dH1 . feats[b] . 1[H1 > 0]; dfeats[b] = sum_n dH1 . H1 / feats[b]
// It replaces head_bwd + 2 x (cuBLAS sgemm + scale_relu_bwd) and leaves dP3 / dP2 / dP1 (+ the aux tile: S dz | 1) in
// HBM as operand images for K12, which computes every weight and bias gradient from them.  FULL = false is the
// actor update's pass (AGENT:420-427): only d z / d aenc flows back, nothing is stored for K12 and the chain stops
// after B1.
//
// Scaling.  Gradients are small numbers and the operands are two fp16 pieces, so dz is multiplied by a power of two S
// chosen in the prologue from max |dz| (S max|dz| in [256, 512)); dP1 = dH1 . feats also carries the observation
// features' range (thrust up to 1000 N gives features in the hundreds) and uses S1 = S 2^-k with 2^-k max feats in
// (1, 2] (max feats: `fmax`, a device word of the caller, or — `fmax` NULL and `repack` 0 — the 64 partial maxima that
// K10's weight-pack launch leaves behind the image when it prepares this workspace).  Both are exact and are removed
// again here (daenc, dfeats) and in K12's finalize pass.
//
// The per-sample sums over the N quantile rows are warp transposes (lane = row): N must be a power of two <= 32.
// Structure, operand buffers, barriers and the two tile slots are K10's (critic_umma.cu); the weight ring streams
// Wh2^T (2 chunks) and Wh^T (2 N-halves x 2 chunks).
#include "umma.cuh"

namespace asvrl {

constexpr int CB_THREADS = 640;
constexpr int CB_ROWS = 128;
constexpr int CB_W = 256, CB_H = 128;
constexpr int CB_TILE_B = CB_ROWS * 128;
constexpr int CB_CHUNK_B = 2 * CB_TILE_B;
constexpr int CB_WSTAGES = 3;
constexpr int CB_W_CHUNKS = 6;
constexpr int CB_IMAGE_B = CB_W_CHUNKS * CB_CHUNK_B;   // 192 KB
constexpr int CB_FPART = 64;                           // partial maxima of feats kept behind the image (float each)
constexpr float CB_WSCALE = 64.0f, CB_WUNSCALE = 1.0f / 64.0f;

constexpr int CB_OFF_X = 0;
constexpr int CB_OFF_WR = 4 * CB_CHUNK_B;
constexpr int CB_OFF_BAR = CB_OFF_WR + CB_WSTAGES * CB_CHUNK_B;
constexpr int CB_NBAR = 2 * 8 + 2 * CB_WSTAGES;                    // per slot a_full[2] x_free[2] d_full[2] img_full[2]; w_full, w_empty
constexpr int CB_OFF_MISC = CB_OFF_BAR + CB_NBAR * 8;              // tmem slot, abort flag, S, 1 / S
constexpr int CB_OFF_RED = CB_OFF_MISC + 32;                       // 20 floats of the max reduction
constexpr int CB_SMEM_B = CB_OFF_RED + 256 + 1024;
static_assert(CB_SMEM_B <= 232448, "K11 shared memory");

#ifdef CF_TRACE
// timeline of CTA 0 for profiles/probes/k10_trace.py: [writer][event] = (globaltimer << 8) | tag
__device__ unsigned long long cb_trace_buf[4][1024];
__device__ int cb_trace_n[4];
#define CB_TR(writer, tag)                                                                              \
  do {                                                                                                   \
    if (blockIdx.x == 0) {                                                                               \
      const int n_ = cb_trace_n[writer];                                                                 \
      if (n_ < 1024) { cb_trace_buf[writer][n_] = (um_globaltimer() << 8) | (unsigned)(tag); cb_trace_n[writer] = n_ + 1; } \
    }                                                                                                    \
  } while (0)
#else
#define CB_TR(writer, tag) do {} while (0)
#endif

struct CriticBwdArgs {
  const float* dz;                              // [rows]
  const unsigned char *r3_img, *h2_img, *h1_img;
  const float *feats, *aenc, *wo, *fmax;
  const unsigned char* image;                   // packed Wh2^T | Wh^T chunks (critic_bwd_pack_kernel)
  int rows, n_tau, n_tiles;
  int* status;
  float *daenc, *dfeats;                        // [B][128], [B][256]
  unsigned char *dp3_img, *dp2_img, *dp1_img, *aux_img;
  float* scales;                                // {1 / S, 1 / S1, sum dz}
};

// chunk 0, 1: B[n = i][k = j - 64 c] = Wh2[j][i]; chunk 2 + 2 nh + c: B[n][k] = Wh[64 c + k][128 nh + n]
__global__ void __launch_bounds__(256)
critic_bwd_pack_kernel(const float* __restrict__ wh, const float* __restrict__ wh2, unsigned char* __restrict__ image,
                       const float* __restrict__ fmax) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  // the 64 partial maxima of feats behind the image (written by K10's pack launch when it prepares this workspace):
  // a stand-alone pack leaves the caller's value, or the documented default (2^8, i.e. S1 = S 2^-8)
  if (t < CB_FPART) reinterpret_cast<float*>(image + CB_IMAGE_B)[t] = t ? 0.0f : (fmax ? __ldg(fmax) : 256.0f);
  if (t >= CB_W_CHUNKS * CB_H * 8) return;
  const int c = t / (CB_H * 8), n = (t / 8) % CB_H, oc = t & 7;
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = oc * 8 + e;
    if (c < 2) v[e] = wh2[(size_t)(64 * c + k) * CB_H + n];
    else v[e] = wh[(size_t)(64 * ((c - 2) & 1) + k) * CB_W + 128 * ((c - 2) >> 1) + n];
    v[e] *= CB_WSCALE;
  }
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) um_split2(v[2 * i], v[2 * i + 1], h[i], l[i]);
  unsigned char* tile = image + (size_t)c * CB_CHUNK_B + um_sw128_off(n, oc);
  *reinterpret_cast<uint4*>(tile) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(tile + CB_TILE_B) = make_uint4(l[0], l[1], l[2], l[3]);
}

// 8 halves (one 16-byte octet) of the hi piece + the lo piece -> 8 floats
__device__ __forceinline__ void cb_octet(const uint4 h, const uint4 l, float (&x)[8]) {
  const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&hw[i]));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&lw[i]));
    x[2 * i] = a.x + b.x;
    x[2 * i + 1] = a.y + b.y;
  }
}

template <bool FULL, int LOGG>
__global__ void __launch_bounds__(CB_THREADS, 1)
critic_backward_kernel(const CriticBwdArgs a) {
  extern __shared__ __align__(1024) unsigned char cb_raw[];
  unsigned char* base = cb_raw + ((1024u - (um_smem(cb_raw) & 1023u)) & 1023u);
  const uint32_t sb = um_smem(base);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  const uint32_t bar0 = sb + CB_OFF_BAR;
  auto bar_a_full = [&](int s, int x) { return bar0 + 8u * (8 * s + x); };
  auto bar_x_free = [&](int s, int x) { return bar0 + 8u * (8 * s + 2 + x); };
  auto bar_d_full = [&](int s, int d) { return bar0 + 8u * (8 * s + 4 + d); };
  auto bar_img = [&](int s, int x) { return bar0 + 8u * (8 * s + 6 + x); };
  auto bar_w_full = [&](int st) { return bar0 + 8u * (16 + st); };
  auto bar_w_empty = [&](int st) { return bar0 + 8u * (16 + CB_WSTAGES + st); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base + CB_OFF_MISC);
  volatile uint32_t* abort_flag = reinterpret_cast<volatile uint32_t*>(base + CB_OFF_MISC + 4);
  float* s_scale = reinterpret_cast<float*>(base + CB_OFF_MISC + 8);   // S, 1 / S
  float* s_red = reinterpret_cast<float*>(base + CB_OFF_RED);

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      for (int x = 0; x < 2; ++x) {
        um_mbar_init(bar_a_full(s, x), FULL ? 1 : 8);
        um_mbar_init(bar_x_free(s, x), 1);
        um_mbar_init(bar_img(s, x), 1);
      }
      for (int d = 0; d < 2; ++d) um_mbar_init(bar_d_full(s, d), 1);
    }
    for (int st = 0; st < CB_WSTAGES; ++st) {
      um_mbar_init(bar_w_full(st), 1);
      um_mbar_init(bar_w_empty(st), 1);
    }
    *abort_flag = 0u;
    um_mbar_init_fence();
  }
  if (tid == 0) CB_TR(3, 0);
  if (warp == 2) um_tmem_alloc<512>(um_smem(tmem_slot));
  __syncthreads();   // barriers initialised
  // the R3 chunks of each slot's first tile start their way into the operand buffers under the dz reduction
  if ((tid == 128 || tid == 384) && (int)blockIdx.x + (tid == 384 ? (int)gridDim.x : 0) < a.n_tiles) {
    const int s = tid == 384 ? 1 : 0;   // the slots' elected threads: warp 4 / warp 12, lane 0
    const size_t tile = (size_t)blockIdx.x + (size_t)s * gridDim.x;
    for (int x = 0; x < 2; ++x) {
      um_mbar_arrive_expect_tx(bar_img(s, x), CB_CHUNK_B);
      um_bulk_g2s(sb + CB_OFF_X + (uint32_t)(2 * s + x) * CB_CHUNK_B, a.r3_img + (tile * 2 + x) * CB_CHUNK_B, CB_CHUNK_B,
                  bar_img(s, x));
    }
  }
  {
    // S from max |dz| over the whole batch (every CTA computes the same value)
    float mx = 0.0f, sm = 0.0f;
    if ((a.rows & 3) == 0 && (reinterpret_cast<uintptr_t>(a.dz) & 15) == 0) {
      const float4* d4 = reinterpret_cast<const float4*>(a.dz);
#pragma unroll 4
      for (int i = tid; i < (a.rows >> 2); i += CB_THREADS) {
        const float4 d = __ldg(d4 + i);
        mx = fmaxf(fmaxf(mx, fmaxf(fabsf(d.x), fabsf(d.y))), fmaxf(fabsf(d.z), fabsf(d.w)));
        sm += (d.x + d.y) + (d.z + d.w);     // d loss / d bo (ACIQN:480): the same fixed order in every launch
      }
    } else {
      for (int i = tid; i < a.rows; i += CB_THREADS) {
        const float d = __ldg(a.dz + i);
        mx = fmaxf(mx, fabsf(d));
        sm += d;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      sm += __shfl_xor_sync(0xffffffffu, sm, o);
    }
    if (lane == 0) { s_red[warp] = mx; s_red[CB_THREADS / 32 + warp] = sm; }
    if (FULL && warp == 0) {
      // max feats: the caller's word, else the partial maxima the forward's pack launch left behind the weight image
      const float* fp = reinterpret_cast<const float*>(a.image + CB_IMAGE_B);
      float fm = a.fmax ? __ldg(a.fmax) : fmaxf(__ldg(fp + lane), __ldg(fp + 32 + lane));
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) fm = fmaxf(fm, __shfl_xor_sync(0xffffffffu, fm, o));
      if (lane == 0) s_red[2 * (CB_THREADS / 32)] = fm;
    }
  }
  um_tc_before();
  __syncthreads();
  um_tc_after();
  if (tid == 0) {
    float mx = 0.0f;
    for (int w = 0; w < CB_THREADS / 32; ++w) mx = fmaxf(mx, s_red[w]);
    float S = 1.0f;
    if (mx > 0.0f && mx < 3.0e38f) {
      int e;
      (void)frexpf(mx, &e);                 // mx = f 2^e, f in [0.5, 1)
      const int p = min(max(9 - e, -100), 100);
      S = ldexpf(1.0f, p);                  // S mx in [256, 512)
    }
    float F = 1.0f / 256.0f;   // S1 / S
    if (FULL) {
      const float fm = s_red[2 * (CB_THREADS / 32)];
      F = 1.0f;
      if (fm > 2.0f && fm < 3.0e38f) {
        int e;
        (void)frexpf(fm, &e);               // fm = f 2^e, f in [0.5, 1): fm 2^-(e-1) in [1, 2)
        F = ldexpf(1.0f, -min(e - 1, 100));
      }
    }
    s_scale[0] = S;
    s_scale[1] = 1.0f / S;
    s_scale[2] = F;
    if (FULL && blockIdx.x == 0) {
      float sum = 0.0f;
      for (int w = 0; w < CB_THREADS / 32; ++w) sum += s_red[CB_THREADS / 32 + w];
      a.scales[0] = 1.0f / S; a.scales[1] = 1.0f / (S * F); a.scales[2] = sum;
    }
  }
  __syncthreads();
  if (tid == 0) CB_TR(3, 1);
  const uint32_t tmem = *tmem_slot;
  const float S = s_scale[0], invS = s_scale[1], F1 = s_scale[2];

  const int n_my = ((int)blockIdx.x < a.n_tiles) ? (a.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto wait = [&](uint32_t bar, uint32_t parity, int code) { (void)um_mbar_wait(bar, parity, abort_flag, a.status, code); };

  if (warp == 0) {
    // =============== W producer ===============
    if (lane == 0) {
      int pos = 0;
      auto emit = [&](int first, int n) {
        for (int c = first; c < first + n; ++c, ++pos) {
          const int st = pos % CB_WSTAGES;
          if (pos >= CB_WSTAGES) wait(bar_w_empty(st), (uint32_t)((pos / CB_WSTAGES) - 1) & 1u, UM_TIMEOUT_RING);
          um_mbar_arrive_expect_tx(bar_w_full(st), CB_CHUNK_B);
          um_bulk_g2s(sb + CB_OFF_WR + st * CB_CHUNK_B, a.image + (size_t)c * CB_CHUNK_B, CB_CHUNK_B, bar_w_full(st));
        }
      };
      for (int i = 0; i < n_my; i += 2) {
        const bool two = i + 1 < n_my;
        emit(0, 2); if (two) emit(0, 2);
        if (FULL) { emit(2, 4); if (two) emit(2, 4); }
      }
    }
  } else if (warp == 1) {
    // =============== MMA issuer ===============
    if (lane == 0) {
      constexpr uint32_t idesc = um_idesc_f16(CB_ROWS, CB_H);
      int pos = 0;
      auto xaddr = [&](int s, int x) { return sb + CB_OFF_X + (uint32_t)(2 * s + x) * CB_CHUNK_B; };
      auto chunk = [&](uint32_t ax, uint32_t d_tmem, bool first) {
        const int st = pos % CB_WSTAGES;
        wait(bar_w_full(st), (uint32_t)(pos / CB_WSTAGES) & 1u, UM_TIMEOUT_RING);
        um_tc_after();
        const uint32_t ah = ax, al = ax + CB_TILE_B;
        const uint32_t wh = sb + CB_OFF_WR + st * CB_CHUNK_B, wl = wh + CB_TILE_B;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint32_t o = ks * 32;
          um_mma_f16(d_tmem, um_desc_sw128(al + o), um_desc_sw128(wh + o), idesc, (first && ks == 0) ? 0u : 1u);
          um_mma_f16(d_tmem, um_desc_sw128(ah + o), um_desc_sw128(wl + o), idesc, 1u);
          um_mma_f16(d_tmem, um_desc_sw128(ah + o), um_desc_sw128(wh + o), idesc, 1u);
        }
        um_commit(bar_w_empty(st));
        ++pos;
      };
      // phase of the MMA uses of a slot's buffers: FULL two per tile (dP3, dP2), else one (dP3)
      auto Sa = [&](int s, int k) {   // dH2 = dP3 Wh2 into columns [0,128)
        const uint32_t ph = (uint32_t)(FULL ? 2 * k : k) & 1u;
        wait(bar_a_full(s, 0), ph, UM_TIMEOUT_INPUT);
        wait(bar_a_full(s, 1), ph, UM_TIMEOUT_INPUT);
        um_tc_after();
        const uint32_t d = tmem + (uint32_t)(256 * s);
        CB_TR(0, 16 * s + 1);
        chunk(xaddr(s, 0), d, true);
        chunk(xaddr(s, 1), d, false);
        CB_TR(0, 16 * s + 2);
        um_commit(bar_x_free(s, 0));
        um_commit(bar_x_free(s, 1));
        um_commit(bar_d_full(s, 0));
      };
      auto Sb = [&](int s, int k) {   // dH1 = dP2 Wh into columns [0,256): dH2 has been read by then
        const uint32_t ph = (uint32_t)(2 * k + 1) & 1u;
        wait(bar_a_full(s, 0), ph, UM_TIMEOUT_INPUT);
        wait(bar_a_full(s, 1), ph, UM_TIMEOUT_INPUT);
        um_tc_after();
        const uint32_t d = tmem + (uint32_t)(256 * s);
        CB_TR(0, 16 * s + 3);
        chunk(xaddr(s, 0), d, true);
        chunk(xaddr(s, 1), d, false);
        chunk(xaddr(s, 0), d + 128u, true);
        chunk(xaddr(s, 1), d + 128u, false);
        CB_TR(0, 16 * s + 4);
        um_commit(bar_x_free(s, 0));
        um_commit(bar_x_free(s, 1));
        um_commit(bar_d_full(s, 1));
      };
      for (int i = 0; i < n_my; i += 2) {
        const bool two = i + 1 < n_my;
        const int k = i >> 1;
        Sa(0, k); if (two) Sa(1, k);
        if (FULL) { Sb(0, k); if (two) Sb(1, k); }
      }
    }
  } else if (warp >= 4) {
    // =============== epilogue warps ===============
    const int ew = warp - 4, s = ew >> 3, e = ew & 7, q = e & 3, hh = e >> 2;
    const int r = 32 * q + lane, r7 = r & 7;
    const uint32_t row_off = (uint32_t)((r >> 3) * 1024 + r7 * 128);
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(256 * s);
    const uint32_t xbase = sb + CB_OFF_X + (uint32_t)(2 * s) * CB_CHUNK_B;
    const bool elected = (e == 0 && lane == 0);
    const int grp_bar = 1 + s;
    const float4* wo4 = reinterpret_cast<const float4*>(a.wo);
    constexpr int CNT = 32 >> LOGG;                       // columns per lane after the segmented reduction
    const int col_l = (lane & ((1 << LOGG) - 1)) * CNT;   // first of them

    // Use `ut` of tile k of this slot takes buffer ut & 1.  begin(): wait for the buffer's previous MMA reader and
    // bulk-store reader, then ONE thread brings the saved image chunk the step consumes into that very buffer with a
    // bulk copy (img_full).  Every thread later reads its own row's octets from it and overwrites them in place with
    // the dP chunk: the image never goes through the L1 as 128-byte-strided 16-byte loads (32 wavefronts per request,
    // which clogged the load / shuffle pipe: profiles/probes/k11_trace.py).
    auto begin = [&](int ut, int k, const unsigned char* img, bool issued = false) -> uint32_t {
      const int x = ut & 1;
      if (FULL) {
        if (ut == 2 || ut == 3) wait(bar_x_free(s, x), 0u, UM_TIMEOUT_RING);        // phase 2k
        else if (ut == 4 || ut == 5) wait(bar_x_free(s, x), 1u, UM_TIMEOUT_RING);   // phase 2k + 1
        if (elected) um_bulk_wait_read<0>();   // every bulk store out of the buffers has read its source
        um_bar_sync(grp_bar, 256);
      } else if (ut < 2 && k > 0) {
        wait(bar_x_free(s, x), (uint32_t)(k - 1) & 1u, UM_TIMEOUT_RING);
      }
      const uint32_t xa = xbase + (uint32_t)x * CB_CHUNK_B;
      if (elected && !issued) {
        um_mbar_arrive_expect_tx(bar_img(s, x), CB_CHUNK_B);
        um_bulk_g2s(xa, img, CB_CHUNK_B, bar_img(s, x));
      }
      return xa;
    };
    // the image chunk of use `ut` has landed; this thread's four octets (hi piece, then lo piece)
    auto image = [&](int ut, uint32_t xa, uint4 (&g)[8]) {
      wait(bar_img(s, ut & 1), (uint32_t)(ut >> 1) & 1u, UM_TIMEOUT_INPUT);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const uint32_t so = (uint32_t)(((4 * hh + t) ^ r7) << 4);
        g[t] = um_lds128(xa + row_off + so);
        g[4 + t] = um_lds128(xa + row_off + CB_TILE_B + so);
      }
    };
    auto publish = [&](int ut, uint32_t xa, unsigned char* img) {
      um_fence_async_smem();
      um_tc_before();
      if (FULL) {
        um_bar_sync(grp_bar, 256);
        if (elected) {
          um_bulk_s2g(img, xa, CB_CHUNK_B);
          um_bulk_commit();
          if (ut < 4) um_mbar_arrive(bar_a_full(s, ut & 1));
        }
      } else {
        __syncwarp();
        if (lane == 0) um_mbar_arrive(bar_a_full(s, ut & 1));
      }
    };

    for (int i = s; i < n_my; i += 2) {
      const int k = i >> 1;
      const uint32_t kpar = (uint32_t)k & 1u;
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const long m = (long)tile * CB_ROWS + r;
      const bool live = m < (long)a.rows;
      const int b = (int)((live ? m : (long)a.rows - 1) / a.n_tau);
      const float4* frow = reinterpret_cast<const float4*>(a.feats + (size_t)b * CB_W);
      const float4* arow = reinterpret_cast<const float4*>(a.aenc + (size_t)b * CB_H);
      const float dzs = live ? S * __ldg(a.dz + m) : 0.0f;
      uint32_t xa[2];

      if (elected) CB_TR(1 + s, 1);
      // ---- B0: dP3 = S dz . wo . 1[R3 > 0] ----
      xa[0] = begin(0, k, a.r3_img + ((size_t)tile * 2 + 0) * CB_CHUNK_B, k == 0);   // k == 0: issued in the prologue
      xa[1] = begin(1, k, a.r3_img + ((size_t)tile * 2 + 1) * CB_CHUNK_B, k == 0);
      if (FULL && hh == 0) {   // aux tile for K12: column 0 = S dz, column 1 = 1 (octets 0 and 1 of the row, hi | lo)
        unsigned char* ax = a.aux_img + (size_t)tile * CB_CHUNK_B + row_off;
        uint32_t dh, dl;
        um_split2(dzs, live ? 1.0f : 0.0f, dh, dl);
        *reinterpret_cast<uint4*>(ax + ((0 ^ r7) << 4)) = make_uint4(dh, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(ax + ((1 ^ r7) << 4)) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(ax + CB_TILE_B + ((0 ^ r7) << 4)) = make_uint4(dl, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(ax + CB_TILE_B + ((1 ^ r7) << 4)) = make_uint4(0u, 0u, 0u, 0u);
      }
      for (int j = 0; j < 2; ++j) {
        uint4 g8[8];
        image(j, xa[j], g8);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int oc = 4 * hh + t, c4 = (64 * j + 8 * oc) / 4;
          const uint32_t so = (uint32_t)((oc ^ r7) << 4);
          float r3[8];
          cb_octet(g8[t], g8[4 + t], r3);
          const float4 wa = __ldg(wo4 + c4), wb = __ldg(wo4 + c4 + 1);
          const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
          float h[8];
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) h[kk] = r3[kk] > 0.0f ? dzs * w[kk] : 0.0f;
          um_put8(xa[j] + row_off + so, CB_TILE_B, h);
        }
        publish(j, xa[j], FULL ? a.dp3_img + ((size_t)tile * 2 + j) * CB_CHUNK_B : nullptr);
        if (elected) CB_TR(1 + s, 2 + j);
      }

      // ---- B1: dH2 -> dP2 = dH2 . aenc . 1[H2 > 0]; daenc = sum_n dH2 . H2 / aenc ----
      wait(bar_d_full(s, 0), kpar, UM_TIMEOUT_ACC);
      if (elected) CB_TR(1 + s, 4);
      xa[0] = begin(2, k, a.h2_img + ((size_t)tile * 2 + 0) * CB_CHUNK_B);
      xa[1] = begin(3, k, a.h2_img + ((size_t)tile * 2 + 1) * CB_CHUNK_B);
      for (int j = 0; j < 2; ++j) {
        um_tc_after();
        uint32_t v[32];
        um_tmem_ld32(lane_addr + (uint32_t)(64 * j + 32 * hh), v);
        uint4 g8[8];
        image(2 + j, xa[j], g8);
        um_tmem_ld_wait();
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int oc = 4 * hh + t, c4 = (64 * j + 8 * oc) / 4;
          const uint32_t so = (uint32_t)((oc ^ r7) << 4);
          float h2[8];
          cb_octet(g8[t], g8[4 + t], h2);
          const float4 fa = __ldg(arow + c4), fb = __ldg(arow + c4 + 1);
          const float f[8] = {fa.x, fa.y, fa.z, fa.w, fb.x, fb.y, fb.z, fb.w};
          float h[8];
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const float g = __uint_as_float(v[8 * t + kk]) * CB_WUNSCALE;   // S dH2
            h[kk] = h2[kk] > 0.0f ? g * f[kk] : 0.0f;
            v[8 * t + kk] = __float_as_uint(g * h2[kk]);   // in place: the summand of d aenc
          }
          if (FULL) um_put8(xa[j] + row_off + so, CB_TILE_B, h);
        }
        if (FULL) publish(2 + j, xa[j], a.dp2_img + ((size_t)tile * 2 + j) * CB_CHUNK_B);
        if (elected) CB_TR(1 + s, 5 + j);
        float p[32];
#pragma unroll
        for (int kk = 0; kk < 32; ++kk) p[kk] = __uint_as_float(v[kk]);
        um_seg_reduce<LOGG>(p, lane);
        if (live) {
          const int c0 = 64 * j + 32 * hh + col_l;
#pragma unroll
          for (int i2 = 0; i2 < CNT; ++i2) {
            const float ae = __ldg(a.aenc + (size_t)b * CB_H + c0 + i2);
            a.daenc[(size_t)b * CB_H + c0 + i2] = ae > 0.0f ? p[i2] * invS / ae : 0.0f;
          }
        }
      }
      if (!FULL) {
        um_tc_before();
        um_bar_sync(grp_bar, 256);   // every thread has read the H2 chunks: the next tile may overwrite the buffers
      }

      // ---- B2: dH1 -> dP1 = dH1 . feats . 1[H1 > 0] (scale S1); dfeats = sum_n dH1 . H1 / feats ----
      if (FULL) {
        if (elected) CB_TR(1 + s, 7);
        wait(bar_d_full(s, 1), kpar, UM_TIMEOUT_ACC);
        if (elected) CB_TR(1 + s, 8);
        xa[0] = begin(4, k, a.h1_img + ((size_t)tile * 4 + 0) * CB_CHUNK_B);
        xa[1] = begin(5, k, a.h1_img + ((size_t)tile * 4 + 1) * CB_CHUNK_B);
        for (int j = 0; j < 4; ++j) {
          um_tc_after();
          const uint32_t xj = xa[j & 1];
          uint32_t v[32];
          um_tmem_ld32(lane_addr + (uint32_t)(64 * j + 32 * hh), v);
          uint4 g8[8];
          image(4 + j, xj, g8);
          um_tmem_ld_wait();
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int oc = 4 * hh + t, c4 = (64 * j + 8 * oc) / 4;
            const uint32_t so = (uint32_t)((oc ^ r7) << 4);
            float h1[8];
            cb_octet(g8[t], g8[4 + t], h1);
            const float4 fa = __ldg(frow + c4), fb = __ldg(frow + c4 + 1);
            const float f[8] = {fa.x, fa.y, fa.z, fa.w, fb.x, fb.y, fb.z, fb.w};
            float h[8];
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
              const float g = __uint_as_float(v[8 * t + kk]) * CB_WUNSCALE;   // S dH1
              h[kk] = h1[kk] > 0.0f ? g * (f[kk] * F1) : 0.0f;
              v[8 * t + kk] = __float_as_uint(g * h1[kk]);   // in place: the summand of d feats
            }
            um_put8(xj + row_off + so, CB_TILE_B, h);
          }
          publish(4 + j, xj, a.dp1_img + ((size_t)tile * 4 + j) * CB_CHUNK_B);
          if (elected) CB_TR(1 + s, 9 + j);
          // the chunk after next goes into this buffer: start its image copy before the per-sample reduction
          if (j < 2) xa[j & 1] = begin(6 + j, k, a.h1_img + ((size_t)tile * 4 + j + 2) * CB_CHUNK_B);
          float p[32];
#pragma unroll
          for (int kk = 0; kk < 32; ++kk) p[kk] = __uint_as_float(v[kk]);
          um_seg_reduce<LOGG>(p, lane);
          if (live) {
            const int c0 = 64 * j + 32 * hh + col_l;
#pragma unroll
            for (int i2 = 0; i2 < CNT; ++i2) {
              const float fe = __ldg(a.feats + (size_t)b * CB_W + c0 + i2);
              a.dfeats[(size_t)b * CB_W + c0 + i2] = fe > 0.0f ? p[i2] * invS / fe : 0.0f;
            }
          }
        }
      }
    }
    if (FULL && elected) um_bulk_wait<0>();
  }

  um_tc_before();
  __syncthreads();
  if (warp == 2) {
    um_tc_after();
    um_tmem_free<512>(tmem);
  }
}

static __device__ int cb_status_word = 0;

template <bool FULL>
static void cb_launch(int logg, int grid, cudaStream_t st, const CriticBwdArgs& a) {
  switch (logg) {
    case 0: critic_backward_kernel<FULL, 0><<<grid, CB_THREADS, CB_SMEM_B, st>>>(a); break;
    case 1: critic_backward_kernel<FULL, 1><<<grid, CB_THREADS, CB_SMEM_B, st>>>(a); break;
    case 2: critic_backward_kernel<FULL, 2><<<grid, CB_THREADS, CB_SMEM_B, st>>>(a); break;
    case 3: critic_backward_kernel<FULL, 3><<<grid, CB_THREADS, CB_SMEM_B, st>>>(a); break;
    case 4: critic_backward_kernel<FULL, 4><<<grid, CB_THREADS, CB_SMEM_B, st>>>(a); break;
    default: critic_backward_kernel<FULL, 5><<<grid, CB_THREADS, CB_SMEM_B, st>>>(a); break;
  }
}
template <bool FULL, int LOGG>
static cudaError_t cb_attr() {
  return cudaFuncSetAttribute(critic_backward_kernel<FULL, LOGG>, cudaFuncAttributeMaxDynamicSharedMemorySize, CB_SMEM_B);
}

}  // namespace asvrl

using namespace asvrl;

extern "C" long asvrl_critic_backward_workspace(void) { return (long)CB_IMAGE_B + CB_FPART * (long)sizeof(float); }

extern "C" int asvrl_critic_backward(const float* dz, const void* r3_img, const void* h2_img, const void* h1_img,
                                     const float* feats, const float* aenc, const float* wh, const float* wh2,
                                     const float* wo, const float* fmax, int batch, int n_tau, int repack, void* workspace,
                                     float* daenc,
                                     float* dfeats, void* dp3_img, void* dp2_img, void* dp1_img, void* aux_img,
                                     float* scales, void* stream) {
  if (!dz || !r3_img || !h2_img || !feats || !aenc || !wh || !wh2 || !wo || !workspace || !daenc)
    return fail(ASVRL_E_BADARG, "asvrl_critic_backward: null pointer");
  const bool full = dfeats || dp3_img || dp2_img || dp1_img || aux_img || scales;
  if (full && !(h1_img && dfeats && dp3_img && dp2_img && dp1_img && aux_img && scales))
    return fail(ASVRL_E_BADARG, "asvrl_critic_backward: h1_img, dfeats, the three dP images, aux and scales go together");
  int logg = -1;
  for (int l = 0; l <= 5; ++l)
    if (n_tau == (1 << l)) logg = l;
  if (batch < 1 || logg < 0 || (long)batch * n_tau > 0x7FFFFFFFL - CB_ROWS)
    return fail(ASVRL_E_BADARG, "asvrl_critic_backward: n_tau must be a power of two <= 32 (B=%d N=%d)", batch, n_tau);
  if (((uintptr_t)workspace | (uintptr_t)r3_img | (uintptr_t)h2_img | (uintptr_t)h1_img | (uintptr_t)dp3_img |
       (uintptr_t)dp2_img | (uintptr_t)dp1_img | (uintptr_t)aux_img) & 127)
    return fail(ASVRL_E_BADARG, "asvrl_critic_backward: workspace and images must be 128-byte aligned");
  if (((uintptr_t)feats | (uintptr_t)aenc | (uintptr_t)wo) & 15)
    return fail(ASVRL_E_BADARG, "asvrl_critic_backward: feats, aenc and wo must be 16-byte aligned");
  static int n_sm = 0;
  cudaError_t e;
  if (n_sm == 0) {
    int dev = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess || (e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess)
      return fail((int)e, "asvrl_critic_backward: %s", cudaGetErrorString(e));
    e = cudaSuccess;
#define CB_ATTR(L) if (e == cudaSuccess) e = cb_attr<true, L>(); if (e == cudaSuccess) e = cb_attr<false, L>();
    CB_ATTR(0) CB_ATTR(1) CB_ATTR(2) CB_ATTR(3) CB_ATTR(4) CB_ATTR(5)
#undef CB_ATTR
    if (e != cudaSuccess) { n_sm = 0; return fail((int)e, "asvrl_critic_backward: %d B shared memory: %s", CB_SMEM_B, cudaGetErrorString(e)); }
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (repack) {
    const int n = CB_W_CHUNKS * CB_H * 8;
    critic_bwd_pack_kernel<<<(n + 255) / 256, 256, 0, st>>>(wh, wh2, (unsigned char*)workspace, fmax);
    if (int rc = check_launch("asvrl_critic_backward (pack)")) return rc;
  }
  CriticBwdArgs a;
  a.dz = dz;
  a.r3_img = (const unsigned char*)r3_img; a.h2_img = (const unsigned char*)h2_img; a.h1_img = (const unsigned char*)h1_img;
  a.feats = feats; a.aenc = aenc; a.wo = wo; a.fmax = fmax;
  a.image = (const unsigned char*)workspace;
  a.rows = batch * n_tau; a.n_tau = n_tau; a.n_tiles = (a.rows + CB_ROWS - 1) / CB_ROWS;
  int* status = nullptr;
  if (cudaGetSymbolAddress((void**)&status, cb_status_word) != cudaSuccess) status = nullptr;
  a.status = status;
  a.daenc = daenc; a.dfeats = dfeats;
  a.dp3_img = (unsigned char*)dp3_img; a.dp2_img = (unsigned char*)dp2_img; a.dp1_img = (unsigned char*)dp1_img;
  a.aux_img = (unsigned char*)aux_img;
  a.scales = scales;
  const int grid = a.n_tiles < n_sm ? a.n_tiles : n_sm;
  if (full) cb_launch<true>(logg, grid, st, a); else cb_launch<false>(logg, grid, st, a);
  return check_launch("asvrl_critic_backward");
}

#ifdef CF_TRACE
extern "C" __attribute__((visibility("default"))) int asvrl_critic_backward_trace(unsigned long long* events, int* counts) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(events, cb_trace_buf, sizeof(unsigned long long) * 4 * 1024);
  cudaMemcpyFromSymbol(counts, cb_trace_n, sizeof(int) * 4);
  int zero[4] = {0, 0, 0, 0};
  cudaMemcpyToSymbol(cb_trace_n, zero, sizeof zero);
  return 0;
}
#endif

extern "C" int asvrl_critic_backward_status(int* status_out) {
  if (!status_out) return fail(ASVRL_E_BADARG, "asvrl_critic_backward_status: null pointer");
  cudaError_t e = cudaMemcpyFromSymbol(status_out, cb_status_word, sizeof(int));
  if (e != cudaSuccess) return fail((int)e, "asvrl_critic_backward_status: %s", cudaGetErrorString(e));
  return 0;
}
