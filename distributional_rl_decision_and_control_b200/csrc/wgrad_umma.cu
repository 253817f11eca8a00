// wgrad_umma.cu — K12: the weight and bias gradients of the IQN critic's quantile trunk on the tcgen05 tensor cores,
// straight from the operand images K10 (forward, SAVE mode) and K11 (backward chain) left in HBM.
//
// Replaces the autograd backward of rfarl/rfarl/policy/AC_IQN_model.py:453-480 for the parameters (reference:
// loss.backward() at rfarl/rfarl/agent.py:413 / 425), i.e. three split-K cuBLAS products with K = B*N, three column
// sums and the head-weight product:
//   dWh2 = dP3^T H2   dbh2 = dP3^T 1   dwo = R3^T dz        (job 3)
//   dWh  = dP2^T H1   dbh  = dP2^T 1                        (job 2)
//   dWc  = dP1^T cos  dbc  = dP1^T 1                        (job 1)
// The contraction runs over the B*N quantile rows, which are the ROWS of every image tile, so both operands are
// read MN-major (the tile's 128-byte rows run along M / N, K advances by rows): the bytes K10 / K11 wrote as
// K-major A operands are consumed unchanged.  The column sums ride along as a 16-wide product with the `aux`
// tile K11 writes per 128 rows (column 0 = S dz, column 1 = 1).  fp16 two-piece operands, three products,
// fp32 accumulation in TMEM for the whole launch: one partial per CTA, summed in a fixed order by
// wgrad_finalize_kernel (deterministic), which also removes K11's power-of-two scales.
//
// Mapping.  The grid is split in three groups of CTAs, one per job; a CTA walks its job's tiles (stride = group
// size), 64 rows (four K steps of 16) per pipeline stage: warp 0 streams the stage's operand pieces with
// cp.async.bulk (14 copies of 8 KB), warp 1 issues the MMAs, warps 4-7 write the accumulators at the end.
#include "umma.cuh"

namespace asvrl {

constexpr int WG_THREADS = 256;
constexpr int WG_CHUNK_B = 32768;                 // image chunk: hi | lo tile of 128 rows x 128 B
constexpr int WG_TILE_B = 16384;
constexpr int WG_PIECE_B = 8192;                  // 64 rows of one fp16 piece
constexpr int WG_SLOT_B = 2 * WG_PIECE_B;         // hi | lo half-tile of one chunk
constexpr int WG_SLOTS = 7;
constexpr int WG_STAGE_B = WG_SLOTS * WG_SLOT_B;  // 112 KB
constexpr int WG_STAGES = 2;
constexpr int WG_OFF_BAR = WG_STAGES * WG_STAGE_B;
constexpr int WG_OFF_MISC = WG_OFF_BAR + 8 * (2 * WG_STAGES + 1);
constexpr int WG_SMEM_B = WG_OFF_MISC + 16 + 1024;
constexpr int WG_PART_LD = 272;                   // floats per accumulator row in the partial buffer
static_assert(WG_SMEM_B <= 232448, "K12 shared memory");

struct WgradArgs {
  const unsigned char *dp3, *dp2, *dp1, *h2, *h1, *cosi, *r3, *aux;   // images [tile][2 | 2 | 4 | 2 | 4 | 1 | 2 | 1][32 KB]
  int n_tiles, n2, n3, n1;                                            // CTAs of job 2, 3, 1 (in this order in the grid)
  float* part;                                                        // [CTA][128][WG_PART_LD]
  int* status;
};

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_kernel(const WgradArgs a) {
  extern __shared__ __align__(1024) unsigned char wg_raw[];
  unsigned char* base = wg_raw + ((1024u - (um_smem(wg_raw) & 1023u)) & 1023u);
  const uint32_t sb = um_smem(base);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t bar0 = sb + WG_OFF_BAR;
  auto bar_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_empty = [&](int s) { return bar0 + 8u * (WG_STAGES + s); };
  const uint32_t bar_done = bar0 + 8u * (2 * WG_STAGES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base + WG_OFF_MISC);
  volatile uint32_t* abort_flag = reinterpret_cast<volatile uint32_t*>(base + WG_OFF_MISC + 4);

  if (tid == 0) {
    for (int s = 0; s < WG_STAGES; ++s) { um_mbar_init(bar_full(s), 1); um_mbar_init(bar_empty(s), 1); }
    um_mbar_init(bar_done, 1);
    *abort_flag = 0u;
    um_mbar_init_fence();
  }
  if (warp == 2) um_tmem_alloc<512>(um_smem(tmem_slot));
  um_tc_before();
  __syncthreads();
  um_tc_after();
  const uint32_t tmem = *tmem_slot;

  const int bid = (int)blockIdx.x;
  int job, li, nj;
  if (bid < a.n2) { job = 2; li = bid; nj = a.n2; }
  else if (bid < a.n2 + a.n3) { job = 3; li = bid - a.n2; nj = a.n3; }
  else { job = 1; li = bid - a.n2 - a.n3; nj = a.n1; }
  const int n_my = li < a.n_tiles ? (a.n_tiles - li + nj - 1) / nj : 0;
  const int n_slots = job == 1 ? 6 : 7;
  auto wait = [&](uint32_t bar, uint32_t parity, int code) { (void)um_mbar_wait(bar, parity, abort_flag, a.status, code); };

  if (warp == 0) {
    if (lane == 0) {
      // slot q of a stage <- chunk ci[q] of image img[q] (cpt[q] chunks per tile)
      const unsigned char* img[WG_SLOTS];
      int cpt[WG_SLOTS], ci[WG_SLOTS];
      auto set = [&](int q, const unsigned char* p, int per_tile, int idx) { img[q] = p; cpt[q] = per_tile; ci[q] = idx; };
      if (job == 2) {
        set(0, a.dp2, 2, 0); set(1, a.dp2, 2, 1);
        for (int c = 0; c < 4; ++c) set(2 + c, a.h1, 4, c);
        set(6, a.aux, 1, 0);
      } else if (job == 3) {
        set(0, a.dp3, 2, 0); set(1, a.dp3, 2, 1); set(2, a.h2, 2, 0); set(3, a.h2, 2, 1);
        set(4, a.r3, 2, 0); set(5, a.r3, 2, 1); set(6, a.aux, 1, 0);
      } else {
        for (int c = 0; c < 4; ++c) set(c, a.dp1, 4, c);
        set(4, a.cosi, 1, 0); set(5, a.aux, 1, 0); set(6, a.aux, 1, 0);
      }
      int it = 0;
      for (int i = 0; i < n_my; ++i) {
        const size_t tile = (size_t)(li + i * nj);
        for (int half = 0; half < 2; ++half, ++it) {
          const int s = it % WG_STAGES;
          if (it >= WG_STAGES) wait(bar_empty(s), (uint32_t)((it / WG_STAGES) - 1) & 1u, UM_TIMEOUT_RING);
          um_mbar_arrive_expect_tx(bar_full(s), (uint32_t)(n_slots * WG_SLOT_B));
          for (int q = 0; q < n_slots; ++q) {
            const unsigned char* src = img[q] + (tile * cpt[q] + ci[q]) * WG_CHUNK_B + half * WG_PIECE_B;
            const uint32_t dst = sb + s * WG_STAGE_B + q * WG_SLOT_B;
            um_bulk_g2s(dst, src, WG_PIECE_B, bar_full(s));                              // hi rows [64 half, +64)
            um_bulk_g2s(dst + WG_PIECE_B, src + WG_TILE_B, WG_PIECE_B, bar_full(s));     // lo
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t i256 = um_idesc_f16_mn(128, 256), i128 = um_idesc_f16_mn(128, 128);
      constexpr uint32_t i64 = um_idesc_f16_mn(128, 64), i16 = um_idesc_f16_mn(128, 16);
      // three products of the two-piece split, small terms first; operand = hi piece address, lo piece + 8 KB
      auto mma3 = [&](uint32_t d, uint32_t pa, uint32_t pb, uint32_t idesc, bool fresh) {
        const uint64_t ah = um_desc_sw128_mn(pa, WG_SLOT_B), al = um_desc_sw128_mn(pa + WG_PIECE_B, WG_SLOT_B);
        const uint64_t bh = um_desc_sw128_mn(pb, WG_SLOT_B), bl = um_desc_sw128_mn(pb + WG_PIECE_B, WG_SLOT_B);
        um_mma_f16(d, al, bh, idesc, fresh ? 0u : 1u);
        um_mma_f16(d, ah, bl, idesc, 1u);
        um_mma_f16(d, ah, bh, idesc, 1u);
      };
      int it = 0;
      for (int i = 0; i < n_my; ++i)
        for (int half = 0; half < 2; ++half, ++it) {
          const int s = it % WG_STAGES;
          wait(bar_full(s), (uint32_t)(it / WG_STAGES) & 1u, UM_TIMEOUT_RING);
          um_tc_after();
          const uint32_t st = sb + s * WG_STAGE_B;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {   // 16 rows = 2048 B per K step
            const bool fresh = it == 0 && ks == 0;
            const uint32_t o = ks * 2048;
            auto slot = [&](int q) { return st + q * WG_SLOT_B + o; };
            if (job == 2) {
              mma3(tmem + 0, slot(0), slot(2), i256, fresh);
              mma3(tmem + 256, slot(0), slot(6), i16, fresh);
            } else if (job == 3) {
              mma3(tmem + 0, slot(0), slot(2), i128, fresh);
              mma3(tmem + 128, slot(0), slot(6), i16, fresh);
              mma3(tmem + 144, slot(4), slot(6), i16, fresh);
            } else {
              mma3(tmem + 0, slot(0), slot(4), i64, fresh);
              mma3(tmem + 64, slot(0), slot(5), i16, fresh);
              mma3(tmem + 80, slot(2), slot(4), i64, fresh);
              mma3(tmem + 144, slot(2), slot(5), i16, fresh);
            }
          }
          um_commit(bar_empty(s));
        }
      um_commit(bar_done);
    }
  } else if (warp >= 4) {
    // accumulators -> this CTA's partial (thread = accumulator row)
    const int q = warp - 4, j = 32 * q + lane;
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16);
    float* dst = a.part + ((size_t)bid * 128 + j) * WG_PART_LD;
    const int ncols = job == 2 ? 272 : 160;
    if (n_my > 0) {
      wait(bar_done, 0u, UM_TIMEOUT_ACC);
      um_tc_after();
    }
    for (int c0 = 0; c0 < ncols; c0 += 16) {
      uint32_t v[16];
      if (n_my > 0) {
        um_tmem_ld16(lane_addr + (uint32_t)c0, v);
        um_tmem_ld_wait();
      } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = 0u;
      }
#pragma unroll
      for (int k = 0; k < 16; k += 4)
        *reinterpret_cast<uint4*>(dst + c0 + k) = make_uint4(v[k], v[k + 1], v[k + 2], v[k + 3]);
    }
  }

  um_tc_before();
  __syncthreads();
  if (warp == 2) {
    um_tc_after();
    um_tmem_free<512>(tmem);
  }
}

// out = scale * sum over the job's CTAs, fixed order.  scales: {1 / S, 1 / S1} (device)
__global__ void __launch_bounds__(256)
wgrad_finalize_kernel(const float* __restrict__ part, int n2, int n3, int n1, const float* __restrict__ scales,
                      float* __restrict__ d_wh, float* __restrict__ d_bh, float* __restrict__ d_wh2,
                      float* __restrict__ d_bh2, float* __restrict__ d_wo, float* __restrict__ d_wc,
                      float* __restrict__ d_bc) {
  const int t = blockIdx.x * 256 + threadIdx.x;
  int c0, nc, j, col, which;
  float* out;
  int e = t;
  if (e < 128 * 256) { c0 = 0; nc = n2; j = e >> 8; col = e & 255; out = d_wh + e; which = 0; }
  else if ((e -= 128 * 256) < 128) { c0 = 0; nc = n2; j = e; col = 257; out = d_bh + e; which = 0; }
  else if ((e -= 128) < 128 * 128) { c0 = n2; nc = n3; j = e >> 7; col = e & 127; out = d_wh2 + e; which = 0; }
  else if ((e -= 128 * 128) < 128) { c0 = n2; nc = n3; j = e; col = 129; out = d_bh2 + e; which = 0; }
  else if ((e -= 128) < 128) { c0 = n2; nc = n3; j = e; col = 144; out = d_wo + e; which = 0; }
  else if ((e -= 128) < 256 * 64) { c0 = n2 + n3; nc = n1; const int k = e >> 6; j = k & 127; col = 80 * (k >> 7) + (e & 63); out = d_wc + e; which = 1; }
  else if ((e -= 256 * 64) < 256) { c0 = n2 + n3; nc = n1; j = e & 127; col = 80 * (e >> 7) + 65; out = d_bc + e; which = 1; }
  else return;
  const float* p = part + ((size_t)c0 * 128 + j) * WG_PART_LD + col;
  float s = 0.0f;
  for (int c = 0; c < nc; ++c) s += p[(size_t)c * 128 * WG_PART_LD];
  *out = s * scales[which];
}

static __device__ int wg_status_word = 0;

}  // namespace asvrl

using namespace asvrl;

// floats of the partial buffer for `n_tiles` 128-row tiles
extern "C" long asvrl_critic_wgrad_workspace(int n_tiles) {
  const int per = n_tiles < 50 ? (n_tiles < 1 ? 1 : n_tiles) : 50;
  return (long)3 * per * 128 * WG_PART_LD;
}

extern "C" int asvrl_critic_wgrad(const void* dp3_img, const void* dp2_img, const void* dp1_img, const void* aux_img,
                                  const void* cos_img, const void* h1_img, const void* h2_img, const void* r3_img,
                                  const float* scales, int batch, int n_tau, float* workspace, float* d_wc, float* d_bc,
                                  float* d_wh, float* d_bh, float* d_wh2, float* d_bh2, float* d_wo, void* stream) {
  if (!dp3_img || !dp2_img || !dp1_img || !aux_img || !cos_img || !h1_img || !h2_img || !r3_img || !scales || !workspace ||
      !d_wc || !d_bc || !d_wh || !d_bh || !d_wh2 || !d_bh2 || !d_wo)
    return fail(ASVRL_E_BADARG, "asvrl_critic_wgrad: null pointer");
  if (batch < 1 || n_tau < 1) return fail(ASVRL_E_BADARG, "asvrl_critic_wgrad: bad sizes (B=%d N=%d)", batch, n_tau);
  if (((uintptr_t)dp3_img | (uintptr_t)dp2_img | (uintptr_t)dp1_img | (uintptr_t)aux_img | (uintptr_t)cos_img |
       (uintptr_t)h1_img | (uintptr_t)h2_img | (uintptr_t)r3_img) & 127)
    return fail(ASVRL_E_BADARG, "asvrl_critic_wgrad: images must be 128-byte aligned");
  static int n_sm = 0;
  cudaError_t e;
  if (n_sm == 0) {
    int dev = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess || (e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess)
      return fail((int)e, "asvrl_critic_wgrad: %s", cudaGetErrorString(e));
    if ((e = cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_B)) != cudaSuccess) {
      n_sm = 0;
      return fail((int)e, "asvrl_critic_wgrad: %d B shared memory: %s", WG_SMEM_B, cudaGetErrorString(e));
    }
  }
  WgradArgs a;
  a.dp3 = (const unsigned char*)dp3_img; a.dp2 = (const unsigned char*)dp2_img; a.dp1 = (const unsigned char*)dp1_img;
  a.h2 = (const unsigned char*)h2_img; a.h1 = (const unsigned char*)h1_img; a.cosi = (const unsigned char*)cos_img;
  a.r3 = (const unsigned char*)r3_img; a.aux = (const unsigned char*)aux_img;
  a.n_tiles = (int)(((long)batch * n_tau + 127) / 128);
  const int per = a.n_tiles < 50 ? a.n_tiles : 50;   // 3 x 50 > 148 SMs: the tail CTAs run as a short second wave
  a.n2 = a.n3 = a.n1 = per;
  if (3 * per > n_sm && n_sm >= 3) { a.n2 = n_sm - 2 * (n_sm / 3); a.n3 = a.n1 = n_sm / 3; }
  a.part = workspace;
  int* status = nullptr;
  if (cudaGetSymbolAddress((void**)&status, wg_status_word) != cudaSuccess) status = nullptr;
  a.status = status;
  cudaStream_t st = (cudaStream_t)stream;
  wgrad_kernel<<<a.n2 + a.n3 + a.n1, WG_THREADS, WG_SMEM_B, st>>>(a);
  if (int rc = check_launch("asvrl_critic_wgrad")) return rc;
  const int n_out = 128 * 256 + 128 + 128 * 128 + 128 + 128 + 256 * 64 + 256;
  wgrad_finalize_kernel<<<(n_out + 255) / 256, 256, 0, st>>>(workspace, a.n2, a.n3, a.n1, scales, d_wh, d_bh, d_wh2, d_bh2,
                                                            d_wo, d_wc, d_bc);
  return check_launch("asvrl_critic_wgrad (finalize)");
}

extern "C" int asvrl_critic_wgrad_status(int* status_out) {
  if (!status_out) return fail(ASVRL_E_BADARG, "asvrl_critic_wgrad_status: null pointer");
  cudaError_t e = cudaMemcpyFromSymbol(status_out, wg_status_word, sizeof(int));
  if (e != cudaSuccess) return fail((int)e, "asvrl_critic_wgrad_status: %s", cudaGetErrorString(e));
  return 0;
}
