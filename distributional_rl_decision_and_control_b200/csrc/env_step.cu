// env_step.cu — K1: one batched MarineNavEnv3.step for thousands of environments (sm_100a).
//
// Replaces, for every env of the batch in ONE launch (reference paths under rfarl/rfarl/):
//   envs/marinenav/env.py:240-333        MarineNavEnv3.step
//   envs/marinenav/env.py:458-501        get_velocity / compute_speed (vortex current)
//   envs/marinenav/vehicles/wamv.py:204-279  Robot.update_state / compute_motion
//   envs/marinenav/vehicles/wamv.py:436-529  Robot.perception_output (+ 281-322 helpers)
//   envs/marinenav/vehicles/wamv.py:324-423  COLREGs zone tests and turn angle
//   policy/trainer.py:168-170            deactivation, when ASVRL_M_AUTO_DEACTIVATE
//
// Mapping.  A 128-thread CTA owns EPB consecutive environments (EPB*RP <= 32 ASVs):
//   phase 0  the CTA stages the envs' buoy / vortex tables and pre-step poses in shared memory;
//   phase A  warp 0, one lane per ASV, integrates the 3-DOF model for the 10 sub-steps (pose in
//            fp64 registers so the progress reward does not cancel); SoA loads and
//            stores are coalesced over the robot-fastest layout;
//   phase A' warps 1-3 CONCURRENTLY draw the Philox sensor noise (state independent) for the
//            pairs that can end up in range, hiding the serial integration chain;
//   phase B  the (ego, candidate) pairs are flattened over all threads: detection / collision
//            predicates in fp64, body-frame observation and sort key into shared memory;
//   phase B2 every detected pair computes its stable rank among its ego's keys (top-5);
//   phase C  COLREGs per (ego, kept object) in parallel, then reward / done / info per ego;
//   phase D  the CTA writes the padded observation rows back with coalesced stores.
// The kernel is instruction-issue bound (about 1e4 thread instructions per 280 algorithmic
// bytes); HBM traffic is one pass over state in, state + observations out.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "env_consts.cuh"
#include "env_reset.cuh"

namespace asvrl {

struct StatePtrs {
  float *x, *y, *theta, *velr_x, *velr_y, *velr_r, *vel_x, *vel_y, *vel_r, *thrust_l, *thrust_r;
  const float *goal_x, *goal_y;
  uint32_t* flags;
  const int32_t *n_robots, *n_obs, *n_cores;
  int32_t* ep_steps;
  const float4* obstacles;
  const float4* cores;
};

struct OutPtrs {
  float *obs_self, *obs_objects, *obs_mask, *reward;
  uint32_t* status;
  float* phi;
  int32_t* top_ids;
  unsigned long long* detected;
  uint32_t* env_status;
  uint8_t* ended;
};

constexpr int TPB = 128;  // 4 warps per CTA; a CTA owns <= 32 ASVs (EPB envs)

// extra arguments of the fused step (asvrl_env_step_autoreset): K1, then in the same launch the
// CTA re-generates its envs whose episode just ended (K1b's warp routine) and re-observes them
struct FusedArgs {
  ResetConsts rk;
  const int32_t *want_robots, *want_cores, *want_obs;
  const float* min_sg;
  uint64_t observe_seed;
  float *term_self, *term_objects, *term_mask;  // terminal observation rows of ended envs (nullable)
  int32_t* episode_idx;  // [E] episodes generated so far per env: the next scenario's RNG counter is idx + 1
  // pre-generated next episodes (nullable): stage state + stage_tag[e] = episode number it holds
  ResetState stage;
  int32_t* stage_tag;
  const float *stage_self, *stage_objects, *stage_mask;  // staged FIRST observation rows (nullable)
  int n_workers;         // leading CTAs of the grid that only refill the stage
  // any variant: nullable per-env RNG step of the sensor noise (replaces `step` for that env)
  const int32_t* step_arr;
};

__device__ __forceinline__ int ld_acquire(const int32_t* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int32_t* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Refill worker (leading CTAs of the fused grid): every warp scans its slice of the envs for a
// stage that does not hold the env's NEXT episode (tag != episode_idx + 1) and re-generates ONE of
// them with K1b's warp routine, off the critical path of the step CTAs that run beside it.  The
// scenario is a pure function of (seed, env, episode number), so it does not matter who generates
// it or when: a step CTA that finds no valid stage generates the same episode inline.
__device__ __forceinline__ void refill_worker(const FusedArgs& fa, char* smem, uint64_t seed, int E, int RP,
                                              int OP, int CP, int64_t env_offset) {
  if (!fa.stage_tag) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_warps = fa.n_workers * (TPB / 32), gw = blockIdx.x * (TPB / 32) + warp;
  const int chunk = (E + n_warps - 1) / n_warps;
  const int lo = gw * chunk, hi = min(E, lo + chunk);
  float4* sg = (float4*)smem + warp * (RP + CP + OP);
  float4* cores = sg + RP;
  float4* obs = cores + CP;
  for (int base = lo; base < hi; base += 32) {
    const int e = base + lane;
    const bool need = e < hi && ld_acquire(fa.stage_tag + e) != __ldcg(fa.episode_idx + e) + 1;
    const unsigned bal = __ballot_sync(0xffffffffu, need);
    if (!bal) continue;
    const int pick = base + __ffs(bal) - 1;
    const int ep = __ldcg(fa.episode_idx + pick) + 1;
    reset_env_warp(fa.rk, fa.stage, pick, env_offset + pick, fa.want_robots[pick], fa.want_cores[pick],
                   fa.want_obs[pick], fa.min_sg[pick], seed, (uint64_t)ep, RP, OP, CP, sg, cores, obs, lane);
    __threadfence();
    __syncwarp();
    if (lane == 0) st_release(fa.stage_tag + pick, ep);
    break;  // one env per warp per launch keeps the worker shorter than the step it hides behind
  }
}


// shared-memory layout: byte offsets computed once on the host and passed by value, so the
// kernel prologue is a handful of 32-bit adds instead of per-thread 64-bit size arithmetic
struct SmemLayout {
  uint32_t px, py, key, obst, cores, c, s, vx, vy, vz, Tl, Tr, gx, gy, rew, ox, oy, osp, cand,
      st_obj, st_self, st_mask, cphi, fl0, fl, det_lo, det_hi, plist, dlist, cnt, el_of, r_of, chit,
      nr, no, nc, ep, ok, ended, epn, pre_ep, pre_tag, rs_sg, rs_core, rs_obs, rs_cand, rs_okm, crx, cry;
  uint32_t bytes;
  // pair-loop stepping without integer division: TPB = q*S + rem
  int step_a, step_sl, step3_a, step3_sl;
};

struct Smem {
  double *px, *py, *key;
  float *crx, *cry;  // [CP][A] core - pose offsets at the start of the step (phase A only)
  float *c, *s, *vx, *vy, *vz, *Tl, *Tr, *gx, *gy, *rew, *ox, *oy, *osp;
  float *cand, *st_obj, *st_self, *st_mask, *cphi;
  uint32_t *fl0, *fl, *det_lo, *det_hi;
  uint16_t *plist, *dlist;  // compacted (ego << 6 | slot): pairs worth a noise draw / detected pairs
  int* cnt;                 // [0] entries of plist, [1] entries of dlist, [2] noise work ticket
  uint8_t *el_of, *r_of, *chit;
  int *nr, *no, *nc, *ep, *ok, *ended, *epn;
  int *pre_ep, *pre_tag;      // fused tail: next episode number / stage readiness, read in phase 0
  float4 *obst, *cores;
  float4 *rs_sg, *rs_core, *rs_obs, *rs_cand;  // fused auto-reset scratch: accepted sets, round candidates
  unsigned* rs_okm;
};

// constexpr: the shape-specialised kernel instantiation evaluates it at compile time, so every
// shared-memory offset becomes an immediate
static constexpr SmemLayout make_layout(int EPB, int RP, int OP, int CP, int tpb, bool fused = false) {
  SmemLayout L{};
  const uint32_t A = (uint32_t)EPB * RP, S = OP + RP, SK = S | 1;
  uint32_t off = 0;
  auto take = [&](uint32_t n, uint32_t elem) { uint32_t o = off; off += (n * elem + 15u) & ~15u; return o; };
  L.px = take(A, 8); L.py = take(A, 8); L.key = take(A * SK, 8);
  L.obst = take(EPB * (OP > 0 ? OP : 1), 16); L.cores = take(EPB * (CP > 0 ? CP : 1), 16);
  L.c = take(A, 4); L.s = take(A, 4); L.vx = take(A, 4); L.vy = take(A, 4); L.vz = take(A, 4);
  L.Tl = take(A, 4); L.Tr = take(A, 4); L.gx = take(A, 4); L.gy = take(A, 4); L.rew = take(A, 4);
  L.ox = take(A, 4); L.oy = take(A, 4); L.osp = take(A, 4);
  L.cand = take(A * S * 5, 4);  // prefetched noise first, then the candidate observation (same thread)
  L.st_obj = take(A * 25, 4); L.st_self = take(A * 7, 4); L.st_mask = take(A * 5, 4);
  L.cphi = take(A * 5, 4);
  L.fl0 = take(A, 4); L.fl = take(A, 4); L.det_lo = take(A, 4); L.det_hi = take(A, 4);
  L.plist = take(A * S, 2); L.dlist = take(A * S, 2); L.cnt = take(4, 4);
  L.el_of = take(A, 1); L.r_of = take(A, 1); L.chit = take(A * 5, 1);
  L.nr = take(EPB, 4); L.no = take(EPB, 4); L.nc = take(EPB, 4); L.ep = take(EPB, 4); L.ok = take(EPB, 4);
  L.ended = take(EPB, 4); L.epn = take(EPB, 4);
  L.pre_ep = take(EPB, 4); L.pre_tag = take(EPB, 4);
  const uint32_t f = fused ? 1u : 0u;
  L.rs_sg = take(f * RP, 16); L.rs_core = take(f * CP, 16); L.rs_obs = take(f * OP, 16);
  L.rs_cand = take(f * tpb, 16); L.rs_okm = take(f * 8, 4);
  // phase A's offset tables are dead before phase B writes the keys: alias them when they fit
  if ((uint32_t)CP * A * 8 <= A * SK * 8) { L.crx = L.key; L.cry = L.key + (uint32_t)CP * A * 4; }
  else { L.crx = take((uint32_t)CP * A, 4); L.cry = take((uint32_t)CP * A, 4); }
  // a refill worker CTA carves per-warp accepted sets from the base of the same allocation
  const uint32_t worker_need = fused ? (uint32_t)(tpb / 32) * (RP + CP + OP) * 16u : 0u;
  L.bytes = off > worker_need ? off : worker_need;
  L.step_a = tpb / (int)S; L.step_sl = tpb % (int)S;
  L.step3_a = (tpb - 32) / (int)S; L.step3_sl = (tpb - 32) % (int)S;
  return L;
}

__device__ __forceinline__ Smem carve(char* base, const SmemLayout& L) {
  Smem m;
  m.px = (double*)(base + L.px); m.py = (double*)(base + L.py); m.key = (double*)(base + L.key);
  m.crx = (float*)(base + L.crx); m.cry = (float*)(base + L.cry);
  m.obst = (float4*)(base + L.obst); m.cores = (float4*)(base + L.cores);
  m.c = (float*)(base + L.c); m.s = (float*)(base + L.s); m.vx = (float*)(base + L.vx);
  m.vy = (float*)(base + L.vy); m.vz = (float*)(base + L.vz); m.Tl = (float*)(base + L.Tl);
  m.Tr = (float*)(base + L.Tr); m.gx = (float*)(base + L.gx); m.gy = (float*)(base + L.gy);
  m.rew = (float*)(base + L.rew); m.ox = (float*)(base + L.ox); m.oy = (float*)(base + L.oy);
  m.osp = (float*)(base + L.osp); m.cand = (float*)(base + L.cand);
  m.st_obj = (float*)(base + L.st_obj); m.st_self = (float*)(base + L.st_self);
  m.st_mask = (float*)(base + L.st_mask); m.cphi = (float*)(base + L.cphi);
  m.fl0 = (uint32_t*)(base + L.fl0); m.fl = (uint32_t*)(base + L.fl);
  m.det_lo = (uint32_t*)(base + L.det_lo); m.det_hi = (uint32_t*)(base + L.det_hi);
  m.plist = (uint16_t*)(base + L.plist); m.dlist = (uint16_t*)(base + L.dlist);
  m.cnt = (int*)(base + L.cnt);
  m.el_of = (uint8_t*)(base + L.el_of); m.r_of = (uint8_t*)(base + L.r_of);
  m.chit = (uint8_t*)(base + L.chit);
  m.nr = (int*)(base + L.nr); m.no = (int*)(base + L.no); m.nc = (int*)(base + L.nc);
  m.ep = (int*)(base + L.ep); m.ok = (int*)(base + L.ok); m.ended = (int*)(base + L.ended); m.epn = (int*)(base + L.epn);
  m.pre_ep = (int*)(base + L.pre_ep); m.pre_tag = (int*)(base + L.pre_tag);
  m.rs_sg = (float4*)(base + L.rs_sg); m.rs_core = (float4*)(base + L.rs_core);
  m.rs_obs = (float4*)(base + L.rs_obs); m.rs_cand = (float4*)(base + L.rs_cand);
  m.rs_okm = (unsigned*)(base + L.rs_okm);
  return m;
}

// pair index p -> (ego a, slot sl) walked without division
struct PairIt {
  int p, a, sl;
  __device__ __forceinline__ PairIt(int p0, int S) : p(p0), a(p0 / S), sl(p0 - (p0 / S) * S) {}
  __device__ __forceinline__ void advance(int dp, int da, int dsl, int S) {
    p += dp; a += da; sl += dsl;
    if (sl >= S) { sl -= S; ++a; }
  }
};

// ENV:458-501.  All cores contribute (the reference's `continue` is a no-op).
// tangent * speed = (-ry, rx) * dir * speed / d, and speed / d is Gamma/(2 pi r^2) inside the
// core (solid body) and Gamma/(2 pi d^2) outside: no square root, one reciprocal per core.
// The offsets core - p are formed in fp64 from the fp64 pose: the 1/d^2 field amplifies the fp32
// rounding of a ~50 m coordinate by 1/d near a core.
// Inside the 10 sub-steps the fp64 subtraction is hoisted: (core - pose0) is rounded to fp32 once
// per (ASV, core) (table rx0/ry0, stride A) and the displacement since the step started, (dx, dy) =
// fl32(pose - pose0) <= 1.5 m, is subtracted in fp32; the field is singular only for d < core_r
// where it is replaced by the solid-body law, so the error on f stays below 1e-6 relative.
__device__ __forceinline__ void current_at(const EnvConsts& k, const float4* cores, int nc,
                                           const float* rx0, const float* ry0, int A, float dx, float dy,
                                           float& cx, float& cy) {
  float ax = 0.0f, ay = 0.0f;
  const float r2 = k.core_r * k.core_r;
#pragma unroll 4
  for (int i = 0; i < nc; ++i) {
    const float4 co = cores[i];
    const float rx = rx0[i * A] - dx, ry = ry0[i * A] - dy;
    const float d2 = rx * rx + ry * ry;
    const float g = co.w * co.z;  // signed circulation: +clockwise, -counter-clockwise (ENV:483-487)
    // ENV:497-501.  d2 > core_r^2 on this branch, so the reciprocal needs none of __fdividef's range guards
    // (six predicated instructions per core on the integration's serial chain); same bits for normal operands
    float rcp;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcp) : "f"(d2));
    const float f = d2 <= r2 ? g * k.solid_k : (g * k.inv_2pi) * rcp;
    ax -= ry * f;
    ay += rx * f;
  }
  cx = ax + k.drift_vx;
  cy = ay + k.drift_vy;
}

__device__ __forceinline__ float wrap_to_pi(float a) {  // WAMV:425-434
  const float PI = 3.14159265358979f;
  while (a < -PI) a += 2.0f * PI;
  while (a >= PI) a -= 2.0f * PI;
  return a;
}

// WAMV:398-423 for one kept object o = [px,py,vx,vy,r] (ego frame); w = ego velocity in ego frame
__device__ __forceinline__ bool colregs_one(const EnvConsts& k, const float* o, float w0, float w1,
                                            float& phi) {
  const float so2 = o[2] * o[2] + o[3] * o[3];
  if (sqrtf(so2) < k.min_speed) return false;                    // WAMV:400-402
  const float se = sqrtf(w0 * w0 + w1 * w1);
  if (se < k.min_speed) return false;                            // WAMV:404-407
  const float inv = rsqrtf(so2);
  const float ca = o[2] * inv, sa = o[3] * inv;                  // R(atan2(vy,vx)), WAMV:328-330
  const float epx = -(ca * o[0] + sa * o[1]), epy = -(-sa * o[0] + ca * o[1]);  // WAMV:334
  const float evx = ca * w0 + sa * w1, evy = -sa * w0 + ca * w1;                // WAMV:339
  const float ang = atan2f(evy, evx);
  const float PI = 3.14159265358979f;
  const bool tri = (epy - k.lc_y1) > k.lc_grad * (epx - k.lc_x1);  // WAMV:348-351
  const bool left = epx >= k.lc_x0 && epx <= k.lc_x1 && epy >= k.lc_y0 && epy <= 0.0f && !tri &&
                    ang >= 0.25f * PI && ang <= 0.75f * PI;        // WAMV:345-357
  const bool head = epx >= 0.0f && epx <= k.head_on_x && epy >= -k.head_on_half_y &&
                    epy <= k.head_on_half_y && fabsf(ang) > 0.75f * PI;  // WAMV:365-372
  if (!(left || head)) return false;
  // WAMV:379-396; NaN (dist < base_1) propagates and fails phi > 0 like the reference
  const float dist = sqrtf(o[0] * o[0] + o[1] * o[1]);
  const float base_1 = o[4] + 1.0f;
  const float add_1 = asinf(base_1 / dist);
  const float add_2 = atan2f(k.robot_r, sqrtf(dist * dist - base_1 * base_1));
  phi = wrap_to_pi(wrap_to_pi(atan2f(o[1], o[0]) + add_1 + add_2) - atan2f(w1, w0));
  return phi > 0.0f;                                             // WAMV:421
}

// candidate slot -> (live, true position, radius, stale velocity) from the shared tables
struct Cand {
  bool live;
  double px, py, rt;
  float vx, vy;
};
__device__ __forceinline__ Cand candidate(const EnvConsts& k, const Smem& m, int el, int r, int sl,
                                          int RP, int OP) {
  Cand c;
  c.vx = 0.0f; c.vy = 0.0f; c.px = 0.0; c.py = 0.0; c.rt = 0.0;
  if (sl < OP) {  // static obstacles, WAMV:465-483
    c.live = sl < m.no[el];
    const float4 ob = m.obst[el * OP + sl];
    c.px = ob.x; c.py = ob.y; c.rt = ob.z;
  } else {  // other robots still on the map, WAMV:486-510
    const int j = sl - OP, b = el * RP + j;
    c.live = j != r && !(m.fl0[b] & ASVRL_F_DEACTIVATED);
    c.px = m.px[b]; c.py = m.py[b]; c.vx = m.vx[b]; c.vy = m.vy[b]; c.rt = k.robot_r_d;
  }
  return c;
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_l2_span(const void* p, int bytes) {
  const char* c = (const char*)p;
  for (int o = 0; o < bytes; o += 128) prefetch_l2(c + o);
  prefetch_l2(c + bytes - 1);
}
// everything take_stage_batched will read for env e (one thread issues ~25 non-blocking prefetches)
__device__ __forceinline__ void prefetch_stage(const FusedArgs& fa, int e, int RP, int OP, int CP) {
  const ResetState& sg = fa.stage;
  const size_t g = (size_t)e * RP;
  prefetch_l2_span(sg.x + g, RP * 4); prefetch_l2_span(sg.y + g, RP * 4); prefetch_l2_span(sg.theta + g, RP * 4);
  prefetch_l2_span(sg.goal_x + g, RP * 4); prefetch_l2_span(sg.goal_y + g, RP * 4);
  if (OP > 0) prefetch_l2_span(sg.obstacles + (size_t)e * OP, OP * 16);
  if (CP > 0) prefetch_l2_span(sg.cores + (size_t)e * CP, CP * 16);
  prefetch_l2(sg.n_robots + e); prefetch_l2(sg.n_obs + e); prefetch_l2(sg.n_cores + e);
  if (fa.stage_self) {
    prefetch_l2_span(fa.stage_self + g * 7, RP * 28);
    prefetch_l2_span(fa.stage_objects + g * 25, RP * 100);
    prefetch_l2_span(fa.stage_mask + g * 5, RP * 20);
  }
}


// The fused tail's copy of a staged episode into the live state (what take_stage does) and of its
// staged first observation rows into `out`, with all loads issued before the first store.
__device__ __forceinline__ void take_stage_batched(const ResetState& st, const FusedArgs& fa, const OutPtrs& out,
                                                   int e, int RP, int OP, int CP, int tid) {
  static_assert(RS_MAXR <= TPB && RS_MAXO <= TPB && RS_MAXC <= TPB, "one table row per thread");
  constexpr int IT_O = (RS_MAXR * 25 + TPB - 1) / TPB, IT_S = (RS_MAXR * 7 + TPB - 1) / TPB,
                IT_M = (RS_MAXR * 5 + TPB - 1) / TPB;
  const ResetState& sg = fa.stage;
  const size_t g = (size_t)e * RP + tid, b0 = (size_t)e * RP;
  float rx = 0, ry = 0, rth = 0, rgx = 0, rgy = 0;
  float4 ob = make_float4(0, 0, 0, 0), co = make_float4(0, 0, 0, 0);
  int cnt = 0;
  float fo[IT_O], fs[IT_S], fm[IT_M];
  if (tid < RP) {
    rx = __ldcg(sg.x + g); ry = __ldcg(sg.y + g); rth = __ldcg(sg.theta + g);
    rgx = __ldcg(sg.goal_x + g); rgy = __ldcg(sg.goal_y + g);
  }
  if (tid < OP) ob = __ldcg(sg.obstacles + (size_t)e * OP + tid);
  if (tid < CP) co = __ldcg(sg.cores + (size_t)e * CP + tid);
  if (tid < 3) cnt = __ldcg((tid == 0 ? sg.n_robots : tid == 1 ? sg.n_obs : sg.n_cores) + e);
  if (fa.stage_self) {
#pragma unroll
    for (int q = 0; q < IT_O; ++q) { const int i = tid + q * TPB; fo[q] = i < RP * 25 ? __ldcg(fa.stage_objects + b0 * 25 + i) : 0.0f; }
#pragma unroll
    for (int q = 0; q < IT_S; ++q) { const int i = tid + q * TPB; fs[q] = i < RP * 7 ? __ldcg(fa.stage_self + b0 * 7 + i) : 0.0f; }
#pragma unroll
    for (int q = 0; q < IT_M; ++q) { const int i = tid + q * TPB; fm[q] = i < RP * 5 ? __ldcg(fa.stage_mask + b0 * 5 + i) : 0.0f; }
  }
  if (tid < RP) {
    st.x[g] = rx; st.y[g] = ry; st.theta[g] = rth; st.goal_x[g] = rgx; st.goal_y[g] = rgy;
    st.velr_x[g] = 0; st.velr_y[g] = 0; st.velr_r[g] = 0;
    st.vel_x[g] = 0; st.vel_y[g] = 0; st.vel_r[g] = 0;
    st.thrust_l[g] = 0; st.thrust_r[g] = 0;
    st.flags[g] = 0;
  }
  if (tid < OP) st.obstacles[(size_t)e * OP + tid] = ob;
  if (tid < CP) st.cores[(size_t)e * CP + tid] = co;
  if (tid < 3) (tid == 0 ? st.n_robots : tid == 1 ? st.n_obs : st.n_cores)[e] = cnt;
  if (tid == 3) st.ep_steps[e] = 0;
  if (fa.stage_self) {
#pragma unroll
    for (int q = 0; q < IT_O; ++q) { const int i = tid + q * TPB; if (i < RP * 25) out.obs_objects[b0 * 25 + i] = fo[q]; }
#pragma unroll
    for (int q = 0; q < IT_S; ++q) { const int i = tid + q * TPB; if (i < RP * 7) out.obs_self[b0 * 7 + i] = fs[q]; }
#pragma unroll
    for (int q = 0; q < IT_M; ++q) { const int i = tid + q * TPB; if (i < RP * 5) out.obs_mask[b0 * 5 + i] = fm[q]; }
  }
}


// SHAPE = 1 is the instantiation for BASELINE config 2 (8 ASVs x 10 buoys x 8 vortex cores, 4 envs per
// CTA): the table sizes and the shared-memory layout are compile-time constants, which folds the
// per-use `base + offset` arithmetic of the 45 shared arrays (7.6 % of the generic kernel's
// instructions, profiles/r01_k1_v12_source.txt), the index divisions and the short loops' bounds.
// SHAPE = 0 takes every size at run time.  Same code, same results.
// SHAPE = 1: 8 vortex cores; SHAPE = 2: none (the reference's shipped curricula have num_cores = 0).
constexpr int C2_RP = 8, C2_OP = 10, C2_EPB = 4;
constexpr int shape_cores(int shape) { return shape == 2 ? 0 : 8; }

// 0 = generic instantiation; ASVRL_K1_GENERIC=1 in the environment forces it (A/B timing, tests).
static int pick_shape(int max_robots, int max_obs, int max_cores, int EPB) {
  static const bool generic = [] { const char* v = getenv("ASVRL_K1_GENERIC"); return v && v[0] == '1'; }();
  if (generic || max_robots != C2_RP || max_obs != C2_OP || EPB != C2_EPB) return 0;
  return max_cores == shape_cores(1) ? 1 : (max_cores == shape_cores(2) ? 2 : 0);
}

// -DK1_TRACE (debug builds, profiles/probes/k1_trace.py): SM clock of the first step CTA at the phase boundaries
#ifdef K1_TRACE
__device__ unsigned long long k1_trace_buf[32];
#define K1_TR(slot, thread) do { if ((int)blockIdx.x == (FUSED ? fa.n_workers : 0) && threadIdx.x == (thread) && !second) \
    k1_trace_buf[slot] = clock64(); } while (0)
#else
#define K1_TR(slot, thread) do { } while (0)
#endif

template <bool INJECT, bool FUSED, int SHAPE = 0>
__global__ void __launch_bounds__(TPB, 7)
env_step_kernel(const __grid_constant__ EnvConsts k, const __grid_constant__ SmemLayout L_,
                StatePtrs st, const void* __restrict__ actions,
                const float* __restrict__ noise, const uint8_t* __restrict__ env_mask, uint64_t seed,
                uint64_t step, OutPtrs out, int E, int RP_, int OP_, int CP_, int64_t env_offset,
                uint32_t mode, int EPB_, const uint64_t* __restrict__ counter_dev,
                const __grid_constant__ FusedArgs fa) {
  extern __shared__ __align__(16) char smem_raw[];
  constexpr bool SP = SHAPE != 0;
  constexpr int C2_CP = shape_cores(SHAPE);
  const int RP = SP ? C2_RP : RP_, OP = SP ? C2_OP : OP_, CP = SP ? C2_CP : CP_, EPB = SP ? C2_EPB : EPB_;
  if (FUSED && (int)blockIdx.x < fa.n_workers) {
    refill_worker(fa, smem_raw, seed, E, RP, OP, CP, env_offset);
    return;
  }
  constexpr SmemLayout LS = make_layout(C2_EPB, C2_RP, C2_OP, C2_CP, TPB, FUSED);
  const Smem m = SP ? carve(smem_raw, LS) : carve(smem_raw, L_);
  const int step3_a = SP ? LS.step3_a : L_.step3_a, step3_sl = SP ? LS.step3_sl : L_.step3_sl;
  const int tid = threadIdx.x;
  const int A = EPB * RP, S = OP + RP, SK = S | 1, NP = A * S;
  const int e0 = ((int)blockIdx.x - (FUSED ? fa.n_workers : 0)) * EPB;
  bool dyn = mode & ASVRL_M_DYNAMICS;
  if (counter_dev) step += *counter_dev;  // graph-replay friendly RNG counter (include/asvrl.h)
  // FUSED: pass 0 is the step itself; pass 1 (only in CTAs that own an env whose episode just
  // ended) re-observes the re-generated envs: no dynamics, its own noise stream, observation
  // rows only.  `second` is CTA-uniform.
  bool second = false;
  for (;;) {

  // ---- phase 0: per-env tables and the pre-step poses (for the noise prefetch filter).  Every
  // global load of the prologue — counts, buoy / vortex tables, the state row, the fused tail's
  // episode number and stage tag — is issued before the first dependent use, so a CTA pays ONE
  // DRAM round trip with cold data instead of one per dependent stage. ----
  K1_TR(0, 0);
  int c_nr = 0, c_no = 0, c_nc = 0, c_ep = 0, c_epn = 0, c_epi = 0, c_tag = -1;
  bool c_v = false;
  if (tid < EPB) {
    const int e = e0 + tid;
    c_v = second ? m.ended[tid] != 0 : (e < E && (!env_mask || env_mask[e]));
    if (c_v) {
      c_nr = st.n_robots[e]; c_no = st.n_obs[e]; c_nc = st.n_cores[e]; c_ep = st.ep_steps[e];
      if (!second && fa.step_arr) c_epn = fa.step_arr[e];
    }
  }
  const int n_tab_o = min(EPB, E - e0) * OP, n_tab_c = min(EPB, E - e0) * CP;  // tables are contiguous over the CTA's envs
  float4 t_ob = make_float4(0, 0, 0, 0), t_co = make_float4(0, 0, 0, 0);
  if (tid < n_tab_o) t_ob = st.obstacles[(size_t)e0 * OP + tid];
  if (tid < n_tab_c) t_co = st.cores[(size_t)e0 * CP + tid];
  // one lane per ASV: the whole state row, one DRAM round trip ahead of the integration that consumes it
  float p_x = 0, p_y = 0, p_th = 0, vr0 = 0, vr1 = 0, vr2 = 0, v0 = 0, v1 = 0, v2 = 0, Tl = 0, Tr = 0,
        gx = 0, gy = 0;
  float2 act = make_float2(0.0f, 0.0f);
  int act_id = 0, my_nr = 0;
  uint32_t f_in = (uint32_t)ASVRL_F_DEACTIVATED;  // absent == not on the map
  bool my_v = false;
  if (tid < A) {
    const int el = tid / RP, r = tid - el * RP;
    const int e = e0 + el;
    my_v = second ? m.ended[el] != 0 : (e < E && (!env_mask || env_mask[e]));
    if (my_v) {
      const size_t g = (size_t)e * RP + r;
      my_nr = st.n_robots[e];
      p_x = st.x[g]; p_y = st.y[g]; p_th = st.theta[g];
      vr0 = st.velr_x[g]; vr1 = st.velr_y[g]; vr2 = st.velr_r[g];
      v0 = st.vel_x[g]; v1 = st.vel_y[g]; v2 = st.vel_r[g];
      Tl = st.thrust_l[g]; Tr = st.thrust_r[g];
      gx = st.goal_x[g]; gy = st.goal_y[g];
      f_in = st.flags[g];
      if (dyn) {
        if (mode & ASVRL_M_DISCRETE) act_id = ((const int32_t*)actions)[g];
        else act = ((const float2*)actions)[g];
      }
    }
  }
  if (FUSED && !second && tid < EPB && e0 + tid < E) {  // last of the batch: an acquire orders later loads behind it
    c_epi = fa.episode_idx[e0 + tid];
    if (fa.stage_tag) c_tag = ld_acquire(fa.stage_tag + e0 + tid);
  }
  // ---- first uses ----
  if (tid < EPB) {
    m.ok[tid] = c_v;
    m.nr[tid] = min(c_nr, RP); m.no[tid] = min(c_no, OP); m.nc[tid] = min(c_nc, CP);
    m.ep[tid] = c_ep;
    if (!second && fa.step_arr) m.epn[tid] = c_epn;
    if (FUSED && !second) {
      const bool ready = fa.stage_tag && c_tag == c_epi + 1;
      m.pre_ep[tid] = c_epi + 1;
      m.pre_tag[tid] = ready;
      // an env whose clock runs out in this step is certain to end: pull its staged episode into L2
      // now, 20 us before the tail copies it
      if (ready && c_v && dyn && c_ep >= k.max_ep) prefetch_stage(fa, e0 + tid, RP, OP, CP);
    }
  }
  if (tid < EPB * OP) m.obst[tid] = t_ob;
  if (tid < EPB * CP) m.cores[tid] = t_co;
  for (int i = tid + TPB; i < EPB * OP; i += TPB)  // shapes with more than 128 table rows per CTA
    m.obst[i] = i < n_tab_o ? st.obstacles[(size_t)e0 * OP + i] : make_float4(0, 0, 0, 0);
  for (int i = tid + TPB; i < EPB * CP; i += TPB)
    m.cores[i] = i < n_tab_c ? st.cores[(size_t)e0 * CP + i] : make_float4(0, 0, 0, 0);
  if (tid < A) {
    const int el = tid / RP, r = tid - el * RP;
    if (!(my_v && r < min(my_nr, RP))) {  // absent robot (or masked env): zero state, not on the map
      p_x = p_y = p_th = vr0 = vr1 = vr2 = v0 = v1 = v2 = Tl = Tr = gx = gy = 0.0f;
      act = make_float2(0.0f, 0.0f); act_id = 0;
      f_in = (uint32_t)ASVRL_F_DEACTIVATED;
    }
    m.ox[tid] = p_x;
    m.oy[tid] = p_y;
    m.osp[tid] = sqrtf(v0 * v0 + v1 * v1);
    m.fl0[tid] = f_in;
    m.el_of[tid] = (uint8_t)el;
    m.r_of[tid] = (uint8_t)r;
    m.det_lo[tid] = 0; m.det_hi[tid] = 0;
  }
  K1_TR(1, 0);
  if (tid < 4) m.cnt[tid] = 0;
  if (!second) {  // CTAs whose envs are all masked out leave at once (the masked re-observation after resets)
    if (!__syncthreads_or(tid < EPB && c_v)) return;
  } else {
    __syncthreads();
  }

  K1_TR(2, 0);
  K1_TR(3, 32);
  // ---- phase P: compact the (ego, slot) pairs that can possibly end up detected into plist, so
  // the expensive per-pair work below (Philox, Box-Muller, von Mises, fp64 geometry) runs with
  // full warps instead of the ~20 % of lanes a flat pair loop keeps busy.  The filter uses the
  // pre-step poses plus a travel allowance (it must only be a superset; phase B decides).  With
  // injected noise (tests) the noise is unbounded, so every live pair is kept. ----
  // Warps 1..3 run it while warp 0 is already integrating (phase A), then meet on barrier 2 and
  // signal warp 0 through barrier 1 (96 arrivals + warp 0's sync).
  if (tid >= 32) {
  for (PairIt it(tid - 32, S); it.p - (tid & 31) < NP; it.advance(TPB - 32, step3_a, step3_sl, S)) {
    bool keep = false;
    const int a = it.a, sl = it.sl;
    if (it.p < NP && !(m.fl0[a] & ASVRL_F_DEACTIVATED)) {
      const int el = m.el_of[a], r = m.r_of[a];
      float qx, qy, rt, travel = 0.5f * m.osp[a] + 1.5f;
      bool live;
      if (sl < OP) {
        live = sl < m.no[el];
        const float4 ob = m.obst[el * OP + sl];
        qx = ob.x; qy = ob.y; rt = ob.z;
      } else {
        const int j = sl - OP, b = el * RP + j;
        live = j != r && !(m.fl0[b] & ASVRL_F_DEACTIVATED);
        qx = m.ox[b]; qy = m.oy[b]; rt = k.robot_r; travel += 0.5f * m.osp[b] + 1.5f;
      }
      if (live) {
        const float dx = qx - m.ox[a], dy = qy - m.oy[a];
        const float far = k.detect_bound + rt + (dyn ? travel : 0.0f);
        keep = INJECT || dx * dx + dy * dy <= far * far;
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (bal) {
      const int lane = tid & 31;
      int base = 0;
      if (lane == 0) base = atomicAdd(&m.cnt[0], __popc(bal));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (keep) m.plist[base + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)((a << 6) | sl);
    }
  }
  K1_TR(4, 32);
  asm volatile("bar.sync 2, 96;" ::: "memory");
  asm volatile("bar.arrive 1, 128;" ::: "memory");
  }

  if (tid < 32) {
    // ---- phase A (warp 0): dynamics, one lane per ASV ----
    if (tid < A) {
      const int el = tid / RP, r = tid - el * RP;
      const int e = e0 + el;
      const bool present = m.ok[el] && r < m.nr[el];
      const size_t g = (size_t)e * RP + r;
      double x = p_x, y = p_y, th = p_th;
      const uint32_t f = m.fl0[tid];
      float rew = 0.0f;
      if (present && dyn && !(f & ASVRL_F_DEACTIVATED)) {
        // thrust increments in fp64: T + a*500 cancels when the thruster reverses
        double dl, dr;
        if (mode & ASVRL_M_DISCRETE) {  // WAMV:146-147, 225: id = 5*iL + iR
          const int a = min(max(act_id, 0), 24);
          dl = k.dl[a / 5];
          dr = k.dr[a % 5];
        } else {  // WAMV:220-223
          dl = (double)act.x * k.act_scale;
          dr = (double)act.y * k.act_scale;
        }
        const float4* cores = m.cores + el * CP;
        const int nc = m.nc[el];
        float* rx0 = m.crx + tid;
        float* ry0 = m.cry + tid;
        for (int i = 0; i < nc; ++i) {  // (core - pose0) in fp64, rounded once
          const float4 co = cores[i];
          rx0[i * A] = (float)((double)co.x - x);
          ry0[i * A] = (float)((double)co.y - y);
        }
        K1_TR(5, 0);
        const double x_start = x, y_start = y;
        double dxg = (double)gx - x, dyg = (double)gy - y;
        const double dis_before = sqrt(dxg * dxg + dyg * dyg);  // ENV:254
        const double TWO_PI = 6.283185307179586;
        for (int idx = 0; idx < k.n_sub; ++idx) {  // ENV:257-260
          float cx, cy;
          current_at(k, cores, nc, rx0, ry0, A, (float)(x - x_start), (float)(y - y_start), cx, cy);
          v0 = vr0 + cx; v1 = vr1 + cy; v2 = vr2;  // WAMV:201-202
          x += (double)(v0 * k.dt);                // WAMV:207-210
          y += (double)(v1 * k.dt);
          th += (double)(v2 * k.dt);
          while (th < 0.0) th += TWO_PI;           // WAMV:213-216
          while (th >= TWO_PI) th -= TWO_PI;
          if (idx == 0) {                          // WAMV:218-229
            Tl = fminf(fmaxf((float)((double)Tl + dl), k.min_thrust), k.max_thrust);
            Tr = fminf(fmaxf((float)((double)Tr + dr), k.min_thrust), k.max_thrust);
          }
          // compute_motion, WAMV:233-279
          float s, c;
          sincosf((float)th, &s, &c);
          float u_r = c * vr0 + s * vr1, v_r = -s * vr0 + c * vr1;
          const float u = c * v0 + s * v1, v = -s * v0 + c * v1;
          float rr = v2;
          const float a = k.yDotV * v_r + k.yDotR * rr;
          const float au = fabsf(u_r), av = fabsf(v_r), ar = fabsf(rr);
          const float n0 = (-k.xU - k.xUU * au) * u_r + a * rr;
          const float n1 = (-k.yV - k.yVV * av - k.yRV * ar) * v_r +
                           (-k.xDotU * u_r - k.yR - k.yVR * av - k.yRR * ar) * rr;
          const float n2 = -a * u_r + (k.xDotU * u_r - k.nV - k.nVV * av - k.nRV * ar) * v_r +
                           (-k.nR - k.nVR * av - k.nRR * ar) * rr;
          const float Fxl = Tl * k.cl, Fyl = Tl * k.sl, Fxr = Tr * k.cr, Fyr = Tr * k.sr;
          const float b0 = k.m * rr * v - n0 + (Fxl + Fxr);
          const float b1 = -k.m * rr * u - n1 + (Fyl + Fyr);
          const float b2 = -n2 + (Fxl * k.half_w - Fyl * k.half_l - Fxr * k.half_w - Fyr * k.half_l);
          u_r += (k.Ainv[0] * b0 + k.Ainv[1] * b1 + k.Ainv[2] * b2) * k.dt;
          v_r += (k.Ainv[3] * b0 + k.Ainv[4] * b1 + k.Ainv[5] * b2) * k.dt;
          rr += (k.Ainv[6] * b0 + k.Ainv[7] * b1 + k.Ainv[8] * b2) * k.dt;
          vr0 = c * u_r - s * v_r;
          vr1 = s * u_r + c * v_r;
          vr2 = rr;
        }
        K1_TR(6, 0);
        dxg = (double)gx - x; dyg = (double)gy - y;
        const double dis_after = sqrt(dxg * dxg + dyg * dyg);
        rew = k.timestep_penalty + (float)(dis_before - dis_after);  // ENV:271-277
        st.x[g] = (float)x; st.y[g] = (float)y; st.theta[g] = (float)th;
        st.velr_x[g] = vr0; st.velr_y[g] = vr1; st.velr_r[g] = vr2;
        st.vel_x[g] = v0; st.vel_y[g] = v1; st.vel_r[g] = v2;
        st.thrust_l[g] = Tl; st.thrust_r[g] = Tr;
      }
      // heading of the pose everything below reads, with the fp32 rounding residual folded in
      const float thf = (float)th;
      float s, c;
      sincosf(thf, &s, &c);
      const float dth = (float)(th - (double)thf);
      m.px[tid] = x; m.py[tid] = y;
      m.c[tid] = c - s * dth; m.s[tid] = s + c * dth;
      m.vx[tid] = v0; m.vy[tid] = v1; m.vz[tid] = v2;
      m.Tl[tid] = Tl; m.Tr[tid] = Tr; m.gx[tid] = gx; m.gy[tid] = gy;
      m.rew[tid] = rew;
      uint32_t fnew = f;
      if (present && !(f & ASVRL_F_DEACTIVATED)) {  // check_reach_goal, WAMV:170-172, 462
        const double dxg = (double)gx - x, dyg = (double)gy - y;
        if (dxg * dxg + dyg * dyg <= k.goal_dis_d * k.goal_dis_d) fnew |= ASVRL_F_REACH_GOAL;
      }
      m.fl[tid] = fnew;
    }
    K1_TR(7, 0);
    asm volatile("bar.sync 1, 128;" ::: "memory");  // plist / cnt[0] of warps 1..3 are complete
  }
  if (!INJECT) {
    // ---- phase A' (warps 1..3 at once, warp 0 when its integration is done): the sensor noise
    // does not depend on the state, so it is drawn for the compacted pairs while warp 0 walks
    // the serial integration chain.  Warps take 32 list entries at a time from a ticket. ----
    const int lane = tid & 31;
    const int n_pairs = m.cnt[0];
    for (;;) {
      int first = 0;
      if (lane == 0) first = atomicAdd(&m.cnt[2], 32);
      first = __shfl_sync(0xffffffffu, first, 0);
      if (first >= n_pairs) break;
      const int i = first + lane;
      if (i < n_pairs) {
        const int pr = m.plist[i], a = pr >> 6, sl = pr & 63;
        const int el = m.el_of[a];
        const NoiseDraw nd = sensor_noise(k, seed, ((FUSED && second) || fa.step_arr) ? (uint64_t)m.epn[el] : step,
                                          env_offset + e0 + el, (uint32_t)m.r_of[a], (uint32_t)sl);
        float* np = m.cand + (a * S + sl) * 5;
#pragma unroll
        for (int q = 0; q < 5; ++q) np[q] = nd.n[q];
      }
    }
  }
  K1_TR(8, 0);
  K1_TR(9, 32);
  __syncthreads();
  const int n_pairs = m.cnt[0];
  K1_TR(10, 0);

  // ---- phase B: the compacted pairs, one per thread, with the post-step poses: detection and
  // collision predicates in fp64, body-frame observation and sort key into shared memory;
  // detected pairs are compacted again (dlist) for the ranking pass ----
  for (int i0 = tid & ~31; i0 < n_pairs; i0 += TPB) {
    const int i = i0 + (tid & 31);
    bool det = false;
    int pr = 0;
    if (i < n_pairs) {
      pr = m.plist[i];
      const int a = pr >> 6, sl = pr & 63;
      const int p = a * S + sl;
      const int el = m.el_of[a], r = m.r_of[a];
      const uint32_t fe = m.fl0[a];
      const Cand cd = candidate(k, m, el, r, sl, RP, OP);  // live by construction of plist
      const double tx = cd.px - m.px[a], ty = cd.py - m.py[a];  // true offset
      const double t2 = tx * tx + ty * ty;
      bool maybe = true;
      if (!INJECT) {  // the in-kernel generator's noise is bounded: skip hopeless candidates
        const float far = k.detect_bound + (float)cd.rt;
        maybe = (float)t2 <= far * far;
      }
      if (maybe) {
        const float* np = INJECT ? noise + ((size_t)e0 * RP * S + p) * 5 : m.cand + p * 5;
        float n[5];
#pragma unroll
        for (int q = 0; q < 5; ++q) n[q] = np[q];
        const double qx = tx + (double)n[0], qy = ty + (double)n[1];  // WAMV:27-30
        const double orr = k.r_mean_ratio * cd.rt + k.r_noise_k * (double)n[4] * cd.rt;  // WAMV:37-40
        const double d2 = qx * qx + qy * qy, lim = k.perc_range + orr;
        det = d2 <= lim * lim;  // WAMV:296
        const float c = m.c[a], s = m.s[a];
        const float fx = (float)qx, fy = (float)qy;
        const float bx = c * fx + s * fy, by = -s * fx + c * fy;  // WAMV:305-322
        if (det && k.angle_limited) {  // WAMV:299-301 (vacuous at the default 2 pi)
          const float ang = atan2f(by, bx);
          det = !((double)ang < -k.half_angle || (double)ang > k.half_angle);
        }
        if (det) {
          atomicOr(sl < 32 ? &m.det_lo[a] : &m.det_hi[a], 1u << (sl & 31));
          if (!(fe & ASVRL_F_COLLISION)) {  // true geometry, WAMV:281-291, 475-476
            const double reach = cd.rt + k.robot_r_d;
            if (t2 <= reach * reach) atomicOr(&m.fl[a], (uint32_t)ASVRL_F_COLLISION);
          }
          const float ovx = cd.vx + n[2], ovy = cd.vy + n[3];  // WAMV:32-35
          float* o = m.cand + p * 5;
          o[0] = bx; o[1] = by;
          o[2] = c * ovx + s * ovy; o[3] = -s * ovx + c * ovy;
          o[4] = (float)orr;
          m.key[a * SK + sl] = sqrt(d2) - orr - k.robot_r_d;  // WAMV:515, 288
        }
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, det);
    if (bal) {
      const int lane = tid & 31;
      int base = 0;
      if (lane == 0) base = atomicAdd(&m.cnt[1], __popc(bal));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (det) m.dlist[base + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)pr;
    }
  }
  for (int i = tid; i < A * 25; i += TPB) m.st_obj[i] = 0.0f;
  for (int i = tid; i < A * 5; i += TPB) { m.st_mask[i] = 0.0f; m.chit[i] = 0; }
  __syncthreads();

  K1_TR(11, 0);
  // ---- phase B2: rank of every detected candidate among its ego's keys (stable: ties keep slot
  // order, like heapq.nsmallest, WAMV:513-515); the 5 smallest go to the staged observation.
  // Only detected slots have keys, and only their bits are walked. ----
  const int n_det = m.cnt[1];
  for (int i = tid; i < n_det; i += TPB) {
    const int pr = m.dlist[i], a = pr >> 6, sl = pr & 63;
    const double* keys = m.key + a * SK;
    const double kk = keys[sl];
    int rank = 0;
    uint32_t lo = m.det_lo[a], hi = m.det_hi[a];
    while (lo) { const int q = __ffs(lo) - 1; lo &= lo - 1; const double kq = keys[q]; rank += (kq < kk || (kq == kk && q < sl)) ? 1 : 0; }
    while (hi) { const int q = 32 + __ffs(hi) - 1; hi &= hi - 1; const double kq = keys[q]; rank += (kq < kk || (kq == kk && q < sl)) ? 1 : 0; }
    if (rank < ASVRL_MAX_OBJ) {
      const float* cd = m.cand + (a * S + sl) * 5;
      float* so = m.st_obj + a * 25 + rank * 5;
      float o5[5];
#pragma unroll
      for (int q = 0; q < 5; ++q) { o5[q] = cd[q]; so[q] = o5[q]; }
      m.st_mask[a * 5 + rank] = 1.0f;
      if (out.top_ids && !second) out.top_ids[((size_t)e0 * RP + a) * 5 + rank] = sl;
      // COLREGs test of the kept object (WAMV:398-423) by the thread that placed it: no (ego, rank) pass of its own
      // (160 items on 128 threads were two rounds) and one CTA barrier fewer on the CTA's serial chain
      const float c = m.c[a], s = m.s[a];
      const float w0 = c * m.vx[a] + s * m.vy[a], w1 = -s * m.vx[a] + c * m.vy[a];
      float phi = 0.0f;
      const bool hit = colregs_one(k, o5, w0, w1, phi);
      m.chit[a * 5 + rank] = hit ? 1 : 0;
      m.cphi[a * 5 + rank] = phi;
    }
  }
  __syncthreads();

  K1_TR(13, 0);
  // ---- phases C2 and D side by side.  Warp 0, one lane per ego: self observation, reward / done / info, then its
  // own rows and the env-level reductions (ballots over the env's lanes); warps 1..3 meanwhile store the object rows
  // and masks, complete since the barrier above.  No CTA barrier between the two any more. ----
  const int n_env = min(EPB, E - e0);
  const size_t a0 = (size_t)e0 * RP;
  const int n_asv = n_env * RP;
  int my_end = 0;
  if (tid >= 32) {
    const int t = tid - 32;
    constexpr int T3 = TPB - 32;
    if (!env_mask && !second) {
      for (int i = t; i < n_asv * 25; i += T3) out.obs_objects[a0 * 25 + i] = m.st_obj[i];
      for (int i = t; i < n_asv * 5; i += T3) out.obs_mask[a0 * 5 + i] = m.st_mask[i];
    } else {
      for (int el = 0; el < n_env; ++el) {
        if (!m.ok[el]) continue;
        const size_t b0 = a0 + (size_t)el * RP;
        const int l0 = el * RP;
        for (int i = t; i < RP * 25; i += T3) out.obs_objects[b0 * 25 + i] = m.st_obj[l0 * 25 + i];
        for (int i = t; i < RP * 5; i += T3) out.obs_mask[b0 * 5 + i] = m.st_mask[l0 * 5 + i];
      }
    }
  } else {
    uint32_t f_fin = 0;
    bool pres = false;
    if (tid < A) {
      const int a = tid, el = a / RP, r = a - el * RP;
      const int e = e0 + el;
      const size_t g = (size_t)e * RP + r;
      float* ss = m.st_self + a * 7;
#pragma unroll
      for (int q = 0; q < 7; ++q) ss[q] = 0.0f;
      if (m.ok[el]) {
        const bool present = r < m.nr[el];
        const uint32_t f0 = m.fl0[a];
        uint32_t f = m.fl[a], status = 0;
        float reward = 0.0f, phi = 0.0f;
        if (!present) {
          status = (uint32_t)ASVRL_INFO_ABSENT << ASVRL_S_INFO_SHIFT;
          f = 0;
        } else if (f0 & ASVRL_F_DEACTIVATED) {  // WAMV:437-438, ENV:295-304
          const int info = (f0 & ASVRL_F_COLLISION) ? ASVRL_INFO_DEACT_COLLISION
                           : (f0 & ASVRL_F_REACH_GOAL) ? ASVRL_INFO_DEACT_GOAL : ASVRL_INFO_DEACT_INVALID;
          status = ASVRL_S_DONE | ASVRL_S_WAS_DEACTIVATED | ((uint32_t)info << ASVRL_S_INFO_SHIFT);
        } else {
          const float c = m.c[a], s = m.s[a];
          const float dgx = (float)((double)m.gx[a] - m.px[a]), dgy = (float)((double)m.gy[a] - m.py[a]);
          ss[0] = c * dgx + s * dgy; ss[1] = -s * dgx + c * dgy;  // WAMV:443-453
          ss[2] = c * m.vx[a] + s * m.vy[a]; ss[3] = -s * m.vx[a] + c * m.vy[a];
          ss[4] = m.vz[a]; ss[5] = m.Tl[a]; ss[6] = m.Tr[a];
          bool apply = false;  // first hit in sorted order wins, WAMV:517-521
#pragma unroll
          for (int t = 0; t < ASVRL_MAX_OBJ; ++t)
            if (!apply && m.chit[a * 5 + t]) { apply = true; phi = m.cphi[a * 5 + t]; }
          if (apply) status |= ASVRL_S_COLREGS;
          int info = ASVRL_INFO_NORMAL;
          if (dyn) {  // ENV:306-328
            reward = m.rew[a];
            if (apply) reward += k.colregs_penalty * phi;
            if (m.ep[el] >= k.max_ep) { status |= ASVRL_S_DONE; info = ASVRL_INFO_TOO_LONG; }
            else if (f & ASVRL_F_COLLISION) { reward += k.collision_penalty; status |= ASVRL_S_DONE; info = ASVRL_INFO_COLLISION; }
            else if (f & ASVRL_F_REACH_GOAL) { reward += k.goal_reward; status |= ASVRL_S_DONE; info = ASVRL_INFO_REACH_GOAL; }
          }
          status |= (uint32_t)info << ASVRL_S_INFO_SHIFT;
          if ((mode & ASVRL_M_AUTO_DEACTIVATE) && (f & (ASVRL_F_COLLISION | ASVRL_F_REACH_GOAL)))
            f |= ASVRL_F_DEACTIVATED;  // TRAINER:168-170
          if (f != f0) st.flags[g] = f;
          m.fl[a] = f;
        }
        f_fin = f; pres = present;
        if (present) {
          if (f & ASVRL_F_COLLISION) status |= ASVRL_S_COLLISION;
          if (f & ASVRL_F_REACH_GOAL) status |= ASVRL_S_REACH_GOAL;
        }
        if (!second) {  // reward / status describe the transition, not the fresh episode
        out.reward[g] = reward;
        out.status[g] = status;
        out.phi[g] = phi;
        if (out.top_ids) {
          const int kept = (int)(m.st_mask[a * 5] + m.st_mask[a * 5 + 1] + m.st_mask[a * 5 + 2] +
                                 m.st_mask[a * 5 + 3] + m.st_mask[a * 5 + 4]);
          for (int t = kept; t < 5; ++t) out.top_ids[g * 5 + t] = -1;
        }
        if (out.detected)
          out.detected[g] = (unsigned long long)m.det_lo[a] | ((unsigned long long)m.det_hi[a] << 32);
        }
      }
    }
    __syncwarp();
    K1_TR(14, 0);
    if (!env_mask && !second) {
      for (int i = tid; i < n_asv * 7; i += 32) out.obs_self[a0 * 7 + i] = m.st_self[i];
    } else {
      for (int el = 0; el < n_env; ++el) {
        if (!m.ok[el]) continue;
        const size_t b0 = a0 + (size_t)el * RP;
        for (int i = tid; i < RP * 7; i += 32) out.obs_self[b0 * 7 + i] = m.st_self[el * RP * 7 + i];
      }
    }
    // env-level reductions over the env's present robots (ENV:178-200)
    const unsigned b_pr = __ballot_sync(0xffffffffu, pres);
    const unsigned b_de = __ballot_sync(0xffffffffu, pres && (f_fin & ASVRL_F_DEACTIVATED));
    const unsigned b_co = __ballot_sync(0xffffffffu, pres && (f_fin & ASVRL_F_COLLISION));
    const unsigned b_go = __ballot_sync(0xffffffffu, pres && (f_fin & ASVRL_F_REACH_GOAL));
    if (tid < A) {
      const int el = tid / RP, r = tid - el * RP;
      if (r == 0) {
        if (el < n_env && m.ok[el] && !second) {
          const int e = e0 + el;
          const unsigned msk = (RP >= 32 ? 0xffffffffu : ((1u << RP) - 1u)) << (el * RP);
          const unsigned P = b_pr & msk;
          const bool all_deact = (b_de & msk) == P, any_coll = (b_co & msk) != 0u, all_goal = (b_go & msk) == P;
          const uint32_t es = (all_deact ? 1u : 0u) | (m.ep[el] >= k.max_ep ? 2u : 0u) |
                              (any_coll ? 4u : 0u) | (all_goal ? 8u : 0u);
          out.env_status[e] = es;
          if (out.ended) out.ended[e] = (es & 3u) ? 1 : 0;
          if (dyn) st.ep_steps[e] = m.ep[el] + 1;  // ENV:330
          my_end = (es & 3u) ? 1 : 0;
        }
        if (FUSED && !second) m.ended[el] = my_end;
      }
    }
  }
  {
    K1_TR(15, 0);
    if (!FUSED || second) return;
    // ---- fused tail: episodes that just ended (TRAINER:172, 240-245) ----
    if (!__syncthreads_or(my_end)) return;
    if (fa.term_self) {  // the terminal observation rows stay available to the replay push
      for (int el = 0; el < n_env; ++el) {
        if (!m.ended[el]) continue;
        const size_t b0 = a0 + (size_t)el * RP;
        const int l0 = el * RP;
        for (int i = tid; i < RP * 25; i += TPB) fa.term_objects[b0 * 25 + i] = m.st_obj[l0 * 25 + i];
        for (int i = tid; i < RP * 7; i += TPB) fa.term_self[b0 * 7 + i] = m.st_self[l0 * 7 + i];
        for (int i = tid; i < RP * 5; i += TPB) fa.term_mask[b0 * 5 + i] = m.st_mask[l0 * 5 + i];
      }
    }
    {  // K1b's routine, run by the whole CTA for each ended env in turn
      const ResetState rs = {st.x, st.y, st.theta, st.velr_x, st.velr_y, st.velr_r, st.vel_x, st.vel_y,
                             st.vel_r, st.thrust_l, st.thrust_r, (float*)st.goal_x, (float*)st.goal_y,
                             st.flags, (int32_t*)st.n_robots, (int32_t*)st.n_obs, (int32_t*)st.n_cores,
                             st.ep_steps, (float4*)st.obstacles, (float4*)st.cores};
      // thread el decides for its env: 2 = the stage holds its next episode, 1 = generate it here.
      // Episode number and tag were read in phase 0; only a stage that was not ready then is
      // looked at again (the side stream may have published it meanwhile).
      if (tid < EPB && m.ended[tid]) {
        const int e = e0 + tid;
        const int ep = m.pre_ep[tid];
        m.epn[tid] = ep;
        if (m.pre_tag[tid] || (fa.stage_tag && ld_acquire(fa.stage_tag + e) == ep)) m.ended[tid] = 2;
      }
      __syncthreads();
      static_assert(TPB == 128, "reset_env_cta resolves rounds of 128 candidates");
      const CtaScratch sc = {m.rs_cand, m.rs_okm};
      for (int el = 0; el < n_env; ++el) {
        const int how = m.ended[el];
        if (!how) continue;
        const int e = e0 + el;
        if (how == 2) {
          // scenario AND first observation of the staged episode: every load first, then every
          // store, so the copy costs one DRAM round trip and not one per array
          take_stage_batched(rs, fa, out, e, RP, OP, CP, tid);
        } else
          reset_env_cta(fa.rk, rs, e, env_offset + e, fa.want_robots[e], fa.want_cores[e], fa.want_obs[e],
                        fa.min_sg[e], seed, (uint64_t)m.epn[el], RP, OP, CP, m.rs_sg, m.rs_core, m.rs_obs,
                        sc, tid);
      }
      __syncthreads();  // the stage has been read: only now may a worker see the bumped counter
      int observe = 0;
      if (tid < EPB && m.ended[tid]) {
        __threadfence();
        fa.episode_idx[e0 + tid] = m.epn[tid];
        if (fa.stage_self && m.ended[tid] == 2) m.ended[tid] = 0;  // observed already
        observe = m.ended[tid];
      }
      if (!__syncthreads_or(observe)) return;
      second = true;
      dyn = false;
      mode = 0;
      seed = fa.observe_seed;  // pass 1 draws its noise at step = the env's episode number (m.epn)
    }
    __syncthreads();
  }
  }  // pass loop
}

// fills the [E][RP][OP+RP][5] table the NULL-noise path of K1 uses (same device function)
__global__ void env_noise_kernel(const __grid_constant__ EnvConsts k, uint64_t seed, uint64_t step,
                                 int E, int RP, int OP, int64_t env_offset, float* __restrict__ out) {
  const int S = OP + RP;
  const size_t n = (size_t)E * RP * S;
  for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < n;
       p += (size_t)gridDim.x * blockDim.x) {
    const int sl = (int)(p % S);
    const size_t a = p / S;
    const int r = (int)(a % RP);
    const int64_t e = (int64_t)(a / RP);
    NoiseDraw nd = sensor_noise(k, seed, step, env_offset + e, (uint32_t)r, (uint32_t)sl);
#pragma unroll
    for (int q = 0; q < 5; ++q) out[p * 5 + q] = nd.n[q];
  }
}

static void invert3(const double G[3][3], double inv[3][3], double* det_out) {
  double c00 = G[1][1] * G[2][2] - G[1][2] * G[2][1];
  double c01 = G[1][2] * G[2][0] - G[1][0] * G[2][2];
  double c02 = G[1][0] * G[2][1] - G[1][1] * G[2][0];
  double det = G[0][0] * c00 + G[0][1] * c01 + G[0][2] * c02;
  *det_out = det;
  inv[0][0] = c00 / det; inv[0][1] = (G[0][2] * G[2][1] - G[0][1] * G[2][2]) / det;
  inv[0][2] = (G[0][1] * G[1][2] - G[0][2] * G[1][1]) / det;
  inv[1][0] = c01 / det; inv[1][1] = (G[0][0] * G[2][2] - G[0][2] * G[2][0]) / det;
  inv[1][2] = (G[0][2] * G[1][0] - G[0][0] * G[1][2]) / det;
  inv[2][0] = c02 / det; inv[2][1] = (G[0][1] * G[2][0] - G[0][0] * G[2][1]) / det;
  inv[2][2] = (G[0][0] * G[1][1] - G[0][1] * G[1][0]) / det;
}

int make_env_consts(const asvrl_env_params* p, EnvConsts* k) {
  if (!p || !k) return fail(ASVRL_E_BADARG, "null params");
  if (p->n_substeps < 0 || p->dt <= 0.0) return fail(ASVRL_E_BADARG, "bad dt / n_substeps");
  const double PI = 3.14159265358979323846;
  k->dt = (float)p->dt; k->n_sub = p->n_substeps; k->max_ep = p->max_episode_steps;
  k->min_thrust = (float)p->min_thrust; k->max_thrust = (float)p->max_thrust;
  const double hold = p->dt * p->n_substeps;
  k->act_scale = p->thrust_rate * hold;
  for (int i = 0; i < 5; ++i) {
    k->dl[i] = p->thrust_change_l[i] * hold;
    k->dr[i] = p->thrust_change_r[i] * hold;
  }
  k->m = (float)p->m; k->xDotU = (float)p->xDotU; k->yDotV = (float)p->yDotV; k->yDotR = (float)p->yDotR;
  k->xU = (float)p->xU; k->xUU = (float)p->xUU; k->yV = (float)p->yV; k->yVV = (float)p->yVV;
  k->yR = (float)p->yR; k->yRV = (float)p->yRV; k->yVR = (float)p->yVR; k->yRR = (float)p->yRR;
  k->nR = (float)p->nR; k->nRR = (float)p->nRR; k->nV = (float)p->nV; k->nVV = (float)p->nVV;
  k->nRV = (float)p->nRV; k->nVR = (float)p->nVR;
  k->cl = (float)cos(p->left_pos); k->sl = (float)sin(p->left_pos);
  k->cr = (float)cos(p->right_pos); k->sr = (float)sin(p->right_pos);
  k->half_w = (float)(p->width / 2); k->half_l = (float)(p->length / 2);
  {  // pinv(A) = (A^T A)^-1 A^T, WAMV:267-271
    double A[3][3] = {{p->m - p->xDotU, 0, 0}, {0, p->m - p->yDotV, -p->yDotR}, {0, -p->nDotV, p->Izz - p->nDotR}};
    double G[3][3], Gi[3][3], det;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) { G[i][j] = 0; for (int q = 0; q < 3; ++q) G[i][j] += A[q][i] * A[q][j]; }
    invert3(G, Gi, &det);
    if (!(fabs(det) > 0.0)) return fail(ASVRL_E_BADARG, "singular M_RB + M_A");
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) { double v = 0; for (int q = 0; q < 3; ++q) v += Gi[i][q] * A[j][q]; k->Ainv[i * 3 + j] = (float)v; }
  }
  k->core_r = (float)p->core_r; k->solid_k = (float)(1.0 / (2 * PI * p->core_r * p->core_r));
  k->inv_2pi = (float)(1.0 / (2 * PI)); k->drift_vx = (float)p->drift_vx; k->drift_vy = (float)p->drift_vy;
  k->perc_range = p->perc_range; k->half_angle = 0.5 * p->perc_angle;
  k->angle_limited = p->perc_angle < 2 * PI ? 1 : 0;
  k->r_mean_ratio = p->r_mean_ratio; k->r_noise_k = (1.0 - p->r_mean_ratio) / PI;
  k->robot_r_d = p->robot_r; k->goal_dis_d = p->goal_dis; k->robot_r = (float)p->robot_r;
  k->pos_std = (float)p->pos_std; k->vel_std = (float)p->vel_std;
  k->vm_uniform = p->r_kappa < 1e-5 ? 1 : 0;
  if (!k->vm_uniform) {
    double kap = p->r_kappa, rr = 1.0 + sqrt(1.0 + 4.0 * kap * kap);
    double rho = (rr - sqrt(2.0 * rr)) / (2.0 * kap);
    k->vm_kappa = (float)kap; k->vm_s = (float)((1.0 + rho * rho) / (2.0 * rho));
  } else { k->vm_kappa = 0; k->vm_s = 1; }
  // Box-Muller on 24-bit uniforms: |n| <= sqrt(-2 ln 2^-25) sigma < 5.9 sigma per axis; the
  // noisy radius never exceeds the true one (r_mean_ratio + (1-r_mean_ratio) <= 1).
  k->detect_bound = (float)(p->perc_range + 5.9 * 1.4143 * fabs(p->pos_std) + 1e-3);
  k->head_on_x = (float)p->head_on_x; k->head_on_half_y = (float)(0.5 * p->head_on_y);
  k->lc_x0 = (float)p->lc_x0; k->lc_x1 = (float)p->lc_x1; k->lc_y0 = (float)p->lc_y0; k->lc_y1 = (float)p->lc_y1;
  k->lc_grad = (float)(p->lc_y1 / p->lc_x1); k->min_speed = (float)p->colregs_min_speed;
  k->timestep_penalty = (float)p->timestep_penalty; k->colregs_penalty = (float)p->colregs_penalty;
  k->collision_penalty = (float)p->collision_penalty; k->goal_reward = (float)p->goal_reward;
  return 0;
}

}  // namespace asvrl

using namespace asvrl;

static int pick_epb(int RP) { int e = 32 / RP; return e < 1 ? 1 : e; }

extern "C" int asvrl_env_step(const asvrl_env_params* params, const asvrl_env_state* state,
                              const void* actions, const float* noise, const uint8_t* env_mask,
                              uint64_t seed, uint64_t step, const asvrl_env_out* out, int n_envs,
                              int max_robots, int max_obs, int max_cores, int64_t env_offset,
                              uint32_t mode, const uint64_t* counter_dev, const int32_t* step_per_env,
                              void* stream) {
  if (!params || !state || !out) return fail(ASVRL_E_BADARG, "asvrl_env_step: null struct");
  if (n_envs < 0 || max_robots < 1 || max_obs < 0 || max_cores < 0)
    return fail(ASVRL_E_BADARG, "asvrl_env_step: bad sizes");
  if (max_robots > 32 || max_obs + max_robots > ASVRL_MAX_SLOTS)
    return fail(ASVRL_E_LIMIT, "asvrl_env_step: max_robots <= 32 and max_obs + max_robots <= %d", ASVRL_MAX_SLOTS);
  if ((mode & ASVRL_M_DYNAMICS) && !actions) return fail(ASVRL_E_BADARG, "asvrl_env_step: actions required");
  if (n_envs == 0) return 0;
  EnvConsts k;
  int rc = make_env_consts(params, &k);
  if (rc) return rc;
  StatePtrs st = {state->x, state->y, state->theta, state->velr_x, state->velr_y, state->velr_r,
                  state->vel_x, state->vel_y, state->vel_r, state->thrust_l, state->thrust_r,
                  state->goal_x, state->goal_y, state->flags, state->n_robots, state->n_obs,
                  state->n_cores, state->ep_steps, (const float4*)state->obstacles,
                  (const float4*)state->cores};
  OutPtrs o = {out->obs_self, out->obs_objects, out->obs_mask, out->reward, out->status, out->phi,
               out->top_ids, (unsigned long long*)out->detected, out->env_status, out->ended};
  if (!o.obs_self || !o.obs_objects || !o.obs_mask || !o.reward || !o.status || !o.phi || !o.env_status)
    return fail(ASVRL_E_BADARG, "asvrl_env_step: null output");
  if ((max_obs > 0 && !st.obstacles) || (max_cores > 0 && !st.cores))
    return fail(ASVRL_E_BADARG, "asvrl_env_step: null obstacle / core table");
  const int EPB = pick_epb(max_robots);
  const SmemLayout L = make_layout(EPB, max_robots, max_obs, max_cores, TPB);
  const size_t smem = L.bytes;
  const int grid = (n_envs + EPB - 1) / EPB;
  cudaStream_t s = (cudaStream_t)stream;
  const int shape = pick_shape(max_robots, max_obs, max_cores, EPB);
  auto kern = noise ? env_step_kernel<true, false>
                    : (shape == 1 ? env_step_kernel<false, false, 1>
                                  : (shape == 2 ? env_step_kernel<false, false, 2> : env_step_kernel<false, false>));
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "asvrl_env_step: %zu B shared memory: %s", smem, cudaGetErrorString(e));
  }
  FusedArgs fa;
  memset(&fa, 0, sizeof fa);
  fa.step_arr = step_per_env;
  kern<<<grid, TPB, smem, s>>>(k, L, st, actions, noise, env_mask, seed, step, o, n_envs, max_robots, max_obs,
                               max_cores, env_offset, mode, EPB, counter_dev, fa);
  return check_launch("asvrl_env_step");
}

extern "C" int asvrl_env_step_autoreset(const asvrl_env_params* params, const asvrl_env_state* state,
                                        const void* actions, uint64_t seed, uint64_t step,
                                        const asvrl_env_out* out, const asvrl_env_autoreset* ar,
                                        int n_envs, int max_robots, int max_obs, int max_cores,
                                        int64_t env_offset, uint32_t mode,
                                        const uint64_t* counter_dev, void* stream) {
  if (!params || !state || !out || !ar) return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: null struct");
  if (!ar->reset_params || !ar->want_robots || !ar->want_cores || !ar->want_obs || !ar->min_start_goal_dis ||
      !ar->episode_idx)
    return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: null auto-reset argument");
  if ((ar->term_self || ar->term_objects || ar->term_mask) &&
      !(ar->term_self && ar->term_objects && ar->term_mask))
    return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: term_* must be all set or all NULL");
  if (n_envs < 0 || max_robots < 1 || max_obs < 0 || max_cores < 0)
    return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: bad sizes");
  if (max_robots > 32 || max_obs + max_robots > ASVRL_MAX_SLOTS || max_robots > RS_MAXR ||
      max_obs > RS_MAXO || max_cores > RS_MAXC)
    return fail(ASVRL_E_LIMIT, "asvrl_env_step_autoreset: max_robots <= 32, max_obs + max_robots <= %d, max_cores <= %d",
                ASVRL_MAX_SLOTS, RS_MAXC);
  if (!(mode & ASVRL_M_DYNAMICS) || !actions)
    return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: needs ASVRL_M_DYNAMICS and actions");
  if (n_envs == 0) return 0;
  EnvConsts k;
  int rc = make_env_consts(params, &k);
  if (rc) return rc;
  StatePtrs st = {state->x, state->y, state->theta, state->velr_x, state->velr_y, state->velr_r,
                  state->vel_x, state->vel_y, state->vel_r, state->thrust_l, state->thrust_r,
                  state->goal_x, state->goal_y, state->flags, state->n_robots, state->n_obs,
                  state->n_cores, state->ep_steps, (const float4*)state->obstacles,
                  (const float4*)state->cores};
  OutPtrs o = {out->obs_self, out->obs_objects, out->obs_mask, out->reward, out->status, out->phi,
               out->top_ids, (unsigned long long*)out->detected, out->env_status, out->ended};
  if (!o.obs_self || !o.obs_objects || !o.obs_mask || !o.reward || !o.status || !o.phi || !o.env_status)
    return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: null output");
  if ((max_obs > 0 && !st.obstacles) || (max_cores > 0 && !st.cores))
    return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: null obstacle / core table");
  FusedArgs fa;
  fa.rk = make_reset_consts(ar->reset_params);
  fa.want_robots = ar->want_robots; fa.want_cores = ar->want_cores; fa.want_obs = ar->want_obs;
  fa.min_sg = ar->min_start_goal_dis;
  fa.observe_seed = ar->observe_seed;
  fa.episode_idx = ar->episode_idx;
  fa.stage_tag = nullptr;
  fa.n_workers = 0;
  fa.step_arr = nullptr;
  fa.stage_self = fa.stage_objects = fa.stage_mask = nullptr;
  memset(&fa.stage, 0, sizeof fa.stage);
  if (ar->stage) {
    const asvrl_env_state* g = ar->stage;
    if (!ar->stage_tag) return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: stage needs stage_tag");
    if (!g->x || !g->y || !g->theta || !g->velr_x || !g->velr_y || !g->velr_r || !g->vel_x || !g->vel_y ||
        !g->vel_r || !g->thrust_l || !g->thrust_r || !g->goal_x || !g->goal_y || !g->flags || !g->n_robots ||
        !g->n_obs || !g->n_cores || !g->ep_steps || (max_obs > 0 && !g->obstacles) || (max_cores > 0 && !g->cores))
      return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: incomplete stage state");
    fa.stage = ResetState{g->x, g->y, g->theta, g->velr_x, g->velr_y, g->velr_r, g->vel_x, g->vel_y, g->vel_r,
                          g->thrust_l, g->thrust_r, g->goal_x, g->goal_y, g->flags, g->n_robots, g->n_obs,
                          g->n_cores, g->ep_steps, (float4*)g->obstacles, (float4*)g->cores};
    fa.stage_tag = ar->stage_tag;
    if (ar->stage_self || ar->stage_objects || ar->stage_mask) {
      if (!(ar->stage_self && ar->stage_objects && ar->stage_mask))
        return fail(ASVRL_E_BADARG, "asvrl_env_step_autoreset: stage_self/objects/mask must be all set or all NULL");
      fa.stage_self = ar->stage_self; fa.stage_objects = ar->stage_objects; fa.stage_mask = ar->stage_mask;
    }
  }
  fa.term_self = ar->term_self; fa.term_objects = ar->term_objects; fa.term_mask = ar->term_mask;
  const int EPB = pick_epb(max_robots);
  const SmemLayout L = make_layout(EPB, max_robots, max_obs, max_cores, TPB, true);
  const size_t smem = L.bytes;
  const int n_step = (n_envs + EPB - 1) / EPB;
  const int shape = pick_shape(max_robots, max_obs, max_cores, EPB);
  auto kern = shape == 1 ? env_step_kernel<false, true, 1>
                         : (shape == 2 ? env_step_kernel<false, true, 2> : env_step_kernel<false, true>);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "asvrl_env_step_autoreset: %zu B shared memory: %s", smem, cudaGetErrorString(e));
  }
  if (fa.stage_tag && ar->refill_in_launch) {
    // refill workers take the CTA slots the last wave of step CTAs leaves free (4..16 of them)
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TPB, smem);
    const int slots = sms * (per_sm > 0 ? per_sm : 1);
    int spare = slots - n_step % slots;
    if (spare == slots) spare = 0;
    fa.n_workers = spare < 4 ? 4 : (spare > 16 ? 16 : spare);
    const int max_useful = (n_envs + TPB / 32 - 1) / (TPB / 32);
    if (fa.n_workers > max_useful) fa.n_workers = max_useful;
  }
  const int grid = n_step + fa.n_workers;
  kern<<<grid, TPB, smem, (cudaStream_t)stream>>>(k, L, st, actions, nullptr, nullptr, seed, step, o, n_envs,
                                                  max_robots, max_obs, max_cores, env_offset, mode, EPB,
                                                  counter_dev, fa);
  return check_launch("asvrl_env_step_autoreset");
}

extern "C" int asvrl_env_noise(const asvrl_env_params* params, uint64_t seed, uint64_t step,
                               int n_envs, int max_robots, int max_obs, int64_t env_offset,
                               float* noise_out, void* stream) {
  if (!params || !noise_out) return fail(ASVRL_E_BADARG, "asvrl_env_noise: null argument");
  EnvConsts k;
  int rc = make_env_consts(params, &k);
  if (rc) return rc;
  size_t n = (size_t)n_envs * max_robots * (max_obs + max_robots);
  if (n == 0) return 0;
  int grid = (int)((n + 255) / 256);
  if (grid > 148 * 16) grid = 148 * 16;
  env_noise_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(k, seed, step, n_envs, max_robots, max_obs,
                                                           env_offset, noise_out);
  return check_launch("asvrl_env_noise");
}

#ifdef K1_TRACE
extern "C" __attribute__((visibility("default"))) int asvrl_env_step_trace(unsigned long long* out) {
  cudaDeviceSynchronize();
  return (int)cudaMemcpyFromSymbol(out, asvrl::k1_trace_buf, sizeof(unsigned long long) * 32);
}
#endif
