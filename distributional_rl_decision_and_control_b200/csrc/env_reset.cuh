// env_reset.cuh — warp-level scenario generation shared by K1b (env_reset.cu) and the fused
// auto-reset tail of K1 (env_step.cu).  See env_reset.cu for the algorithm and its reference lines.
#pragma once
#include <math.h>

#include "common.cuh"

namespace asvrl {

struct ResetConsts {
  float map_w, map_h, start_margin, obs_margin, clear_r, core_r, v_rel_max, p, v_min, v_max,
      obs_r_min, obs_r_max;
  int max_tries;
};

struct ResetState {
  float *x, *y, *theta, *velr_x, *velr_y, *velr_r, *vel_x, *vel_y, *vel_r, *thrust_l, *thrust_r,
      *goal_x, *goal_y;
  uint32_t* flags;
  int32_t *n_robots, *n_obs, *n_cores, *ep_steps;
  float4 *obstacles, *cores;
};

inline ResetConsts make_reset_consts(const asvrl_reset_params* p) {
  ResetConsts k = {(float)p->map_w, (float)p->map_h, (float)p->start_margin, (float)p->obs_margin,
                   (float)p->clear_r, (float)p->core_r, (float)p->v_rel_max, (float)p->p,
                   (float)p->v_min, (float)p->v_max, (float)p->obs_r_min, (float)p->obs_r_max,
                   p->max_tries};
  return k;
}

constexpr int RS_WARPS = 4;
constexpr int RS_MAXR = 32, RS_MAXO = 64, RS_MAXC = 64;

__device__ __forceinline__ float lerp_u(float lo, float hi, float u) { return lo + (hi - lo) * u; }


__device__ __forceinline__ float dist2f(float ax, float ay, float bx, float by) {
  const float dx = ax - bx, dy = ay - by;
  return sqrtf(dx * dx + dy * dy);
}

// squared distance: the acceptance tests below compare d^2 with the squared bound (no square root,
// no dependent chain); they decide like the reference's `norm(...) <= bound` except within one
// ulp of the bound, which the parity tests' float64 margin already excludes
__device__ __forceinline__ float sqd(float ax, float ay, float bx, float by) {
  const float dx = ax - bx, dy = ay - by;
  return dx * dx + dy * dy;
}

// ENV:368-374: start too close to an accepted start, or goal to an accepted goal
__device__ __forceinline__ bool sg_conflict(const ResetConsts& k, float4 c, float4 o) {
  const float c2 = k.clear_r * k.clear_r;
  return (sqd(o.x, o.y, c.x, c.y) <= c2) | (sqd(o.z, o.w, c.z, c.w) <= c2);
}

// ENV:396-416: core c = (x, y, dir, Gamma) against an accepted core o
__device__ __forceinline__ bool core_conflict(const ResetConsts& k, float4 c, float4 o) {
  const float TWO_PI = 6.28318530717959f;
  const float dis = dist2f(o.x, o.y, c.x, c.y);
  const float bi = o.w / (TWO_PI * k.v_rel_max), bj = c.w / (TWO_PI * k.v_rel_max);
  const float gl = fmaxf(o.w, c.w), gs = fminf(o.w, c.w);
  const float v1 = gl / (TWO_PI * (dis - 2.0f * k.core_r)), v2 = gs / (TWO_PI * k.core_r);
  return o.z == c.z ? dis < bi + bj : v1 > k.p * v2;
}

// ENV:448-454: obstacle c = (x, y, r) overlapping an accepted obstacle o
__device__ __forceinline__ bool obs_conflict(float4 c, float4 o) {
  const float rr = o.z + c.z;
  return sqd(o.x, o.y, c.x, c.y) <= rr * rr;
}

__device__ __forceinline__ float4 shfl4(float4 v, int src) {
  const unsigned FULL = 0xffffffffu;
  return make_float4(__shfl_sync(FULL, v.x, src), __shfl_sync(FULL, v.y, src),
                     __shfl_sync(FULL, v.z, src), __shfl_sync(FULL, v.w, src));
}

// Resolve one round of 32 candidates in order.  `okm` = lanes that pass against the set accepted
// before the round; CONFLICT(c, accepted) strikes later candidates.  Returns the new count.
template <class Conflict>
__device__ __forceinline__ int resolve_round(unsigned okm, float4 c, float4* set, int n, int want,
                                             int lane, Conflict conflict) {
  const unsigned FULL = 0xffffffffu;
  while (okm && n < want) {
    const int f = __ffs(okm) - 1;
    const float4 cf = shfl4(c, f);
    if (lane == f) set[n] = c;
    ++n;
    okm &= ~((2u << f) - 1u);                       // candidates up to f are decided
    okm &= ~__ballot_sync(FULL, conflict(c, cf));   // later ones must also clear the newcomer
  }
  __syncwarp();
  return n;
}

// One shared, rolled Philox for the scenario generator: it runs in a single warp, mostly with a cold
// instruction cache (once per episode), so code bytes cost more than a call and a loop counter.
static __device__ __noinline__ uint4 philox_reset(uint64_t seed, uint64_t episode, int64_t env, uint32_t cls,
                                           uint32_t t) {
  uint4 c = make_uint4((t & 255u) | ((t >> 8) << 8) | (cls << 16), (uint32_t)env, (uint32_t)episode,
                       ((uint32_t)(episode >> 32) & 0x00FFFFFFu) | ((uint32_t)STREAM_RESET << 24));
  uint2 k = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll 1
  for (int round = 0; round < 10; ++round) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// ---- candidate t of each class (Philox stream 2, counters as documented in env_reset.cu) and the
// part of its acceptance test that does not involve the class's own accepted set ----
__device__ __forceinline__ float4 robot_candidate(const ResetConsts& k, uint64_t seed, uint64_t episode,
                                                  int64_t ge, int t) {  // ENV:108-110
  const uint4 w = philox_reset(seed, episode, ge, 0u, (uint32_t)t);
  const float lo = k.start_margin;
  return make_float4(lerp_u(lo, k.map_w - lo, u01(w.x)), lerp_u(lo, k.map_h - lo, u01(w.y)),
                     lerp_u(lo, k.map_w - lo, u01(w.z)), lerp_u(lo, k.map_h - lo, u01(w.w)));
}
__device__ __forceinline__ bool robot_ok(float4 c, float min_dis) {  // ENV:361
  return !(sqd(c.z, c.w, c.x, c.y) < min_dis * min_dis);
}

__device__ __forceinline__ float4 core_candidate(const ResetConsts& k, uint64_t seed, uint64_t episode,
                                                 int64_t ge, int t) {  // ENV:125-129
  const float TWO_PI = 6.28318530717959f;
  const uint4 w = philox_reset(seed, episode, ge, 1u, (uint32_t)t);
  const float dir = (w.z >> 31) ? 1.0f : -1.0f;  // binomial(1, 0.5), ENV:127
  return make_float4(k.map_w * u01(w.x), k.map_h * u01(w.y), dir,
                     TWO_PI * k.core_r * lerp_u(k.v_min, k.v_max, u01(w.w)));
}
__device__ __forceinline__ bool core_ok(const ResetConsts& k, float4 c, const float4* sg, int n_r) {
  bool bad = (c.x - k.core_r < 0.0f) | (c.x + k.core_r > k.map_w);   // ENV:381
  bad |= (c.y - k.core_r < 0.0f) | (c.y + k.core_r > k.map_w);       // ENV:383 (sic: width)
  const float lim = k.core_r + k.clear_r, lim2 = lim * lim;
#pragma unroll 4
  for (int i = 0; i < n_r; ++i) {                                    // ENV:386-394
    const float4 o = sg[i];
    bad |= (sqd(c.x, c.y, o.x, o.y) < lim2) | (sqd(c.x, c.y, o.z, o.w) < lim2);
  }
  return !bad;
}

__device__ __forceinline__ float4 obs_candidate(const ResetConsts& k, uint64_t seed, uint64_t episode,
                                                int64_t ge, int t) {  // ENV:153-155
  const uint4 w = philox_reset(seed, episode, ge, 2u, (uint32_t)t);
  const float lo = k.obs_margin;
  return make_float4(lerp_u(lo, k.map_w - lo, u01(w.x)), lerp_u(lo, k.map_h - lo, u01(w.y)),
                     lerp_u(k.obs_r_min, k.obs_r_max, u01(w.z)), 0.0f);
}
__device__ __forceinline__ bool obs_ok(const ResetConsts& k, float4 c, const float4* sg, int n_r,
                                       const float4* cores, int n_c) {
  const float cx = c.x, cy = c.y, rr = c.z;
  bool bad = (cx - rr < 0.0f) | (cx + rr > k.map_w) | (cy - rr < 0.0f) | (cy + rr > k.map_h);
  const float lim = rr + k.clear_r, lim2 = lim * lim, lc = k.core_r + rr, lc2 = lc * lc;
#pragma unroll 4
  for (int i = 0; i < n_r; ++i) {  // ENV:428-436
    const float4 o = sg[i];
    bad |= (sqd(cx, cy, o.x, o.y) < lim2) | (sqd(cx, cy, o.z, o.w) < lim2);
  }
#pragma unroll 4
  for (int i = 0; i < n_c; ++i)    // ENV:439-445
    bad |= sqd(cores[i].x, cores[i].y, cx, cy) <= lc2;
  return !bad;
}

// write the fresh episode: reset_robot ENV:166-176, Robot.reset_state WAMV:177-193
__device__ __forceinline__ void write_episode(const ResetState& st, int e, int64_t ge, uint64_t seed,
                                              uint64_t episode, int RP, int OP, int CP, const float4* sg,
                                              int n_r, const float4* cores, int n_c, const float4* obs,
                                              int n_o, int tid, int nthreads) {
  const float TWO_PI = 6.28318530717959f;
  for (int r = tid; r < RP; r += nthreads) {
    const size_t g = (size_t)e * RP + r;
    const bool on = r < n_r;
    float4 c = on ? sg[r] : make_float4(0, 0, 0, 0);
    float th = 0.0f;
    if (on) {  // init_theta ~ U(0, 2 pi), ENV:171
      uint4 w = philox_reset(seed, episode, ge, 3u, (uint32_t)r << 8);
      th = fminf(TWO_PI * u01(w.x), 6.2831850f);
    }
    st.x[g] = c.x; st.y[g] = c.y; st.theta[g] = th;
    st.velr_x[g] = 0; st.velr_y[g] = 0; st.velr_r[g] = 0;
    st.vel_x[g] = 0; st.vel_y[g] = 0; st.vel_r[g] = 0;  // no cores exist yet at ENV:175
    st.thrust_l[g] = 0; st.thrust_r[g] = 0;
    st.goal_x[g] = c.z; st.goal_y[g] = c.w;
    st.flags[g] = 0;
  }
  for (int i = tid; i < OP; i += nthreads) st.obstacles[(size_t)e * OP + i] = i < n_o ? obs[i] : make_float4(0, 0, 0, 0);
  for (int i = tid; i < CP; i += nthreads) st.cores[(size_t)e * CP + i] = i < n_c ? cores[i] : make_float4(0, 0, 0, 0);
  if (tid == 0) {
    st.n_robots[e] = n_r; st.n_obs[e] = n_o; st.n_cores[e] = n_c;
    st.ep_steps[e] = 0;  // ENV:96
  }
}

// the pre-generated episode of env e moves from the stage to the live state (all `nthreads` threads)
__device__ __forceinline__ void take_stage(const ResetState& st, const ResetState& sg, int e, int RP, int OP,
                                           int CP, int tid, int nthreads) {
  for (int r = tid; r < RP; r += nthreads) {
    const size_t g = (size_t)e * RP + r;
    st.x[g] = __ldcg(sg.x + g); st.y[g] = __ldcg(sg.y + g); st.theta[g] = __ldcg(sg.theta + g);
    st.goal_x[g] = __ldcg(sg.goal_x + g); st.goal_y[g] = __ldcg(sg.goal_y + g);
    st.velr_x[g] = 0; st.velr_y[g] = 0; st.velr_r[g] = 0;
    st.vel_x[g] = 0; st.vel_y[g] = 0; st.vel_r[g] = 0;
    st.thrust_l[g] = 0; st.thrust_r[g] = 0;
    st.flags[g] = 0;
  }
  for (int i = tid; i < OP; i += nthreads) st.obstacles[(size_t)e * OP + i] = __ldcg(sg.obstacles + (size_t)e * OP + i);
  for (int i = tid; i < CP; i += nthreads) st.cores[(size_t)e * CP + i] = __ldcg(sg.cores + (size_t)e * CP + i);
  if (tid == 0) {
    st.n_robots[e] = __ldcg(sg.n_robots + e); st.n_obs[e] = __ldcg(sg.n_obs + e); st.n_cores[e] = __ldcg(sg.n_cores + e);
    st.ep_steps[e] = 0;
  }
}

// One warp re-generates env `e` (global index `ge` for the RNG counters) and writes the fresh episode
// to the state arrays.  sg / cores / obs: per-warp shared scratch of >= RP / CP / OP float4.
__device__ __forceinline__ void reset_env_warp(const ResetConsts& k, const ResetState& st, int e, int64_t ge,
                                               int want_r_in, int want_c_in, int want_o_in, float min_dis,
                                               uint64_t seed, uint64_t episode, int RP, int OP, int CP,
                                               float4* sg, float4* cores, float4* obs, int lane) {
  const unsigned FULL = 0xffffffffu;
  // ---- robots: ENV:105-120, check_start_and_goal ENV:358-376 ----
  const int want_r = min(want_r_in, RP);
  int n_r = 0;
  for (int base = 0; base < k.max_tries && n_r < want_r; base += 32) {
    const int t = base + lane;
    const float4 c = robot_candidate(k, seed, episode, ge, t);
    bool ok = t < k.max_tries && robot_ok(c, min_dis);
#pragma unroll 4
    for (int i = 0; i < n_r; ++i) ok &= !sg_conflict(k, c, sg[i]);
    n_r = resolve_round(__ballot_sync(FULL, ok), c, sg, n_r, want_r, lane,
                        [&](float4 a, float4 b) { return sg_conflict(k, a, b); });
  }
  // ---- vortex cores: ENV:122-136, check_core ENV:378-418 ----
  const int want_c = min(want_c_in, CP);
  int n_c = 0;
  for (int base = 0; base < k.max_tries && n_c < want_c; base += 32) {
    const int t = base + lane;
    const float4 c = core_candidate(k, seed, episode, ge, t);
    bool ok = t < k.max_tries && core_ok(k, c, sg, n_r);
#pragma unroll 4
    for (int i = 0; i < n_c; ++i) ok &= !core_conflict(k, c, cores[i]);  // ENV:396-416
    n_c = resolve_round(__ballot_sync(FULL, ok), c, cores, n_c, want_c, lane,
                        [&](float4 a, float4 b) { return core_conflict(k, a, b); });
  }
  // ---- static obstacles: ENV:150-162, check_obstacle ENV:420-456 ----
  const int want_o = min(want_o_in, OP);
  int n_o = 0;
  for (int base = 0; base < k.max_tries && n_o < want_o; base += 32) {
    const int t = base + lane;
    const float4 c = obs_candidate(k, seed, episode, ge, t);
    bool ok = t < k.max_tries && obs_ok(k, c, sg, n_r, cores, n_c);
#pragma unroll 4
    for (int i = 0; i < n_o; ++i) ok &= !obs_conflict(c, obs[i]);  // ENV:448-454
    n_o = resolve_round(__ballot_sync(FULL, ok), c, obs, n_o, want_o, lane,
                        [&](float4 a, float4 b) { return obs_conflict(a, b); });
  }
  __syncwarp();
  write_episode(st, e, ge, seed, episode, RP, OP, CP, sg, n_r, cores, n_c, obs, n_o, lane, 32);
}

// ---- the same algorithm run by a whole 128-thread CTA (the fused tail of K1, where one env of
// the CTA has ended and all four warps are free): rounds of 128 candidates, resolved in candidate
// order by warp 0 (lane l owns candidates 4l..4l+3 of the round).  Identical results: a candidate
// is kept iff it passes against everything accepted before it. ----
struct CtaScratch {
  float4* cand;     // [128] the round's candidates
  unsigned* okm;    // [4] per-warp ballots, [4] = accepted count after the round
};

template <class Gen, class Fixed, class Conflict>
__device__ __forceinline__ int place_class_cta(int want, int max_tries, float4* set, const CtaScratch& sc,
                                               int tid, Gen gen, Fixed fixed_ok, Conflict conflict) {
  const unsigned FULL = 0xffffffffu;
  const int lane = tid & 31, warp = tid >> 5;
  int n = 0;
  for (int base = 0; base < max_tries && n < want; base += 128) {
    const int t = base + tid;
    const float4 c = gen(t);
    bool ok = t < max_tries && fixed_ok(c);
#pragma unroll 4
    for (int i = 0; i < n; ++i) ok &= !conflict(c, set[i]);
    sc.cand[tid] = c;
    const unsigned b = __ballot_sync(FULL, ok);
    if (lane == 0) sc.okm[warp] = b;
    __syncthreads();
    if (warp == 0) {
      unsigned mine = (sc.okm[lane >> 3] >> (4 * (lane & 7))) & 15u;
      float4 c4[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) c4[q] = sc.cand[4 * lane + q];
      while (n < want) {
        const unsigned bal = __ballot_sync(FULL, mine != 0u);
        if (!bal) break;
        const int fl = __ffs(bal) - 1;
        const int fq = __ffs(__shfl_sync(FULL, mine, fl)) - 1;
        const float4 cf = sc.cand[4 * fl + fq];
        if (lane == 0) set[n] = cf;
        ++n;
        if (lane < fl) mine = 0u;
        else if (lane == fl) mine &= ~((2u << fq) - 1u);  // candidates up to the accepted one are decided
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (((mine >> q) & 1u) && conflict(c4[q], cf)) mine &= ~(1u << q);
      }
      if (lane == 0) sc.okm[4] = (unsigned)n;
    }
    __syncthreads();
    n = (int)sc.okm[4];
  }
  return n;
}

// all 128 threads of the CTA call this with CTA-uniform arguments
__device__ __forceinline__ void reset_env_cta(const ResetConsts& k, const ResetState& st, int e, int64_t ge,
                                              int want_r_in, int want_c_in, int want_o_in, float min_dis,
                                              uint64_t seed, uint64_t episode, int RP, int OP, int CP,
                                              float4* sg, float4* cores, float4* obs, const CtaScratch& sc,
                                              int tid) {
  const int n_r = place_class_cta(
      min(want_r_in, RP), k.max_tries, sg, sc, tid,
      [&](int t) { return robot_candidate(k, seed, episode, ge, t); },
      [&](float4 c) { return robot_ok(c, min_dis); },
      [&](float4 a, float4 b) { return sg_conflict(k, a, b); });
  const int n_c = place_class_cta(
      min(want_c_in, CP), k.max_tries, cores, sc, tid,
      [&](int t) { return core_candidate(k, seed, episode, ge, t); },
      [&](float4 c) { return core_ok(k, c, sg, n_r); },
      [&](float4 a, float4 b) { return core_conflict(k, a, b); });
  const int n_o = place_class_cta(
      min(want_o_in, OP), k.max_tries, obs, sc, tid,
      [&](int t) { return obs_candidate(k, seed, episode, ge, t); },
      [&](float4 c) { return obs_ok(k, c, sg, n_r, cores, n_c); },
      [&](float4 a, float4 b) { return obs_conflict(a, b); });
  write_episode(st, e, ge, seed, episode, RP, OP, CP, sg, n_r, cores, n_c, obs, n_o, tid, 128);
  __syncthreads();
}

}  // namespace asvrl
