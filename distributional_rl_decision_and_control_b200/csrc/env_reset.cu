// env_reset.cu — K1b: masked per-env scenario re-generation (sm_100a).
//
// Replaces MarineNavEnv3.reset (rfarl/rfarl/envs/marinenav/env.py:72-176) with the acceptance
// predicates of check_start_and_goal / check_core / check_obstacle (env.py:358-456).
//
// One warp per environment that needs a reset.  The reference's loop "draw a candidate, keep it
// if it passes the checks against everything kept so far, give up after 500 draws" is sequential
// only through the accepted set, so the warp evaluates 32 consecutive candidates of the
// counter-indexed sequence at once against the set accepted so far, then resolves the round in
// candidate order with ballots: the first survivor is accepted, later survivors that conflict with
// it are struck, and so on — a few shuffles per acceptance instead of a new Philox round.  A
// candidate is kept iff it passes against everything accepted before it, so the result is exactly
// the sequential algorithm applied to the Philox candidate sequence.
// Counter use (stream 2): ego field = class (0 robots, 1 cores, 2 obstacles, 3 theta0),
// candidate t -> slot = t >> 8, block = t & 255; theta0 of robot r -> slot = r.
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

constexpr int RS_WARPS = 4;
constexpr int RS_MAXR = 32, RS_MAXO = 64, RS_MAXC = 64;

__device__ __forceinline__ float lerp_u(float lo, float hi, float u) { return lo + (hi - lo) * u; }


__device__ __forceinline__ float dist2f(float ax, float ay, float bx, float by) {
  const float dx = ax - bx, dy = ay - by;
  return sqrtf(dx * dx + dy * dy);
}

// ENV:368-374: start too close to an accepted start, or goal to an accepted goal
__device__ __forceinline__ bool sg_conflict(const ResetConsts& k, float4 c, float4 o) {
  return dist2f(o.x, o.y, c.x, c.y) <= k.clear_r || dist2f(o.z, o.w, c.z, c.w) <= k.clear_r;
}

// ENV:396-416: core c = (x, y, dir, Gamma) against an accepted core o
__device__ __forceinline__ bool core_conflict(const ResetConsts& k, float4 c, float4 o) {
  const float TWO_PI = 6.28318530717959f;
  const float dis = dist2f(o.x, o.y, c.x, c.y);
  if (o.z == c.z) {
    const float bi = o.w / (TWO_PI * k.v_rel_max), bj = c.w / (TWO_PI * k.v_rel_max);
    return dis < bi + bj;
  }
  const float gl = fmaxf(o.w, c.w), gs = fminf(o.w, c.w);
  const float v1 = gl / (TWO_PI * (dis - 2.0f * k.core_r)), v2 = gs / (TWO_PI * k.core_r);
  return v1 > k.p * v2;
}

// ENV:448-454: obstacle c = (x, y, r) overlapping an accepted obstacle o
__device__ __forceinline__ bool obs_conflict(float4 c, float4 o) {
  return dist2f(o.x, o.y, c.x, c.y) <= o.z + c.z;
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

__global__ void __launch_bounds__(RS_WARPS * 32)
env_reset_kernel(const __grid_constant__ ResetConsts k, ResetState st,
                 const uint8_t* __restrict__ reset_mask, const int32_t* __restrict__ want_robots,
                 const int32_t* __restrict__ want_cores, const int32_t* __restrict__ want_obs,
                 const float* __restrict__ min_sg, uint64_t seed, uint64_t episode, int E, int RP,
                 int OP, int CP, int64_t env_offset, const uint64_t* __restrict__ counter_dev) {
  __shared__ float4 s_sg[RS_WARPS][RS_MAXR];    // start.xy, goal.xy
  __shared__ float4 s_core[RS_WARPS][RS_MAXC];  // x, y, +-1, Gamma
  __shared__ float4 s_obs[RS_WARPS][RS_MAXO];   // x, y, r, 0
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int e = blockIdx.x * RS_WARPS + warp;
  if (e >= E) return;
  if (reset_mask && !reset_mask[e]) return;
  if (counter_dev) episode += *counter_dev;
  const int64_t ge = env_offset + e;
  const unsigned FULL = 0xffffffffu;
  const float TWO_PI = 6.28318530717959f;
  float4* sg = s_sg[warp];
  float4* cores = s_core[warp];
  float4* obs = s_obs[warp];

  // ---- robots: ENV:105-120, check_start_and_goal ENV:358-376 ----
  const int want_r = min(want_robots[e], RP);
  const float min_dis = min_sg[e];
  int n_r = 0;
  for (int base = 0; base < k.max_tries && n_r < want_r; base += 32) {
    const int t = base + lane;
    uint4 w = philox_at(seed, STREAM_RESET, episode, ge, 0u, (uint32_t)t >> 8, (uint32_t)t & 255u);
    const float lo = k.start_margin;
    float4 c = make_float4(lerp_u(lo, k.map_w - lo, u01(w.x)), lerp_u(lo, k.map_h - lo, u01(w.y)),
                           lerp_u(lo, k.map_w - lo, u01(w.z)), lerp_u(lo, k.map_h - lo, u01(w.w)));
    bool ok = t < k.max_tries;
    ok = ok && !(dist2f(c.z, c.w, c.x, c.y) < min_dis);  // ENV:361
    for (int i = 0; i < n_r && ok; ++i) ok = !sg_conflict(k, c, sg[i]);
    n_r = resolve_round(__ballot_sync(FULL, ok), c, sg, n_r, want_r, lane,
                        [&](float4 a, float4 b) { return sg_conflict(k, a, b); });
  }

  // ---- vortex cores: ENV:122-136, check_core ENV:378-418 ----
  const int want_c = min(want_cores[e], CP);
  int n_c = 0;
  for (int base = 0; base < k.max_tries && n_c < want_c; base += 32) {
    const int t = base + lane;
    uint4 w = philox_at(seed, STREAM_RESET, episode, ge, 1u, (uint32_t)t >> 8, (uint32_t)t & 255u);
    const float cx = k.map_w * u01(w.x), cy = k.map_h * u01(w.y);
    const float dir = (w.z >> 31) ? 1.0f : -1.0f;                    // binomial(1, 0.5), ENV:127
    const float gam = TWO_PI * k.core_r * lerp_u(k.v_min, k.v_max, u01(w.w));  // ENV:128-129
    const float4 c = make_float4(cx, cy, dir, gam);
    bool ok = t < k.max_tries;
    ok = ok && !(cx - k.core_r < 0.0f || cx + k.core_r > k.map_w);   // ENV:381
    ok = ok && !(cy - k.core_r < 0.0f || cy + k.core_r > k.map_w);   // ENV:383 (sic: width)
    for (int i = 0; i < n_r && ok; ++i) {                            // ENV:386-394
      const float4 o = sg[i];
      ok = !(dist2f(cx, cy, o.x, o.y) < k.core_r + k.clear_r) && !(dist2f(cx, cy, o.z, o.w) < k.core_r + k.clear_r);
    }
    for (int i = 0; i < n_c && ok; ++i) ok = !core_conflict(k, c, cores[i]);  // ENV:396-416
    n_c = resolve_round(__ballot_sync(FULL, ok), c, cores, n_c, want_c, lane,
                        [&](float4 a, float4 b) { return core_conflict(k, a, b); });
  }

  // ---- static obstacles: ENV:150-162, check_obstacle ENV:420-456 ----
  const int want_o = min(want_obs[e], OP);
  int n_o = 0;
  for (int base = 0; base < k.max_tries && n_o < want_o; base += 32) {
    const int t = base + lane;
    uint4 w = philox_at(seed, STREAM_RESET, episode, ge, 2u, (uint32_t)t >> 8, (uint32_t)t & 255u);
    const float lo = k.obs_margin;
    const float cx = lerp_u(lo, k.map_w - lo, u01(w.x)), cy = lerp_u(lo, k.map_h - lo, u01(w.y));
    const float rr = lerp_u(k.obs_r_min, k.obs_r_max, u01(w.z));
    const float4 c = make_float4(cx, cy, rr, 0.0f);
    bool ok = t < k.max_tries;
    ok = ok && !(cx - rr < 0.0f || cx + rr > k.map_w) && !(cy - rr < 0.0f || cy + rr > k.map_h);
    for (int i = 0; i < n_r && ok; ++i) {  // ENV:428-436
      const float4 o = sg[i];
      ok = !(dist2f(cx, cy, o.x, o.y) < rr + k.clear_r) && !(dist2f(cx, cy, o.z, o.w) < rr + k.clear_r);
    }
    for (int i = 0; i < n_c && ok; ++i)    // ENV:439-445
      ok = !(dist2f(cores[i].x, cores[i].y, cx, cy) <= k.core_r + rr);
    for (int i = 0; i < n_o && ok; ++i) ok = !obs_conflict(c, obs[i]);  // ENV:448-454
    n_o = resolve_round(__ballot_sync(FULL, ok), c, obs, n_o, want_o, lane,
                        [&](float4 a, float4 b) { return obs_conflict(a, b); });
  }
  __syncwarp();

  // ---- write the fresh episode: reset_robot ENV:166-176, Robot.reset_state WAMV:177-193 ----
  for (int r = lane; r < RP; r += 32) {
    const size_t g = (size_t)e * RP + r;
    const bool on = r < n_r;
    float4 c = on ? sg[r] : make_float4(0, 0, 0, 0);
    float th = 0.0f;
    if (on) {  // init_theta ~ U(0, 2 pi), ENV:171
      uint4 w = philox_at(seed, STREAM_RESET, episode, ge, 3u, (uint32_t)r, 0u);
      th = fminf(TWO_PI * u01(w.x), 6.2831850f);
    }
    st.x[g] = c.x; st.y[g] = c.y; st.theta[g] = th;
    st.velr_x[g] = 0; st.velr_y[g] = 0; st.velr_r[g] = 0;
    st.vel_x[g] = 0; st.vel_y[g] = 0; st.vel_r[g] = 0;  // no cores exist yet at ENV:175
    st.thrust_l[g] = 0; st.thrust_r[g] = 0;
    st.goal_x[g] = c.z; st.goal_y[g] = c.w;
    st.flags[g] = 0;
  }
  for (int i = lane; i < OP; i += 32) st.obstacles[(size_t)e * OP + i] = i < n_o ? obs[i] : make_float4(0, 0, 0, 0);
  for (int i = lane; i < CP; i += 32) st.cores[(size_t)e * CP + i] = i < n_c ? cores[i] : make_float4(0, 0, 0, 0);
  if (lane == 0) {
    st.n_robots[e] = n_r; st.n_obs[e] = n_o; st.n_cores[e] = n_c;
    st.ep_steps[e] = 0;  // ENV:96
  }
}

}  // namespace asvrl

using namespace asvrl;

extern "C" void asvrl_reset_default_params(asvrl_reset_params* p) {
  if (!p) return;
  p->map_w = 55.0; p->map_h = 55.0; p->start_margin = 2.0; p->obs_margin = 5.0;
  p->clear_r = 10.0; p->core_r = 0.5; p->v_rel_max = 1.0; p->p = 0.8;
  p->v_min = 3.0; p->v_max = 3.0; p->obs_r_min = 1.0; p->obs_r_max = 1.0;
  p->max_tries = 500; p->reserved = 0;
}

extern "C" int asvrl_env_reset(const asvrl_reset_params* params, const asvrl_env_state* state,
                               const uint8_t* reset_mask, const int32_t* want_robots,
                               const int32_t* want_cores, const int32_t* want_obs,
                               const float* min_start_goal_dis, uint64_t seed, uint64_t episode,
                               int n_envs, int max_robots, int max_obs, int max_cores,
                               int64_t env_offset, const uint64_t* counter_dev, void* stream) {
  if (!params || !state || !want_robots || !want_cores || !want_obs || !min_start_goal_dis)
    return fail(ASVRL_E_BADARG, "asvrl_env_reset: null argument");
  if (n_envs < 0 || max_robots < 1 || max_obs < 0 || max_cores < 0)
    return fail(ASVRL_E_BADARG, "asvrl_env_reset: bad sizes");
  if (max_robots > RS_MAXR || max_obs > RS_MAXO || max_cores > RS_MAXC)
    return fail(ASVRL_E_LIMIT, "asvrl_env_reset: max_robots <= %d, max_obs <= %d, max_cores <= %d", RS_MAXR, RS_MAXO, RS_MAXC);
  if ((max_obs > 0 && !state->obstacles) || (max_cores > 0 && !state->cores))
    return fail(ASVRL_E_BADARG, "asvrl_env_reset: null obstacle / core table");
  if (n_envs == 0) return 0;
  ResetConsts k = {(float)params->map_w, (float)params->map_h, (float)params->start_margin,
                   (float)params->obs_margin, (float)params->clear_r, (float)params->core_r,
                   (float)params->v_rel_max, (float)params->p, (float)params->v_min,
                   (float)params->v_max, (float)params->obs_r_min, (float)params->obs_r_max,
                   params->max_tries};
  ResetState st = {state->x, state->y, state->theta, state->velr_x, state->velr_y, state->velr_r,
                   state->vel_x, state->vel_y, state->vel_r, state->thrust_l, state->thrust_r,
                   state->goal_x, state->goal_y, state->flags, state->n_robots, state->n_obs,
                   state->n_cores, state->ep_steps, (float4*)state->obstacles, (float4*)state->cores};
  const int grid = (n_envs + RS_WARPS - 1) / RS_WARPS;
  env_reset_kernel<<<grid, RS_WARPS * 32, 0, (cudaStream_t)stream>>>(
      k, st, reset_mask, want_robots, want_cores, want_obs, min_start_goal_dis, seed, episode,
      n_envs, max_robots, max_obs, max_cores, env_offset, counter_dev);
  return check_launch("asvrl_env_reset");
}
