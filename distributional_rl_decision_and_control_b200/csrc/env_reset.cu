// env_reset.cu — K1b: masked per-env scenario re-generation (sm_100a).
//
// Replaces MarineNavEnv3.reset (rfarl/rfarl/envs/marinenav/env.py:72-176) with the acceptance
// predicates of check_start_and_goal / check_core / check_obstacle (env.py:358-456).
//
// One warp per environment that needs a reset.  The reference's loop "draw a candidate, keep it
// if it passes the checks against everything kept so far, give up after 500 draws" is sequential
// only through the accepted set, so the warp evaluates 32 consecutive candidates of the
// counter-indexed sequence at once and takes the FIRST passing one (ballot + ffs); candidates
// after it are re-evaluated against the grown set in the next round.  The result is exactly the
// sequential algorithm applied to the Philox candidate sequence.
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
  for (int base = 0; base < k.max_tries && n_r < want_r;) {
    const int t = base + lane;
    uint4 w = philox_at(seed, STREAM_RESET, episode, ge, 0u, (uint32_t)t >> 8, (uint32_t)t & 255u);
    const float lo = k.start_margin;
    float4 c = make_float4(lerp_u(lo, k.map_w - lo, u01(w.x)), lerp_u(lo, k.map_h - lo, u01(w.y)),
                           lerp_u(lo, k.map_w - lo, u01(w.z)), lerp_u(lo, k.map_h - lo, u01(w.w)));
    bool ok = t < k.max_tries;
    float dx = c.z - c.x, dy = c.w - c.y;
    ok = ok && !(sqrtf(dx * dx + dy * dy) < min_dis);  // ENV:361
    for (int i = 0; i < n_r && ok; ++i) {
      float4 o = sg[i];
      float sx = o.x - c.x, sy = o.y - c.y, gx = o.z - c.z, gy = o.w - c.w;
      ok = !(sqrtf(sx * sx + sy * sy) <= k.clear_r) && !(sqrtf(gx * gx + gy * gy) <= k.clear_r);  // ENV:368-374
    }
    const unsigned bal = __ballot_sync(FULL, ok);
    if (bal == 0) { base += 32; continue; }
    const int first = __ffs(bal) - 1;
    if (lane == first) sg[n_r] = c;
    __syncwarp();
    ++n_r;
    base += first + 1;
  }

  // ---- vortex cores: ENV:122-136, check_core ENV:378-418 ----
  const int want_c = min(want_cores[e], CP);
  int n_c = 0;
  for (int base = 0; base < k.max_tries && n_c < want_c;) {
    const int t = base + lane;
    uint4 w = philox_at(seed, STREAM_RESET, episode, ge, 1u, (uint32_t)t >> 8, (uint32_t)t & 255u);
    const float cx = k.map_w * u01(w.x), cy = k.map_h * u01(w.y);
    const float dir = (w.z >> 31) ? 1.0f : -1.0f;                    // binomial(1, 0.5), ENV:127
    const float gam = TWO_PI * k.core_r * lerp_u(k.v_min, k.v_max, u01(w.w));  // ENV:128-129
    bool ok = t < k.max_tries;
    ok = ok && !(cx - k.core_r < 0.0f || cx + k.core_r > k.map_w);   // ENV:381
    ok = ok && !(cy - k.core_r < 0.0f || cy + k.core_r > k.map_w);   // ENV:383 (sic: width)
    for (int i = 0; i < n_r && ok; ++i) {                            // ENV:386-394
      float4 o = sg[i];
      float sx = cx - o.x, sy = cy - o.y, gx = cx - o.z, gy = cy - o.w;
      ok = !(sqrtf(sx * sx + sy * sy) < k.core_r + k.clear_r) && !(sqrtf(gx * gx + gy * gy) < k.core_r + k.clear_r);
    }
    for (int i = 0; i < n_c && ok; ++i) {                            // ENV:396-416
      float4 o = cores[i];
      float dx = o.x - cx, dy = o.y - cy, dis = sqrtf(dx * dx + dy * dy);
      if (o.z == dir) {
        float bi = o.w / (TWO_PI * k.v_rel_max), bj = gam / (TWO_PI * k.v_rel_max);
        ok = !(dis < bi + bj);
      } else {
        float gl = fmaxf(o.w, gam), gs = fminf(o.w, gam);
        float v1 = gl / (TWO_PI * (dis - 2.0f * k.core_r)), v2 = gs / (TWO_PI * k.core_r);
        ok = !(v1 > k.p * v2);
      }
    }
    const unsigned bal = __ballot_sync(FULL, ok);
    if (bal == 0) { base += 32; continue; }
    const int first = __ffs(bal) - 1;
    if (lane == first) cores[n_c] = make_float4(cx, cy, dir, gam);
    __syncwarp();
    ++n_c;
    base += first + 1;
  }

  // ---- static obstacles: ENV:150-162, check_obstacle ENV:420-456 ----
  const int want_o = min(want_obs[e], OP);
  int n_o = 0;
  for (int base = 0; base < k.max_tries && n_o < want_o;) {
    const int t = base + lane;
    uint4 w = philox_at(seed, STREAM_RESET, episode, ge, 2u, (uint32_t)t >> 8, (uint32_t)t & 255u);
    const float lo = k.obs_margin;
    const float cx = lerp_u(lo, k.map_w - lo, u01(w.x)), cy = lerp_u(lo, k.map_h - lo, u01(w.y));
    const float rr = lerp_u(k.obs_r_min, k.obs_r_max, u01(w.z));
    bool ok = t < k.max_tries;
    ok = ok && !(cx - rr < 0.0f || cx + rr > k.map_w) && !(cy - rr < 0.0f || cy + rr > k.map_h);
    for (int i = 0; i < n_r && ok; ++i) {  // ENV:428-436
      float4 o = sg[i];
      float sx = cx - o.x, sy = cy - o.y, gx = cx - o.z, gy = cy - o.w;
      ok = !(sqrtf(sx * sx + sy * sy) < rr + k.clear_r) && !(sqrtf(gx * gx + gy * gy) < rr + k.clear_r);
    }
    for (int i = 0; i < n_c && ok; ++i) {  // ENV:439-445
      float dx = cores[i].x - cx, dy = cores[i].y - cy;
      ok = !(sqrtf(dx * dx + dy * dy) <= k.core_r + rr);
    }
    for (int i = 0; i < n_o && ok; ++i) {  // ENV:448-454
      float dx = obs[i].x - cx, dy = obs[i].y - cy;
      ok = !(sqrtf(dx * dx + dy * dy) <= obs[i].z + rr);
    }
    const unsigned bal = __ballot_sync(FULL, ok);
    if (bal == 0) { base += 32; continue; }
    const int first = __ffs(bal) - 1;
    if (lane == first) obs[n_o] = make_float4(cx, cy, rr, 0.0f);
    __syncwarp();
    ++n_o;
    base += first + 1;
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
