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
#include "env_reset.cuh"

namespace asvrl {

__global__ void __launch_bounds__(RS_WARPS * 32)
env_reset_kernel(const __grid_constant__ ResetConsts k, ResetState st,
                 const uint8_t* __restrict__ reset_mask, const int32_t* __restrict__ want_robots,
                 const int32_t* __restrict__ want_cores, const int32_t* __restrict__ want_obs,
                 const float* __restrict__ min_sg, uint64_t seed, uint64_t episode, int E, int RP,
                 int OP, int CP, int64_t env_offset, const uint64_t* __restrict__ counter_dev,
                 const int32_t* __restrict__ episode_arr) {
  __shared__ float4 s_sg[RS_WARPS][RS_MAXR];    // start.xy, goal.xy
  __shared__ float4 s_core[RS_WARPS][RS_MAXC];  // x, y, +-1, Gamma
  __shared__ float4 s_obs[RS_WARPS][RS_MAXO];   // x, y, r, 0
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int e = blockIdx.x * RS_WARPS + warp;
  if (e >= E) return;
  if (reset_mask && !reset_mask[e]) return;
  if (counter_dev) episode += *counter_dev;
  if (episode_arr) episode = (uint64_t)episode_arr[e];  // per-env episode numbers (staging)
  reset_env_warp(k, st, e, env_offset + e, want_robots[e], want_cores[e], want_obs[e], min_sg[e], seed,
                 episode, RP, OP, CP, s_sg[warp], s_core[warp], s_obs[warp], lane);
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
                               int64_t env_offset, const uint64_t* counter_dev,
                               const int32_t* episode_per_env, void* stream) {
  if (!params || !state || !want_robots || !want_cores || !want_obs || !min_start_goal_dis)
    return fail(ASVRL_E_BADARG, "asvrl_env_reset: null argument");
  if (n_envs < 0 || max_robots < 1 || max_obs < 0 || max_cores < 0)
    return fail(ASVRL_E_BADARG, "asvrl_env_reset: bad sizes");
  if (max_robots > RS_MAXR || max_obs > RS_MAXO || max_cores > RS_MAXC)
    return fail(ASVRL_E_LIMIT, "asvrl_env_reset: max_robots <= %d, max_obs <= %d, max_cores <= %d", RS_MAXR, RS_MAXO, RS_MAXC);
  if ((max_obs > 0 && !state->obstacles) || (max_cores > 0 && !state->cores))
    return fail(ASVRL_E_BADARG, "asvrl_env_reset: null obstacle / core table");
  if (n_envs == 0) return 0;
  ResetConsts k = make_reset_consts(params);
  ResetState st = {state->x, state->y, state->theta, state->velr_x, state->velr_y, state->velr_r,
                   state->vel_x, state->vel_y, state->vel_r, state->thrust_l, state->thrust_r,
                   state->goal_x, state->goal_y, state->flags, state->n_robots, state->n_obs,
                   state->n_cores, state->ep_steps, (float4*)state->obstacles, (float4*)state->cores};
  const int grid = (n_envs + RS_WARPS - 1) / RS_WARPS;
  env_reset_kernel<<<grid, RS_WARPS * 32, 0, (cudaStream_t)stream>>>(
      k, st, reset_mask, want_robots, want_cores, want_obs, min_start_goal_dis, seed, episode,
      n_envs, max_robots, max_obs, max_cores, env_offset, counter_dev, episode_per_env);
  return check_launch("asvrl_env_reset");
}

// ---- staging of next episodes on a side stream (asvrl_env_stage_refill) ----
namespace asvrl {

// need[e] = the stage of env e does not hold its next episode; ep[e] = that episode's number
__global__ void stage_begin_kernel(const int32_t* __restrict__ episode_idx, const int32_t* stage_tag,
                                   uint8_t* __restrict__ need, int32_t* __restrict__ ep, int E) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const int n = __ldcg(episode_idx + e) + 1;
  int tag;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(tag) : "l"(stage_tag + e) : "memory");
  need[e] = tag != n;
  ep[e] = n;
}

// the data kernels before this one in the stream have completed: publish the tags
__global__ void stage_publish_kernel(const uint8_t* __restrict__ need, const int32_t* __restrict__ ep,
                                     int32_t* stage_tag, int E) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E || !need[e]) return;
  __threadfence();
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(stage_tag + e), "r"(ep[e]) : "memory");
}

}  // namespace asvrl

extern "C" int asvrl_env_stage_refill(const asvrl_env_params* params, const asvrl_reset_params* rparams,
                                      const asvrl_env_state* stage, const asvrl_env_out* stage_out,
                                      const int32_t* want_robots, const int32_t* want_cores,
                                      const int32_t* want_obs, const float* min_start_goal_dis,
                                      const int32_t* episode_idx, int32_t* stage_tag, uint8_t* need_ws,
                                      int32_t* episode_ws, uint64_t seed, uint64_t observe_seed,
                                      int n_envs, int max_robots, int max_obs, int max_cores,
                                      int64_t env_offset, void* stream) {
  if (!params || !rparams || !stage || !stage_out || !episode_idx || !stage_tag || !need_ws || !episode_ws)
    return fail(ASVRL_E_BADARG, "asvrl_env_stage_refill: null argument");
  if (n_envs <= 0) return n_envs == 0 ? 0 : fail(ASVRL_E_BADARG, "asvrl_env_stage_refill: bad sizes");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = (n_envs + 255) / 256;
  stage_begin_kernel<<<grid, 256, 0, s>>>(episode_idx, stage_tag, need_ws, episode_ws, n_envs);
  int rc = check_launch("asvrl_env_stage_refill(begin)");
  if (rc) return rc;
  rc = asvrl_env_reset(rparams, stage, need_ws, want_robots, want_cores, want_obs, min_start_goal_dis, seed,
                       0, n_envs, max_robots, max_obs, max_cores, env_offset, nullptr, episode_ws, stream);
  if (rc) return rc;
  rc = asvrl_env_step(params, stage, nullptr, nullptr, need_ws, observe_seed, 0, stage_out, n_envs, max_robots,
                      max_obs, max_cores, env_offset, 0u, nullptr, episode_ws, stream);
  if (rc) return rc;
  stage_publish_kernel<<<grid, 256, 0, s>>>(need_ws, episode_ws, stage_tag, n_envs);
  return check_launch("asvrl_env_stage_refill(publish)");
}
