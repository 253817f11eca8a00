/*
 * asvrl.h — C-ABI of libasvrl.so, the B200 (sm_100a) hot path of
 * pszenher/Distributional_RL_Decision_and_Control's `rfarl` package.
 *
 * Nothing like this boundary exists in the reference (pure Python);
 * each entry point below cites the reference code it replaces (paths relative to
 * the reference root, abbreviations as in SURVEY.md):
 *   ENV   = rfarl/rfarl/envs/marinenav/env.py
 *   WAMV  = rfarl/rfarl/envs/marinenav/vehicles/wamv.py
 *   AGENT = rfarl/rfarl/agent.py
 *   ACIQN = rfarl/rfarl/policy/AC_IQN_model.py
 *   IQN   = rfarl/rfarl/policy/IQN_model.py
 *   RB    = rfarl/rfarl/utils/replay_buffer.py
 *
 * Conventions
 *  - every function returns int: 0 = ok, >0 = a cudaError_t, <0 = ASVRL_E_* below;
 *    no exceptions, no aborts, nothing is printed; asvrl_last_error_string() explains.
 *  - all data pointers are DEVICE pointers on the current device, owned by the caller
 *    (PyTorch's allocator in the shipped host code); the library allocates nothing
 *    persistent and keeps no mutable global state besides the per-thread error string.
 *  - kernels are enqueued on `stream` (a cudaStream_t passed as void*) and the call
 *    returns without synchronising.
 *  - the CPU oracle (oracle/asv_oracle.h, test infrastructure only) mirrors these structs
 *    with HOST pointers and float64 arrays, and shares asvrl_env_params verbatim.
 */
#ifndef ASVRL_H_
#define ASVRL_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define ASVRL_API __attribute__((visibility("default")))
#else
#define ASVRL_API
#endif

#define ASVRL_VERSION 200 /* 0.2.0: round 2 (K9 - K13, asvrl_c51_greedy layout arguments, static-scratch tower backward) */

/* fixed by the network input shape ((1,7),(1,5,5),(1,5)): WAMV:14, AGENT:20-22 */
#define ASVRL_MAX_OBJ 5
#define ASVRL_SELF_DIM 7
#define ASVRL_OBJ_DIM 5
/* noise draws per candidate object: N px, N py, N vx, N vy, vonMises r  (WAMV:27-40) */
#define ASVRL_NOISE_DIM 5
/* hard limit on candidate slots (max_obs + max_robots) per ego: detected mask is 64 bit */
#define ASVRL_MAX_SLOTS 64

/* library error codes (negative) */
#define ASVRL_E_BADARG (-1)
#define ASVRL_E_LIMIT (-2)  /* a size exceeds a compiled limit */
#define ASVRL_E_NODEV (-3)

/* per-ASV sticky flags word (state.flags), WAMV:130-132 */
#define ASVRL_F_COLLISION 1u
#define ASVRL_F_REACH_GOAL 2u
#define ASVRL_F_DEACTIVATED 4u

/* per-ASV status word written by a step (out.status):
 *   bit 0      done                      ENV:290-328
 *   bit 1      collision   (sticky, after this step's perception)
 *   bit 2      reach_goal  (sticky, after this step's perception)
 *   bit 3      deactivated (value the step STARTED with; the transition of a robot that was
 *              already deactivated is not stored, TRAINER:157-159)
 *   bit 4      apply_COLREGs             WAMV:517-521
 *   bits 8..11 info code below           ENV:291-328
 */
#define ASVRL_S_DONE 1u
#define ASVRL_S_COLLISION 2u
#define ASVRL_S_REACH_GOAL 4u
#define ASVRL_S_WAS_DEACTIVATED 8u
#define ASVRL_S_COLREGS 16u
#define ASVRL_S_INFO_SHIFT 8
#define ASVRL_INFO_NORMAL 0
#define ASVRL_INFO_COLLISION 1
#define ASVRL_INFO_REACH_GOAL 2
#define ASVRL_INFO_TOO_LONG 3
#define ASVRL_INFO_DEACT_COLLISION 4
#define ASVRL_INFO_DEACT_GOAL 5
#define ASVRL_INFO_DEACT_INVALID 6 /* ENV:303 RuntimeError case */
#define ASVRL_INFO_ABSENT 7        /* padding slot r >= n_robots[env] */

/* step mode bits */
#define ASVRL_M_DYNAMICS 1u        /* 0 = perception only (ENV:341-356 as used by reset, ENV:164) */
#define ASVRL_M_DISCRETE 2u        /* actions are int32 ids into the 5x5 thrust table, WAMV:146-147,225 */
#define ASVRL_M_AUTO_DEACTIVATE 4u /* fold TRAINER:168-170 into the step (batched API) */

/* All scalar parameters of the path, float64 as in the reference.  The CUDA library rounds
 * them to fp32 once per launch.  One parameter set per batch (the reference allows per-robot
 * hull parameters through reset_with_eval_config, ENV:551-607; every shipped config uses the
 * defaults, and the host façade rejects heterogeneous robots). */
typedef struct asvrl_env_params {
  /* integration, WAMV:46-47 */
  double dt;
  int32_t n_substeps;
  int32_t max_episode_steps; /* ENV:312 (1000) */
  /* hull, WAMV:51-54,79,84-87 */
  double length, width, robot_r, goal_dis;
  double min_thrust, max_thrust;
  double thrust_rate;           /* continuous action scale, N/s per unit action (1000), WAMV:222-223 */
  double thrust_change_l[5];    /* discrete deltas N/s, WAMV:86 */
  double thrust_change_r[5];    /* WAMV:87 */
  double left_pos, right_pos;   /* thruster angles, WAMV:98-99,137-138 (0 in training) */
  /* rigid body + hydrodynamic derivatives, WAMV:103-125 */
  double m, Izz;
  double xDotU, yDotV, yDotR, nDotR, nDotV;
  double xU, xUU, yV, yVV, yR, yRV, yVR, yRR, nR, nRR, nV, nVV, nRV, nVR;
  /* perception, WAMV:12-25 */
  double perc_range, perc_angle;
  double pos_std, vel_std, r_kappa, r_mean_ratio;
  /* COLREGs zones, WAMV:73-76,400-406 */
  double head_on_x, head_on_y;
  double lc_x0, lc_x1, lc_y0, lc_y1; /* left-crossing x in [-9,12], y in [-17,0], cut at y=-7 */
  double colregs_min_speed;          /* 0.5 */
  /* rewards, ENV:47-50 */
  double timestep_penalty, colregs_penalty, collision_penalty, goal_reward;
  /* current field, ENV:35 */
  double core_r;
  /* extension terms with NO reference counterpart: uniform wind/wave drift
   * velocity added to the current, world frame.  Must be 0 for parity; default 0. */
  double drift_vx, drift_vy;
} asvrl_env_params;

/* Structure-of-arrays environment state in HBM.  E = n_envs, RP = max_robots (robot index
 * fastest), OP = max_obs, CP = max_cores.  The 13 per-ASV floats are the persistent attributes
 * of WAMV.Robot (WAMV:92-101,128-129); `vel_*` is the STALE sea-floor velocity. */
typedef struct asvrl_env_state {
  float* x;        /* [E][RP] */
  float* y;
  float* theta;
  float* velr_x;   /* velocity_r, world frame, relative to water */
  float* velr_y;
  float* velr_r;
  float* vel_x;    /* velocity, world frame, over ground (stale by one sub-step) */
  float* vel_y;
  float* vel_r;
  float* thrust_l;
  float* thrust_r;
  float* goal_x;
  float* goal_y;
  uint32_t* flags;    /* [E][RP] ASVRL_F_* */
  int32_t* n_robots;  /* [E] */
  int32_t* n_obs;     /* [E] */
  int32_t* n_cores;   /* [E] */
  int32_t* ep_steps;  /* [E] episode_timesteps, ENV:64,330 */
  float* obstacles;   /* [E][OP][4] = x, y, r, 0         ENV:16-22 */
  float* cores;       /* [E][CP][4] = x, y, +1 clockwise / -1 counter-clockwise, Gamma   ENV:7-14 */
} asvrl_env_state;

/* Everything one step produces besides the in-place state update. */
typedef struct asvrl_env_out {
  float* obs_self;     /* [E][RP][7]      WAMV:443-453 */
  float* obs_objects;  /* [E][RP][5][5]   WAMV:513-515, zero padded (RB:64) */
  float* obs_mask;     /* [E][RP][5]      1 = object present (RB:65) */
  float* reward;       /* [E][RP]         ENV:271-277,310-323 */
  uint32_t* status;    /* [E][RP]         ASVRL_S_* */
  float* phi;          /* [E][RP]         COLREGs turn angle, valid when ASVRL_S_COLREGS */
  int32_t* top_ids;    /* [E][RP][5]      candidate slot of each kept object (obstacle i -> i,
                                          robot j -> OP + j), -1 = empty; nullable */
  uint64_t* detected;  /* [E][RP]         bit s = candidate slot s detected (observed_obs /
                                          observed_objs, WAMV:473,500); nullable */
  uint32_t* env_status;/* [E]             bit0 all deactivated (ENV:178-184), bit1 timeout
                                          (ep_steps BEFORE this step >= max), bit2 any collision
                                          (ENV:194-200), bit3 all reach goal (ENV:186-192);
                                          low two bits != 0 <=> episode over (TRAINER:172) */
  uint8_t* ended;      /* [E]             1 iff (env_status & 3) != 0; nullable.  Feeds
                                          asvrl_env_reset(reset_mask) / asvrl_env_step(env_mask)
                                          directly so the auto-reset needs no host round trip */
} asvrl_env_out;

/* Counter-based RNG contract (Philox4x32-10, Salmon et al. SC'11), shared by every kernel and
 * restated by the oracle so that streams are reproducible and invariant to the GPU count:
 *   key = (seed lo, seed hi)
 *   ctr = (block | slot<<8 | ego<<16,  global env index,  step lo,  step hi (24 bit) | stream<<24)
 *   stream: 1 sensor noise, 2 reset, 3 tau, 4 replay sampling, 5 exploration (asvrl_actor_head: the
 *           "env" word is the row index; words 0..2 = explore?, action 0, action 1)
 *   uniform(word) = ((word >> 8) + 0.5) * 2^-24
 * Sensor noise of (env, ego, slot, step): block 0 -> two Box-Muller pairs (n_px,n_py)*pos_std,
 * (n_vx,n_vy)*vel_std; blocks 1..8 -> Best-Fisher von Mises attempts, two per block
 * ((U,V) = words (0,1) then (2,3), sign = bit 0 of the V word). */

ASVRL_API int asvrl_version(void);
ASVRL_API const char* asvrl_last_error_string(void);
/* fills *p with the reference defaults (WAMV:45-126, ENV:33-54) */
ASVRL_API void asvrl_env_default_params(asvrl_env_params* p);

/* K1 — replaces MarineNavEnv3.step (ENV:240-333) incl. Robot.update_state/compute_motion
 * (WAMV:204-279), get_velocity (ENV:458-501), get_observations/perception_output
 * (ENV:341-356, WAMV:436-529), COLREGs (WAMV:324-423) and reward/done (ENV:271-328).
 *   actions   [E][RP][2] float (continuous) or [E][RP] int32 (ASVRL_M_DISCRETE); ignored for
 *             deactivated robots (the reference passes None, TRAINER:112-114)
 *   noise     nullable [E][RP][OP+RP][5] float injected sensor noise in draw order
 *             (n_px, n_py, n_vx, n_vy already scaled by std; raw von Mises sample w); slot of
 *             obstacle i is i, of robot j is OP+j.  NULL = in-kernel Philox stream above.
 *   env_mask  nullable [E] uint8: envs with 0 are skipped entirely (nothing read or written);
 *             used to observe only the freshly reset envs (ENV:164).
 *   env_offset  global index of local env 0 (multi-GPU sharding keeps streams invariant)
 *   counter_dev nullable DEVICE uint64: the RNG step used is `step + *counter_dev`.  Lets a
 *             captured CUDA graph be replayed with fresh random streams (the host bumps the
 *             device word inside the graph); the same convention holds for asvrl_env_reset
 *             (episode), asvrl_replay_sample (counter) and asvrl_iqn_tau_cos (counter).
 * Table sizes max_robots = 8, max_obs = 10, max_cores = 8 or 0 (BASELINE config 2) run a
 * kernel instantiation with those sizes as compile-time constants; results are bit-identical to
 * the run-time-sized one, which ASVRL_K1_GENERIC=1 in the process environment (read once) forces. */
ASVRL_API int asvrl_env_step(const asvrl_env_params* params, const asvrl_env_state* state,
                             const void* actions, const float* noise, const uint8_t* env_mask,
                             uint64_t seed, uint64_t step, const asvrl_env_out* out, int n_envs,
                             int max_robots, int max_obs, int max_cores, int64_t env_offset,
                             uint32_t mode, const uint64_t* counter_dev,
                             const int32_t* step_per_env /* nullable [E]: RNG step of env e's sensor
                                                            noise instead of `step` */,
                             void* stream);

/* The noise table the NULL-noise path of K1 draws: float [E][RP][OP+RP][5] (test / debug aid;
 * same device function as K1). */
ASVRL_API int asvrl_env_noise(const asvrl_env_params* params, uint64_t seed, uint64_t step,
                              int n_envs, int max_robots, int max_obs, int64_t env_offset,
                              float* noise_out, void* stream);

/* scenario-generation constants of MarineNavEnv3 (ENV:33-41, 109-110, 126-129, 154-155) */
typedef struct asvrl_reset_params {
  double map_w, map_h;        /* 55 x 55 */
  double start_margin;        /* starts / goals in [2, w-2] x [2, h-2]   ENV:109-110 */
  double obs_margin;          /* obstacle centres in [5, w-5] x [5, h-5] ENV:154 */
  double clear_r;             /* 10   ENV:40 */
  double core_r;              /* 0.5  ENV:35 */
  double v_rel_max, p;        /* 1.0, 0.8  ENV:36-37 */
  double v_min, v_max;        /* vortex edge speed range [3,3]  ENV:38 */
  double obs_r_min, obs_r_max;/* [1,1]  ENV:39 */
  int32_t max_tries;          /* 500  ENV:107,124,152 */
  int32_t reserved;
} asvrl_reset_params;
ASVRL_API void asvrl_reset_default_params(asvrl_reset_params* p);

/* K1b — replaces MarineNavEnv3.reset (ENV:72-176) for every env with reset_mask != 0
 * (NULL = all): rejection-samples starts/goals, cores, obstacles with the reference's
 * acceptance rules (ENV:358-456) over a counter-indexed candidate sequence (candidate t of a
 * class is accepted iff it passes the checks against everything accepted before it, exactly
 * the sequential loop of the reference; at most max_tries candidates per class), zeroes
 * velocity/thrust, draws theta0 (ENV:166-176).  Counts requested per env come from
 * want_robots/want_cores/want_obs [E] (the curriculum lookup ENV:75-87 is done by the host).
 * Perception of the fresh state is a following asvrl_env_step(mode=0, env_mask=reset_mask). */
ASVRL_API int asvrl_env_reset(const asvrl_reset_params* params, const asvrl_env_state* state,
                              const uint8_t* reset_mask, const int32_t* want_robots,
                              const int32_t* want_cores, const int32_t* want_obs,
                              const float* min_start_goal_dis /* [E] */, uint64_t seed,
                              uint64_t episode, int n_envs, int max_robots, int max_obs,
                              int max_cores, int64_t env_offset, const uint64_t* counter_dev,
                              const int32_t* episode_per_env /* nullable [E]: replaces `episode` */,
                              void* stream);

/* Auto-reset arguments of the fused step: what the Trainer does between two steps of one env
 * (TRAINER:172, 240-245: `if end_episode: states, _, _ = train_env.reset()`). */
typedef struct asvrl_env_autoreset {
  const asvrl_reset_params* reset_params;
  const int32_t* want_robots;          /* [E] requested counts of the NEXT episode (ENV:75-87) */
  const int32_t* want_cores;           /* [E] */
  const int32_t* want_obs;             /* [E] */
  const float* min_start_goal_dis;     /* [E] */
  int32_t* episode_idx;                /* [E] in/out: episodes generated so far for each env.  The
                                          scenario RNG counter of env e's next episode (the
                                          `episode` argument of asvrl_env_reset) is episode_idx[e]+1,
                                          and so is the `step` of its first observation's noise */
  uint64_t observe_seed;               /* seed of the sensor-noise stream of the first observation */
  float* term_self;                    /* nullable [E][RP][7]: rows of envs with ended=1 receive */
  float* term_objects;                 /* nullable [E][RP][5][5]  the TERMINAL observation (the  */
  float* term_mask;                    /* nullable [E][RP][5]     replay row's next state)       */
  const asvrl_env_state* stage;        /* nullable: a second, caller-owned state block of the same
                                          shape holding PRE-GENERATED next episodes */
  int32_t* stage_tag;                  /* [E] with stage: episode number the stage of env e holds
                                          (0 = none; zero-initialise, and zero again whenever
                                          want_* / min_start_goal_dis / reset_params change) */
  const float* stage_self;             /* nullable, with stage: the staged episodes' FIRST        */
  const float* stage_objects;          /*   observation rows ([E][RP][7], [E][RP][5][5],          */
  const float* stage_mask;             /*   [E][RP][5]) as asvrl_env_stage_refill writes them     */
  int32_t refill_in_launch;            /* != 0: a few extra CTAs of this launch re-generate consumed
                                          stages (scenario only); 0: the caller runs
                                          asvrl_env_stage_refill, e.g. on a side stream */
  int32_t reserved;
} asvrl_env_autoreset;

/* K1 + K1b + first observation in ONE launch.  For every env, equal bit for bit to
 *   asvrl_env_step(mode, env_mask=NULL, noise=NULL)                       (writes out, ended)
 * and then, for each env e with out->ended[e] != 0, with n = ++episode_idx[e]:
 *   asvrl_env_reset(reset_mask={e}, seed, episode=n)
 *   asvrl_env_step(mode=0, env_mask={e}, seed=ar->observe_seed, step=n)
 * except that reward / status / phi / env_status / ended keep describing the TRANSITION for the
 * envs that ended (the call sequence overwrites them with the fresh episode's), and the terminal
 * observation rows of those envs are kept in ar->term_*.  out->obs_* hold, for every env, the
 * observation to act on next.
 * Scenario generation is a pure function of (seed, env, episode number), which is what lets it
 * leave the critical path: with `stage`, a few extra CTAs of the same launch pre-generate the
 * NEXT episode of every env whose stage was consumed (K1b's warp routine, concurrently with the
 * step CTAs), and a CTA whose env ends just copies the staged scenario and re-runs the perception
 * phases for it.  An env whose stage is not ready (first steps, bursts) is generated in place by
 * its CTA — same result, longer tail.  Requires ASVRL_M_DYNAMICS; counter_dev is added to `step`. */
ASVRL_API int asvrl_env_step_autoreset(const asvrl_env_params* params, const asvrl_env_state* state,
                                       const void* actions, uint64_t seed, uint64_t step,
                                       const asvrl_env_out* out, const asvrl_env_autoreset* ar,
                                       int n_envs, int max_robots, int max_obs, int max_cores,
                                       int64_t env_offset, uint32_t mode,
                                       const uint64_t* counter_dev, void* stream);

/* Re-generates, on `stream`, the stage of every env whose stage does not hold its next episode
 * (stage_tag[e] != episode_idx[e] + 1): K1b into `stage` with episode number episode_idx[e] + 1,
 * the perception-only K1 pass over it into `stage_out` (noise step = that number, seed =
 * observe_seed), then publishes stage_tag.  Safe to run CONCURRENTLY with
 * asvrl_env_step_autoreset on another stream: tags are written with release / read with acquire
 * semantics, a stage is only rewritten while its tag is stale, and an env that ends while its
 * stage is not ready is generated in place with the same result.
 *   need_ws uint8[E], episode_ws int32[E]: scratch owned by the caller. */
ASVRL_API int asvrl_env_stage_refill(const asvrl_env_params* params, const asvrl_reset_params* rparams,
                                     const asvrl_env_state* stage, const asvrl_env_out* stage_out,
                                     const int32_t* want_robots, const int32_t* want_cores,
                                     const int32_t* want_obs, const float* min_start_goal_dis,
                                     const int32_t* episode_idx, int32_t* stage_tag, uint8_t* need_ws,
                                     int32_t* episode_ws, uint64_t seed, uint64_t observe_seed,
                                     int n_envs, int max_robots, int max_obs, int max_cores,
                                     int64_t env_offset, void* stream);

/* Replay rows are dense float32[ASVRL_ROW] records: RB:22-69 + AGENT:360-368,387-392 */
#define ASVRL_ROW_SELF 0     /* 7  */
#define ASVRL_ROW_OBJ 7      /* 25 */
#define ASVRL_ROW_NOBJ 32    /* 1: number of valid objects (mask = first n ones) */
#define ASVRL_ROW_ACT 33     /* 2: continuous action, or (float)id, 0 */
#define ASVRL_ROW_REW 35     /* 1 */
#define ASVRL_ROW_NSELF 36   /* 7  */
#define ASVRL_ROW_NOBJ_F 43  /* 25 */
#define ASVRL_ROW_NNOBJ 68   /* 1 */
#define ASVRL_ROW_DONE 69    /* 1 */
#define ASVRL_ROW 70

/* K2a — replaces ReplayBuffer.add over a whole batched step (RB:22-24, TRAINER:157-166):
 * appends one row per ASV whose status has neither ASVRL_S_WAS_DEACTIVATED nor
 * ASVRL_INFO_ABSENT, in (env, robot) order (deterministic), at the ring position given by the
 * device counters head_size[0] = rows ever written, head_size[1] = rows valid = min(total,
 * capacity).  workspace: uint32[ceil(n_asv/64) + 4], zero-initialised once by the caller and
 * left clean by every call.  After asvrl_env_step_autoreset the next state of an ASV whose env
 * ended (ended[i / robots_per_env] != 0) is read from term_* instead of next_* (all four
 * nullable together: NULL = always next_*). */
ASVRL_API int asvrl_replay_push(float* rows, int64_t capacity, unsigned long long* head_size,
                                const float* obs_self, const float* obs_objects,
                                const float* obs_mask, const void* actions, int discrete,
                                const float* reward, const float* next_self,
                                const float* next_objects, const float* next_mask,
                                const uint32_t* status, int n_asv, uint32_t* workspace,
                                const float* term_self, const float* term_objects,
                                const float* term_mask, const uint8_t* ended, int robots_per_env,
                                void* stream);

/* K2b — replaces ReplayBuffer.sample + state_batch + Agent.state_to_tensor (RB:26-69,
 * AGENT:387-392): draws `batch` DISTINCT uniform row indices (a keyed Feistel permutation of
 * [0,size) walked from 0, i.e. sampling without replacement like random.sample, RB:28) or
 * takes them from `indices` (int64[batch], nullable) and gathers into dense tensors
 * self [B][7], obj [B][5][5], mask [B][5], act [B][2], rew [B], done [B].  Requires
 * size >= 1 (and size >= batch for distinct indices; otherwise indices repeat modulo size). */
ASVRL_API int asvrl_replay_sample(const float* rows, const unsigned long long* head_size,
                                  int64_t capacity, const int64_t* indices, uint64_t seed,
                                  uint64_t counter, int batch, float* self_b, float* obj_b,
                                  float* mask_b, float* act_b, float* rew_b, float* nself_b,
                                  float* nobj_b, float* nmask_b, float* done_b,
                                  int64_t* indices_out /* nullable */,
                                  const uint64_t* counter_dev /* nullable */, void* stream);

/* K3 — replaces Critic.calc_cos / IQN_Policy.calc_cos (ACIQN:410-426, IQN:56-72):
 * tau = U(0,1) * cvar (Philox stream 3 keyed on seed, counter) unless tau_in is given;
 * cos_out[b][n][i] = cosf(tau * float32(pi*i)), i < n_cos.  tau_out [B][N] (nullable),
 * cos_out [B][N][n_cos]. */
ASVRL_API int asvrl_iqn_tau_cos(const float* tau_in /* nullable */, uint64_t seed,
                                uint64_t counter, float cvar, int batch, int n_tau, int n_cos,
                                float* tau_out, float* cos_out,
                                const uint64_t* counter_dev /* nullable */, void* stream);

/* K4 — replaces the loss core of Agent.train_AC_IQN / train_IQN (AGENT:401-412, 455-467,
 * 701-707) and its autograd backward:
 *   T_bj = r_b + gamma (1 - d_b) Z'_bj ; delta_bij = T_bj - Z_bi ;
 *   loss_b = (1/N') sum_j sum_i |tau_bi - 1[delta<0]| * huber_kappa(delta) / kappa
 *   loss   = mean_b loss_b ;  grad_z_bi = dloss/dZ_bi.
 * z [B][N], tau [B][N], z_next [B][N'], r [B], done [B]; outputs loss_per_sample [B],
 * grad_z [B][N], loss [1].  workspace: one zero-initialised uint32 (left zero on completion). */
ASVRL_API int asvrl_qhuber_fwd_bwd(const float* z, const float* tau, const float* z_next,
                                   const float* r, const float* done, float gamma, float kappa,
                                   int batch, int n, int n_next, float* loss_per_sample,
                                   float* grad_z, float* loss, uint32_t* workspace, void* stream);

/* K5 — the element-wise / reduction glue of the IQN critic's quantile trunk (forward and the
 * autograd backward of ACIQN:453-460, 470-474, 476-480); the dense layers between the calls stay
 * plain GEMMs.  Rows m = b * n_tau + n.  dim: multiple of 4 with (dim / 4) dividing 256.
 *   scale_relu fwd: out[m][d] = scale[b][d] * relu(pre[m][d] + bias[d])
 *   scale_relu bwd: dpre[m][d] = dout[m][d] * scale[b][d] * 1[pre+bias > 0]   (nullable)
 *                   dscale[b][d] = sum_n dout[m][d] * relu(pre[m][d] + bias[d]) (nullable)
 *                   dbias[d] = sum_m dpre[m][d]                                (nullable)
 *   head fwd:       z[m] = sum_d relu(pre[m][d] + bias[d]) * w[d] + b_out[0]
 *   head bwd:       dpre[m][d] = dz[m] * w[d] * 1[pre+bias > 0];  dw[d] = sum_m dz[m] relu(pre+bias);
 *                   dbias[d] = sum_m dpre[m][d];  db_out[0] = sum_m dz[m]   (the three nullable together)
 * partials: caller-owned scratch of asvrl_iqn_*_bwd_partials() floats; counter: one zero-initialised
 * uint32 (left zero).  Reductions are summed in a fixed order (deterministic). */
ASVRL_API int asvrl_iqn_scale_relu_fwd(const float* pre, const float* bias, const float* scale,
                                       int batch, int n_tau, int dim, float* out, void* stream);
ASVRL_API int asvrl_iqn_scale_relu_bwd(const float* dout, const float* pre, const float* bias,
                                       const float* scale, int batch, int n_tau, int dim,
                                       float* dpre /* nullable */, float* dscale /* nullable */,
                                       float* dbias /* nullable */, float* partials, uint32_t* counter,
                                       void* stream);
ASVRL_API long asvrl_iqn_scale_relu_bwd_partials(int batch, int dim);
ASVRL_API int asvrl_iqn_head_fwd(const float* pre, const float* bias, const float* w,
                                 const float* b_out /* nullable */, int rows, int dim, float* z,
                                 void* stream);
ASVRL_API int asvrl_iqn_head_bwd(const float* dz, const float* pre, const float* bias, const float* w,
                                 int rows, int dim, float* dpre, float* dw, float* dbias, float* db_out,
                                 float* partials, uint32_t* counter, void* stream);
ASVRL_API long asvrl_iqn_head_bwd_partials(int rows, int dim);

/* K6 — the observation tower of the actor and the critic (ACIQN:284-308, 428-451) in one launch:
 *   feats[b][j]                  = relu(w_self[j] . self_obs[b] + b_self[j])                 j < self_feat
 *   feats[b][self_feat+k*OF+f]   = mask[b][k] >= 0.5 ? relu(w_obj[f] . objects[b][k] + b_obj[f]) : 0
 * self_obs [B][self_dim], objects [B][max_obj][obj_dim], mask [B][max_obj], feats [B][self_feat +
 * max_obj*obj_feat]; self_dim, obj_dim <= 8.  The backward returns the four parameter gradients
 * (observations are data); partials: asvrl_tower_bwd_partials() floats, counter: one zeroed uint32. */
ASVRL_API int asvrl_tower_fwd(const float* self_obs, const float* objects, const float* mask,
                              const float* w_self, const float* b_self, const float* w_obj,
                              const float* b_obj, int batch, int self_dim, int obj_dim, int max_obj,
                              int self_feat, int obj_feat, float* feats, void* stream);
ASVRL_API int asvrl_tower_bwd(const float* dfeats, const float* feats, const float* self_obs,
                              const float* objects, int batch, int self_dim, int obj_dim, int max_obj,
                              int self_feat, int obj_feat, float* dw_self, float* db_self, float* dw_obj,
                              float* db_obj, float* partials, uint32_t* counter, void* stream);
ASVRL_API long asvrl_tower_bwd_partials(int batch, int max_obj, int self_feat, int obj_feat);
/* asvrl_tower_bwd without caller scratch and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and results as asvrl_tower_bwd minus `partials` / `counter`: the
 * per-CTA partial sums and the arrival ticket live in a file-scope device array of the library (295 168 bytes per GPU
 * that loads it, held until exit; its device address is resolved once per device with cudaGetSymbolAddress).
 * Limit: self_feat + obj_feat <= ASVRL_TOWER_BWD_STATIC_MAX_FEATURES, beyond it ASVRL_E_LIMIT.  The call performs no
 * cudaMalloc / cudaFree / library allocation.  Every call shares the one array, and the ticket is cleared in stream
 * order before each launch: calls that overlap on the GPU (different streams or threads) RACE — serialise them. */
#define ASVRL_TOWER_BWD_STATIC_MAX_FEATURES 128
ASVRL_API int asvrl_tower_bwd_static(const float* dfeats, const float* feats, const float* self_obs,
                                     const float* objects, int batch, int self_dim, int obj_dim, int max_obj,
                                     int self_feat, int obj_feat, float* dw_self, float* db_self, float* dw_obj,
                                     float* db_obj, void* stream);

/* K7 — replaces torch.nn.utils.clip_grad_norm_(params, max_norm) followed by Adam.step()
 * (AGENT:414-416, 424-427 and the other agents' twins) with two launches over a list of tensors:
 *   total = ||g||_2 * grad_scale;  c = min(1, max_norm / (total + 1e-6));  g' = g * grad_scale * c
 *   m <- m + (1-b1)(g'-m);  v <- b2 v + (1-b2) g'^2;  t <- t+1
 *   p <- p - lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
 * t is kept per tensor like torch (a tensor left out of a call keeps its count); norm_out: nullable
 * device float (total); scratch: ASVRL_OPT_SCRATCH_FLOATS caller-owned floats; counter: one
 * zero-initialised uint32 (left zero).  grad_scale = 1/world_size when the gradients hold an all-reduced SUM.  grad is not
 * modified (torch scales .grad in place; nothing on the path reads it afterwards). */
#define ASVRL_OPT_MAX_TENSORS 64
#define ASVRL_OPT_SCRATCH_FLOATS 320
typedef struct asvrl_adam_tensors {
  int n;
  float* param[ASVRL_OPT_MAX_TENSORS];
  const float* grad[ASVRL_OPT_MAX_TENSORS];
  float* exp_avg[ASVRL_OPT_MAX_TENSORS];
  float* exp_avg_sq[ASVRL_OPT_MAX_TENSORS];
  float* step[ASVRL_OPT_MAX_TENSORS];   /* device float per tensor: its step count t, incremented here */
  long long numel[ASVRL_OPT_MAX_TENSORS];
} asvrl_adam_tensors;
ASVRL_API int asvrl_clip_adam_step(const asvrl_adam_tensors* tensors, float max_norm, float lr,
                                   float beta1, float beta2, float eps, float grad_scale,
                                   float* norm_out /* nullable */, float* scratch, uint32_t* counter,
                                   void* stream);

/* K7, multi-GPU: the same update with the all-reduce of the gradient bucket done by the kernels
 * themselves over NVLink peer memory (one-shot, summed in rank order: bit-identical replicas)
 * instead of an NCCL call.  bucket[r]: rank r's symmetric buffer, peer-mapped into this process
 * (e.g. torch.distributed._symmetric_memory), laid out [half 0 | half 1 | >= 16 uint32 flag words],
 * half_floats floats per half, zero-initialised before the first call on any rank.  grad_sum:
 * local scratch of the summed gradient (>= total floats); epoch: device uint32, zero-initialised,
 * advanced by every call (all ranks must call in lock step).  A peer that never arrives makes the
 * wait give up after ~2^26 polls (environment ASVRL_P2P_TIMEOUT_POLLS overrides) and norm_out read NaN (no
 * hang); the rank that gave up raises flag word ASVRL_MAX_PEERS (sticky) in EVERY rank's buffer, and a rank
 * that finds it set skips the step too — from then on every step is skipped on every rank, so replicas never
 * drift apart silently.  A skipped step (here and in asvrl_clip_adam_step: NaN / inf norm) leaves parameters,
 * moments and step counts untouched and adds 1 to scratch[ASVRL_OPT_SCRATCH_FLOATS - 1]. */
#define ASVRL_MAX_PEERS 8
typedef struct asvrl_peer_buckets {
  int world, rank;
  float* bucket[ASVRL_MAX_PEERS];
  long long half_floats;
} asvrl_peer_buckets;
ASVRL_API int asvrl_clip_adam_step_p2p(const asvrl_adam_tensors* tensors, const asvrl_peer_buckets* peers,
                                       float* grad_sum, uint32_t* epoch, float max_norm, float lr,
                                       float beta1, float beta2, float eps, float* norm_out /* nullable */,
                                       float* scratch, uint32_t* counter, void* stream);

/* Collection path (no gradient), AC-IQN actor: from the pre-activation of the last hidden layer to the
 * epsilon-greedy action of every ASV in one pass (ACIQN:318-321, AGENT:207-225):
 *   a[m][j] = atan_scale * atan(sum_d relu(pre[m][d] + bias[d]) w[j][d] + b_out[j]),  j < 2; with
 *   probability eps (eps_dev, a device float, overrides eps when not NULL) a = lo + (hi - lo) u,
 *   u from Philox stream 5 keyed on (seed, counter [+ *counter_dev], row).
 * pre [rows][dim], w [2][dim], action_lo / action_hi: HOST float[2]; actions [rows][2]. */
ASVRL_API int asvrl_actor_head(const float* pre, const float* bias, const float* w, const float* b_out,
                               float atan_scale, const float* eps_dev /* nullable */, float eps,
                               const float* action_lo, const float* action_hi, uint64_t seed,
                               uint64_t counter, const uint64_t* counter_dev /* nullable */, int rows,
                               int dim, float* actions, void* stream);

/* K8 — out = [relu](x * W^T [+ bias]): a dense layer of the policy MLPs (ACIQN:315-316) in the nn.Linear
 * layout, x [rows][in_dim], w [out_dim][in_dim], bias [out_dim] nullable, out [rows][out_dim], fp32 in
 * and out.  Computed on the tcgen05 tensor cores as a three-pass TF32 split (x_lo W_hi + x_hi W_lo +
 * x_hi W_hi, fp32 accumulation in TMEM): fp32-GEMM-level accuracy (~3e-7 relative), not TF32's 3e-4.
 * in_dim a multiple of 32; out_dim a multiple of 16, <= 256; pointers 16-byte aligned.  Forward only
 * (the collection path, torch.no_grad). */
ASVRL_API int asvrl_linear_tf32x3(const float* x, const float* w, const float* bias, int relu, long rows,
                                  int in_dim, int out_dim, float* out, void* stream);
/* Synchronous read of K8's device status word: 0, or a non-zero code when a bounded wait on the tensor pipe
 * expired in some launch since the library was loaded (that launch's output is unspecified; nothing traps). */
ASVRL_API int asvrl_linear_tf32x3_status(int* status_out);

/* K9 — the whole collection policy forward of the AC-IQN / DDPG actor for `rows` ASVs in ONE persistent launch:
 * encoders + mask + concat (ACIQN:284-308), hidden_layer / hidden_layer_2 with ReLU (ACIQN:315-316) on the tcgen05
 * tensor cores (fp16 two-piece split, three products, fp32 accumulation in TMEM: fp32-GEMM-level accuracy), output
 * layer, (2/pi) atan (ACIQN:318-321) and the epsilon-greedy choice of AGENT:207-225 (Philox stream 5, the same draw
 * as asvrl_actor_head).  The network shape is the reference's fixed one (AGENT:20-26): self 7 -> 56, 5 objects x
 * 5 -> 40, 256 -> 128 -> 128 -> 2; weights in the nn.Linear layout.  self_obs [rows][7], objects [rows][5][5],
 * mask [rows][5] (>= 0.5 = present), actions [rows][2].  `workspace`: asvrl_policy_forward_workspace() bytes,
 * 128-byte aligned, holds the packed weight image; `repack` != 0 rebuilds it from w1 / w2 first (needed after every
 * optimiser step; one small launch).  Bounded waits: a pipeline fault leaves a non-zero code in the status word
 * read by asvrl_policy_forward_status (synchronous) instead of hanging or trapping. */
ASVRL_API long asvrl_policy_forward_workspace(void);
ASVRL_API int asvrl_policy_forward(const float* self_obs, const float* objects, const float* mask,
                                   const float* w_self, const float* b_self, const float* w_obj,
                                   const float* b_obj, const float* w1, const float* b1, const float* w2,
                                   const float* b2, const float* w_out, const float* b_out, float atan_scale,
                                   const float* eps_dev /* nullable */, float eps, const float* action_lo,
                                   const float* action_hi, uint64_t seed, uint64_t counter,
                                   const uint64_t* counter_dev /* nullable */, long rows, int repack,
                                   void* workspace, float* actions, void* stream);
/* The trunk of asvrl_policy_forward alone — tower -> 256 -> 128 -> 128 (ReLU) — with the second hidden layer written to
 * hidden [rows][128] (16-byte aligned) instead of the actor's head: the front end that DQN_model.py:50-77, SAC's actor
 * (SAC_model.py Actor.forward) and Rainbow's two noisy streams (Rainbow_model.py:100-135) share; their own output
 * layers follow.  hidden_img (128-byte aligned, ceil(rows / 128) * 65536 bytes): the same values as a tcgen05 operand
 * image — per 128-row tile two chunks (columns 0-63, 64-127) of fp16 hi | lo pieces in the 128-byte-swizzled K-major
 * layout, rows beyond `rows` zero — which asvrl_c51_head consumes by bulk copy.  At least one of the two outputs.
 * Same kernel, workspace and status word as asvrl_policy_forward. */
ASVRL_API int asvrl_policy_hidden(const float* self_obs, const float* objects, const float* mask, const float* w_self,
                                  const float* b_self, const float* w_obj, const float* b_obj, const float* w1,
                                  const float* b1, const float* w2, const float* b2, long rows, int repack,
                                  void* workspace, float* hidden /* nullable */, void* hidden_img /* nullable */,
                                  void* stream);
ASVRL_API int asvrl_policy_forward_status(int* status_out);

/* K10 — the IQN critic's quantile trunk forward (ACIQN:410-426 calc_cos, ACIQN:453-480 the three dense layers with the
 * feature / action-encoding products and the 128 -> 1 head; with aenc == NULL the trunk of IQN:74-86 up to the last
 * hidden layer) for all batch * n_tau quantile rows in ONE persistent launch on the tcgen05 tensor cores (fp16
 * two-piece split, three products, fp32 accumulation in TMEM: fp32-GEMM-level accuracy).  The cosine table is
 * generated inside the kernel's operand loader; no activation leaves the SM.
 *   tau_in [B*N] nullable: NULL draws tau = u01 * cvar from Philox stream 3 exactly as asvrl_iqn_tau_cos does
 *   (same seed / counter / counter_dev meaning); tau_out [B*N] nullable; feats [B][256] (the observation tower's
 *   output), aenc [B][128] nullable; wc [256][64], wh [128][256], wh2 [128][128], wo [128], bo [1] and the biases in
 *   the nn.Linear layout; z [B*N].  head = 1 (IQN's act, AGENT:227-250; n_tau = 32, aenc = wo = bo = NULL, no images):
 *   z [B][128] = mean over the 32 quantile rows of the last hidden layer instead — the action values are the 128 -> 25
 *   output layer applied to it, which is linear, so the mean commutes.  `workspace`: asvrl_critic_forward_workspace() bytes, 128-byte aligned, holds the
 *   packed weight image; `repack` != 0 rebuilds it from wc / wh / wh2 first (one small launch), and with
 *   `bwd_workspace` (asvrl_critic_backward_workspace() bytes) the same launch packs asvrl_critic_backward's image
 *   and the partial maxima of feats behind it, so that call can pass repack = 0 and fmax = NULL.
 *   cos_img / h1_img / h2_img / r3_img: all NULL (inference), or all set (the differentiable pass): the operand images
 *   the backward kernels read — per 128-row tile 1 / 4 / 2 / 2 chunks of asvrl_critic_forward_chunk_bytes() bytes
 *   (fp16 hi | lo pieces of cos, H1 = feats * relu(P1 + bc), H2 = aenc * relu(P2 + bh), R3 = relu(P3 + bh2), in the
 *   128-byte-swizzled K-major tile layout the MMAs consume), 128-byte aligned.
 * Bounded waits: a pipeline fault leaves a non-zero code in the status word read by asvrl_critic_forward_status
 * (synchronous) instead of hanging or trapping.  An activation beyond fp16's range (65 504) yields NaN quantiles. */
ASVRL_API long asvrl_critic_forward_workspace(void);
ASVRL_API long asvrl_critic_forward_chunk_bytes(void);
ASVRL_API int asvrl_critic_forward(const float* tau_in /* nullable */, uint64_t seed, uint64_t counter,
                                   const uint64_t* counter_dev /* nullable */, float cvar, const float* feats,
                                   const float* aenc /* nullable */, const float* wc, const float* bc,
                                   const float* wh, const float* bh, const float* wh2, const float* bh2,
                                   const float* wo, const float* bo, int batch, int n_tau, int head, int repack,
                                   void* workspace, void* bwd_workspace /* nullable */,
                                   float* tau_out /* nullable */, float* z, void* cos_img,
                                   void* h1_img, void* h2_img, void* r3_img, void* stream);
ASVRL_API int asvrl_critic_forward_status(int* status_out);

/* K12 — weight and bias gradients of the critic trunk (the parameter half of the autograd backward of ACIQN:453-480)
 * on the tcgen05 tensor cores, from operand images: dWh2 = dP3^T H2, dbh2 = dP3^T 1, dwo = R3^T dz, dWh = dP2^T H1,
 * dbh = dP2^T 1, dWc = dP1^T cos, dbc = dP1^T 1, contraction over the batch * n_tau quantile rows.
 *   cos_img / h1_img / h2_img / r3_img: as written by asvrl_critic_forward; dp3_img / dp2_img / dp1_img / aux_img
 *   ([tile][2 | 2 | 4 | 1] chunks) and scales[2] = {1/S, 1/S1} (device) as written by asvrl_critic_backward: dP3, dP2
 *   and aux column 0 (dz) carry the power-of-two factor S, dP1 carries S1; the outputs are unscaled.
 *   workspace: asvrl_critic_wgrad_workspace(n_tiles) floats (per-CTA partial accumulators, summed in a fixed order).
 *   Outputs in the nn.Linear layout: d_wc [256][64], d_bc [256], d_wh [128][256], d_bh [128], d_wh2 [128][128],
 *   d_bh2 [128], d_wo [128]. */
ASVRL_API long asvrl_critic_wgrad_workspace(int n_tiles);
ASVRL_API int asvrl_critic_wgrad(const void* dp3_img, const void* dp2_img, const void* dp1_img, const void* aux_img,
                                 const void* cos_img, const void* h1_img, const void* h2_img, const void* r3_img,
                                 const float* scales, int batch, int n_tau, float* workspace, float* d_wc,
                                 float* d_bc, float* d_wh, float* d_bh, float* d_wh2, float* d_bh2, float* d_wo,
                                 void* stream);
ASVRL_API int asvrl_critic_wgrad_status(int* status_out);

/* K11 — the backward chain of the critic trunk (the activation half of the autograd backward of ACIQN:453-480) for all
 * batch * n_tau quantile rows in ONE persistent tcgen05 launch: dz -> dP3 -> dH2 = dP3 Wh2 -> dP2, daenc -> dH1 = dP2 Wh
 * -> dP1, dfeats, reading the operand images asvrl_critic_forward saved.  n_tau: a power of two <= 32.
 *   dz [B*N]; feats [B][256], aenc [B][128], wo [128], wh [128][256], wh2 [128][128] as in asvrl_critic_forward;
 *   workspace: asvrl_critic_backward_workspace() bytes (packed Wh2^T | Wh^T image; `repack` != 0 rebuilds it);
 *   daenc [B][128] = d loss / d aenc; fmax (device, nullable): max over feats, sets dP1's power-of-two scale.  NULL:
 *   with `repack` != 0 the scale is 2^-8; with `repack` == 0 it comes from the 64 partial maxima of feats that
 *   asvrl_critic_forward's weight-pack launch wrote behind the image when it was given this buffer as `bwd_workspace`.
 *   Full pass: dfeats [B][256], dp3_img / dp2_img / dp1_img / aux_img ([tile][2 | 2 | 4 | 1] chunks, 128-byte aligned)
 *   and scales[3] (device: {1/S, 1/S1} for asvrl_critic_wgrad, then sum dz = d loss / d bo) — all set; or all NULL (with h1_img NULL allowed): the actor
 *   update's pass (AGENT:420-427), which needs d z / d aenc only and stops after dP2. */
ASVRL_API long asvrl_critic_backward_workspace(void);
ASVRL_API int asvrl_critic_backward(const float* dz, const void* r3_img, const void* h2_img,
                                    const void* h1_img /* nullable */, const float* feats, const float* aenc,
                                    const float* wh, const float* wh2, const float* wo,
                                    const float* fmax /* nullable */, int batch, int n_tau, int repack,
                                    void* workspace, float* daenc, float* dfeats /* nullable */,
                                    void* dp3_img, void* dp2_img, void* dp1_img, void* aux_img, float* scales,
                                    void* stream);
ASVRL_API int asvrl_critic_backward_status(int* status_out);

/* Greedy action of Rainbow's dueling C51 head (Rainbow_model.py:127-135 + agent.py:235-246 act_rainbow) for `rows`
 * ASVs in one launch: q = v + a - mean_a(a), softmax over the atoms, expectation over the support, first argmax.
 *   zv [rows][atoms], za [rows][n_actions][atoms]: the two output layers' products WITHOUT bias; bv [atoms],
 *   ba [n_actions][atoms]: their (noisy) biases; support [atoms]; action [rows] int32; q_out [rows][n_actions]
 *   nullable (the expected values).  n_actions <= 32, atoms <= 64.  Layout of the products (0 = dense): zv_ld = row
 *   stride of zv in floats (>= atoms); za may be stored in blocks of `za_actions_per_block` actions — block k is
 *   [rows][za_ld] with its actions' atoms packed at the front, blocks `za_block_stride` floats apart — which is what
 *   asvrl_linear_tf32x3 produces when the advantage layer is evaluated 5 actions (255 -> 256 outputs) at a time. */
ASVRL_API int asvrl_c51_greedy(const float* zv, const float* za, const float* bv, const float* ba, const float* support,
                               long rows, int n_actions, int atoms, int zv_ld, int za_actions_per_block, int za_ld,
                               long za_block_stride, int* action, float* q_out /* nullable */, void* stream);

/* K13 — Rainbow's dueling C51 head for `rows` ASVs in ONE persistent tcgen05 launch (+ a weight-pack launch): the
 * 128 -> n_actions * atoms advantage layer, the dueling combination q = v + a - mean_a(a) (Rainbow_model.py:127-135),
 * softmax over the atoms, expectation over the support and the greedy action (agent.py:235-246), without the
 * (rows, n_actions, atoms) product ever reaching HBM (accumulated in TMEM two actions at a time).
 *   h_img: the advantage stream's second hidden layer as the operand image asvrl_policy_hidden writes (128-byte aligned);
 *   zv [rows][zv_ld]: the value stream's output product WITHOUT bias (zv_ld >= 64, a multiple of 4; rows 16-byte aligned);
 *   wa [n_actions * atoms][128], ba [n_actions * atoms], bv [atoms]: the (noisy) output layers' parameters; support [atoms];
 *   workspace: asvrl_c51_head_workspace() bytes, 128-byte aligned; action [rows] int32; q_out [rows][n_actions] nullable.
 *   n_actions <= 32, atoms <= 64, hidden width 128.  Bounded waits report through asvrl_c51_head_status. */
ASVRL_API long asvrl_c51_head_workspace(void);
ASVRL_API int asvrl_c51_head(const void* h_img, const float* zv, int zv_ld, const float* wa, const float* ba,
                             const float* bv, const float* support, long rows, int n_actions, int atoms, void* workspace,
                             int* action, float* q_out /* nullable */, void* stream);
ASVRL_API int asvrl_c51_head_status(int* status_out);

/* Rainbow's prioritised n-step memory: the batched append of one env step (the reference appends one transition per
 * call, rfarl/rfarl/policy/replay_memory_rainbow.py:125-131; here one row per ASV, lane = env * R + robot) in ONE
 * launch, graph-replay safe.  rows [slots][lanes][36] = [self 7 | objects 25 | n_obj | action | reward | nonterminal],
 * tstep [slots][lanes] (the lane's episode clock, 0 cuts an n-step window), prio [slots * lanes] (max_priority for an
 * active ASV, 0 otherwise), lane_t [lanes]; head / filled / count: the ring's device words (int64[1] each), advanced by
 * the last CTA; scratch: two zero-initialised uint32 (left zero).  status: the env step's status words. */
ASVRL_API int asvrl_rainbow_append(const float* obs_self, const float* obs_objects, const float* obs_mask,
                                   const int* actions, const float* reward, const uint32_t* status, int lanes, long slots,
                                   long capacity, float* rows, int* tstep, float* prio, int* lane_t,
                                   const float* max_priority, long long* head, long long* filled, long long* count,
                                   uint32_t* scratch, void* stream);

/* The two narrow layers of the AC-IQN update, each direction in one launch (batch = the learner's sampled states; the
 * backward kernels keep per-CTA partial column sums in `partials` — asvrl_narrow_bwd_partials(dim) floats — and the last
 * CTA, by the zero-initialised ticket `counter` which it leaves at zero, adds them in a fixed order: deterministic).
 * 32 <= dim <= 256, a divisor of 1024.
 *   action encoder (ACIQN:399, 472): aenc [B][dim] = relu(wa a + ba), actions [B][2], wa [dim][2], ba [dim];
 *     backward: dwa [dim][2], dba [dim], dact [B][2] (nullable) from daenc [B][dim] and the saved aenc.
 *   actor head (ACIQN:318-321): y [B][2] = w3 h + b3, actions = atan_scale * atan(y); h [B][dim], w3 [2][dim], b3 [2];
 *     backward: dh [B][dim] (nullable), dw3 [2][dim], db3 [2] from dact [B][2] and the saved y, h. */
ASVRL_API int asvrl_action_encoder_fwd(const float* actions, const float* wa, const float* ba, int batch, int dim,
                                       float* aenc, void* stream);
ASVRL_API int asvrl_action_encoder_bwd(const float* daenc, const float* aenc, const float* actions, const float* wa,
                                       int batch, int dim, float* dwa, float* dba, float* dact /* nullable */,
                                       float* partials, uint32_t* counter, void* stream);
ASVRL_API long asvrl_narrow_bwd_partials(int dim);
ASVRL_API int asvrl_atan_head_fwd(const float* h, const float* w3, const float* b3, float atan_scale, int batch, int dim,
                                  float* y, float* actions, void* stream);
ASVRL_API int asvrl_atan_head_bwd(const float* dact, const float* y, const float* h, const float* w3, float atan_scale,
                                  int batch, int dim, float* dh /* nullable */, float* dw3, float* db3, float* partials,
                                  uint32_t* counter, void* stream);

/* out[0] = -mean(x[0 .. n)) in one launch and a fixed order: the actor loss "-Q.mean()" of AGENT:423 (its gradient is
 * the constant -1 / n, so the learner needs no autograd node for it).  n <= 2^24. */
ASVRL_API int asvrl_neg_mean(const float* x, long n, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ASVRL_H_ */
