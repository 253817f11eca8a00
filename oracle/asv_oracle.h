/*
 * asv_oracle.h — CPU oracle of the rfarl hot path (TEST INFRASTRUCTURE, not product).
 *
 * A float64 restatement, in plain C, of the reference's Python algorithm for the path
 * libasvrl.so accelerates.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it; the product package never does.
 *
 * Parity status: PINNED.  oracle/gen_golden.py drives the unmodified Python reference
 * (imported from /root/reference in the build container) on seeded states with injected
 * sensor noise and commits the input/output vectors under tests/golden/;
 * tests/test_oracle_golden.py checks this restatement against every one of them.
 *
 * Reference citations (paths relative to the reference root):
 *   ENV   = rfarl/rfarl/envs/marinenav/env.py
 *   WAMV  = rfarl/rfarl/envs/marinenav/vehicles/wamv.py
 *   AGENT = rfarl/rfarl/agent.py
 *
 * Shapes and semantics mirror include/asvrl.h (same params struct, same status / flag
 * bits, same noise-slot layout) with every float array widened to double and HOST pointers.
 */
#ifndef ASV_ORACLE_H_
#define ASV_ORACLE_H_

#include <stdint.h>
#include "../include/asvrl.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_env_state {
  double *x, *y, *theta;             /* [E][RP] */
  double *velr_x, *velr_y, *velr_r;  /* velocity_r  WAMV:95 */
  double *vel_x, *vel_y, *vel_r;     /* velocity (stale)  WAMV:96, 206 */
  double *thrust_l, *thrust_r;
  double *goal_x, *goal_y;
  uint32_t* flags;                   /* ASVRL_F_* */
  int32_t *n_robots, *n_obs, *n_cores, *ep_steps; /* [E] */
  double* obstacles;                 /* [E][OP][4] x,y,r,0 */
  double* cores;                     /* [E][CP][4] x,y,+1 cw/-1 ccw,Gamma */
} oracle_env_state;

typedef struct oracle_env_out {
  double* obs_self;     /* [E][RP][7] */
  double* obs_objects;  /* [E][RP][5][5] */
  double* obs_mask;     /* [E][RP][5] */
  double* reward;       /* [E][RP] */
  uint32_t* status;     /* [E][RP] ASVRL_S_* */
  double* phi;          /* [E][RP] */
  int32_t* top_ids;     /* [E][RP][5] */
  uint64_t* detected;   /* [E][RP] */
  uint32_t* env_status; /* [E] */
  /* smallest |margin| of any threshold predicate evaluated for this ASV: the parity
   * tests demand bit-exact flags wherever the margin exceeds a stated epsilon
   * ("Bit-exact flags from fp32 state").  nullable. */
  double* margin_perc;    /* [E][RP] detection / collision / goal / top-5 key gaps (metres) */
  double* margin_colregs; /* [E][RP] speed gates, zone bounds, sector tests, phi sign */
} oracle_env_out;

/* ENV:240-333 for a batch of environments.  actions: double[E][RP][2] or int32[E][RP]
 * (ASVRL_M_DISCRETE).  noise: double[E][RP][OP+RP][5] or NULL (Philox stream, see
 * oracle_env_noise).  n_threads > 1 splits the env range over OpenMP threads. */
int oracle_env_step(const asvrl_env_params* params, const oracle_env_state* state,
                    const void* actions, const double* noise, uint64_t seed, uint64_t step,
                    const oracle_env_out* out, int n_envs, int max_robots, int max_obs,
                    int max_cores, int64_t env_offset, uint32_t mode, int n_threads);

/* the sensor-noise table the NULL-noise path uses: double[E][RP][OP+RP][5] */
int oracle_env_noise(const asvrl_env_params* params, uint64_t seed, uint64_t step, int n_envs,
                     int max_robots, int max_obs, int64_t env_offset, double* noise_out);

/* ENV:458-501 at one point: cores double[C][4]; out[2] */
void oracle_current(const asvrl_env_params* params, const double* cores, int n_cores, double x,
                    double y, double* out);

/* AGENT:401-412, 701-707 + analytic backward; all double. */
int oracle_qhuber(const double* z, const double* tau, const double* z_next, const double* r,
                  const double* done, double gamma, double kappa, int batch, int n, int n_next,
                  double* loss_per_sample, double* grad_z, double* loss);

/* ACIQN:410-426: cos_out[b][n][i] = cos(tau[b][n] * (float)(pi*i)) evaluated in double on
 * the float32-rounded pi*i table the reference builds (ACIQN:397). */
int oracle_tau_cos(const double* tau, int batch, int n_tau, int n_cos, double* cos_out);

/* Scenario generation (ENV:72-176).  The three acceptance predicates restate
 * check_start_and_goal / check_core / check_obstacle (ENV:358-456) and are pinned against the
 * reference's own methods by tests/golden/reset_checks.npz.  Each returns 1 = accept and lowers
 * *margin (nullable) to the smallest |distance to a threshold| it evaluated. */
int oracle_check_start_goal(const asvrl_reset_params* rp, const double* sg, int n,
                            const double* cand, double min_dis, double* margin);
int oracle_check_core(const asvrl_reset_params* rp, const double* sg, int n_r, const double* cores,
                      int n_c, const double* cand, double* margin);
int oracle_check_obstacle(const asvrl_reset_params* rp, const double* sg, int n_r,
                          const double* cores, int n_c, const double* obs, int n_o,
                          const double* cand, double* margin);
/* the reference's sequential rejection loops applied to the Philox candidate sequence of
 * include/asvrl.h (stream 2); margin_out [E] nullable */
int oracle_env_reset(const asvrl_reset_params* rp, const oracle_env_state* st,
                     const uint8_t* reset_mask, const int32_t* want_robots,
                     const int32_t* want_cores, const int32_t* want_obs, const double* min_sg,
                     uint64_t seed, uint64_t episode, int n_envs, int max_robots, int max_obs,
                     int max_cores, int64_t env_offset, double* margin_out);

/* Philox4x32-10 block, exposed so the tests can pin the generator against the published
 * Random123 known-answer vectors. */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
