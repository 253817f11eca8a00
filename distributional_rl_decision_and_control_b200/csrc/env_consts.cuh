// env_consts.cuh — per-launch constant block of the env kernels, derived on the host from
// asvrl_env_params (float64) once per call and passed by value (__grid_constant__).
#pragma once
#include "common.cuh"

namespace asvrl {

struct EnvConsts {
  // integration, WAMV:46-47
  float dt;
  int n_sub;
  int max_ep;
  // thrust, WAMV:84-87, 218-229 (deltas already multiplied by dt*N)
  float min_thrust, max_thrust;
  double act_scale, dl[5], dr[5];
  // rigid body + hydrodynamics, WAMV:103-125, 242-248
  float m, xDotU, yDotV, yDotR;
  float xU, xUU, yV, yVV, yR, yRV, yVR, yRR, nR, nRR, nV, nVV, nRV, nVR;
  // propulsion geometry, WAMV:251-264
  float cl, sl, cr, sr, half_w, half_l;
  // (A^T A)^-1 A^T with A = M_RB + M_A, row major, WAMV:152-156, 267-271 (float64 on the host)
  float Ainv[9];
  // current field, ENV:497-501
  float core_r, solid_k /* 1/(2 pi r^2) */, inv_2pi, drift_vx, drift_vy;
  // perception thresholds evaluated in float64 so flags are bit-exact from identical inputs
  double perc_range, half_angle, r_mean_ratio, r_noise_k /* (1-ratio)/pi */, robot_r_d, goal_dis_d;
  int angle_limited;
  float robot_r;
  // in-kernel sensor noise, WAMV:22-40
  float pos_std, vel_std, vm_kappa, vm_s;
  int vm_uniform;  // kappa < 1e-5: numpy's uniform branch
  float detect_bound;  // range + largest possible |position noise| of the in-kernel generator
  // COLREGs, WAMV:73-76, 344-377, 400-406
  float head_on_x, head_on_half_y, lc_x0, lc_x1, lc_y0, lc_y1, lc_grad, min_speed;
  // rewards, ENV:47-50
  float timestep_penalty, colregs_penalty, collision_penalty, goal_reward;
};

// fills *k; returns 0 or ASVRL_E_BADARG (singular inertia matrix, bad sizes)
int make_env_consts(const asvrl_env_params* p, EnvConsts* k);

struct NoiseDraw {
  float n[5];
};

// 5 sensor-noise draws for (env, ego, slot, step): two Box-Muller pairs and one Best-Fisher
// von Mises(0, kappa) sample (the algorithm numpy's legacy RandomState.vonmises uses,
// SURVEY.md Appendix C).  Attempt a of Philox block b>=1 uses words (2a, 2a+1) as (U, V) and
// bit 0 of the V word as the sign; 16 attempts at most.
__device__ __forceinline__ NoiseDraw sensor_noise(const EnvConsts& k, uint64_t seed, uint64_t step,
                                                  int64_t env, uint32_t ego, uint32_t slot) {
  NoiseDraw d;
  uint4 w = philox_at(seed, STREAM_SENSOR, step, env, ego, slot, 0u);
  // fast intrinsics: the draws are noise, only their distribution matters (abs. error ~1e-6 sigma)
  const float TWO_PI = 6.28318530717959f;
  float r0 = sqrtf(-2.0f * __logf(u01(w.x)));
  float r1 = sqrtf(-2.0f * __logf(u01(w.z)));
  float s0, c0, s1, c1;
  __sincosf(TWO_PI * (u01(w.y) - 0.5f), &s0, &c0);  // angle in (-pi, pi): MUFU range
  __sincosf(TWO_PI * (u01(w.w) - 0.5f), &s1, &c1);
  s0 = -s0; c0 = -c0; s1 = -s1; c1 = -c1;            // sin/cos(2 pi u) = -sin/cos(2 pi u - pi)
  d.n[0] = k.pos_std * r0 * c0;
  d.n[1] = k.pos_std * r0 * s0;
  d.n[2] = k.vel_std * r1 * c1;
  d.n[3] = k.vel_std * r1 * s1;
  if (k.vm_uniform) {
    w = philox_at(seed, STREAM_SENSOR, step, env, ego, slot, 1u);
    d.n[4] = 3.14159265358979f * (2.0f * u01(w.x) - 1.0f);
    return d;
  }
  float W = 1.0f;
  uint32_t sign_word = 0;
  bool accepted = false;
  for (uint32_t block = 1; block <= 8 && !accepted; ++block) {
    w = philox_at(seed, STREAM_SENSOR, step, env, ego, slot, block);
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      if (accepted) break;
      uint32_t wu = a == 0 ? w.x : w.z, wv = a == 0 ? w.y : w.w;
      float U = u01(wu), V = u01(wv);
      float Z = __cosf(3.14159265358979f * U);
      W = __fdividef(1.0f + k.vm_s * Z, k.vm_s + Z);
      float Y = k.vm_kappa * (k.vm_s - W);
      sign_word = wv;
      accepted = (Y * (2.0f - Y) - V >= 0.0f) || (__logf(__fdividef(Y, V)) + 1.0f - Y >= 0.0f);
    }
  }
  float res = acosf(W);
  d.n[4] = (sign_word & 1u) ? res : -res;
  return d;
}

}  // namespace asvrl
