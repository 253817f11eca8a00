/*
 * asv_oracle.c — float64 CPU oracle of the rfarl env step and IQN loss.
 * TEST INFRASTRUCTURE ONLY (see asv_oracle.h).  Every function cites the reference lines
 * it restates; it is written from the reference's behaviour, in C, with no code taken
 * from it.  Pinned against the Python reference by tests/golden/ (oracle/gen_golden.py).
 */
#include "asv_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------------------------ */
/* Counter-based RNG (no reference counterpart: the reference uses numpy's sequential    */
/* MT19937; value parity is by noise injection, distribution parity by statistics).      */
/* Published algorithm: Philox4x32-10, Salmon et al., SC'11 (Random123).                 */
/* ------------------------------------------------------------------------------------ */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int round = 0; round < 10; ++round) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static double u01(uint32_t w) { return ((double)(w >> 8) + 0.5) * (1.0 / 16777216.0); }

/* Stream layout shared with the CUDA library (documented in include/asvrl.h):
 * key = seed; ctr = (block | slot<<8 | ego<<16, global env, step lo, step hi(24) | stream<<24) */
#define STREAM_SENSOR 1u

static void sensor_block(uint64_t seed, uint64_t step, int64_t env, int ego, int slot,
                         uint32_t block, uint32_t w[4]) {
  uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  uint32_t ctr[4] = {block | ((uint32_t)slot << 8) | ((uint32_t)ego << 16), (uint32_t)env,
                     (uint32_t)step, ((uint32_t)(step >> 32) & 0x00FFFFFFu) | (STREAM_SENSOR << 24)};
  oracle_philox4x32_10(ctr, key, w);
}

/* 5 sensor-noise draws for (env, ego, slot, step): 4 scaled normals (Box-Muller) and one raw
 * von Mises(0, kappa) sample by the Best-Fisher rejection scheme that numpy's legacy
 * RandomState.vonmises uses (WAMV:27-40 are the call sites). */
static void sensor_noise(const asvrl_env_params* p, uint64_t seed, uint64_t step, int64_t env,
                         int ego, int slot, double n[5]) {
  uint32_t w[4];
  sensor_block(seed, step, env, ego, slot, 0u, w);
  double r0 = sqrt(-2.0 * log(u01(w[0]))), a0 = 2.0 * M_PI * u01(w[1]);
  double r1 = sqrt(-2.0 * log(u01(w[2]))), a1 = 2.0 * M_PI * u01(w[3]);
  n[0] = p->pos_std * r0 * cos(a0);
  n[1] = p->pos_std * r0 * sin(a0);
  n[2] = p->vel_std * r1 * cos(a1);
  n[3] = p->vel_std * r1 * sin(a1);
  double kappa = p->r_kappa;
  if (kappa < 1e-5) { /* numpy: uniform on the circle */
    sensor_block(seed, step, env, ego, slot, 1u, w);
    n[4] = M_PI * (2.0 * u01(w[0]) - 1.0);
    return;
  }
  double rr = 1.0 + sqrt(1.0 + 4.0 * kappa * kappa);
  double rho = (rr - sqrt(2.0 * rr)) / (2.0 * kappa);
  double s = (1.0 + rho * rho) / (2.0 * rho);
  double W = 1.0;
  uint32_t sign_word = 0;
  int accepted = 0;
  for (uint32_t block = 1; block <= 8 && !accepted; ++block) {
    sensor_block(seed, step, env, ego, slot, block, w);
    for (int a = 0; a < 2 && !accepted; ++a) {
      double U = u01(w[2 * a]), V = u01(w[2 * a + 1]);
      double Z = cos(M_PI * U);
      W = (1.0 + s * Z) / (s + Z);
      double Y = kappa * (s - W);
      sign_word = w[2 * a + 1];
      if (Y * (2.0 - Y) - V >= 0.0 || log(Y / V) + 1.0 - Y >= 0.0) accepted = 1;
    }
  }
  double res = acos(W);
  n[4] = (sign_word & 1u) ? res : -res;
}

int oracle_env_noise(const asvrl_env_params* p, uint64_t seed, uint64_t step, int n_envs,
                     int max_robots, int max_obs, int64_t env_offset, double* noise_out) {
  int S = max_obs + max_robots;
  for (int e = 0; e < n_envs; ++e)
    for (int r = 0; r < max_robots; ++r)
      for (int sl = 0; sl < S; ++sl)
        sensor_noise(p, seed, step, env_offset + e, r, sl,
                     noise_out + (((size_t)e * max_robots + r) * S + sl) * 5);
  return 0;
}

/* ------------------------------------------------------------------------------------ */
/* Current field, ENV:458-501.  The `continue` at ENV:473-478 only continues the inner    */
/* loop, so every core contributes; the KD-tree fixes only the sum  */
/* order, which we ignore (differences are O(1e-16)).                                     */
/* ------------------------------------------------------------------------------------ */
void oracle_current(const asvrl_env_params* p, const double* cores, int n_cores, double x,
                    double y, double* out) {
  double vx = 0.0, vy = 0.0;
  for (int k = 0; k < n_cores; ++k) {
    const double* c = cores + 4 * k;
    double rx = c[0] - x, ry = c[1] - y;
    double d = sqrt(rx * rx + ry * ry);
    double ux = rx / d, uy = ry / d; /* ENV:481-482 */
    double tx, ty;
    if (c[2] > 0.0) { tx = -uy; ty = ux; }  /* clockwise: [[0,-1],[1,0]]  ENV:483-484 */
    else            { tx = uy;  ty = -ux; } /* ENV:485-486 */
    double speed = (d <= p->core_r) ? c[3] / (2.0 * M_PI * p->core_r * p->core_r) * d
                                    : c[3] / (2.0 * M_PI * d); /* ENV:497-501 */
    vx += tx * speed;
    vy += ty * speed;
  }
  out[0] = vx + p->drift_vx; /* drift_* are the zero-default extension terms (no reference) */
  out[1] = vy + p->drift_vy;
}

/* pseudo-inverse application acc = (A^T A)^-1 A^T b, WAMV:267-271 */
static void solve_acc(const asvrl_env_params* p, const double b[3], double acc[3]) {
  double A[3][3] = {{p->m - p->xDotU, 0.0, 0.0},
                    {0.0, p->m - p->yDotV, -p->yDotR},
                    {0.0, -p->nDotV, p->Izz - p->nDotR}}; /* WAMV:152-156 */
  double G[3][3], Atb[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) {
      G[i][j] = 0.0;
      for (int k = 0; k < 3; ++k) G[i][j] += A[k][i] * A[k][j];
    }
    Atb[i] = 0.0;
    for (int k = 0; k < 3; ++k) Atb[i] += A[k][i] * b[k];
  }
  double c00 = G[1][1] * G[2][2] - G[1][2] * G[2][1];
  double c01 = G[1][2] * G[2][0] - G[1][0] * G[2][2];
  double c02 = G[1][0] * G[2][1] - G[1][1] * G[2][0];
  double det = G[0][0] * c00 + G[0][1] * c01 + G[0][2] * c02;
  double inv[3][3] = {
      {c00 / det, (G[0][2] * G[2][1] - G[0][1] * G[2][2]) / det, (G[0][1] * G[1][2] - G[0][2] * G[1][1]) / det},
      {c01 / det, (G[0][0] * G[2][2] - G[0][2] * G[2][0]) / det, (G[0][2] * G[1][0] - G[0][0] * G[1][2]) / det},
      {c02 / det, (G[0][1] * G[2][0] - G[0][0] * G[2][1]) / det, (G[0][0] * G[1][1] - G[0][1] * G[1][0]) / det}};
  for (int i = 0; i < 3; ++i) acc[i] = inv[i][0] * Atb[0] + inv[i][1] * Atb[1] + inv[i][2] * Atb[2];
}

typedef struct robot_t {
  double x, y, th, vr[3], v[3], Tl, Tr, gx, gy;
} robot_t;

/* Robot.compute_motion, WAMV:233-279 */
static void compute_motion(const asvrl_env_params* p, robot_t* q) {
  double c = cos(q->th), s = sin(q->th);
  double u_r = c * q->vr[0] + s * q->vr[1], v_r = -s * q->vr[0] + c * q->vr[1]; /* WAMV:235 */
  double u = c * q->v[0] + s * q->v[1], v = -s * q->v[0] + c * q->v[1];         /* WAMV:236 */
  double r = q->v[2];                                                            /* WAMV:241 */
  double a = p->yDotV * v_r + p->yDotR * r; /* C_A[0][2], WAMV:243 */
  double au = fabs(u_r), av = fabs(v_r), ar = fabs(r);
  /* N V_r with N = C_A + D + D_n, WAMV:243-248 */
  double n0 = (-p->xU - p->xUU * au) * u_r + a * r;
  double n1 = (-p->yV - p->yVV * av - p->yRV * ar) * v_r +
              (-p->xDotU * u_r - p->yR - p->yVR * av - p->yRR * ar) * r;
  double n2 = -a * u_r + (p->xDotU * u_r - p->nV - p->nVV * av - p->nRV * ar) * v_r +
              (-p->nR - p->nVR * av - p->nRR * ar) * r;
  /* propulsion, WAMV:251-264 */
  double Fxl = q->Tl * cos(p->left_pos), Fyl = q->Tl * sin(p->left_pos);
  double Fxr = q->Tr * cos(p->right_pos), Fyr = q->Tr * sin(p->right_pos);
  double Fx = Fxl + Fxr, Fy = Fyl + Fyr;
  double Mn = Fxl * p->width / 2 - Fyl * p->length / 2 - Fxr * p->width / 2 - Fyr * p->length / 2;
  /* b = -C_RB V - N V_r + tau_p, WAMV:242, 270 */
  double b[3] = {p->m * r * v - n0 + Fx, -p->m * r * u - n1 + Fy, -n2 + Mn};
  double acc[3];
  solve_acc(p, b, acc);
  u_r += acc[0] * p->dt; /* WAMV:274 */
  v_r += acc[1] * p->dt;
  r += acc[2] * p->dt;
  q->vr[0] = c * u_r - s * v_r; /* WAMV:277-279 */
  q->vr[1] = s * u_r + c * v_r;
  q->vr[2] = r;
}

static double clip(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* Robot.update_state, WAMV:204-231 */
static void update_state(const asvrl_env_params* p, robot_t* q, const double cur[2], int first,
                         double dl, double dr) {
  q->v[0] = q->vr[0] + cur[0]; /* WAMV:201-202 */
  q->v[1] = q->vr[1] + cur[1];
  q->v[2] = q->vr[2];
  q->x += q->v[0] * p->dt; /* WAMV:207-210 */
  q->y += q->v[1] * p->dt;
  q->th += q->v[2] * p->dt;
  while (q->th < 0.0) q->th += 2 * M_PI; /* WAMV:213-216 */
  while (q->th >= 2 * M_PI) q->th -= 2 * M_PI;
  if (first) { /* WAMV:218-229; dl, dr are N/s */
    q->Tl = clip(q->Tl + dl * p->dt * p->n_substeps, p->min_thrust, p->max_thrust);
    q->Tr = clip(q->Tr + dr * p->dt * p->n_substeps, p->min_thrust, p->max_thrust);
  }
  compute_motion(p, q);
}

static double wrap_to_pi(double a) { /* WAMV:425-434 */
  while (a < -M_PI) a += 2 * M_PI;
  while (a >= M_PI) a -= 2 * M_PI;
  return a;
}

static void upd_margin(double* m, double v) {
  v = fabs(v);
  if (v < *m) *m = v;
}

/* check_apply_COLREGs + helpers, WAMV:324-423.  o = [px,py,vx,vy,r] in the ego frame,
 * w = ego velocity over ground in the ego frame.  Returns 1 and sets *phi_out on a hit.
 * phi is written whenever compute_COLREGs_turn_angle ran (self.phi persists, WAMV:396). */
static int colregs_one(const asvrl_env_params* p, const double o[5], const double w[2],
                       double* phi_out, double* margin) {
  double so = sqrt(o[2] * o[2] + o[3] * o[3]);
  upd_margin(margin, so - p->colregs_min_speed);
  if (so < p->colregs_min_speed) return 0; /* WAMV:400-402 */
  double se = sqrt(w[0] * w[0] + w[1] * w[1]);
  upd_margin(margin, se - p->colregs_min_speed);
  if (se < p->colregs_min_speed) return 0; /* WAMV:404-407 */
  double al = atan2(o[3], o[2]);           /* WAMV:328 */
  double ca = cos(al), sa = sin(al);
  double epx = -(ca * o[0] + sa * o[1]), epy = -(-sa * o[0] + ca * o[1]); /* WAMV:334 */
  double evx = ca * w[0] + sa * w[1], evy = -sa * w[0] + ca * w[1];       /* WAMV:339 */
  double ang = atan2(evy, evx);
  /* left-crossing zone, WAMV:344-362 */
  int lx = epx >= p->lc_x0 && epx <= p->lc_x1;
  int ly = epy >= p->lc_y0 && epy <= 0.0;
  double x_diff = epx - p->lc_x1, y_diff = epy - p->lc_y1; /* lc_y1 = -7 (front edge) */
  double grad = p->lc_y1 / p->lc_x1;
  int tri = y_diff > grad * x_diff;
  int la = ang >= M_PI / 4 && ang <= 3 * M_PI / 4;
  int left = lx && ly && !tri && la;
  /* head-on zone, WAMV:364-377 */
  int hx = epx >= 0.0 && epx <= p->head_on_x;
  int hy = epy >= -0.5 * p->head_on_y && epy <= 0.5 * p->head_on_y;
  int ha = fabs(ang) > 3 * M_PI / 4;
  int head = hx && hy && ha;
  upd_margin(margin, epx - p->lc_x0); upd_margin(margin, epx - p->lc_x1);
  upd_margin(margin, epy - p->lc_y0); upd_margin(margin, epy);
  upd_margin(margin, y_diff - grad * x_diff);
  upd_margin(margin, (ang - M_PI / 4) * se); upd_margin(margin, (ang - 3 * M_PI / 4) * se);
  upd_margin(margin, (fabs(ang) - 3 * M_PI / 4) * se);
  upd_margin(margin, epx - p->head_on_x);
  upd_margin(margin, epy + 0.5 * p->head_on_y); upd_margin(margin, epy - 0.5 * p->head_on_y);
  if (!(left || head)) return 0;
  /* compute_COLREGs_turn_angle, WAMV:379-396 (NaN when dist < base_1, then phi > 0 is False) */
  double ego_ang = atan2(w[1], w[0]);
  double obj_ang = atan2(o[1], o[0]);
  double base_1 = o[4] + 1.0;
  double dist = sqrt(o[0] * o[0] + o[1] * o[1]);
  double add_1 = asin(base_1 / dist);
  double tangent = sqrt(dist * dist - base_1 * base_1);
  double add_2 = atan2(p->robot_r, tangent);
  double phi = wrap_to_pi(wrap_to_pi(obj_ang + add_1 + add_2) - ego_ang);
  *phi_out = phi;
  upd_margin(margin, dist - base_1);
  if (phi == phi) {
    upd_margin(margin, phi);
    upd_margin(margin, M_PI - fabs(phi)); /* wrap discontinuity */
  }
  return phi > 0.0 ? 1 : 0; /* WAMV:421 */
}

typedef struct cand_t {
  double o[5];
  double key;
  int slot;
} cand_t;

static void step_one_env(const asvrl_env_params* p, const oracle_env_state* st, const void* actions,
                         const double* noise, uint64_t seed, uint64_t step,
                         const oracle_env_out* out, int e, int RP, int OP, int CP,
                         int64_t env_offset, uint32_t mode) {
  const int S = OP + RP;
  const size_t base = (size_t)e * RP;
  int nr = st->n_robots[e], no = st->n_obs[e], nc = st->n_cores[e];
  const double* obst = st->obstacles + (size_t)e * OP * 4;
  const double* cores = st->cores ? st->cores + (size_t)e * CP * 4 : NULL;
  int ep_before = st->ep_steps[e];
  robot_t rb[64];
  uint32_t fl[64], fl0[64]; /* fl0 = flags the step STARTED with: who is absent from the map */
  double rew[64];
  if (RP > 64) return;

  /* ---- ENV:247-277: execute actions ---- */
  for (int i = 0; i < nr; ++i) {
    size_t k = base + i;
    robot_t* q = &rb[i];
    q->x = st->x[k]; q->y = st->y[k]; q->th = st->theta[k];
    q->vr[0] = st->velr_x[k]; q->vr[1] = st->velr_y[k]; q->vr[2] = st->velr_r[k];
    q->v[0] = st->vel_x[k]; q->v[1] = st->vel_y[k]; q->v[2] = st->vel_r[k];
    q->Tl = st->thrust_l[k]; q->Tr = st->thrust_r[k];
    q->gx = st->goal_x[k]; q->gy = st->goal_y[k];
    fl[i] = fl0[i] = st->flags[k];
    rew[i] = 0.0;
    if (!(mode & ASVRL_M_DYNAMICS) || (fl[i] & ASVRL_F_DEACTIVATED)) continue;
    double dl, dr;
    if (mode & ASVRL_M_DISCRETE) { /* WAMV:146-147, 225: index = 5*iL + iR */
      int a = ((const int32_t*)actions)[k];
      dl = p->thrust_change_l[a / 5];
      dr = p->thrust_change_r[a % 5];
    } else { /* WAMV:220-223 */
      dl = ((const double*)actions)[2 * k] * p->thrust_rate;
      dr = ((const double*)actions)[2 * k + 1] * p->thrust_rate;
    }
    double dx = q->gx - q->x, dy = q->gy - q->y;
    double dis_before = sqrt(dx * dx + dy * dy); /* ENV:254 */
    for (int idx = 0; idx < p->n_substeps; ++idx) { /* ENV:257-260 */
      double cur[2];
      oracle_current(p, cores, nc, q->x, q->y, cur);
      update_state(p, q, cur, idx == 0, dl, dr);
    }
    dx = q->gx - q->x; dy = q->gy - q->y;
    double dis_after = sqrt(dx * dx + dy * dy);
    rew[i] += p->timestep_penalty;      /* ENV:274 */
    rew[i] += dis_before - dis_after;   /* ENV:277 */
  }

  /* ---- ENV:288, 341-356 + WAMV:436-529: perception of every active robot ---- */
  int all_deact = 1, any_coll = 0, all_goal = 1;
  for (int i = 0; i < RP; ++i) {
    size_t k = base + i;
    double* o_self = out->obs_self + k * 7;
    double* o_obj = out->obs_objects + k * 25;
    double* o_mask = out->obs_mask + k * 5;
    memset(o_self, 0, 7 * sizeof(double));
    memset(o_obj, 0, 25 * sizeof(double));
    memset(o_mask, 0, 5 * sizeof(double));
    out->reward[k] = 0.0;
    out->phi[k] = 0.0;
    if (out->top_ids) for (int t = 0; t < 5; ++t) out->top_ids[k * 5 + t] = -1;
    if (out->detected) out->detected[k] = 0;
    if (out->margin_perc) out->margin_perc[k] = INFINITY;
    if (out->margin_colregs) out->margin_colregs[k] = INFINITY;
    if (i >= nr) {
      out->status[k] = (uint32_t)ASVRL_INFO_ABSENT << ASVRL_S_INFO_SHIFT;
      continue;
    }
    robot_t* q = &rb[i];
    uint32_t f = fl[i];
    uint32_t status = 0;
    if (f & ASVRL_F_DEACTIVATED) { /* WAMV:437-438, ENV:295-304 */
      int info = (f & ASVRL_F_COLLISION) ? ASVRL_INFO_DEACT_COLLISION
                 : (f & ASVRL_F_REACH_GOAL) ? ASVRL_INFO_DEACT_GOAL : ASVRL_INFO_DEACT_INVALID;
      status = ASVRL_S_DONE | ASVRL_S_WAS_DEACTIVATED | ((uint32_t)info << ASVRL_S_INFO_SHIFT);
      if (f & ASVRL_F_COLLISION) status |= ASVRL_S_COLLISION;
      if (f & ASVRL_F_REACH_GOAL) status |= ASVRL_S_REACH_GOAL;
      out->status[k] = status;
      if (f & ASVRL_F_COLLISION) any_coll = 1;
      if (!(f & ASVRL_F_REACH_GOAL)) all_goal = 0;
      continue;
    }
    double mperc = INFINITY, mcol = INFINITY;
    double c = cos(q->th), s = sin(q->th);
    /* project_to_robot_frame as the reference evaluates it: R_rw p + (-R_rw t), WAMV:313-319 */
    double trx = -(c * q->x + s * q->y), try_ = -(-s * q->x + c * q->y);
    double w[2] = {c * q->v[0] + s * q->v[1], -s * q->v[0] + c * q->v[1]}; /* WAMV:445 */
    o_self[0] = c * q->gx + s * q->gy + trx; /* WAMV:448 */
    o_self[1] = -s * q->gx + c * q->gy + try_;
    o_self[2] = w[0]; o_self[3] = w[1]; o_self[4] = q->v[2];
    o_self[5] = q->Tl; o_self[6] = q->Tr; /* WAMV:450-453 */
    { /* check_reach_goal, WAMV:170-172 */
      double dx = q->gx - q->x, dy = q->gy - q->y;
      double dg = sqrt(dx * dx + dy * dy);
      upd_margin(&mperc, dg - p->goal_dis);
      if (dg <= p->goal_dis) f |= ASVRL_F_REACH_GOAL;
    }
    cand_t cand[ASVRL_MAX_SLOTS];
    int ncand = 0;
    uint64_t det = 0;
    for (int sl = 0; sl < S; ++sl) {
      double px, py, vx, vy, rt;
      if (sl < OP) { /* static obstacles, WAMV:465-483 */
        if (sl >= no) continue;
        px = obst[4 * sl]; py = obst[4 * sl + 1]; rt = obst[4 * sl + 2];
        vx = 0.0; vy = 0.0;
      } else { /* other robots, WAMV:486-510 */
        int j = sl - OP;
        if (j >= nr || j == i || (fl0[j] & ASVRL_F_DEACTIVATED)) continue;
        px = rb[j].x; py = rb[j].y; vx = rb[j].v[0]; vy = rb[j].v[1]; rt = p->robot_r;
      }
      double n[5];
      if (noise) memcpy(n, noise + ((k * S) + sl) * 5, sizeof n);
      else sensor_noise(p, seed, step, env_offset + e, i, sl, n);
      double opx = px + n[0], opy = py + n[1];                 /* WAMV:27-30 */
      double ovx = vx + n[2], ovy = vy + n[3];                 /* WAMV:32-35 */
      double orr = p->r_mean_ratio * rt + (1 - p->r_mean_ratio) * n[4] / M_PI * rt; /* WAMV:37-40 */
      double qx = c * opx + s * opy + trx, qy = -s * opx + c * opy + try_;          /* WAMV:294 */
      double dq = sqrt(qx * qx + qy * qy);
      upd_margin(&mperc, dq - (p->perc_range + orr));
      if (dq > p->perc_range + orr) continue; /* WAMV:296 */
      double ang = atan2(qy, qx);
      if (ang < -0.5 * p->perc_angle || ang > 0.5 * p->perc_angle) continue; /* WAMV:299-301 */
      det |= (uint64_t)1 << sl;
      if (!(f & ASVRL_F_COLLISION)) { /* WAMV:475-476, 281-291: true geometry */
        double d = sqrt((q->x - px) * (q->x - px) + (q->y - py) * (q->y - py)) - rt - p->robot_r;
        upd_margin(&mperc, d);
        if (d <= 0.0) f |= ASVRL_F_COLLISION;
      }
      cand_t* cd = &cand[ncand++];
      cd->o[0] = qx; cd->o[1] = qy;
      cd->o[2] = c * ovx + s * ovy; cd->o[3] = -s * ovx + c * ovy; /* WAMV:480 */
      cd->o[4] = orr;
      cd->key = sqrt(qx * qx + qy * qy) - orr - p->robot_r; /* WAMV:515, 288 */
      cd->slot = sl;
    }
    /* heapq.nsmallest(5, ...) == first 5 of a stable sort, WAMV:513-515 */
    for (int a = 1; a < ncand; ++a) {
      cand_t t = cand[a];
      int b = a - 1;
      while (b >= 0 && cand[b].key > t.key) { cand[b + 1] = cand[b]; --b; }
      cand[b + 1] = t;
    }
    int keep = ncand < ASVRL_MAX_OBJ ? ncand : ASVRL_MAX_OBJ;
    for (int a = 0; a + 1 < ncand && a < ASVRL_MAX_OBJ; ++a) {
      double gap = cand[a + 1].key - cand[a].key;
      if (gap > 0.0) upd_margin(&mperc, gap);
    }
    for (int a = 0; a < keep; ++a) {
      memcpy(o_obj + 5 * a, cand[a].o, 5 * sizeof(double));
      o_mask[a] = 1.0;
      if (out->top_ids) out->top_ids[k * 5 + a] = cand[a].slot;
    }
    /* WAMV:517-521 */
    int apply = 0;
    double phi = 0.0;
    for (int a = 0; a < keep && !apply; ++a) apply = colregs_one(p, cand[a].o, w, &phi, &mcol);
    if (apply) status |= ASVRL_S_COLREGS;
    out->phi[k] = apply ? phi : 0.0;
    if (out->detected) out->detected[k] = det;
    if (f & ASVRL_F_COLLISION) status |= ASVRL_S_COLLISION;
    if (f & ASVRL_F_REACH_GOAL) status |= ASVRL_S_REACH_GOAL;

    /* ---- ENV:306-328 reward / done / info ---- */
    int info = ASVRL_INFO_NORMAL;
    if (mode & ASVRL_M_DYNAMICS) {
      double r = rew[i];
      if (apply) r += p->colregs_penalty * phi; /* ENV:232-238, 310 */
      if (ep_before >= p->max_episode_steps) { status |= ASVRL_S_DONE; info = ASVRL_INFO_TOO_LONG; }
      else if (f & ASVRL_F_COLLISION) { r += p->collision_penalty; status |= ASVRL_S_DONE; info = ASVRL_INFO_COLLISION; }
      else if (f & ASVRL_F_REACH_GOAL) { r += p->goal_reward; status |= ASVRL_S_DONE; info = ASVRL_INFO_REACH_GOAL; }
      out->reward[k] = r;
    }
    status |= (uint32_t)info << ASVRL_S_INFO_SHIFT;
    out->status[k] = status;
    if (out->margin_perc) out->margin_perc[k] = mperc;
    if (out->margin_colregs) out->margin_colregs[k] = mcol;
    /* TRAINER:168-170 folded in for the batched API */
    if ((mode & ASVRL_M_AUTO_DEACTIVATE) && (f & (ASVRL_F_COLLISION | ASVRL_F_REACH_GOAL)))
      f |= ASVRL_F_DEACTIVATED;
    fl[i] = f;
    if (f & ASVRL_F_COLLISION) any_coll = 1;
    if (!(f & ASVRL_F_REACH_GOAL)) all_goal = 0;
  }
  /* ---- write back state ---- */
  for (int i = 0; i < nr; ++i) {
    size_t k = base + i;
    robot_t* q = &rb[i];
    st->x[k] = q->x; st->y[k] = q->y; st->theta[k] = q->th;
    st->velr_x[k] = q->vr[0]; st->velr_y[k] = q->vr[1]; st->velr_r[k] = q->vr[2];
    st->vel_x[k] = q->v[0]; st->vel_y[k] = q->v[1]; st->vel_r[k] = q->v[2];
    st->thrust_l[k] = q->Tl; st->thrust_r[k] = q->Tr;
    st->flags[k] = fl[i];
    if (!(fl[i] & ASVRL_F_DEACTIVATED)) all_deact = 0;
  }
  if (nr == 0) { all_deact = 1; all_goal = 1; }
  uint32_t es = 0;
  if (all_deact) es |= 1u;                            /* ENV:178-184 */
  if (ep_before >= p->max_episode_steps) es |= 2u;    /* TRAINER:172 */
  if (any_coll) es |= 4u;                             /* ENV:194-200 */
  if (all_goal) es |= 8u;                             /* ENV:186-192 */
  out->env_status[e] = es;
  if (mode & ASVRL_M_DYNAMICS) st->ep_steps[e] = ep_before + 1; /* ENV:330 */
}

typedef struct step_job_t {
  const asvrl_env_params* params; const oracle_env_state* state; const void* actions;
  const double* noise; uint64_t seed, step; const oracle_env_out* out;
  int e0, e1, RP, OP, CP; int64_t env_offset; uint32_t mode;
} step_job_t;

static void* step_worker(void* arg) {
  step_job_t* j = (step_job_t*)arg;
  for (int e = j->e0; e < j->e1; ++e)
    step_one_env(j->params, j->state, j->actions, j->noise, j->seed, j->step, j->out, e, j->RP,
                 j->OP, j->CP, j->env_offset, j->mode);
  return NULL;
}

int oracle_env_step(const asvrl_env_params* params, const oracle_env_state* state,
                    const void* actions, const double* noise, uint64_t seed, uint64_t step,
                    const oracle_env_out* out, int n_envs, int max_robots, int max_obs,
                    int max_cores, int64_t env_offset, uint32_t mode, int n_threads) {
  if (max_robots > 64 || max_obs + max_robots > ASVRL_MAX_SLOTS) return ASVRL_E_LIMIT;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  if (n_threads > n_envs) n_threads = n_envs > 0 ? n_envs : 1;
  /* envs are independent (the reference fans trials out over processes, TRAIN:131-140):
   * contiguous env ranges, one pthread each */
  step_job_t jobs[256];
  pthread_t tid[256];
  for (int t = 0; t < n_threads; ++t) {
    step_job_t j = {params, state, actions, noise, seed, step, out,
                    (int)((int64_t)n_envs * t / n_threads), (int)((int64_t)n_envs * (t + 1) / n_threads),
                    max_robots, max_obs, max_cores, env_offset, mode};
    jobs[t] = j;
  }
  if (n_threads == 1) { step_worker(&jobs[0]); return 0; }
  for (int t = 0; t < n_threads; ++t) pthread_create(&tid[t], NULL, step_worker, &jobs[t]);
  for (int t = 0; t < n_threads; ++t) pthread_join(tid[t], NULL);
  return 0;
}

/* AGENT:401-412 (loss core), AGENT:701-707 (Huber), analytic gradient of SURVEY.md B6 */
int oracle_qhuber(const double* z, const double* tau, const double* z_next, const double* r,
                  const double* done, double gamma, double kappa, int batch, int n, int n_next,
                  double* loss_per_sample, double* grad_z, double* loss) {
  double total = 0.0;
  for (int b = 0; b < batch; ++b) {
    double lb = 0.0;
    for (int i = 0; i < n; ++i) {
      double zi = z[(size_t)b * n + i], ti = tau[(size_t)b * n + i];
      double g = 0.0;
      for (int j = 0; j < n_next; ++j) {
        double T = r[b] + gamma * z_next[(size_t)b * n_next + j] * (1.0 - done[b]); /* AGENT:401 */
        double d = T - zi;                                                           /* AGENT:406 */
        double ad = fabs(d);
        double h = ad <= kappa ? 0.5 * d * d : kappa * (ad - 0.5 * kappa);           /* AGENT:705 */
        double wgt = fabs(ti - (d < 0.0 ? 1.0 : 0.0));                               /* AGENT:409 */
        lb += wgt * h / kappa;
        double dh = ad <= kappa ? d : (d > 0 ? kappa : -kappa);
        g += wgt * dh / kappa;
      }
      grad_z[(size_t)b * n + i] = -g / ((double)batch * n_next);
    }
    lb /= n_next; /* sum over i (dim 1), mean over j (dim 2)  AGENT:411 */
    loss_per_sample[b] = lb;
    total += lb;
  }
  *loss = total / batch; /* AGENT:412 */
  return 0;
}

int oracle_tau_cos(const double* tau, int batch, int n_tau, int n_cos, double* cos_out) {
  for (size_t t = 0; t < (size_t)batch * n_tau; ++t)
    for (int i = 0; i < n_cos; ++i) {
      float pi_i = (float)(M_PI * i); /* torch.FloatTensor([np.pi*i ...]), ACIQN:397 */
      float arg = (float)tau[t] * pi_i; /* float32 product, as torch evaluates ACIQN:424 */
      cos_out[t * n_cos + i] = cos((double)arg);
    }
  return 0;
}

/* ------------------------------------------------------------------------------------ */
/* Scenario generation, ENV:72-176 with the acceptance predicates of ENV:358-456.         */
/* The three predicates are pinned against the reference's own check_* methods by          */
/* tests/golden/reset_checks.npz; the generator applies them, in the reference's order,    */
/* to the counter-indexed Philox candidate sequence documented in include/asvrl.h          */
/* (stream 2; ego field = class, candidate t -> slot = t >> 8, block = t & 255).           */
/* Each predicate returns 1 = accept and lowers *margin to the smallest |distance to a     */
/* threshold| it evaluated (metres), so the parity tests can skip razor-edge cases.        */
/* ------------------------------------------------------------------------------------ */
#define STREAM_RESET 2u

static void marg(double* m, double v) {
  if (!m) return;
  v = fabs(v);
  if (v < *m) *m = v;
}

/* ENV:358-376.  sg = accepted robots [n][4] = start.xy, goal.xy; cand = start.xy, goal.xy */
int oracle_check_start_goal(const asvrl_reset_params* rp, const double* sg, int n,
                            const double* cand, double min_dis, double* margin) {
  double dx = cand[2] - cand[0], dy = cand[3] - cand[1];
  double d = sqrt(dx * dx + dy * dy);
  marg(margin, d - min_dis);
  if (d < min_dis) return 0; /* ENV:361 */
  for (int i = 0; i < n; ++i) {
    const double* o = sg + 4 * i;
    double ds = hypot(o[0] - cand[0], o[1] - cand[1]);
    marg(margin, ds - rp->clear_r);
    if (ds <= rp->clear_r) return 0; /* ENV:366-369 */
    double dg = hypot(o[2] - cand[2], o[3] - cand[3]);
    marg(margin, dg - rp->clear_r);
    if (dg <= rp->clear_r) return 0; /* ENV:371-374 */
  }
  return 1;
}

/* ENV:378-418.  cores [n_c][4] = x, y, +1 clockwise / -1, Gamma; cand likewise */
int oracle_check_core(const asvrl_reset_params* rp, const double* sg, int n_r, const double* cores,
                      int n_c, const double* cand, double* margin) {
  double r = rp->core_r;
  marg(margin, cand[0] - r); marg(margin, cand[0] + r - rp->map_w);
  if (cand[0] - r < 0.0 || cand[0] + r > rp->map_w) return 0; /* ENV:381 */
  marg(margin, cand[1] - r); marg(margin, cand[1] + r - rp->map_w);
  if (cand[1] - r < 0.0 || cand[1] + r > rp->map_w) return 0; /* ENV:383 compares with width */
  for (int i = 0; i < n_r; ++i) {
    const double* o = sg + 4 * i;
    double ds = hypot(cand[0] - o[0], cand[1] - o[1]);
    marg(margin, ds - (r + rp->clear_r));
    if (ds < r + rp->clear_r) return 0; /* ENV:389-391 */
    double dg = hypot(cand[0] - o[2], cand[1] - o[3]);
    marg(margin, dg - (r + rp->clear_r));
    if (dg < r + rp->clear_r) return 0; /* ENV:392-394 */
  }
  for (int i = 0; i < n_c; ++i) {
    const double* o = cores + 4 * i;
    double dx = o[0] - cand[0], dy = o[1] - cand[1];
    double dis = sqrt(dx * dx + dy * dy);
    if ((o[2] > 0.0) == (cand[2] > 0.0)) { /* same spin, ENV:401-407 */
      double bi = o[3] / (2.0 * M_PI * rp->v_rel_max), bj = cand[3] / (2.0 * M_PI * rp->v_rel_max);
      marg(margin, dis - (bi + bj));
      if (dis < bi + bj) return 0;
    } else { /* opposite spin, ENV:408-416 */
      double gl = o[3] > cand[3] ? o[3] : cand[3], gs = o[3] > cand[3] ? cand[3] : o[3];
      double v1 = gl / (2.0 * M_PI * (dis - 2.0 * r)), v2 = gs / (2.0 * M_PI * r);
      /* v1 > p v2  <=>  dis - 2r < gl / (2 pi p v2) when dis > 2r; report the metric margin */
      marg(margin, (dis - 2.0 * r) - gl / (2.0 * M_PI * rp->p * v2));
      marg(margin, dis - 2.0 * r);
      if (v1 > rp->p * v2) return 0;
    }
  }
  return 1;
}

/* ENV:420-456.  obs [n_o][4] = x, y, r, 0; cand = x, y, r */
int oracle_check_obstacle(const asvrl_reset_params* rp, const double* sg, int n_r,
                          const double* cores, int n_c, const double* obs, int n_o,
                          const double* cand, double* margin) {
  double r = cand[2];
  marg(margin, cand[0] - r); marg(margin, cand[0] + r - rp->map_w);
  if (cand[0] - r < 0.0 || cand[0] + r > rp->map_w) return 0; /* ENV:423 */
  marg(margin, cand[1] - r); marg(margin, cand[1] + r - rp->map_h);
  if (cand[1] - r < 0.0 || cand[1] + r > rp->map_h) return 0; /* ENV:425 */
  for (int i = 0; i < n_r; ++i) {
    const double* o = sg + 4 * i;
    double ds = hypot(cand[0] - o[0], cand[1] - o[1]);
    marg(margin, ds - (r + rp->clear_r));
    if (ds < r + rp->clear_r) return 0; /* ENV:431-433 */
    double dg = hypot(cand[0] - o[2], cand[1] - o[3]);
    marg(margin, dg - (r + rp->clear_r));
    if (dg < r + rp->clear_r) return 0; /* ENV:434-436 */
  }
  for (int i = 0; i < n_c; ++i) {
    double dis = hypot(cores[4 * i] - cand[0], cores[4 * i + 1] - cand[1]);
    marg(margin, dis - (rp->core_r + r));
    if (dis <= rp->core_r + r) return 0; /* ENV:439-445 */
  }
  for (int i = 0; i < n_o; ++i) {
    double dis = hypot(obs[4 * i] - cand[0], obs[4 * i + 1] - cand[1]);
    marg(margin, dis - (obs[4 * i + 2] + r));
    if (dis <= obs[4 * i + 2] + r) return 0; /* ENV:448-454 */
  }
  return 1;
}

static void reset_block(uint64_t seed, uint64_t episode, int64_t env, uint32_t cls, uint32_t t,
                        int theta, uint32_t w[4]) {
  uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  uint32_t slot = theta ? t : (t >> 8), block = theta ? 0u : (t & 255u);
  uint32_t ctr[4] = {block | (slot << 8) | (cls << 16), (uint32_t)env, (uint32_t)episode,
                     ((uint32_t)(episode >> 32) & 0x00FFFFFFu) | (STREAM_RESET << 24)};
  oracle_philox4x32_10(ctr, key, w);
}

static double lerp_u(double lo, double hi, double u) { return lo + (hi - lo) * u; }

int oracle_env_reset(const asvrl_reset_params* rp, const oracle_env_state* st,
                     const uint8_t* reset_mask, const int32_t* want_robots,
                     const int32_t* want_cores, const int32_t* want_obs, const double* min_sg,
                     uint64_t seed, uint64_t episode, int n_envs, int RP, int OP, int CP,
                     int64_t env_offset, double* margin_out) {
  double* sg = (double*)malloc(sizeof(double) * 4 * (size_t)(RP > 0 ? RP : 1));
  double* cores = (double*)malloc(sizeof(double) * 4 * (size_t)(CP > 0 ? CP : 1));
  double* obs = (double*)malloc(sizeof(double) * 4 * (size_t)(OP > 0 ? OP : 1));
  for (int e = 0; e < n_envs; ++e) {
    if (reset_mask && !reset_mask[e]) continue;
    const int64_t ge = env_offset + e;
    double margin = INFINITY;
    uint32_t w[4];
    /* robots, ENV:105-120: stop after max_tries draws or when the wanted count is reached */
    int want_r = want_robots[e] < RP ? want_robots[e] : RP, n_r = 0;
    for (int t = 0; t < rp->max_tries && n_r < want_r; ++t) {
      reset_block(seed, episode, ge, 0u, (uint32_t)t, 0, w);
      double lo = rp->start_margin;
      double c[4] = {lerp_u(lo, rp->map_w - lo, u01(w[0])), lerp_u(lo, rp->map_h - lo, u01(w[1])),
                     lerp_u(lo, rp->map_w - lo, u01(w[2])), lerp_u(lo, rp->map_h - lo, u01(w[3]))};
      if (oracle_check_start_goal(rp, sg, n_r, c, min_sg[e], &margin)) {
        memcpy(sg + 4 * n_r, c, sizeof c);
        ++n_r;
      }
    }
    /* vortex cores, ENV:122-136 */
    int want_c = want_cores[e] < CP ? want_cores[e] : CP, n_c = 0;
    for (int t = 0; t < rp->max_tries && n_c < want_c; ++t) {
      reset_block(seed, episode, ge, 1u, (uint32_t)t, 0, w);
      double c[4] = {rp->map_w * u01(w[0]), rp->map_h * u01(w[1]), (w[2] >> 31) ? 1.0 : -1.0,
                     2.0 * M_PI * rp->core_r * lerp_u(rp->v_min, rp->v_max, u01(w[3]))};
      if (oracle_check_core(rp, sg, n_r, cores, n_c, c, &margin)) {
        memcpy(cores + 4 * n_c, c, sizeof c);
        ++n_c;
      }
    }
    /* static obstacles, ENV:150-162 */
    int want_o = want_obs[e] < OP ? want_obs[e] : OP, n_o = 0;
    for (int t = 0; t < rp->max_tries && n_o < want_o; ++t) {
      reset_block(seed, episode, ge, 2u, (uint32_t)t, 0, w);
      double lo = rp->obs_margin;
      double c[4] = {lerp_u(lo, rp->map_w - lo, u01(w[0])), lerp_u(lo, rp->map_h - lo, u01(w[1])),
                     lerp_u(rp->obs_r_min, rp->obs_r_max, u01(w[2])), 0.0};
      if (oracle_check_obstacle(rp, sg, n_r, cores, n_c, obs, n_o, c, &margin)) {
        memcpy(obs + 4 * n_o, c, sizeof c);
        ++n_o;
      }
    }
    /* reset_robot ENV:166-176, Robot.reset_state WAMV:177-193 (no cores exist yet at ENV:175,
     * so the initial sea-floor velocity is zero) */
    for (int r = 0; r < RP; ++r) {
      size_t g = (size_t)e * RP + r;
      int on = r < n_r;
      double th = 0.0;
      if (on) {
        reset_block(seed, episode, ge, 3u, (uint32_t)r, 1, w);
        th = 2.0 * M_PI * u01(w[0]); /* ENV:171 */
      }
      st->x[g] = on ? sg[4 * r] : 0.0; st->y[g] = on ? sg[4 * r + 1] : 0.0; st->theta[g] = th;
      st->velr_x[g] = st->velr_y[g] = st->velr_r[g] = 0.0;
      st->vel_x[g] = st->vel_y[g] = st->vel_r[g] = 0.0;
      st->thrust_l[g] = st->thrust_r[g] = 0.0;
      st->goal_x[g] = on ? sg[4 * r + 2] : 0.0; st->goal_y[g] = on ? sg[4 * r + 3] : 0.0;
      st->flags[g] = 0;
    }
    for (int i = 0; i < OP; ++i)
      for (int q = 0; q < 4; ++q) st->obstacles[((size_t)e * OP + i) * 4 + q] = i < n_o ? obs[4 * i + q] : 0.0;
    for (int i = 0; i < CP; ++i)
      for (int q = 0; q < 4; ++q) st->cores[((size_t)e * CP + i) * 4 + q] = i < n_c ? cores[4 * i + q] : 0.0;
    st->n_robots[e] = n_r; st->n_obs[e] = n_o; st->n_cores[e] = n_c;
    st->ep_steps[e] = 0; /* ENV:96 */
    if (margin_out) margin_out[e] = margin;
  }
  free(sg); free(cores); free(obs);
  return 0;
}
