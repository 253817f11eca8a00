// api.cu — version, error string and defaults of libasvrl.so (include/asvrl.h).
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"

namespace asvrl {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) return 0;
  snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(e));
  return (int)e;
}

}  // namespace asvrl

extern "C" int asvrl_version(void) { return ASVRL_VERSION; }

extern "C" const char* asvrl_last_error_string(void) { return asvrl::g_err; }

// WAMV:45-126, ENV:33-54
extern "C" void asvrl_env_default_params(asvrl_env_params* p) {
  if (!p) return;
  memset(p, 0, sizeof *p);
  p->dt = 0.05; p->n_substeps = 10; p->max_episode_steps = 1000;
  p->length = 5.0; p->width = 2.5;
  p->robot_r = 0.5 * sqrt(p->length * p->length + p->width * p->width);
  p->goal_dis = 2.0;
  p->min_thrust = -500.0; p->max_thrust = 1000.0; p->thrust_rate = 1000.0;
  const double change[5] = {0.0, -500.0, -1000.0, 500.0, 1000.0};
  for (int i = 0; i < 5; ++i) { p->thrust_change_l[i] = change[i]; p->thrust_change_r[i] = change[i]; }
  p->m = 400.0; p->Izz = 450.0;
  p->xDotU = 20.0; p->nDotR = -980.0;
  p->xU = -100.0; p->xUU = -150.0; p->yV = -100.0; p->yVV = -150.0;
  p->nR = -980.0; p->nRR = -950.0;
  p->perc_range = 20.0; p->perc_angle = 2.0 * 3.14159265358979323846;
  p->pos_std = 0.05; p->vel_std = 0.05; p->r_kappa = 1.0; p->r_mean_ratio = 0.8;
  p->head_on_x = 17.0; p->head_on_y = 9.0;
  p->lc_x0 = -9.0; p->lc_x1 = 12.0; p->lc_y0 = -17.0; p->lc_y1 = -7.0;
  p->colregs_min_speed = 0.5;
  p->timestep_penalty = -0.1; p->colregs_penalty = -0.1;
  p->collision_penalty = -5.0; p->goal_reward = 10.0;
  p->core_r = 0.5;
}
