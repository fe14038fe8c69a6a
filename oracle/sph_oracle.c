/* TEST INFRASTRUCTURE ONLY -- never linked into, imported or called by the
 * product (swift_subtask_dev_b200/); only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may use it, as the checker.
 *
 * CPU restatement, in plain C, of the reference's SPH neighbour-loop hot path
 * (SWIFT 0.9.0 PlanetSim fork, Minimal and Planetary schemes, 3D cubic spline,
 * gamma = 5/3): top-level grid + octree, task splitting, 13-axis sorts,
 * density self/pair loops, ghost Newton-Raphson with subset redo passes,
 * force self/pair loops, end_force.  Every function cites the reference
 * file:line whose arithmetic it follows.  Compile WITHOUT fast-math and
 * WITHOUT FMA contraction (-O2 -fno-fast-math -ffp-contract=off) so the float
 * sequences are the same IEEE operations the reference build in oracle/_ref
 * executes.
 *
 * Parity pinning: tests/test_oracle.py runs this file and the compiled
 * reference (oracle/_ref/libswiftref_*.so, through ref_harness.c) on the same
 * inputs and requires identical ordered pair lists and fields equal to
 * summation-order round-off; tests/golden/ holds outputs of the reference
 * itself for the GPU box, where /root/reference does not exist. */
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "oracle_api.h"

/* ------------------------------------------------------------------------ */
/* Constants: src/kernel_hydro.h:45-64, 205-241; src/dimension.h:41         */
/* ------------------------------------------------------------------------ */
#define K_GAMMA ((float)(1.825742))
#define K_CONSTANT ((float)(16. * M_1_PI))
#define K_GAMMA_INV ((float)(1. / K_GAMMA))
#define K_GAMMA2 ((float)(K_GAMMA * K_GAMMA))
#define K_GAMMA_INV_DIM ((float)(1. / (K_GAMMA * K_GAMMA * K_GAMMA)))
#define K_GAMMA_INV_DIM_P1 ((float)(1. / (K_GAMMA * K_GAMMA * K_GAMMA * K_GAMMA)))
#define K_ROOT ((float)(0.5f) * K_CONSTANT * K_GAMMA_INV_DIM)
#define DIM_F 3.f
#define DIM_INV_F 0.3333333333f
#define GAMMA_F 1.66666666666666667f
#define GAMMA_M1_F 0.66666666666666667f
#define SPACE_STRETCH 1.10f
#define SPLITSIZE 400
#define SUBSIZE_PAIR 256000000
#define SUBSIZE_SELF 32000
#define MAX_TOP_CELLS 12
#define NUM_TIME_BINS 56

static const float k_coeffs[12] = {3.f, -3.f, 0.f, 0.5f, -1.f, 3.f, -3.f, 1.f, 0.f, 0.f, 0.f, 0.f};

/* src/sort_part.h:42-56 */
static const double axis_dir[13][3] = {
    {5.773502691896258e-01, 5.773502691896258e-01, 5.773502691896258e-01},
    {7.071067811865475e-01, 7.071067811865475e-01, 0.0},
    {5.773502691896258e-01, 5.773502691896258e-01, -5.773502691896258e-01},
    {7.071067811865475e-01, 0.0, 7.071067811865475e-01},
    {1.0, 0.0, 0.0},
    {7.071067811865475e-01, 0.0, -7.071067811865475e-01},
    {5.773502691896258e-01, -5.773502691896258e-01, 5.773502691896258e-01},
    {7.071067811865475e-01, -7.071067811865475e-01, 0.0},
    {5.773502691896258e-01, -5.773502691896258e-01, -5.773502691896258e-01},
    {0.0, 7.071067811865475e-01, 7.071067811865475e-01},
    {0.0, 1.0, 0.0},
    {0.0, 7.071067811865475e-01, -7.071067811865475e-01},
    {0.0, 0.0, 1.0},
};

/* direction index 0..26 -> sort id; indices below 13 are flipped
 * (src/sort_part.h:58-93): by symmetry sid(k) = sid(26-k) = k for k<13. */
static inline int dir_to_sid(int k) { return k < 13 ? k : (k == 13 ? 0 : 26 - k); }
static inline int dir_flipped(int k) { return k < 13; }
static const float sid_scale[13] = {0.1897f, 0.4025f, 0.1897f, 0.4025f, 0.5788f, 0.4025f, 0.1897f,
                                    0.4025f, 0.1897f, 0.4025f, 0.5788f, 0.4025f, 0.5788f};
static inline int sid_is_corner(int sid) { return sid == 0 || sid == 2 || sid == 6 || sid == 8; }

/* ------------------------------------------------------------------------ */
typedef struct {
  double x[3];
  float v[3], a[3];
  float mass, h, u, u_dt, rho;
  float wcount, wcount_dh, rho_dh, div_v, rot_v[3]; /* density sums */
  float f, pressure, soundspeed, v_sig, h_dt, balsara; /* force variables */
  /* SPHENIX: p->viscosity.{alpha, div_v_previous_step, div_v_dt}, p->diffusion.{alpha, laplace_u},
   * p->force.alpha_visc_max_ngb (div_v and v_sig above are its p->viscosity.div_v / .v_sig) */
  float alpha_visc, div_v_prev, div_v_dt, alpha_diff, laplace_u, alpha_visc_max_ngb;
  long long id;
  int mat_id;
  signed char time_bin, min_ngb_time_bin, wakeup; /* wakeup: p->limiter_data.wakeup */
} opart;

typedef struct {
  float d;
  int i;
} osort;

typedef struct {
  double loc[3], width[3], dmin;
  int first, count, parent, progeny[8], depth, split;
  float h_max, h_max_old, h_max_active;
  int active;
  osort *sort[13];
  int *tasks; /* density tasks touching this cell, in creation order */
  int ntasks, captasks;
} ocell;

typedef struct {
  const oracle_params *prm;
  const oracle_eos *eos;
  opart *parts;
  ocell *cells;
  int ncells;
  const oracle_task *tasks;
  int ntasks;
  oracle_pairlog *log;
  int stage; /* 0 density, 1 ghost redo, 2 force, 3 gradient (SPHENIX), 4 time-step limiter */
  float a, H; /* cosmology_init_no_cosmo: a = 1, H = 0 */
  float beta;
  oracle_sphenix *sx; /* SPHENIX only */
} octx;

static inline float fmin_f(float a, float b) { return a < b ? a : b; }
static inline float fmax_f(float a, float b) { return a > b ? a : b; }

static inline void log_pair(octx *c, int slot, long long i, long long j) {
  oracle_pairlog *l = c->log;
  if (l == NULL) return;
  const long long k = l->count[slot]++;
  if (c->prm->record_pairs && l->pairs[slot] != NULL && k < l->capacity[slot]) {
    l->pairs[slot][2 * k] = i;
    l->pairs[slot][2 * k + 1] = j;
  }
}

static inline int part_active(const octx *c, const opart *p) { return p->time_bin <= c->prm->max_active_bin; }

/* ------------------------------------------------------------------------ */
/* kernel_deval: src/kernel_hydro.h:257-284                                  */
/* ------------------------------------------------------------------------ */
static inline void kernel_eval(float u, float *W, float *dW) {
  const float x = u * K_GAMMA_INV;
  const int t = (int)(x * 2.f);
  const int ind = t > 2 ? 2 : t;
  const float *cf = &k_coeffs[ind * 4];
  float w = cf[0] * x + cf[1];
  float dw = cf[0];
  for (int k = 2; k <= 3; k++) {
    dw = dw * x + w;
    w = x * w + cf[k];
  }
  w = fmax_f(w, 0.f);
  dw = fmin_f(dw, 0.f);
  *W = w * K_CONSTANT * K_GAMMA_INV_DIM;
  *dW = dw * K_CONSTANT * K_GAMMA_INV_DIM_P1;
}

/* runner_iact_nonsym_density: src/hydro/Minimal/hydro_iact.h:134-184 and
 * src/hydro/Planetary/hydro_iact.h:136-186 (same arithmetic).  The symmetric
 * runner_iact_density (:52-120) equals this applied in both directions bit
 * for bit (the reference's tests/testSymmetry.c), so only this form exists. */
static inline void iact_gradient(octx *c, float r2, const float dx[3], float hi, opart *pi, const opart *pj);

static inline void iact_density(octx *c, float r2, const float dx[3], float hi, opart *pi, const opart *pj) {
  /* the gradient loop of SPHENIX is FUNCTION = gradient pasted into the same DOSELF1 / DOPAIR1 /
   * DOSUB_*1 walks (src/runner_doiact_hydro.c:47-52): same loops, another interaction */
  if (c->stage == 3) {
    iact_gradient(c, r2, dx, hi, pi, pj);
    return;
  }
  /* likewise the limiter loop: FUNCTION = limiter in walks of the same shape
   * (src/runner_doiact_functions_limiter.h:233-450, 541-680), with "starting" for "active" -- the same
   * comparison time_bin <= max_active_bin (src/active.h:698-717), on the bins the time-step task has
   * just given.  runner_iact_nonsym_limiter, src/timestep_limiter_iact.h:107-117: the NEIGHBOUR is
   * written (accumulate_max_c). */
  if (c->stage == 4) {
    opart *wj = (opart *)pj;
    if (wj->time_bin > pi->time_bin + 2 /* time_bin_neighbour_max_delta_bin */) {
      const signed char w = (signed char)(-pi->time_bin);
      if (w > wj->wakeup) wj->wakeup = w;
    }
    return;
  }
  float wi, wi_dx;
  const float mj = pj->mass;
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float h_inv = 1.f / hi;
  const float ui = r * h_inv;
  kernel_eval(ui, &wi, &wi_dx);
  pi->rho += mj * wi;
  pi->rho_dh -= mj * (DIM_F * wi + ui * wi_dx);
  pi->wcount += wi;
  pi->wcount_dh -= (DIM_F * wi + ui * wi_dx);
  const float faci = mj * wi_dx * r_inv;
  float dv[3], curl[3];
  dv[0] = pi->v[0] - pj->v[0];
  dv[1] = pi->v[1] - pj->v[1];
  dv[2] = pi->v[2] - pj->v[2];
  const float dvdr = dv[0] * dx[0] + dv[1] * dx[1] + dv[2] * dx[2];
  pi->div_v -= faci * dvdr;
  curl[0] = dv[1] * dx[2] - dv[2] * dx[1];
  curl[1] = dv[2] * dx[0] - dv[0] * dx[2];
  curl[2] = dv[0] * dx[1] - dv[1] * dx[0];
  pi->rot_v[0] += faci * curl[0];
  pi->rot_v[1] += faci * curl[1];
  pi->rot_v[2] += faci * curl[2];
  log_pair(c, c->stage, pi->id, pj->id);
}

/* runner_iact_nonsym_gradient: src/hydro/SPHENIX/hydro_iact.h:272-329.  The symmetric runner_iact_gradient
 * (:196-257) is this applied in both directions: the same new_v_sig, and laplace_u of pj with
 * -(u_i - u_j) = (u_j - u_i) exactly. */
static inline void iact_gradient(octx *c, float r2, const float dx[3], float hi, opart *pi, const opart *pj) {
  const float a = c->a, H = c->H;
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float fac_mu = 1.f;
  const float a2_Hubble = a * a * H;
  const float dvdr = (pi->v[0] - pj->v[0]) * dx[0] + (pi->v[1] - pj->v[1]) * dx[1] + (pi->v[2] - pj->v[2]) * dx[2];
  const float dvdr_Hubble = dvdr + a2_Hubble * r2;
  const float omega_ij = fmin_f(dvdr_Hubble, 0.f);
  const float mu_ij = fac_mu * r_inv * omega_ij;
  /* hydro_signal_velocity, SPHENIX/hydro.h:490-498 */
  const float new_v_sig = pi->soundspeed + pj->soundspeed - c->beta * mu_ij;
  pi->v_sig = fmax_f(pi->v_sig, new_v_sig);
  float wi, wi_dx;
  const float ui = r / hi;
  kernel_eval(ui, &wi, &wi_dx);
  const float delta_u_factor = (pi->u - pj->u) * r_inv;
  pi->laplace_u += pj->mass * delta_u_factor * wi_dx / pj->rho;
  pi->alpha_visc_max_ngb = fmax_f(pi->alpha_visc_max_ngb, pj->alpha_visc);
  if (c->sx) {
    const long long k = c->sx->grad_count++;
    if (c->prm->record_pairs && c->sx->grad_pairs != NULL && k < c->sx->grad_capacity) {
      c->sx->grad_pairs[2 * k] = pi->id;
      c->sx->grad_pairs[2 * k + 1] = pj->id;
    }
  }
}

/* runner_iact_nonsym_force of SPHENIX: src/hydro/SPHENIX/hydro_iact.h:488-609 (+ runner_iact_nonsym_timebin) */
static inline void iact_force_sphenix(octx *c, float r2, const float dx[3], float hi, float hj, opart *pi,
                                      const opart *pj) {
  const float a = c->a, H = c->H;
  const float fac_mu = 1.f;
  const float a2_Hubble = a * a * H;
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float mi = pi->mass, mj = pj->mass;
  const float rhoi = pi->rho, rhoj = pj->rho;
  const float pressurei = pi->pressure, pressurej = pj->pressure;
  const float hi_inv = 1.0f / hi;
  const float hi2 = hi_inv * hi_inv;
  const float hid_inv = hi2 * hi2;
  const float xi = r * hi_inv;
  float wi, wi_dx;
  kernel_eval(xi, &wi, &wi_dx);
  const float wi_dr = hid_inv * wi_dx;
  const float hj_inv = 1.0f / hj;
  const float hj2 = hj_inv * hj_inv;
  const float hjd_inv = hj2 * hj2;
  const float xj = r * hj_inv;
  float wj, wj_dx;
  kernel_eval(xj, &wj, &wj_dx);
  const float wj_dr = hjd_inv * wj_dx;
  const float dvdr = (pi->v[0] - pj->v[0]) * dx[0] + (pi->v[1] - pj->v[1]) * dx[1] + (pi->v[2] - pj->v[2]) * dx[2];
  const float dvdr_Hubble = dvdr + a2_Hubble * r2;
  const float omega_ij = fmin_f(dvdr_Hubble, 0.f);
  const float mu_ij = fac_mu * r_inv * omega_ij;
  const float v_sig = pi->soundspeed + pj->soundspeed - c->beta * mu_ij;
  const float f_ij = 1.f - pi->f / mj;
  const float f_ji = 1.f - pj->f / mi;
  const float rho_ij = rhoi + rhoj;
  const float alpha = pi->alpha_visc + pj->alpha_visc;
  const float visc = -0.25f * alpha * v_sig * mu_ij * (pi->balsara + pj->balsara) / rho_ij;
  const float visc_acc_term = 0.5f * visc * (wi_dr * f_ij + wj_dr * f_ji) * r_inv;
  const float P_over_rho2_i = pressurei / (rhoi * rhoi) * f_ij;
  const float P_over_rho2_j = pressurej / (rhoj * rhoj) * f_ji;
  const float sph_acc_term = (P_over_rho2_i * wi_dr + P_over_rho2_j * wj_dr) * r_inv;
  const float acc = sph_acc_term + visc_acc_term;
  pi->a[0] -= mj * acc * dx[0];
  pi->a[1] -= mj * acc * dx[1];
  pi->a[2] -= mj * acc * dx[2];
  const float sph_du_term_i = P_over_rho2_i * dvdr * r_inv * wi_dr;
  const float visc_du_term = 0.5f * visc_acc_term * dvdr_Hubble;
  const float alpha_diff = (pressurei * pi->alpha_diff + pressurej * pj->alpha_diff) / (pressurei + pressurej);
  const float v_diff =
      alpha_diff * 0.5f * (sqrtf(2.f * fabsf(pressurei - pressurej) / rho_ij) + fabsf(fac_mu * r_inv * dvdr_Hubble));
  const float diff_du_term = v_diff * (pi->u - pj->u) * (f_ij * wi_dr / rhoi + f_ji * wj_dr / rhoj);
  const float du_dt_i = sph_du_term_i + visc_du_term + diff_du_term;
  pi->u_dt += du_dt_i * mj;
  pi->h_dt -= mj * dvdr * r_inv / rhoj * wi_dr;
  if (pj->time_bin > 0) pi->min_ngb_time_bin = pi->min_ngb_time_bin < pj->time_bin ? pi->min_ngb_time_bin : pj->time_bin;
  log_pair(c, 2, pi->id, pj->id);
}

/* runner_iact_nonsym_force: src/hydro/Minimal/hydro_iact.h:370-478,
 * src/hydro/Planetary/hydro_iact.h:369-476; plus runner_iact_nonsym_timebin
 * src/timestep_limiter_iact.h:41-55. */
static inline void iact_force(octx *c, float r2, const float dx[3], float hi, float hj, opart *pi, const opart *pj) {
  if (c->prm->scheme == ORACLE_SCHEME_SPHENIX) {
    iact_force_sphenix(c, r2, dx, hi, hj, pi, pj);
    return;
  }
  const float a = c->a, H = c->H;
  const float fac_mu = 1.f; /* pow_three_gamma_minus_five_over_two(a), gamma=5/3: src/adiabatic_index.h:524-526 */
  const float a2_Hubble = a * a * H;
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float mi = pi->mass, mj = pj->mass;
  const float rhoi = pi->rho, rhoj = pj->rho;
  const float pressurei = pi->pressure, pressurej = pj->pressure;
  const float hi_inv = 1.0f / hi;
  const float hi2 = hi_inv * hi_inv;
  const float hid_inv = hi2 * hi2;
  const float xi = r * hi_inv;
  float wi, wi_dx;
  kernel_eval(xi, &wi, &wi_dx);
  const float wi_dr = hid_inv * wi_dx;
  const float hj_inv = 1.0f / hj;
  const float hj2 = hj_inv * hj_inv;
  const float hjd_inv = hj2 * hj2;
  const float xj = r * hj_inv;
  float wj, wj_dx;
  kernel_eval(xj, &wj, &wj_dx);
  const float wj_dr = hjd_inv * wj_dx;
  const float f_ij = 1.f - pi->f / mj;
  const float f_ji = 1.f - pj->f / mi;
  const float P_over_rho2_i = pressurei / (rhoi * rhoi) * f_ij;
  const float P_over_rho2_j = pressurej / (rhoj * rhoj) * f_ji;
  float visc, dvdr, dvdr_visc, v_sig;
  if (c->prm->scheme == ORACLE_SCHEME_PLANETARY) {
    /* Planetary: Hubble term folded into dvdr (:294-296) */
    dvdr = (pi->v[0] - pj->v[0]) * dx[0] + (pi->v[1] - pj->v[1]) * dx[1] + (pi->v[2] - pj->v[2]) * dx[2] +
           a2_Hubble * r2;
    dvdr_visc = dvdr;
    const float omega_ij = fmin_f(dvdr, 0.f);
    const float mu_ij = fac_mu * r_inv * omega_ij;
    v_sig = pi->soundspeed + pj->soundspeed - c->beta * mu_ij;
    const float rho_ij = 0.5f * (rhoi + rhoj);
    visc = -0.25f * v_sig * mu_ij * (pi->balsara + pj->balsara) / rho_ij;
  } else {
    dvdr = (pi->v[0] - pj->v[0]) * dx[0] + (pi->v[1] - pj->v[1]) * dx[1] + (pi->v[2] - pj->v[2]) * dx[2];
    dvdr_visc = dvdr + a2_Hubble * r2;
    const float omega_ij = fmin_f(dvdr_visc, 0.f);
    const float mu_ij = fac_mu * r_inv * omega_ij;
    v_sig = pi->soundspeed + pj->soundspeed - c->beta * mu_ij;
    const float rho_ij = 0.5f * (rhoi + rhoj);
    visc = -0.25f * v_sig * (pi->balsara + pj->balsara) * mu_ij / rho_ij;
  }
  const float visc_acc_term = 0.5f * visc * (wi_dr * f_ij + wj_dr * f_ji) * r_inv;
  const float sph_acc_term = (P_over_rho2_i * wi_dr + P_over_rho2_j * wj_dr) * r_inv;
  const float acc = sph_acc_term + visc_acc_term;
  pi->a[0] -= mj * acc * dx[0];
  pi->a[1] -= mj * acc * dx[1];
  pi->a[2] -= mj * acc * dx[2];
  const float sph_du_term_i = P_over_rho2_i * dvdr * r_inv * wi_dr;
  const float visc_du_term = 0.5f * visc_acc_term * dvdr_visc;
  const float du_dt_i = sph_du_term_i + visc_du_term;
  pi->u_dt += du_dt_i * mj;
  if (c->prm->scheme == ORACLE_SCHEME_PLANETARY)
    pi->h_dt -= mj * dvdr * r_inv / rhoj * wi_dr;
  else
    pi->h_dt -= mj * dvdr * r_inv / rhoj * wi_dr * f_ij;
  pi->v_sig = fmax_f(pi->v_sig, v_sig);
  if (pj->time_bin > 0) pi->min_ngb_time_bin = pi->min_ngb_time_bin < pj->time_bin ? pi->min_ngb_time_bin : pj->time_bin;
  log_pair(c, 2, pi->id, pj->id);
}

/* ------------------------------------------------------------------------ */
/* Equation of state                                                          */
/* ------------------------------------------------------------------------ */
/* Tillotson: src/equation_of_state/planetary/tillotson.h:189-233, 245-304 */
static float til_pressure(float density, float u, const oracle_til *m) {
  const float eta = density / m->rho_0;
  const float eta_sq = eta * eta;
  const float mu = eta - 1.f;
  const float nu = 1.f / eta - 1.f;
  const float w = u / (m->u_0 * eta_sq) + 1.f;
  const float w_inv = 1.f / w;
  float P_c, P_e, P;
  P_c = (m->a + m->b * w_inv) * density * u + m->A * mu + m->B * mu * mu;
  if (eta < m->eta_zero) {
    P_c = 0.f;
  } else if (eta < m->eta_min) {
    P_c *= (eta - m->eta_zero) / (m->eta_min - m->eta_zero);
  }
  P_e = m->a * density * u + (m->b * density * u * w_inv + m->A * mu * expf(-m->beta * nu)) * expf(-m->alpha * nu * nu);
  if ((1.f < eta) || (u < m->u_iv)) {
    P = P_c;
  } else if ((eta < 1.f) && (m->u_cv < u)) {
    P = P_e;
  } else {
    P = ((u - m->u_iv) * P_e + (m->u_cv - u) * P_c) / (m->u_cv - m->u_iv);
  }
  if (P < m->P_min) P = m->P_min;
  return P;
}

static float til_soundspeed(float density, float u, const oracle_til *m) {
  const float rho_0_inv = 1.f / m->rho_0;
  const float eta = density * rho_0_inv;
  const float rho_inv = 1.f / density;
  const float eta_sq = eta * eta;
  const float mu = eta - 1.f;
  const float nu = 1.f / eta - 1.f;
  const float w = u / (m->u_0 * eta_sq) + 1.f;
  const float w_inv = 1.f / w;
  const float w_inv_sq = w_inv * w_inv;
  const float exp_beta = expf(-m->beta * nu);
  const float exp_alpha = expf(-m->alpha * nu * nu);
  float P_c, P_e, c_sq_c, c_sq_e, c_sq;
  P_c = (m->a + m->b * w_inv) * density * u + m->A * mu + m->B * mu * mu;
  if (eta < m->eta_zero) {
    P_c = 0.f;
  } else if (eta < m->eta_min) {
    P_c *= (eta - m->eta_zero) / (m->eta_min - m->eta_zero);
  }
  c_sq_c = P_c * rho_inv * (1.f + m->a + m->b * w_inv) + m->b * (w - 1.f) * w_inv_sq * (2.f * u - P_c * rho_inv) +
           rho_inv * (m->A + m->B * (eta_sq - 1.f));
  P_e = m->a * density * u + (m->b * density * u * w_inv + m->A * mu * exp_beta) * exp_alpha;
  c_sq_e = P_e * rho_inv * (1.f + m->a + m->b * w_inv * exp_alpha) +
           (m->b * density * u * w_inv_sq / eta_sq *
                (rho_inv / m->u_0 * (2.f * u - P_e * rho_inv) + 2.f * m->alpha * nu * w * rho_0_inv) +
            m->A * rho_0_inv * (1.f + mu / eta_sq * (m->beta + 2.f * m->alpha * nu - eta)) * exp_beta) *
               exp_alpha;
  if ((1.f < eta) || (u < m->u_iv)) {
    c_sq = c_sq_c;
  } else if ((eta < 1.f) && (m->u_cv < u)) {
    c_sq = c_sq_e;
  } else {
    c_sq = ((u - m->u_iv) * c_sq_e + (m->u_cv - u) * c_sq_c) / (m->u_cv - m->u_iv);
  }
  c_sq = fmaxf(c_sq, m->A * rho_0_inv);
  return sqrtf(c_sq);
}

/* ---- materials that live in tables --------------------------------------------------
 * find_value_in_monot_incr_array, src/utilities.h:33-57 */
static int tab_find(float x, const float *array, int n) {
  int lo = 0, hi = n;
  while (hi - lo > 1) {
    const int mid = (hi + lo) / 2;
    if (array[mid] <= x) lo = mid; else hi = mid;
  }
  if (x < array[0]) return -1;
  if (array[n - 1] <= x) return n;
  return lo;
}

/* What SESAME_pressure_from_internal_energy and SESAME_soundspeed_from_internal_energy share
 * (src/equation_of_state/planetary/sesame.h:488-544, 613-669).  which = 0: pressure (:546-593), 1: sound speed
 * (:671-720); they differ in the table, in what a non-positive node does, and in whether the "below the minimum u"
 * clamp looks at the raw or at the logged node values.  The u slices are searched in rows clamped into the table
 * (the reference reads rows -1 / num_rho + 1 for a density outside it: undefined). */
static float tab_sesame(const oracle_eos_table *m, float density, float u, int which) {
  if (u <= 0.f) return 0.f;
  const int nT = m->num_T;
  const float log_rho = logf(density);
  const float log_u = logf(u);
  int idx_rho = tab_find(log_rho, m->log_rho, m->num_rho);
  const int r1 = idx_rho < 0 ? 0 : (idx_rho > m->num_rho - 1 ? m->num_rho - 1 : idx_rho);
  const int r2 = idx_rho + 1 < 0 ? 0 : (idx_rho + 1 > m->num_rho - 1 ? m->num_rho - 1 : idx_rho + 1);
  int idx_u_1 = tab_find(log_u, m->log_u + (size_t)r1 * nT, nT);
  int idx_u_2 = tab_find(log_u, m->log_u + (size_t)r2 * nT, nT);
  if (idx_rho <= -1) idx_rho = 0; else if (idx_rho >= m->num_rho) idx_rho = m->num_rho - 2;
  if (idx_u_1 <= -1) idx_u_1 = 0; else if (idx_u_1 >= nT) idx_u_1 = nT - 2;
  if (idx_u_2 <= -1) idx_u_2 = 0; else if (idx_u_2 >= nT) idx_u_2 = nT - 2;
  float intp_rho, intp_u_1, intp_u_2;
  if (m->log_rho[idx_rho + 1] != m->log_rho[idx_rho])
    intp_rho = (log_rho - m->log_rho[idx_rho]) / (m->log_rho[idx_rho + 1] - m->log_rho[idx_rho]);
  else
    intp_rho = 1.f;
  const float *ua = m->log_u + (size_t)idx_rho * nT, *ub = m->log_u + (size_t)(idx_rho + 1) * nT;
  if (ua[idx_u_1 + 1] != ua[idx_u_1]) intp_u_1 = (log_u - ua[idx_u_1]) / (ua[idx_u_1 + 1] - ua[idx_u_1]);
  else intp_u_1 = 1.f;
  if (ub[idx_u_2 + 1] != ub[idx_u_2]) intp_u_2 = (log_u - ub[idx_u_2]) / (ub[idx_u_2 + 1] - ub[idx_u_2]);
  else intp_u_2 = 1.f;
  const float *tab = which == 0 ? m->P : m->c;
  float v1 = tab[(size_t)idx_rho * nT + idx_u_1], v2 = tab[(size_t)idx_rho * nT + idx_u_1 + 1];
  float v3 = tab[(size_t)(idx_rho + 1) * nT + idx_u_2], v4 = tab[(size_t)(idx_rho + 1) * nT + idx_u_2 + 1];
  if (which == 0 && (idx_rho > 0.f) && ((intp_u_1 < 0.f) || (intp_u_2 < 0.f) || (v1 > v2) || (v3 > v4))) {
    intp_u_1 = 0;
    intp_u_2 = 0;
  }
  const int npos = (v1 <= 0.f) + (v2 <= 0.f) + (v3 <= 0.f) + (v4 <= 0.f);
  const float tiny = which == 0 ? m->P_tiny : m->c_tiny;
  if (which == 0) {
    if (npos > 0) {
      if ((npos > 2) || (m->P_tiny == 0.f) || (intp_rho < 0.f) || (intp_u_1 < 0.f) || (intp_u_2 < 0.f)) return 0.f;
    }
  } else {
    if (npos > 2) return m->c_tiny;
    if (npos > 0 && ((intp_rho < 0.f) || (intp_u_1 < 0.f) || (intp_u_2 < 0.f))) return m->c_tiny;
  }
  if (npos > 0) {
    if (v1 <= 0.f) v1 = tiny;
    if (v2 <= 0.f) v2 = tiny;
    if (v3 <= 0.f) v3 = tiny;
    if (v4 <= 0.f) v4 = tiny;
  }
  v1 = logf(v1); v2 = logf(v2); v3 = logf(v3); v4 = logf(v4);
  if (which == 1 && (idx_rho > 0.f) && ((intp_u_1 < 0.f) || (intp_u_2 < 0.f) || (v1 > v2) || (v3 > v4))) {
    intp_u_1 = 0;
    intp_u_2 = 0;
  }
  const float v = (1.f - intp_rho) * ((1.f - intp_u_1) * v1 + intp_u_1 * v2) +
                  intp_rho * ((1.f - intp_u_2) * v3 + intp_u_2 * v4);
  return expf(v);
}

/* HM80_pressure_from_internal_energy, src/equation_of_state/planetary/hm80.h:224-268 */
static float tab_hm80_pressure(const oracle_eos_table *m, float density, float u) {
  if (u <= 0.f) return 0.f;
  const float log_rho = logf(density);
  const float log_u = logf(u);
  int idx_rho = floor((log_rho - m->log_rho_min) * m->inv_log_rho_step);
  int idx_u = floor((log_u - m->log_u_min) * m->inv_log_u_step);
  if (idx_rho <= -1) idx_rho = 0; else if (idx_rho >= m->num_rho - 1) idx_rho = m->num_rho - 2;
  if (idx_u <= -1) idx_u = 0; else if (idx_u >= m->num_T - 1) idx_u = m->num_T - 2;
  const float intp_rho = (log_rho - m->log_rho_min - idx_rho * m->log_rho_step) * m->inv_log_rho_step;
  const float intp_u = (log_u - m->log_u_min - idx_u * m->log_u_step) * m->inv_log_u_step;
  const int nu = m->num_T;
  const float p1 = m->P[(size_t)idx_rho * nu + idx_u], p2 = m->P[(size_t)idx_rho * nu + idx_u + 1];
  const float p3 = m->P[(size_t)(idx_rho + 1) * nu + idx_u], p4 = m->P[(size_t)(idx_rho + 1) * nu + idx_u + 1];
  const float log_P = (1.f - intp_rho) * ((1.f - intp_u) * p1 + intp_u * p2) +
                      intp_rho * ((1.f - intp_u) * p3 + intp_u * p4);
  return expf(log_P);
}

/* the table branch of the planetary dispatcher; returns 0 when mat_id has no table */
static int tab_eval(const oracle_eos *eos, int mat_id, float rho, float u, float *P, float *cs) {
  for (int k = 0; k < eos->n_tab; k++) {
    const oracle_eos_table *m = &eos->tab[k];
    if (m->mat_id != mat_id) continue;
    if (m->kind == 2) {
      *P = tab_hm80_pressure(m, rho, u);
      /* HM80_soundspeed_from_internal_energy, hm80.h:281-300 */
      if (m->bulk_mod != 0) {
        *cs = sqrtf(m->bulk_mod / rho);
      } else {
        *cs = sqrtf(GAMMA_F * (*P) / rho);
        if (*cs <= 0) *cs = sqrtf(GAMMA_F * m->P_min_for_c_min / rho);
      }
    } else {
      *P = tab_sesame(m, rho, u, 0);
      *cs = tab_sesame(m, rho, u, 1);
    }
    return 1;
  }
  return 0;
}

/* Dispatcher: src/equation_of_state/planetary/equation_of_state.h:910-1146
 * (pressure) and :1305-1457 (sound speed): type = mat_id / 100; type 0 is
 * the planetary ideal gas (planetary/ideal_gas.h), type 1 Tillotson. */
static void eos_eval(const octx *c, const opart *p, float *P, float *cs) {
  if (c->prm->scheme != ORACLE_SCHEME_PLANETARY) {
    /* src/equation_of_state/ideal_gas/equation_of_state.h; Minimal/hydro.h:690-693 */
    *P = GAMMA_M1_F * p->u * p->rho;
    *cs = sqrtf(GAMMA_F * (*P) / p->rho);
    return;
  }
  const int type = p->mat_id / 100;
  if (type == 0) {
    /* planetary/ideal_gas.h:132-136, 161-165: P = (gamma-1) u rho, c = sqrt(u gamma (gamma-1)) */
    const float g = c->eos->idg_gamma[p->mat_id % 100];
    *P = (g - 1.f) * p->u * p->rho;
    *cs = sqrtf(p->u * g * (g - 1.f));
    return;
  }
  /* types 2 (HM80), 3 (SESAME), 4 (ANEOS), 9 (custom): tables */
  if (type >= 2 && tab_eval(c->eos, p->mat_id, p->rho, p->u, P, cs)) return;
  const oracle_til *m = NULL;
  for (int k = 0; k < c->eos->n_til; k++)
    if (c->eos->til[k].mat_id == p->mat_id) m = &c->eos->til[k];
  if (m == NULL) {
    fprintf(stderr, "oracle: unknown material id %d\n", p->mat_id);
    abort();
  }
  *P = til_pressure(p->rho, p->u, m);
  *cs = til_soundspeed(p->rho, p->u, m);
}

/* ------------------------------------------------------------------------ */
/* Per-particle hydro.h pieces                                                */
/* ------------------------------------------------------------------------ */
/* hydro_init_part: Minimal/hydro.h:517-528 */
static inline void init_part(opart *p) {
  p->wcount = 0.f;
  p->wcount_dh = 0.f;
  p->rho = 0.f;
  p->rho_dh = 0.f;
  p->div_v = 0.f;
  p->rot_v[0] = p->rot_v[1] = p->rot_v[2] = 0.f;
  p->laplace_u = 0.f; /* SPHENIX/hydro.h:565-566 */
}

/* hydro_end_density: Minimal/hydro.h:543-575 (a2_inv = 1 without cosmology) */
static inline void end_density(opart *p) {
  const float h = p->h;
  const float h_inv = 1.0f / h;
  const float h_inv_dim = h_inv * h_inv * h_inv;
  const float h_inv_dim_plus_one = h_inv_dim * h_inv;
  p->rho += p->mass * K_ROOT;
  p->rho_dh -= DIM_F * p->mass * K_ROOT;
  p->wcount += K_ROOT;
  p->wcount_dh -= DIM_F * K_ROOT;
  p->rho *= h_inv_dim;
  p->rho_dh *= h_inv_dim_plus_one;
  p->wcount *= h_inv_dim;
  p->wcount_dh *= h_inv_dim_plus_one;
  const float rho_inv = 1.f / p->rho;
  const float a_inv2 = 1.f;
  p->rot_v[0] *= h_inv_dim_plus_one * a_inv2 * rho_inv;
  p->rot_v[1] *= h_inv_dim_plus_one * a_inv2 * rho_inv;
  p->rot_v[2] *= h_inv_dim_plus_one * a_inv2 * rho_inv;
  p->div_v *= h_inv_dim_plus_one * a_inv2 * rho_inv;
}

/* hydro_part_has_no_neighbours: Minimal/hydro.h:626-650 */
static inline void no_neighbours(opart *p) {
  const float h = p->h;
  const float h_inv = 1.0f / h;
  const float h_inv_dim = h_inv * h_inv * h_inv;
  p->rho = p->mass * K_ROOT * h_inv_dim;
  p->wcount = K_ROOT * h_inv_dim;
  p->rho_dh = 0.f;
  p->wcount_dh = 0.f;
  p->div_v = 0.f;
  p->rot_v[0] = p->rot_v[1] = p->rot_v[2] = 0.f;
  p->laplace_u = 0.f; /* SPHENIX/hydro.h:789-806 also clears v_sig, which hydro_reset_gradient sets next */
}

/* SPHENIX: what runner_do_ghost does for a converged particle under EXTRA_HYDRO_LOOP
 * (src/runner_ghost.c:1249-1264, 1429-1443): hydro_prepare_gradient + hydro_reset_gradient,
 * SPHENIX/hydro.h:654-733 */
static void prepare_gradient(const octx *c, opart *p) {
  const float fac_B = 1.f; /* cosmo->a_factor_Balsara_eps at a = 1 */
  const float curl_v = sqrtf(p->rot_v[0] * p->rot_v[0] + p->rot_v[1] * p->rot_v[1] + p->rot_v[2] * p->rot_v[2]);
  const float abs_div_v = fabsf(p->div_v);
  float pressure, soundspeed;
  eos_eval(c, p, &pressure, &soundspeed);
  const float balsara = abs_div_v / (abs_div_v + curl_v + 0.0001f * soundspeed * fac_B / p->h);
  const float common_factor = p->h * DIM_INV_F / p->wcount;
  float grad_h_term;
  if (p->h > 0.9999f * c->prm->h_max) {
    grad_h_term = 0.f;
  } else {
    const float grad_W_term = common_factor * p->wcount_dh;
    if (grad_W_term < -0.9999f)
      grad_h_term = 0.f;
    else
      grad_h_term = common_factor * p->rho_dh / (1.f + grad_W_term);
  }
  p->f = grad_h_term;
  p->pressure = pressure;
  p->soundspeed = soundspeed;
  p->balsara = balsara;
  /* hydro_reset_gradient */
  p->v_sig = 2.f * p->soundspeed;
  p->alpha_visc_max_ngb = p->alpha_visc;
}

/* SPHENIX: runner_do_extra_ghost for one active particle (src/runner_ghost.c:1027-1063):
 * hydro_end_gradient (SPHENIX/hydro.h:745-761), hydro_prepare_force (:823-936) with
 * dt_alpha = get_timestep(p->time_bin, time_base), hydro_reset_acceleration (:944-955).
 * The cosmology factors (doubles in struct cosmology) are all 1: a product of two floats formed in
 * double and rounded once is the float product. */
static void extra_ghost_part(const octx *c, opart *p) {
  const oracle_sphenix *sx = c->sx;
  {
    const float h_inv = 1.0f / p->h;
    const float h_inv_dim = h_inv * h_inv * h_inv;
    const float h_inv_dim_plus_one = h_inv_dim * h_inv;
    p->laplace_u *= 2.f * h_inv_dim_plus_one;
  }
  const double dt_d = (p->time_bin <= 0 ? 0.0 : (double)(1LL << (p->time_bin + 1))) * sx->time_base;
  const float dt_alpha = (float)dt_d;
  const float kernel_support_physical = p->h * K_GAMMA;
  const float kernel_support_physical_inv = 1.f / kernel_support_physical;
  const float v_sig_physical = p->v_sig;
  float pressure, soundspeed_physical;
  eos_eval(c, p, &pressure, &soundspeed_physical);
  const float sound_crossing_time_inverse = soundspeed_physical * kernel_support_physical_inv;
  const float div_v_dt = dt_alpha == 0.f ? 0.f : (p->div_v - p->div_v_prev) / dt_alpha;
  const float S = p->div_v < 0.f ? kernel_support_physical * kernel_support_physical * fmax_f(0.f, -1.f * div_v_dt) : 0.f;
  const float soundspeed_square = soundspeed_physical * soundspeed_physical;
  const float alpha_loc = sx->viscosity_alpha_max * S / (soundspeed_square + S);
  if (alpha_loc > p->alpha_visc) {
    p->alpha_visc = alpha_loc;
  } else {
    const float timescale_ratio = dt_alpha * sound_crossing_time_inverse * sx->viscosity_length;
    p->alpha_visc += alpha_loc * timescale_ratio;
    p->alpha_visc /= (1.f + timescale_ratio);
  }
  p->alpha_visc = fmax_f(p->alpha_visc, sx->viscosity_alpha_min);
  p->div_v_prev = p->div_v;
  p->div_v_dt = div_v_dt;
  const float diffusion_timescale_physical_inverse = v_sig_physical * kernel_support_physical_inv;
  const float sqrt_u_inv = 1.f / sqrtf(p->u);
  float alpha_diff_dt = sx->diffusion_beta * kernel_support_physical * p->laplace_u * sqrt_u_inv;
  alpha_diff_dt -= (p->alpha_diff - sx->diffusion_alpha_min) * diffusion_timescale_physical_inverse;
  float new_diffusion_alpha = p->alpha_diff;
  new_diffusion_alpha += alpha_diff_dt * dt_alpha;
  new_diffusion_alpha = fmax_f(new_diffusion_alpha, sx->diffusion_alpha_min);
  const float viscous_diffusion_limit =
      sx->diffusion_alpha_max * (1.f - p->alpha_visc_max_ngb / sx->viscosity_alpha_max);
  new_diffusion_alpha = fmin_f(new_diffusion_alpha, viscous_diffusion_limit);
  p->alpha_diff = new_diffusion_alpha;
  p->a[0] = p->a[1] = p->a[2] = 0.0f;
  p->u_dt = 0.0f;
  p->h_dt = 0.0f;
}

/* hydro_prepare_force + hydro_reset_acceleration:
 * Minimal/hydro.h:669-761, Planetary/hydro.h:667-768 */
static void prepare_force(const octx *c, opart *p) {
  if (c->prm->scheme == ORACLE_SCHEME_SPHENIX) {
    prepare_gradient(c, p);
    return;
  }
  const int planetary = c->prm->scheme == ORACLE_SCHEME_PLANETARY;
  const float fac_Balsara_eps = 1.f; /* a_factor_Balsara_eps at a = 1 */
  const float h_inv = 1.f / p->h;
  const float curl_v = sqrtf(p->rot_v[0] * p->rot_v[0] + p->rot_v[1] * p->rot_v[1] + p->rot_v[2] * p->rot_v[2]);
  const float div_physical_v = p->div_v + DIM_F * c->H;
  const float abs_div_physical_v = fabsf(div_physical_v);
  float pressure, soundspeed;
  eos_eval(c, p, &pressure, &soundspeed);
  const float common_factor = p->h * DIM_INV_F / p->wcount;
  float grad_h_term;
  const float lim = planetary ? 0.999f : 0.9999f;
  if (p->h > lim * c->prm->h_max) {
    grad_h_term = 0.f;
  } else {
    const float grad_W_term = common_factor * p->wcount_dh;
    if (grad_W_term < -0.9999f)
      grad_h_term = 0.f;
    else
      grad_h_term = common_factor * p->rho_dh / (1.f + grad_W_term);
  }
  float balsara;
  if (planetary)
    balsara = c->prm->viscosity_alpha * abs_div_physical_v /
              (abs_div_physical_v + curl_v + 0.0001f * fac_Balsara_eps * soundspeed / p->h);
  else
    balsara = c->prm->viscosity_alpha * abs_div_physical_v /
              (abs_div_physical_v + curl_v + 0.0001f * fac_Balsara_eps * soundspeed * h_inv);
  p->f = grad_h_term;
  p->pressure = pressure;
  p->soundspeed = soundspeed;
  p->balsara = balsara;
  /* reset_acceleration */
  p->a[0] = p->a[1] = p->a[2] = 0.0f;
  p->u_dt = 0.0f;
  p->h_dt = 0.0f;
  p->v_sig = planetary ? p->soundspeed : 2.f * p->soundspeed;
}

/* ------------------------------------------------------------------------ */
/* Sorts: runner_do_hydro_sort, src/runner_sort.c:201-431                    */
/* ------------------------------------------------------------------------ */
static int cmp_sort(const void *a, const void *b) {
  const osort *x = (const osort *)a, *y = (const osort *)b;
  if (x->d < y->d) return -1;
  if (x->d > y->d) return 1;
  return x->i - y->i;
}

/* The key is the f64 dot product of the ABSOLUTE position with the axis,
 * rounded to f32 (:377-387); split cells merge their progeny's arrays
 * (:284-350), i.e. hold the same keys.  Tie order is unspecified in the
 * reference (quicksort) and does not change which pairs interact. */
static const osort *get_sort(octx *c, ocell *cell, int sid) {
  if (cell->sort[sid] != NULL) return cell->sort[sid];
  osort *s = (osort *)malloc(sizeof(osort) * (size_t)(cell->count + 1));
  const opart *parts = &c->parts[cell->first];
  for (int k = 0; k < cell->count; k++) {
    const double px[3] = {parts[k].x[0], parts[k].x[1], parts[k].x[2]};
    s[k].i = k;
    s[k].d = px[0] * axis_dir[sid][0] + px[1] * axis_dir[sid][1] + px[2] * axis_dir[sid][2];
  }
  qsort(s, (size_t)cell->count, sizeof(osort), cmp_sort);
  s[cell->count].d = FLT_MAX;
  s[cell->count].i = 0;
  cell->sort[sid] = s;
  return s;
}

/* ------------------------------------------------------------------------ */
/* space_getsid: src/space_getsid.h:46-84                                    */
/* ------------------------------------------------------------------------ */
static int get_sid(const octx *c, ocell **ci, ocell **cj, double shift[3]) {
  const int periodic = c->prm->periodic;
  double dx[3];
  for (int k = 0; k < 3; k++) {
    dx[k] = (*cj)->loc[k] - (*ci)->loc[k];
    if (periodic && dx[k] < -c->prm->dim[k] / 2)
      shift[k] = c->prm->dim[k];
    else if (periodic && dx[k] > c->prm->dim[k] / 2)
      shift[k] = -c->prm->dim[k];
    else
      shift[k] = 0.0;
    dx[k] += shift[k];
  }
  int dir = 0;
  for (int k = 0; k < 3; k++) dir = 3 * dir + ((dx[k] < 0.0) ? 0 : ((dx[k] > 0.0) ? 2 : 1));
  if (dir_flipped(dir)) {
    ocell *t = *ci;
    *ci = *cj;
    *cj = t;
    for (int k = 0; k < 3; k++) shift[k] = -shift[k];
  }
  return dir_to_sid(dir);
}

/* ------------------------------------------------------------------------ */
/* DOPAIR1 (density): src/runner_doiact_functions_hydro.h:1068-1320           */
/* ------------------------------------------------------------------------ */
static void pair_density(octx *c, ocell *ci, ocell *cj, int sid, const double shift[3]) {
  double rshift = 0.0;
  for (int k = 0; k < 3; k++) rshift += shift[k] * axis_dir[sid][k];
  const osort *sort_i = get_sort(c, ci, sid);
  const osort *sort_j = get_sort(c, cj, sid);
  const double hi_max = ci->h_max * K_GAMMA - rshift;
  const double hj_max = cj->h_max * K_GAMMA;
  const int count_i = ci->count, count_j = cj->count;
  opart *parts_i = &c->parts[ci->first];
  opart *parts_j = &c->parts[cj->first];
  const double di_max = sort_i[count_i - 1].d - rshift;
  const double dj_min = sort_j[0].d;
  const float dx_max = 0.f; /* dx_max_sort of both cells: nothing moved since the sort */

  if (ci->active) {
    for (int pid = count_i - 1; pid >= 0 && sort_i[pid].d + hi_max + dx_max > dj_min; pid--) {
      opart *pi = &parts_i[sort_i[pid].i];
      const float hi = pi->h;
      if (!part_active(c, pi)) continue;
      const double di = sort_i[pid].d + hi * K_GAMMA + dx_max - rshift;
      if (di < dj_min) continue;
      const float hig2 = hi * hi * K_GAMMA2;
      const float pix = pi->x[0] - (cj->loc[0] + shift[0]);
      const float piy = pi->x[1] - (cj->loc[1] + shift[1]);
      const float piz = pi->x[2] - (cj->loc[2] + shift[2]);
      for (int pjd = 0; pjd < count_j && sort_j[pjd].d < di; pjd++) {
        opart *pj = &parts_j[sort_j[pjd].i];
        const float pjx = pj->x[0] - cj->loc[0];
        const float pjy = pj->x[1] - cj->loc[1];
        const float pjz = pj->x[2] - cj->loc[2];
        float dx[3] = {pix - pjx, piy - pjy, piz - pjz};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hig2) iact_density(c, r2, dx, hi, pi, pj);
      }
    }
  }
  if (cj->active) {
    for (int pjd = 0; pjd < count_j && sort_j[pjd].d - hj_max - dx_max < di_max; pjd++) {
      opart *pj = &parts_j[sort_j[pjd].i];
      const float hj = pj->h;
      if (!part_active(c, pj)) continue;
      const double dj = sort_j[pjd].d - hj * K_GAMMA - dx_max + rshift;
      if (dj - rshift > di_max) continue;
      const float hjg2 = hj * hj * K_GAMMA2;
      const float pjx = pj->x[0] - cj->loc[0];
      const float pjy = pj->x[1] - cj->loc[1];
      const float pjz = pj->x[2] - cj->loc[2];
      for (int pid = count_i - 1; pid >= 0 && sort_i[pid].d > dj; pid--) {
        opart *pi = &parts_i[sort_i[pid].i];
        const float pix = pi->x[0] - (cj->loc[0] + shift[0]);
        const float piy = pi->x[1] - (cj->loc[1] + shift[1]);
        const float piz = pi->x[2] - (cj->loc[2] + shift[2]);
        float dx[3] = {pjx - pix, pjy - piy, pjz - piz};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hjg2) iact_density(c, r2, dx, hj, pj, pi);
      }
    }
  }
}

/* DOPAIR1_BRANCH: :1331-1413 */
static void pair_density_branch(octx *c, ocell *ci, ocell *cj) {
  if (ci->count == 0 || cj->count == 0) return;
  if (!ci->active && !cj->active) return;
  double shift[3] = {0.0, 0.0, 0.0};
  const int sid = get_sid(c, &ci, &cj, shift);
  pair_density(c, ci, cj, sid, shift);
}

/* DOSELF1 (density): :2062-2261 */
static void self_density(octx *c, ocell *cell) {
  opart *parts = &c->parts[cell->first];
  const int count = cell->count;
  int *indt = (int *)malloc(sizeof(int) * (size_t)(count + 1));
  int countdt = 0, firstdt = 0;
  for (int k = 0; k < count; k++)
    if (part_active(c, &parts[k])) indt[countdt++] = k;
  for (int pid = 0; pid < count; pid++) {
    opart *pi = &parts[pid];
    const double pix[3] = {pi->x[0], pi->x[1], pi->x[2]};
    const float hi = pi->h;
    const float hig2 = hi * hi * K_GAMMA2;
    if (!part_active(c, pi)) {
      for (int pjd = firstdt; pjd < countdt; pjd++) {
        opart *pj = &parts[indt[pjd]];
        const float hj = pj->h;
        float r2 = 0.0f, dx[3];
        for (int k = 0; k < 3; k++) {
          dx[k] = pj->x[k] - pix[k];
          r2 += dx[k] * dx[k];
        }
        if (r2 < hj * hj * K_GAMMA2) iact_density(c, r2, dx, hj, pj, pi);
      }
    } else {
      firstdt += 1;
      for (int pjd = pid + 1; pjd < count; pjd++) {
        opart *pj = &parts[pjd];
        const float hj = pj->h;
        float r2 = 0.0f, dx[3];
        for (int k = 0; k < 3; k++) {
          dx[k] = pix[k] - pj->x[k];
          r2 += dx[k] * dx[k];
        }
        const int doj = part_active(c, pj) && (r2 < hj * hj * K_GAMMA2);
        const int doi = (r2 < hig2);
        /* limiter loop: doi && doj is the symmetric runner_iact_limiter, which does nothing
         * (src/runner_doiact_functions_limiter.h:655-660, src/timestep_limiter_iact.h:76-102) -- in the
         * sorted pair walk there is no such case, a starting neighbour is written like any other */
        if (c->stage == 4 && doi && doj) continue;
        if (doi) iact_density(c, r2, dx, hi, pi, pj);
        if (doj) {
          float mdx[3] = {-dx[0], -dx[1], -dx[2]};
          iact_density(c, r2, mdx, hj, pj, pi);
        }
      }
    }
  }
  free(indt);
}

/* DOSELF1_BRANCH: :2271-2296 */
static void self_density_branch(octx *c, ocell *cell) {
  if (cell->count == 0) return;
  if (!cell->active) return;
  if (cell->h_max_old * K_GAMMA > cell->dmin) {
    fprintf(stderr, "oracle: Cell smaller than smoothing length\n");
    abort();
  }
  self_density(c, cell);
}

/* ------------------------------------------------------------------------ */
/* DOPAIR2 (force): :1424-1961                                                */
/* ------------------------------------------------------------------------ */
static void pair_force(octx *c, ocell *ci, ocell *cj, int sid, const double shift[3]) {
  double rshift = 0.0;
  for (int k = 0; k < 3; k++) rshift += shift[k] * axis_dir[sid][k];
  const osort *sort_i = get_sort(c, ci, sid);
  const osort *sort_j = get_sort(c, cj, sid);
  const double hi_max = ci->h_max;
  const double hj_max = cj->h_max;
  const int count_i = ci->count, count_j = cj->count;
  opart *parts_i = &c->parts[ci->first];
  opart *parts_j = &c->parts[cj->first];
  const double dx_max = 0.0;
  const double di_max = sort_i[count_i - 1].d;
  const double dj_min = sort_j[0].d;
  const double shift_i[3] = {cj->loc[0] + shift[0], cj->loc[1] + shift[1], cj->loc[2] + shift[2]};
  const double shift_j[3] = {cj->loc[0], cj->loc[1], cj->loc[2]};

  /* active-only copies of the sort lists (:1484-1524) */
  int count_active_i = 0, count_active_j = 0;
  osort *sort_active_i = NULL, *sort_active_j = NULL;
  if (ci->active) {
    sort_active_i = (osort *)malloc(sizeof(osort) * (size_t)(count_i + 1));
    for (int k = 0; k < count_i; k++)
      if (part_active(c, &parts_i[sort_i[k].i])) sort_active_i[count_active_i++] = sort_i[k];
  }
  if (cj->active) {
    sort_active_j = (osort *)malloc(sizeof(osort) * (size_t)(count_j + 1));
    for (int k = 0; k < count_j; k++)
      if (part_active(c, &parts_j[sort_j[k].i])) sort_active_j[count_active_j++] = sort_j[k];
  }

  for (int pid = count_i - 1; pid >= 0 && sort_i[pid].d + hi_max * K_GAMMA + dx_max - rshift > dj_min; pid--) {
    opart *pi = &parts_i[sort_i[pid].i];
    const float hi = pi->h;
    const double di = sort_i[pid].d + hi * K_GAMMA + dx_max - rshift;
    if (di < dj_min) continue;
    const float hig2 = hi * hi * K_GAMMA2;
    const float pix = pi->x[0] - shift_i[0];
    const float piy = pi->x[1] - shift_i[1];
    const float piz = pi->x[2] - shift_i[2];
    if (!part_active(c, pi)) {
      for (int pjd = 0; pjd < count_active_j && sort_active_j[pjd].d < di; pjd++) {
        opart *pj = &parts_j[sort_active_j[pjd].i];
        const float hj = pj->h;
        const float pjx = pj->x[0] - shift_j[0];
        const float pjy = pj->x[1] - shift_j[1];
        const float pjz = pj->x[2] - shift_j[2];
        const float dx[3] = {pjx - pix, pjy - piy, pjz - piz};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hig2) iact_force(c, r2, dx, hj, hi, pj, pi);
      }
    } else {
      for (int pjd = 0; pjd < count_j && sort_j[pjd].d < di; pjd++) {
        opart *pj = &parts_j[sort_j[pjd].i];
        const float hj = pj->h;
        const float pjx = pj->x[0] - shift_j[0];
        const float pjy = pj->x[1] - shift_j[1];
        const float pjz = pj->x[2] - shift_j[2];
        const float dx[3] = {pix - pjx, piy - pjy, piz - pjz};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hig2) {
          iact_force(c, r2, dx, hi, hj, pi, pj);
          if (part_active(c, pj)) {
            const float mdx[3] = {-dx[0], -dx[1], -dx[2]};
            iact_force(c, r2, mdx, hj, hi, pj, pi);
          }
        }
      }
    }
  }

  for (int pjd = 0; pjd < count_j && sort_j[pjd].d - hj_max * K_GAMMA - dx_max < di_max - rshift; pjd++) {
    opart *pj = &parts_j[sort_j[pjd].i];
    const float hj = pj->h;
    const double dj = sort_j[pjd].d - hj * K_GAMMA - dx_max;
    if (dj > di_max - rshift) continue;
    const float hjg2 = hj * hj * K_GAMMA2;
    const float pjx = pj->x[0] - shift_j[0];
    const float pjy = pj->x[1] - shift_j[1];
    const float pjz = pj->x[2] - shift_j[2];
    if (!part_active(c, pj)) {
      for (int pid = count_active_i - 1; pid >= 0 && sort_active_i[pid].d - rshift > dj; pid--) {
        opart *pi = &parts_i[sort_active_i[pid].i];
        const float hi = pi->h;
        const float hig2 = hi * hi * K_GAMMA2;
        const float pix = pi->x[0] - shift_i[0];
        const float piy = pi->x[1] - shift_i[1];
        const float piz = pi->x[2] - shift_i[2];
        const float dx[3] = {pix - pjx, piy - pjy, piz - pjz};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hjg2 && r2 >= hig2) iact_force(c, r2, dx, hi, hj, pi, pj);
      }
    } else {
      for (int pid = count_i - 1; pid >= 0 && sort_i[pid].d - rshift > dj; pid--) {
        opart *pi = &parts_i[sort_i[pid].i];
        const float hi = pi->h;
        const float hig2 = hi * hi * K_GAMMA2;
        const float pix = pi->x[0] - shift_i[0];
        const float piy = pi->x[1] - shift_i[1];
        const float piz = pi->x[2] - shift_i[2];
        const float dx[3] = {pjx - pix, pjy - piy, pjz - piz};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hjg2 && r2 >= hig2) {
          iact_force(c, r2, dx, hj, hi, pj, pi);
          if (part_active(c, pi)) {
            const float mdx[3] = {-dx[0], -dx[1], -dx[2]};
            iact_force(c, r2, mdx, hi, hj, pi, pj);
          }
        }
      }
    }
  }
  free(sort_active_i);
  free(sort_active_j);
}

/* DOPAIR2_BRANCH: :1972-2054 */
static void pair_force_branch(octx *c, ocell *ci, ocell *cj) {
  if (ci->count == 0 || cj->count == 0) return;
  if (!ci->active && !cj->active) return;
  double shift[3] = {0.0, 0.0, 0.0};
  const int sid = get_sid(c, &ci, &cj, shift);
  pair_force(c, ci, cj, sid, shift);
}

/* DOSELF2 (force): :2304-2476 */
static void self_force(octx *c, ocell *cell) {
  opart *parts = &c->parts[cell->first];
  const int count = cell->count;
  int *indt = (int *)malloc(sizeof(int) * (size_t)(count + 1));
  int countdt = 0, firstdt = 0;
  for (int k = 0; k < count; k++)
    if (part_active(c, &parts[k])) indt[countdt++] = k;
  for (int pid = 0; pid < count; pid++) {
    opart *pi = &parts[pid];
    const double pix[3] = {pi->x[0], pi->x[1], pi->x[2]};
    const float hi = pi->h;
    const float hig2 = hi * hi * K_GAMMA2;
    if (!part_active(c, pi)) {
      for (int pjd = firstdt; pjd < countdt; pjd++) {
        opart *pj = &parts[indt[pjd]];
        const float hj = pj->h;
        float r2 = 0.0f, dx[3];
        for (int k = 0; k < 3; k++) {
          dx[k] = pj->x[k] - pix[k];
          r2 += dx[k] * dx[k];
        }
        if (r2 < hig2 || r2 < hj * hj * K_GAMMA2) iact_force(c, r2, dx, hj, hi, pj, pi);
      }
    } else {
      firstdt += 1;
      for (int pjd = pid + 1; pjd < count; pjd++) {
        opart *pj = &parts[pjd];
        const float hj = pj->h;
        float r2 = 0.0f, dx[3];
        for (int k = 0; k < 3; k++) {
          dx[k] = pix[k] - pj->x[k];
          r2 += dx[k] * dx[k];
        }
        if (r2 < hig2 || r2 < hj * hj * K_GAMMA2) {
          iact_force(c, r2, dx, hi, hj, pi, pj);
          if (part_active(c, pj)) {
            const float mdx[3] = {-dx[0], -dx[1], -dx[2]};
            iact_force(c, r2, mdx, hj, hi, pj, pi);
          }
        }
      }
    }
  }
  free(indt);
}

static void self_force_branch(octx *c, ocell *cell) {
  if (cell->count == 0) return;
  if (!cell->active) return;
  if (cell->h_max_old * K_GAMMA > cell->dmin) {
    fprintf(stderr, "oracle: Cell smaller than smoothing length\n");
    abort();
  }
  self_force(c, cell);
}

/* ------------------------------------------------------------------------ */
/* Recursion: DOSUB_PAIR1/2, DOSUB_SELF1/2, :2524-2719; criteria cell.h:761-782 */
/* ------------------------------------------------------------------------ */
static inline int can_recurse_pair(const ocell *c) {
  return c->split && ((K_GAMMA * c->h_max_old + 0.f) < 0.5f * c->dmin);
}
static inline int can_recurse_self(const ocell *c) { return c->split && (K_GAMMA * c->h_max_old < 0.5f * c->dmin); }

/* The progeny pairs of a (ci, cj, sid) pair, role of cell_split_pairs
 * (src/cell.c:62): progeny k sits at offset bit(k,4|2|1)*width/2 along x|y|z
 * (src/space_split.c:223-225); a progeny pair is listed when the two octants
 * touch, i.e. when 2*dir + oct_j - oct_i lies in {-1,0,1} on every axis. */
static int split_pairs(int sid, int out_pid[16], int out_pjd[16]) {
  const int k = 26 - sid; /* un-flipped direction of cj seen from ci */
  const int dir[3] = {k / 9 - 1, (k / 3) % 3 - 1, k % 3 - 1};
  int n = 0;
  for (int pid = 0; pid < 8; pid++)
    for (int pjd = 0; pjd < 8; pjd++) {
      const int bi[3] = {(pid >> 2) & 1, (pid >> 1) & 1, pid & 1};
      const int bj[3] = {(pjd >> 2) & 1, (pjd >> 1) & 1, pjd & 1};
      int ok = 1;
      for (int d = 0; d < 3; d++) {
        const int off = 2 * dir[d] + bj[d] - bi[d];
        if (off < -1 || off > 1) ok = 0;
      }
      if (ok) {
        out_pid[n] = pid;
        out_pjd[n] = pjd;
        n++;
      }
    }
  return n;
}

static void sub_pair(octx *c, ocell *ci, ocell *cj, int force) {
  if (!ci->active && !cj->active) return;
  if (ci->count == 0 || cj->count == 0) return;
  double shift[3];
  const int sid = get_sid(c, &ci, &cj, shift);
  if (can_recurse_pair(ci) && can_recurse_pair(cj)) {
    int pid[16], pjd[16];
    const int n = split_pairs(sid, pid, pjd);
    for (int k = 0; k < n; k++)
      if (ci->progeny[pid[k]] >= 0 && cj->progeny[pjd[k]] >= 0)
        sub_pair(c, &c->cells[ci->progeny[pid[k]]], &c->cells[cj->progeny[pjd[k]]], force);
  } else if (ci->active || cj->active) {
    if (force)
      pair_force_branch(c, ci, cj);
    else
      pair_density_branch(c, ci, cj);
  }
}

static void sub_self(octx *c, ocell *ci, int force) {
  if (ci->count == 0 || !ci->active) return;
  if (can_recurse_self(ci)) {
    for (int k = 0; k < 8; k++)
      if (ci->progeny[k] >= 0) {
        sub_self(c, &c->cells[ci->progeny[k]], force);
        for (int j = k + 1; j < 8; j++)
          if (ci->progeny[j] >= 0) sub_pair(c, &c->cells[ci->progeny[k]], &c->cells[ci->progeny[j]], force);
      }
  } else {
    if (force)
      self_force_branch(c, ci);
    else
      self_density_branch(c, ci);
  }
}

/* ------------------------------------------------------------------------ */
/* Subset loops for the ghost redo: DOPAIR_SUBSET :710-870, DOSELF_SUBSET      */
/* :949-1036, DOPAIR_SUBSET_BRANCH :884-937, DOSUB_SUBSET :2721-2805           */
/* ------------------------------------------------------------------------ */
static void pair_subset(octx *c, ocell *ci, opart *parts_i, const int *ind, int count, ocell *cj, int sid,
                        int flipped, const double shift[3]) {
  const int count_j = cj->count;
  opart *parts_j = &c->parts[cj->first];
  const osort *sort_j = get_sort(c, cj, sid);
  const float dxj = 0.f; /* cj->hydro.dx_max_sort */
  (void)ci;
  for (int pid = 0; pid < count; pid++) {
    opart *pi = &parts_i[ind[pid]];
    const double pix = pi->x[0] - (shift[0]);
    const double piy = pi->x[1] - (shift[1]);
    const double piz = pi->x[2] - (shift[2]);
    const float hi = pi->h;
    const float hig2 = hi * hi * K_GAMMA2;
    if (!flipped) {
      const double di = hi * K_GAMMA + dxj + pix * axis_dir[sid][0] + piy * axis_dir[sid][1] + piz * axis_dir[sid][2];
      for (int pjd = 0; pjd < count_j && sort_j[pjd].d < di; pjd++) {
        opart *pj = &parts_j[sort_j[pjd].i];
        const double pjx = pj->x[0], pjy = pj->x[1], pjz = pj->x[2];
        float dx[3] = {(float)(pix - pjx), (float)(piy - pjy), (float)(piz - pjz)};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hig2) iact_density(c, r2, dx, hi, pi, pj);
      }
    } else {
      const double di = -hi * K_GAMMA - dxj + pix * axis_dir[sid][0] + piy * axis_dir[sid][1] + piz * axis_dir[sid][2];
      for (int pjd = count_j - 1; pjd >= 0 && di < sort_j[pjd].d; pjd--) {
        opart *pj = &parts_j[sort_j[pjd].i];
        const double pjx = pj->x[0], pjy = pj->x[1], pjz = pj->x[2];
        float dx[3] = {(float)(pix - pjx), (float)(piy - pjy), (float)(piz - pjz)};
        const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
        if (r2 < hig2) iact_density(c, r2, dx, hi, pi, pj);
      }
    }
  }
}

static void pair_subset_branch(octx *c, ocell *ci, opart *parts_i, const int *ind, int count, ocell *cj) {
  if (cj->count == 0) return;
  double shift[3] = {0.0, 0.0, 0.0};
  /* note: wraps unconditionally, without looking at s->periodic (:898-904) */
  for (int k = 0; k < 3; k++) {
    if (cj->loc[k] - ci->loc[k] < -c->prm->dim[k] / 2)
      shift[k] = c->prm->dim[k];
    else if (cj->loc[k] - ci->loc[k] > c->prm->dim[k] / 2)
      shift[k] = -c->prm->dim[k];
  }
  int dir = 0;
  for (int k = 0; k < 3; k++)
    dir = 3 * dir + ((cj->loc[k] - ci->loc[k] + shift[k] < 0) ? 0 : (cj->loc[k] - ci->loc[k] + shift[k] > 0) ? 2 : 1);
  const int flipped = dir_flipped(dir);
  const int sid = dir_to_sid(dir);
  pair_subset(c, ci, parts_i, ind, count, cj, sid, flipped, shift);
}

static void self_subset(octx *c, ocell *ci, opart *parts, const int *ind, int count) {
  const int count_i = ci->count;
  opart *parts_j = &c->parts[ci->first];
  for (int pid = 0; pid < count; pid++) {
    opart *pi = &parts[ind[pid]];
    const float pix[3] = {(float)(pi->x[0] - ci->loc[0]), (float)(pi->x[1] - ci->loc[1]),
                          (float)(pi->x[2] - ci->loc[2])};
    const float hi = pi->h;
    const float hig2 = hi * hi * K_GAMMA2;
    for (int pjd = 0; pjd < count_i; pjd++) {
      opart *pj = &parts_j[pjd];
      if (pi == pj) continue;
      const float pjx[3] = {(float)(pj->x[0] - ci->loc[0]), (float)(pj->x[1] - ci->loc[1]),
                            (float)(pj->x[2] - ci->loc[2])};
      float dx[3] = {pix[0] - pjx[0], pix[1] - pjx[1], pix[2] - pjx[2]};
      const float r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
      if (r2 < hig2) iact_density(c, r2, dx, hi, pi, pj);
    }
  }
}

static void sub_subset(octx *c, ocell *ci, opart *parts, const int *ind, int count, ocell *cj) {
  if (!ci->active && (cj == NULL || !cj->active)) return;
  if (ci->count == 0 || (cj != NULL && cj->count == 0)) return;
  /* progeny of ci holding the particles (they all sit in one leaf) */
  int sub = -1;
  if (ci->split) {
    const long pos = (long)(&parts[ind[0]] - c->parts);
    for (int k = 0; k < 8; k++)
      if (ci->progeny[k] >= 0) {
        const ocell *p = &c->cells[ci->progeny[k]];
        if (pos >= p->first && pos < p->first + p->count) {
          sub = ci->progeny[k];
          break;
        }
      }
  }
  if (cj == NULL) {
    if (can_recurse_self(ci)) {
      sub_subset(c, &c->cells[sub], parts, ind, count, NULL);
      for (int j = 0; j < 8; j++)
        if (ci->progeny[j] != sub && ci->progeny[j] >= 0)
          sub_subset(c, &c->cells[sub], parts, ind, count, &c->cells[ci->progeny[j]]);
    } else {
      self_subset(c, ci, parts, ind, count);
    }
  } else {
    if (can_recurse_pair(ci) && can_recurse_pair(cj)) {
      double shift[3] = {0.0, 0.0, 0.0};
      const int sid = get_sid(c, &ci, &cj, shift);
      int pid[16], pjd[16];
      const int n = split_pairs(sid, pid, pjd);
      for (int k = 0; k < n; k++) {
        if (ci->progeny[pid[k]] == sub && sub >= 0 && cj->progeny[pjd[k]] >= 0)
          sub_subset(c, &c->cells[ci->progeny[pid[k]]], parts, ind, count, &c->cells[cj->progeny[pjd[k]]]);
        if (ci->progeny[pid[k]] >= 0 && cj->progeny[pjd[k]] == sub && sub >= 0)
          sub_subset(c, &c->cells[cj->progeny[pjd[k]]], parts, ind, count, &c->cells[ci->progeny[pid[k]]]);
      }
    } else if (ci->active || cj->active) {
      pair_subset_branch(c, ci, parts, ind, count, cj);
    }
  }
}

/* ------------------------------------------------------------------------ */
/* runner_do_ghost: src/runner_ghost.c:1085-1596                              */
/* ------------------------------------------------------------------------ */
static int ghost_failed = 0;

static void ghost(octx *c, ocell *cell) {
  const oracle_params *prm = c->prm;
  const float hydro_h_max = prm->h_max;
  const float hydro_h_min = prm->h_min;
  const float eps = prm->h_tolerance;
  const float hydro_eta_dim = prm->eta_neighbours * prm->eta_neighbours * prm->eta_neighbours;
  const int max_smoothing_iter = prm->max_smoothing_iterations;
  float h_max = cell->h_max;
  float h_max_active = cell->h_max_active;
  if (cell->count == 0) return;
  if (!cell->active) return;

  if (cell->split) {
    for (int k = 0; k < 8; k++)
      if (cell->progeny[k] >= 0) {
        ocell *p = &c->cells[cell->progeny[k]];
        ghost(c, p);
        h_max = fmax_f(h_max, p->h_max);
        h_max_active = fmax_f(h_max_active, p->h_max_active);
      }
  } else {
    opart *parts = &c->parts[cell->first];
    int *pid = (int *)malloc(sizeof(int) * (size_t)cell->count);
    float *h_0 = (float *)malloc(sizeof(float) * (size_t)cell->count);
    float *left = (float *)malloc(sizeof(float) * (size_t)cell->count);
    float *right = (float *)malloc(sizeof(float) * (size_t)cell->count);
    int count = 0, redo = 0;
    for (int k = 0; k < cell->count; k++)
      if (part_active(c, &parts[k])) {
        pid[count] = k;
        h_0[count] = parts[k].h;
        left[count] = 0.f;
        right[count] = hydro_h_max;
        ++count;
      }
    for (int num_reruns = 0; count > 0 && num_reruns < max_smoothing_iter; num_reruns++) {
      redo = 0;
      for (int i = 0; i < count; i++) {
        opart *p = &parts[pid[i]];
        const float h_old = p->h;
        const float h_old_dim = h_old * h_old * h_old;
        const float h_old_dim_minus_one = h_old * h_old;
        float h_new;
        int has_no_neighbours = 0;
        if (p->wcount < 1.e-5 * K_ROOT) {
          has_no_neighbours = 1;
          h_new = 2.f * h_old;
        } else {
          end_density(p);
          if (prm->scheme == ORACLE_SCHEME_SPHENIX) p->div_v += c->H * DIM_F; /* SPHENIX/hydro.h:633-634 */
          const float n_sum = p->wcount * h_old_dim;
          const float n_target = hydro_eta_dim;
          const float f = n_sum - n_target;
          const float f_prime = p->wcount_dh * h_old_dim + DIM_F * p->wcount * h_old_dim_minus_one;
          if (n_sum < n_target)
            left[i] = fmax_f(left[i], h_old);
          else if (n_sum > n_target)
            right[i] = fmin_f(right[i], h_old);
          if (((p->h >= hydro_h_max) && (f < 0.f)) || ((p->h <= hydro_h_min) && (f > 0.f))) {
            prepare_force(c, p);
            continue;
          }
          h_new = h_old - f / (f_prime + FLT_MIN);
          h_new = fmin_f(h_new, 2.f * h_old);
          h_new = fmax_f(h_new, 0.5f * h_old);
          h_new = fmax_f(h_new, left[i]);
          h_new = fmin_f(h_new, right[i]);
        }
        if (fabsf(h_new - h_old) > eps * h_old) {
          if ((h_new == left[i] && h_old == right[i]) || (h_old == left[i] && h_new == right[i])) {
            p->h = cbrtf(0.5f * (left[i] * left[i] * left[i] + right[i] * right[i] * right[i]));
          } else {
            p->h = h_new;
          }
          if (p->h < hydro_h_max && p->h > hydro_h_min) {
            pid[redo] = pid[i];
            h_0[redo] = h_0[i];
            left[redo] = left[i];
            right[redo] = right[i];
            redo += 1;
            init_part(p);
            continue;
          } else if (p->h <= hydro_h_min) {
            p->h = hydro_h_min;
          } else if (p->h >= hydro_h_max) {
            p->h = hydro_h_max;
            if (has_no_neighbours) no_neighbours(p);
          }
        }
        h_max = fmax_f(h_max, p->h);
        h_max_active = fmax_f(h_max_active, p->h);
        prepare_force(c, p);
      }
      count = redo;
      if (count > 0) {
        c->stage = 1;
        if (c->log && num_reruns + 1 > c->log->ghost_sweeps) c->log->ghost_sweeps = num_reruns + 1;
        for (int fi = (int)(cell - c->cells); fi >= 0; fi = c->cells[fi].parent) {
          ocell *finger = &c->cells[fi];
          /* links are prepended when tasks are created, so the list is walked newest first */
          for (int l = finger->ntasks - 1; l >= 0; l--) {
            const oracle_task *t = &c->tasks[finger->tasks[l]];
            ocell *other = NULL;
            if (t->cj >= 0) other = (t->ci == fi) ? &c->cells[t->cj] : &c->cells[t->ci];
            if (t->type == ORACLE_TASK_SELF)
              self_subset(c, finger, parts, pid, count);
            else if (t->type == ORACLE_TASK_PAIR)
              pair_subset_branch(c, finger, parts, pid, count, other);
            else if (t->type == ORACLE_TASK_SUB_SELF)
              sub_subset(c, finger, parts, pid, count, NULL);
            else
              sub_subset(c, finger, parts, pid, count, other);
          }
        }
      }
    }
    if (count) {
      fprintf(stderr, "oracle: Smoothing length failed to converge on %i particles.\n", count);
      ghost_failed = 1;
    }
    free(left);
    free(right);
    free(pid);
    free(h_0);
  }
  cell->h_max = h_max;
  cell->h_max_active = h_max_active;
}


/* ------------------------------------------------------------------------ */
/* Tree: space_regrid (src/space_regrid.c:120-127, 220, 297-301),
 * cell_split (src/cell_split.c:72, 109-110), space_split_recursive
 * (src/space_split.c:199-262)                                               */
/* ------------------------------------------------------------------------ */
typedef struct {
  oracle_cell *cells;
  int n, cap;
} cellvec;

static int push_cell(cellvec *v) {
  if (v->n == v->cap) {
    v->cap = v->cap ? 2 * v->cap : 1024;
    v->cells = (oracle_cell *)realloc(v->cells, sizeof(oracle_cell) * (size_t)v->cap);
  }
  memset(&v->cells[v->n], 0, sizeof(oracle_cell));
  for (int k = 0; k < 8; k++) v->cells[v->n].progeny[k] = -1;
  v->cells[v->n].parent = -1;
  return v->n++;
}

static void split_recursive(cellvec *v, int ci, const double *x, int *order, int *scratch) {
  oracle_cell c = v->cells[ci];
  if (c.count <= SPLITSIZE) return;
  const double pivot[3] = {c.loc[0] + c.width[0] / 2, c.loc[1] + c.width[1] / 2, c.loc[2] + c.width[2] / 2};
  int bucket[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int k = 0; k < c.count; k++) {
    const double *p = &x[3 * (size_t)order[c.first + k]];
    const int bid = (p[0] >= pivot[0]) * 4 + (p[1] >= pivot[1]) * 2 + (p[2] >= pivot[2]);
    bucket[bid]++;
  }
  int off[9];
  off[0] = 0;
  for (int k = 0; k < 8; k++) off[k + 1] = off[k] + bucket[k];
  int fill[8];
  for (int k = 0; k < 8; k++) fill[k] = off[k];
  for (int k = 0; k < c.count; k++) {
    const int o = order[c.first + k];
    const double *p = &x[3 * (size_t)o];
    const int bid = (p[0] >= pivot[0]) * 4 + (p[1] >= pivot[1]) * 2 + (p[2] >= pivot[2]);
    scratch[fill[bid]++] = o;
  }
  memcpy(&order[c.first], scratch, sizeof(int) * (size_t)c.count);
  v->cells[ci].split = 1;
  for (int k = 0; k < 8; k++) {
    if (bucket[k] == 0) continue;
    const int pi = push_cell(v);
    oracle_cell *p = &v->cells[pi];
    for (int d = 0; d < 3; d++) {
      p->loc[d] = c.loc[d];
      p->width[d] = c.width[d] / 2;
    }
    if (k & 4) p->loc[0] += p->width[0];
    if (k & 2) p->loc[1] += p->width[1];
    if (k & 1) p->loc[2] += p->width[2];
    p->first = c.first + off[k];
    p->count = bucket[k];
    p->parent = ci;
    p->depth = c.depth + 1;
    v->cells[ci].progeny[k] = pi;
    split_recursive(v, pi, x, order, scratch);
  }
}

/* Builds the top-level grid and octree.  `order[k]` is the input index of the
 * particle stored at position k; cells index into that order. */
int oracle_build_tree(const oracle_params *prm, int n, const double *x, const float *h, int *order,
                      oracle_cell **cells_out, int *ncells_out, int cdim_out[3]) {
  float h_max = 0.f;
  for (int k = 0; k < n; k++)
    if (h[k] > h_max) h_max = h[k];
  /* hydro h_max is capped by SPH:h_max on read-in */
  if (h_max > prm->h_max) h_max = prm->h_max;
  const double dmax = fmax(prm->dim[0], fmax(prm->dim[1], prm->dim[2]));
  const double cell_min = 0.99 * dmax / MAX_TOP_CELLS;
  int cdim[3];
  for (int d = 0; d < 3; d++) {
    cdim[d] = (int)floor(prm->dim[d] / fmax(h_max * K_GAMMA * SPACE_STRETCH, cell_min));
    if (cdim[d] < 1) return 1;
    if (prm->periodic && cdim[d] < 3) return 2;
    cdim_out[d] = cdim[d];
  }
  double width[3], iwidth[3];
  for (int d = 0; d < 3; d++) {
    width[d] = prm->dim[d] / cdim[d];
    iwidth[d] = 1.0 / width[d];
  }
  const int ntop = cdim[0] * cdim[1] * cdim[2];
  cellvec v = {NULL, 0, 0};
  int *counts = (int *)calloc((size_t)ntop + 1, sizeof(int));
  int *cid = (int *)malloc(sizeof(int) * (size_t)(n + 1));
  for (int k = 0; k < n; k++) {
    int idx[3];
    for (int d = 0; d < 3; d++) {
      idx[d] = (int)(x[3 * (size_t)k + d] * iwidth[d]);
      if (idx[d] < 0) idx[d] = 0;
      if (idx[d] >= cdim[d]) idx[d] = cdim[d] - 1;
    }
    cid[k] = (idx[0] * cdim[1] + idx[1]) * cdim[2] + idx[2];
    counts[cid[k]]++;
  }
  int *first = (int *)malloc(sizeof(int) * (size_t)(ntop + 1));
  first[0] = 0;
  for (int k = 0; k < ntop; k++) first[k + 1] = first[k] + counts[k];
  int *fill = (int *)malloc(sizeof(int) * (size_t)(ntop + 1));
  memcpy(fill, first, sizeof(int) * (size_t)(ntop + 1));
  for (int k = 0; k < n; k++) order[fill[cid[k]]++] = k;
  for (int i = 0; i < cdim[0]; i++)
    for (int j = 0; j < cdim[1]; j++)
      for (int k = 0; k < cdim[2]; k++) {
        const int ci = push_cell(&v);
        oracle_cell *c = &v.cells[ci];
        c->loc[0] = i * width[0];
        c->loc[1] = j * width[1];
        c->loc[2] = k * width[2];
        for (int d = 0; d < 3; d++) c->width[d] = width[d];
        c->first = first[ci];
        c->count = counts[ci];
      }
  int *scratch = (int *)malloc(sizeof(int) * (size_t)(n + 1));
  for (int ci = 0; ci < ntop; ci++) split_recursive(&v, ci, x, order, scratch);
  free(scratch);
  free(fill);
  free(first);
  free(cid);
  free(counts);
  *cells_out = v.cells;
  *ncells_out = v.n;
  return 0;
}

void oracle_free(void *p) { free(p); }

/* ------------------------------------------------------------------------ */
/* Tasks: engine_make_hydroloop_tasks_mapper (src/engine_maketasks.c:4060-4182)
 * followed by scheduler_splittask_hydro (src/scheduler.c:1154-1378); split
 * criteria src/cell.h:899-931.                                               */
/* ------------------------------------------------------------------------ */
typedef struct {
  oracle_task *t;
  int n, cap;
  const oracle_params *prm;
  const oracle_cell *cells;
  const float *cell_hmax;
} taskvec;

static void push_task(taskvec *v, int type, int ci, int cj) {
  if (v->n == v->cap) {
    v->cap = v->cap ? 2 * v->cap : 1024;
    v->t = (oracle_task *)realloc(v->t, sizeof(oracle_task) * (size_t)v->cap);
  }
  v->t[v->n].type = type;
  v->t[v->n].ci = ci;
  v->t[v->n].cj = cj;
  v->n++;
}

static inline double cell_dmin(const oracle_cell *c) { return fmin(c->width[0], fmin(c->width[1], c->width[2])); }
static inline int can_split_task(const taskvec *v, int ci) {
  const oracle_cell *c = &v->cells[ci];
  return c->split && (SPACE_STRETCH * K_GAMMA * v->cell_hmax[ci] < 0.5f * cell_dmin(c));
}

static int task_sid(const taskvec *v, int *ci, int *cj) {
  double dx[3];
  int dir = 0;
  for (int k = 0; k < 3; k++) {
    dx[k] = v->cells[*cj].loc[k] - v->cells[*ci].loc[k];
    if (v->prm->periodic && dx[k] < -v->prm->dim[k] / 2)
      dx[k] += v->prm->dim[k];
    else if (v->prm->periodic && dx[k] > v->prm->dim[k] / 2)
      dx[k] += -v->prm->dim[k];
    dir = 3 * dir + ((dx[k] < 0.0) ? 0 : ((dx[k] > 0.0) ? 2 : 1));
  }
  if (dir_flipped(dir)) {
    const int t = *ci;
    *ci = *cj;
    *cj = t;
  }
  return dir_to_sid(dir);
}

static void split_self_task(taskvec *v, int ci);
static void split_pair_task(taskvec *v, int ci, int cj) {
  if (v->cells[ci].count == 0 || v->cells[cj].count == 0) return;
  const int sid = task_sid(v, &ci, &cj);
  if (can_split_task(v, ci) && can_split_task(v, cj)) {
    const int h_count_i = v->cells[ci].count, h_count_j = v->cells[cj].count;
    const int do_sub = h_count_i * sid_scale[sid] < SUBSIZE_PAIR / h_count_j;
    if (do_sub && !sid_is_corner(sid)) {
      push_task(v, ORACLE_TASK_SUB_PAIR, ci, cj);
    } else {
      int pid[16], pjd[16];
      const int n = split_pairs(sid, pid, pjd);
      for (int k = 0; k < n; k++) {
        const int a = v->cells[ci].progeny[pid[k]], b = v->cells[cj].progeny[pjd[k]];
        if (a >= 0 && b >= 0) split_pair_task(v, a, b);
      }
    }
  } else {
    push_task(v, ORACLE_TASK_PAIR, ci, cj);
  }
}

static void split_self_task(taskvec *v, int ci) {
  if (v->cells[ci].count == 0) return;
  if (can_split_task(v, ci)) {
    if (v->cells[ci].count < SUBSIZE_SELF) {
      push_task(v, ORACLE_TASK_SUB_SELF, ci, -1);
    } else {
      for (int k = 0; k < 8; k++)
        if (v->cells[ci].progeny[k] >= 0) split_self_task(v, v->cells[ci].progeny[k]);
      for (int j = 0; j < 8; j++)
        if (v->cells[ci].progeny[j] >= 0)
          for (int k = j + 1; k < 8; k++)
            if (v->cells[ci].progeny[k] >= 0) split_pair_task(v, v->cells[ci].progeny[j], v->cells[ci].progeny[k]);
    }
  } else {
    push_task(v, ORACLE_TASK_SELF, ci, -1);
  }
}

/* h in cell order (after applying `order`). */
int oracle_build_tasks(const oracle_params *prm, int ncells, const oracle_cell *cells, const int cdim[3],
                       const float *h_sorted, oracle_task **tasks_out, int *ntasks_out) {
  float *hmax = (float *)calloc((size_t)ncells + 1, sizeof(float));
  for (int c = 0; c < ncells; c++) {
    float m = 0.f;
    for (int k = 0; k < cells[c].count; k++) m = fmax_f(m, h_sorted[cells[c].first + k]);
    hmax[c] = m;
  }
  taskvec v = {NULL, 0, 0, prm, cells, hmax};
  for (int i = 0; i < cdim[0]; i++)
    for (int j = 0; j < cdim[1]; j++)
      for (int k = 0; k < cdim[2]; k++) {
        const int cid = (i * cdim[1] + j) * cdim[2] + k;
        if (cells[cid].count == 0) continue;
        split_self_task(&v, cid);
        for (int ii = -1; ii < 2; ii++) {
          int iii = i + ii;
          if (!prm->periodic && (iii < 0 || iii >= cdim[0])) continue;
          iii = (iii + cdim[0]) % cdim[0];
          for (int jj = -1; jj < 2; jj++) {
            int jjj = j + jj;
            if (!prm->periodic && (jjj < 0 || jjj >= cdim[1])) continue;
            jjj = (jjj + cdim[1]) % cdim[1];
            for (int kk = -1; kk < 2; kk++) {
              int kkk = k + kk;
              if (!prm->periodic && (kkk < 0 || kkk >= cdim[2])) continue;
              kkk = (kkk + cdim[2]) % cdim[2];
              const int cjd = (iii * cdim[1] + jjj) * cdim[2] + kkk;
              if (cid >= cjd || cells[cjd].count == 0) continue;
              split_pair_task(&v, cid, cjd);
            }
          }
        }
      }
  free(hmax);
  *tasks_out = v.t;
  *ntasks_out = v.n;
  return 0;
}

/* ------------------------------------------------------------------------ */
/* One step: init -> density -> ghost -> force -> end_force                   */
/* ------------------------------------------------------------------------ */
static void cell_add_task(ocell *c, int t) {
  if (c->ntasks == c->captasks) {
    c->captasks = c->captasks ? 2 * c->captasks : 8;
    c->tasks = (int *)realloc(c->tasks, sizeof(int) * (size_t)c->captasks);
  }
  c->tasks[c->ntasks++] = t;
}

static int run_step(const oracle_params *prm, const oracle_eos *eos, oracle_parts *io, int ncells,
                    const oracle_cell *cd, int ntasks, const oracle_task *td, oracle_pairlog *log, oracle_sphenix *sx);

int oracle_run_step(const oracle_params *prm, const oracle_eos *eos, oracle_parts *io, int ncells,
                    const oracle_cell *cd, int ntasks, const oracle_task *td, oracle_pairlog *log) {
  if (prm->scheme == ORACLE_SCHEME_SPHENIX) return 4; /* the three-loop scheme: oracle_run_step_sphenix */
  return run_step(prm, eos, io, ncells, cd, ntasks, td, log, NULL);
}

int oracle_run_step_sphenix(const oracle_params *prm, oracle_parts *io, int ncells, const oracle_cell *cd, int ntasks,
                            const oracle_task *td, oracle_pairlog *log, oracle_sphenix *sx) {
  if (prm->scheme != ORACLE_SCHEME_SPHENIX || sx == NULL) return 4;
  return run_step(prm, NULL, io, ncells, cd, ntasks, td, log, sx);
}

static int run_step(const oracle_params *prm, const oracle_eos *eos, oracle_parts *io, int ncells,
                    const oracle_cell *cd, int ntasks, const oracle_task *td, oracle_pairlog *log, oracle_sphenix *sx) {
  const int n = io->n;
  octx c;
  memset(&c, 0, sizeof(c));
  c.sx = sx;
  if (sx) sx->grad_count = 0;
  c.prm = prm;
  c.eos = eos;
  c.tasks = td;
  c.ntasks = ntasks;
  c.log = log;
  c.a = 1.f;
  c.H = 0.f;
  c.beta = prm->scheme == ORACLE_SCHEME_PLANETARY ? 4.f : 3.f; /* const_viscosity_beta, hydro_parameters.h:51 */
  c.ncells = ncells;
  c.parts = (opart *)calloc((size_t)n + 1, sizeof(opart));
  c.cells = (ocell *)calloc((size_t)ncells + 1, sizeof(ocell));
  ghost_failed = 0;
  if (log) {
    log->count[0] = log->count[1] = log->count[2] = 0;
    log->ghost_sweeps = 0;
  }
  for (int k = 0; k < n; k++) {
    opart *p = &c.parts[k];
    for (int d = 0; d < 3; d++) {
      p->x[d] = io->x[3 * (size_t)k + d];
      p->v[d] = io->v[3 * (size_t)k + d];
    }
    p->mass = io->mass[k];
    p->h = io->h[k];
    p->u = io->u[k];
    p->mat_id = io->mat_id ? io->mat_id[k] : 0;
    p->id = io->id[k];
    p->time_bin = io->time_bin ? io->time_bin[k] : 1;
    p->min_ngb_time_bin = NUM_TIME_BINS + 1;
    if (sx) {
      p->alpha_visc = sx->in_alpha_visc ? sx->in_alpha_visc[k] : prm->viscosity_alpha;
      p->alpha_diff = sx->in_alpha_diff ? sx->in_alpha_diff[k] : 0.f;
      p->div_v_prev = sx->in_div_v_prev ? sx->in_div_v_prev[k] : 0.f;
    }
    /* only active particles are re-initialised (src/cell_drift.c:342); the others keep
     * what their last active step left in struct part */
    if (part_active(&c, p)) {
      init_part(p);
    } else {
      p->rho = io->in_rho ? io->in_rho[k] : 0.f;
      p->pressure = io->in_pressure ? io->in_pressure[k] : 0.f;
      p->f = io->in_f ? io->in_f[k] : 0.f;
      p->soundspeed = io->in_soundspeed ? io->in_soundspeed[k] : 0.f;
      p->balsara = io->in_balsara ? io->in_balsara[k] : 0.f;
    }
  }
  for (int k = 0; k < ncells; k++) {
    ocell *o = &c.cells[k];
    for (int d = 0; d < 3; d++) {
      o->loc[d] = cd[k].loc[d];
      o->width[d] = cd[k].width[d];
    }
    o->dmin = fmin(o->width[0], fmin(o->width[1], o->width[2]));
    o->first = cd[k].first;
    o->count = cd[k].count;
    o->parent = cd[k].parent;
    o->depth = cd[k].depth;
    o->split = cd[k].split;
    for (int j = 0; j < 8; j++) o->progeny[j] = cd[k].progeny[j];
    float hm = 0.f, hm_old = 0.f;
    int act = 0;
    for (int j = 0; j < o->count; j++) {
      hm = fmax_f(hm, c.parts[o->first + j].h);
      hm_old = fmax_f(hm_old, io->in_h_build ? io->in_h_build[o->first + j] : c.parts[o->first + j].h);
      act |= part_active(&c, &c.parts[o->first + j]);
    }
    o->h_max = o->h_max_active = hm;
    o->h_max_old = hm_old;
    o->active = act;
  }
  for (int t = 0; t < ntasks; t++) {
    cell_add_task(&c.cells[td[t].ci], t);
    if (td[t].cj >= 0) cell_add_task(&c.cells[td[t].cj], t);
  }

  if (io->l_only) goto limiter_loop;

  /* density */
  c.stage = 0;
  for (int t = 0; t < ntasks; t++) {
    ocell *ci = &c.cells[td[t].ci];
    ocell *cj = td[t].cj >= 0 ? &c.cells[td[t].cj] : NULL;
    switch (td[t].type) {
      case ORACLE_TASK_SELF: self_density_branch(&c, ci); break;
      case ORACLE_TASK_PAIR: pair_density_branch(&c, ci, cj); break;
      case ORACLE_TASK_SUB_SELF: sub_self(&c, ci, 0); break;
      default: sub_pair(&c, ci, cj, 0); break;
    }
  }
  for (int k = 0; k < n; k++) {
    const opart *p = &c.parts[k];
    if (io->d_rho) io->d_rho[k] = p->rho;
    if (io->d_rho_dh) io->d_rho_dh[k] = p->rho_dh;
    if (io->d_wcount) io->d_wcount[k] = p->wcount;
    if (io->d_wcount_dh) io->d_wcount_dh[k] = p->wcount_dh;
    if (io->d_div_v) io->d_div_v[k] = p->div_v;
    if (io->d_rot_v)
      for (int d = 0; d < 3; d++) io->d_rot_v[3 * (size_t)k + d] = p->rot_v[d];
  }

  /* ghost */
  for (int k = 0; k < ncells; k++)
    if (c.cells[k].parent < 0) ghost(&c, &c.cells[k]);
  for (int k = 0; k < n; k++) {
    const opart *p = &c.parts[k];
    if (io->g_h) io->g_h[k] = p->h;
    if (io->g_rho) io->g_rho[k] = p->rho;
    if (io->g_f) io->g_f[k] = p->f;
    if (io->g_pressure) io->g_pressure[k] = p->pressure;
    if (io->g_soundspeed) io->g_soundspeed[k] = p->soundspeed;
    if (io->g_balsara) io->g_balsara[k] = p->balsara;
    if (sx && sx->g_div_v) sx->g_div_v[k] = p->div_v;
  }

  if (sx) {
    /* gradient loop (the density walks with runner_iact_nonsym_gradient), then runner_do_extra_ghost */
    c.stage = 3;
    for (int t = 0; t < ntasks; t++) {
      ocell *ci = &c.cells[td[t].ci];
      ocell *cj = td[t].cj >= 0 ? &c.cells[td[t].cj] : NULL;
      switch (td[t].type) {
        case ORACLE_TASK_SELF: self_density_branch(&c, ci); break;
        case ORACLE_TASK_PAIR: pair_density_branch(&c, ci, cj); break;
        case ORACLE_TASK_SUB_SELF: sub_self(&c, ci, 0); break;
        default: sub_pair(&c, ci, cj, 0); break;
      }
    }
    for (int k = 0; k < n; k++) {
      const opart *p = &c.parts[k];
      if (sx->r_v_sig) sx->r_v_sig[k] = p->v_sig;
      if (sx->r_laplace_u) sx->r_laplace_u[k] = p->laplace_u;
      if (sx->r_alpha_visc_max_ngb) sx->r_alpha_visc_max_ngb[k] = p->alpha_visc_max_ngb;
    }
    for (int k = 0; k < n; k++) {
      opart *p = &c.parts[k];
      if (part_active(&c, p)) extra_ghost_part(&c, p);
      if (sx->e_laplace_u) sx->e_laplace_u[k] = p->laplace_u;
      if (sx->e_alpha_visc) sx->e_alpha_visc[k] = p->alpha_visc;
      if (sx->e_alpha_diff) sx->e_alpha_diff[k] = p->alpha_diff;
      if (sx->e_div_v_prev) sx->e_div_v_prev[k] = p->div_v_prev;
      if (sx->e_div_v_dt) sx->e_div_v_dt[k] = p->div_v_dt;
    }
  }

  /* force */
  c.stage = 2;
  for (int t = 0; t < ntasks; t++) {
    ocell *ci = &c.cells[td[t].ci];
    ocell *cj = td[t].cj >= 0 ? &c.cells[td[t].cj] : NULL;
    switch (td[t].type) {
      case ORACLE_TASK_SELF: self_force_branch(&c, ci); break;
      case ORACLE_TASK_PAIR: pair_force_branch(&c, ci, cj); break;
      case ORACLE_TASK_SUB_SELF: sub_self(&c, ci, 1); break;
      default: sub_pair(&c, ci, cj, 1); break;
    }
  }
  for (int k = 0; k < n; k++) {
    opart *p = &c.parts[k];
    /* runner_do_end_hydro_force, src/runner_others.c:618-690: hydro_end_force (Minimal/hydro.h:878-882) */
    if (part_active(&c, p)) p->h_dt *= p->h * DIM_INV_F;
    if (io->f_a)
      for (int d = 0; d < 3; d++) io->f_a[3 * (size_t)k + d] = p->a[d];
    if (io->f_u_dt) io->f_u_dt[k] = p->u_dt;
    if (io->f_h_dt) io->f_h_dt[k] = p->h_dt;
    if (io->f_v_sig) io->f_v_sig[k] = p->v_sig;
    if (io->f_min_ngb_time_bin) io->f_min_ngb_time_bin[k] = p->min_ngb_time_bin;
  }

limiter_loop:
  /* time-step limiter loop: the density-shaped walks once more, after the time-step task (new bins),
   * positions, h and the sorts as they are (src/runner_main.c:205-206, 255-256, 303-304, 349-350) */
  if (io->l_time_bin != NULL && io->l_wakeup != NULL) {
    for (int k = 0; k < n; k++) {
      c.parts[k].time_bin = io->l_time_bin[k];
      c.parts[k].wakeup = -NUM_TIME_BINS; /* time_bin_not_awake, src/timeline.h:48 */
    }
    for (int k = 0; k < ncells; k++) {
      ocell *o = &c.cells[k];
      o->active = 0; /* cell_is_starting_hydro */
      for (int j = 0; j < o->count; j++)
        if (part_active(&c, &c.parts[o->first + j])) o->active = 1;
    }
    c.stage = 4;
    c.log = NULL;
    for (int t = 0; t < ntasks; t++) {
      ocell *ci = &c.cells[td[t].ci];
      ocell *cj = td[t].cj >= 0 ? &c.cells[td[t].cj] : NULL;
      switch (td[t].type) {
        case ORACLE_TASK_SELF: self_density_branch(&c, ci); break;
        case ORACLE_TASK_PAIR: pair_density_branch(&c, ci, cj); break;
        case ORACLE_TASK_SUB_SELF: sub_self(&c, ci, 0); break;
        default: sub_pair(&c, ci, cj, 0); break;
      }
    }
    for (int k = 0; k < n; k++) io->l_wakeup[k] = c.parts[k].wakeup;
  }

  for (int k = 0; k < ncells; k++) {
    for (int s = 0; s < 13; s++) free(c.cells[k].sort[s]);
    free(c.cells[k].tasks);
  }
  free(c.cells);
  free(c.parts);
  return ghost_failed ? 3 : 0;
}

/* ------------------------------------------------------------------------ */
/* Time integration either side of the path              */
/* ------------------------------------------------------------------------ */
#define NUM_TIME_BINS 56                 /* src/timeline.h:36 */
#define MAX_NR_TIMESTEPS (1LL << (NUM_TIME_BINS + 1))
#define MAX_DELTA_BIN 2                  /* time_bin_neighbour_max_delta_bin, timeline.h:51 */

/* src/timeline.h:59-63, 78-83, 107-114, 126-138 */
static inline long long int_timestep(int bin) { return bin <= 0 ? 0 : 1LL << (bin + 1); }
static inline int time_bin_of(long long dti) { return 62 - __builtin_clzll((unsigned long long)dti); }
static inline long long int_time_begin(long long ti_current, int bin) {
  const long long dti = int_timestep(bin);
  return dti == 0 ? 0 : dti * ((ti_current - 1) / dti);
}
static inline long long int_time_end(long long ti_current, int bin) {
  const long long dti = int_timestep(bin);
  if (dti == 0) return 0;
  const long long mod = ti_current % dti;
  return mod == 0 ? ti_current : ti_current - mod + dti;
}

/* src/approx_math.h:35-37 */
static inline float approx_expf(float x) { return 1.f + x * (1.f + x * (0.5f + x * (1.f / 6.f + 1.f / 24.f * x))); }

static void ieos(const oracle_iparams *ip, const oracle_eos *eos, float rho, float u, int mat_id, float *P, float *cs) {
  oracle_params prm;
  memset(&prm, 0, sizeof(prm));
  prm.scheme = ip->scheme;
  octx c;
  memset(&c, 0, sizeof(c));
  c.prm = &prm;
  c.eos = eos;
  opart p;
  memset(&p, 0, sizeof(p));
  p.rho = rho;
  p.u = u;
  p.mat_id = mat_id;
  eos_eval(&c, &p, P, cs);
}

/* drift_part (src/drift.h:186-214) with hydro_predict_extra (Minimal/hydro.h:815-866,
 * Planetary/hydro.h:823-870), then the tail of the particle loop of cell_drift_part
 * (src/cell_drift.c:303-318, 359-360). */
int oracle_drift(const oracle_iparams *ip, const oracle_eos *eos, oracle_istate *s, long long ti_old,
                 float dx_max[2]) {
  const double dt = (double)(ip->ti_current - ti_old) * ip->time_base; /* drift, kick_hydro and therm alike */
  const int planetary = ip->scheme == ORACLE_SCHEME_PLANETARY;
  float dx2_max = 0.f, dx2_max_sort = 0.f;
  for (int k = 0; k < s->n; k++) {
    for (int d = 0; d < 3; d++) {
      s->x[3 * k + d] += s->v_full[3 * k + d] * dt;
      s->v[3 * k + d] += s->a[3 * k + d] * dt;
    }
    const float dt_drift = dt, dt_therm = dt;
    float u = s->u[k];
    u += s->u_dt[k] * dt_therm;
    const float min_u = ip->minimal_internal_energy / 1.0; /* cosmo->a_factor_internal_energy = 1 */
    if (planetary) u = u > min_u ? u : min_u;
    const float h_inv = 1.f / s->h[k];
    const float w1 = s->h_dt[k] * h_inv * dt_drift;
    if (fabsf(w1) < 0.2f)
      s->h[k] *= approx_expf(w1);
    else
      s->h[k] *= expf(w1);
    const float w2 = -3.f * w1; /* hydro_dimension */
    if (fabsf(w2) < 0.2f)
      s->rho[k] *= approx_expf(w2);
    else
      s->rho[k] *= expf(w2);
    if (!planetary) {
      /* entropy floor "none": floor_A = 0, floor_u = 0 * rho^(gamma-1) / (gamma-1) = 0 */
      const float floor_u = 0.f;
      u = u > floor_u ? u : floor_u;
      u = u > min_u ? u : min_u;
    }
    s->u[k] = u;
    float P, cs;
    ieos(ip, eos, s->rho[k], u, s->mat_id ? s->mat_id[k] : 0, &P, &cs);
    s->pressure[k] = P;
    s->soundspeed[k] = cs;
    const float two_cs = 2.f * cs;
    s->v_sig[k] = s->v_sig[k] > two_cs ? s->v_sig[k] : two_cs;
    for (int d = 0; d < 3; d++) {
      const float dx = s->v_full[3 * k + d] * dt;
      s->x_diff[3 * k + d] -= dx;
      s->x_diff_sort[3 * k + d] -= dx;
    }
    /* cell_drift.c:303-304 */
    s->h[k] = s->h[k] < ip->h_max ? s->h[k] : ip->h_max;
    s->h[k] = s->h[k] > ip->h_min ? s->h[k] : ip->h_min;
    const float *xd = &s->x_diff[3 * k], *xs = &s->x_diff_sort[3 * k];
    const float dx2 = xd[0] * xd[0] + xd[1] * xd[1] + xd[2] * xd[2];
    dx2_max = dx2_max > dx2 ? dx2_max : dx2;
    const float dx2s = xs[0] * xs[0] + xs[1] * xs[1] + xs[2] * xs[2];
    dx2_max_sort = dx2_max_sort > dx2s ? dx2_max_sort : dx2s;
  }
  dx_max[0] = sqrtf(dx2_max);
  dx_max[1] = sqrtf(dx2_max_sort);
  return 0;
}

/* kick_part (src/kick.h:244-278) with hydro_kick_extra (Minimal/hydro.h:901-930, Planetary/hydro.h:904-925),
 * driven as runner_do_kick1 / runner_do_kick2 do (src/runner_time_integration.c:141-180, 412-457);
 * kick2 ends with hydro_reset_predicted_values (Minimal/hydro.h:771-796, Planetary/hydro.h:778-803). */
/* one kick_part over dt (all of dt_kick_hydro, dt_kick_therm, dt_kick_corr without cosmology) */
static void kick_one(const oracle_iparams *ip, oracle_istate *s, int k, double dt) {
  const int planetary = ip->scheme == ORACLE_SCHEME_PLANETARY;
  for (int d = 0; d < 3; d++) s->v_full[3 * k + d] += s->a[3 * k + d] * dt;
  const float dt_therm = dt;
  const float delta_u = s->u_dt[k] * dt_therm;
  const float ua = s->u_full[k] + delta_u, ub = 0.5f * s->u_full[k];
  s->u_full[k] = ua > ub ? ua : ub;
  const float min_u = ip->minimal_internal_energy / 1.0;
  const float floor_u = 0.f;
  const float energy_min = planetary ? min_u : (min_u > floor_u ? min_u : floor_u);
  if (s->u_full[k] < energy_min) {
    s->u_full[k] = energy_min;
    s->u_dt[k] = 0.f;
  }
}

/* runner_do_limiter on a leaf (src/runner_time_integration.c:1385-1452) with timestep_limit_part
 * (src/timestep_limiter.h:67-217); kick_get_*_kick_dt without cosmology = (ti_end - ti_beg) * time_base
 * (src/kick.h:43-60). */
int oracle_limiter(const oracle_iparams *ip, const oracle_eos *eos, oracle_istate *s, signed char *wakeup,
                   long long *ti_end_min_out, long long *limited_out) {
  (void)eos;
  const signed char not_awake = -NUM_TIME_BINS;
  long long ti_end_min = MAX_NR_TIMESTEPS, limited = 0;
  for (int k = 0; k < s->n; k++) {
    if (wakeup[k] == not_awake) continue;
    limited++;
    long long ti_end_new;
    if (s->time_bin[k] <= ip->max_active_bin) {
      /* active: only the length of the step that starts now changes */
      s->time_bin[k] = (signed char)(-wakeup[k] + 2);
      ti_end_new = ip->ti_current + int_timestep(s->time_bin[k]);
    } else {
      const int old_bin = s->time_bin[k];
      const int new_bin = (signed char)(-wakeup[k] + 2);
      const long long ti_beg_old = int_time_begin(ip->ti_current, old_bin);
      const long long ti_end_old = int_time_end(ip->ti_current, old_bin);
      const long long dti_old = ti_end_old - ti_beg_old;
      const long long dti_new = int_timestep(new_bin);
      int j = 0;
      while (ti_beg_old + j * dti_new <= ip->ti_current) j++;
      const long long ti_beg_new = ti_beg_old + (j - 1) * dti_new;
      const long long ti_end_half_old = ti_beg_old + dti_old / 2;
      const long long ti_end_half_new = ti_beg_new + dti_new / 2;
      /* undo the kick1 of the old step, kick to the start of the new one */
      kick_one(ip, s, k, -((double)(ti_end_half_old - ti_beg_old) * ip->time_base));
      kick_one(ip, s, k, (double)(ti_beg_new - ti_beg_old) * ip->time_base);
      s->time_bin[k] = (signed char)new_bin;
      if (new_bin > ip->max_active_bin) {
        /* the new step started earlier: its kick1 is due here, the kick task will not see the particle */
        kick_one(ip, s, k, (double)(ti_end_half_new - ti_beg_new) * ip->time_base);
        ti_end_new = ti_beg_new + dti_new;
      } else {
        ti_end_new = ip->ti_current + int_timestep(new_bin);
      }
    }
    wakeup[k] = not_awake;
    ti_end_min = ti_end_new < ti_end_min ? ti_end_new : ti_end_min;
  }
  *ti_end_min_out = ti_end_min;
  *limited_out = limited;
  return 0;
}

int oracle_kick(const oracle_iparams *ip, const oracle_eos *eos, oracle_istate *s, int which) {
  const int planetary = ip->scheme == ORACLE_SCHEME_PLANETARY;
  for (int k = 0; k < s->n; k++) {
    const int bin = s->time_bin[k];
    if (bin > ip->max_active_bin) continue; /* part_is_starting / part_is_active */
    const long long ti_step = int_timestep(bin);
    long long ti_begin;
    if (which == 1)
      ti_begin = int_time_begin(ip->ti_current + 1, bin);
    else
      ti_begin = int_time_begin(ip->ti_current, bin) + ti_step / 2;
    const long long ti_end = ti_begin + ti_step / 2;
    const double dt = (double)(ti_end - ti_begin) * ip->time_base;
    for (int d = 0; d < 3; d++) s->v_full[3 * k + d] += s->a[3 * k + d] * dt;
    const float dt_therm = dt;
    const float delta_u = s->u_dt[k] * dt_therm;
    const float ua = s->u_full[k] + delta_u, ub = 0.5f * s->u_full[k];
    s->u_full[k] = ua > ub ? ua : ub;
    const float min_u = ip->minimal_internal_energy / 1.0;
    const float floor_u = 0.f;
    const float energy_min = planetary ? min_u : (min_u > floor_u ? min_u : floor_u);
    if (s->u_full[k] < energy_min) {
      s->u_full[k] = energy_min;
      s->u_dt[k] = 0.f;
    }
    if (which == 2) {
      for (int d = 0; d < 3; d++) s->v[3 * k + d] = s->v_full[3 * k + d];
      s->u[k] = s->u_full[k];
      float P, cs;
      ieos(ip, eos, s->rho[k], s->u[k], s->mat_id ? s->mat_id[k] : 0, &P, &cs);
      s->pressure[k] = P;
      s->soundspeed[k] = cs;
      const float two_cs = 2.f * cs;
      s->v_sig[k] = s->v_sig[k] > two_cs ? s->v_sig[k] : two_cs;
    }
  }
  return 0;
}

/* get_part_timestep + make_integer_timestep (src/timestep.h:45-82, 141-222) for the active particles,
 * the sync point of the others, as runner_do_timestep does (src/runner_time_integration.c:696-745, 790-800). */
int oracle_timestep(const oracle_iparams *ip, oracle_istate *s, long long *ti_end_min_out, long long *updated_out) {
  const double time_base_inv = 1.0 / ip->time_base;
  long long ti_end_min = MAX_NR_TIMESTEPS, updated = 0;
  int status = 0;
  for (int k = 0; k < s->n; k++) {
    const int old_bin = s->time_bin[k];
    if (old_bin > ip->max_active_bin) {
      const long long ti_end = int_time_end(ip->ti_current, old_bin);
      ti_end_min = ti_end < ti_end_min ? ti_end : ti_end_min;
      continue;
    }
    /* hydro_compute_timestep (Minimal/hydro.h:440-452): cosmo->a and a_factor_sound_speed are doubles (= 1) */
    const float dt_cfl = 2.f * K_GAMMA * ip->cfl_condition * 1.0 * s->h[k] / (1.0 * s->v_sig[k]);
    float new_dt = dt_cfl;
    const float dt_h_change =
        (s->h_dt[k] != 0.0f) ? fabsf(ip->log_max_h_change * s->h[k] / s->h_dt[k]) : FLT_MAX;
    new_dt = new_dt < dt_h_change ? new_dt : dt_h_change;
    new_dt = new_dt < FLT_MAX ? new_dt : FLT_MAX; /* dt_max_RMS_displacement */
    new_dt *= 1.0;                                /* cosmology->time_step_factor */
    new_dt = (double)new_dt < ip->dt_max ? (double)new_dt : ip->dt_max;
    if (new_dt < ip->dt_min) status = 2;
    long long new_dti = (long long)(new_dt * time_base_inv);
    int new_bin = (signed char)time_bin_of(new_dti);
    const int ngb = (s->min_ngb_time_bin ? s->min_ngb_time_bin[k] : NUM_TIME_BINS + 1) + MAX_DELTA_BIN;
    new_bin = (signed char)(new_bin < ngb ? new_bin : ngb);
    new_dti = int_timestep(new_bin);
    const long long current_dti = int_timestep(old_bin);
    const long long ti_end = int_time_end(ip->ti_current, old_bin);
    if (old_bin > 0) new_dti = new_dti < 2 * current_dti ? new_dti : 2 * current_dti;
    long long dti_timeline = MAX_NR_TIMESTEPS;
    while (new_dti < dti_timeline) dti_timeline /= 2LL;
    new_dti = dti_timeline;
    if (new_dti > current_dti) {
      if ((MAX_NR_TIMESTEPS - ti_end) % new_dti > 0) new_dti = current_dti;
    }
    s->time_bin[k] = (signed char)time_bin_of(new_dti);
    updated++;
    const long long t = ip->ti_current + new_dti;
    ti_end_min = t < ti_end_min ? t : ti_end_min;
  }
  *ti_end_min_out = ti_end_min;
  *updated_out = updated;
  return status;
}
