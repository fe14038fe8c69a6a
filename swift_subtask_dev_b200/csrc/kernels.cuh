// sm_100a kernels of the SPH neighbour-loop path.  Arithmetic that decides
// WHICH pairs interact (frame positions, r2, h^2 gamma^2, the sorted-axis cuts,
// the sort keys) is written with explicit round-to-nearest intrinsics
// (__fmul_rn/__fadd_rn/__dmul_rn/__dadd_rn: never contracted into FMAs) and in
// the operand order of the reference, so the neighbour sets are bit-exact:
//   pair      src/runner_doiact_functions_hydro.h:1122-1216, 1222-1316 (density)
//                                                  :1528-1737, 1741-1952 (force)
//   self      :2095-2256 (density), :2337-2471 (force)
//   subset    :745-770, 794-838 (pair), :974-1003 (self)
//   sort key  src/runner_sort.c:377-387
#pragma once
#include <cfloat>
#include <type_traits>

#include "device_types.cuh"

namespace swiftcuda {

__constant__ Consts kc;

constexpr int kWarps = 4;        // warps per CTA of the neighbour-loop kernels (each warp works alone)
constexpr int kBlock = 32 * kWarps;
// resident CTAs per SM the neighbour-loop kernels are compiled for (register cap = 65536 / (kBlock * n))
#ifndef SWIFT_MINB_DENSITY
#define SWIFT_MINB_DENSITY 4
#endif
#ifndef SWIFT_MINB_FORCE
#define SWIFT_MINB_FORCE 4
#endif
constexpr int kNumTimeBins = 56;
constexpr unsigned kFull = 0xffffffffu;
constexpr int kDpFlag = 0x40000000;  // queue entry: dx must be recomputed from the f64 positions
constexpr int kInWin = 0x20000000;   // queue entry (list-recording redo sweep): inside the sorted-axis cut of the CURRENT h
constexpr int kIdxMask = kInWin - 1;
// Neighbour lists of the ghost iteration: the redo sweep that walks the cells also records, per lane,
// every candidate within nl_grow x the search radius as (dx, dy, dz, j) exactly as the reference forms
// dx; the following sweeps re-test that list against the new h (k_density_list) instead of walking.
constexpr int kLD = 96;              // list entries per lane


// ---------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float sq_gamma2(float h) { return __fmul_rn(__fmul_rn(h, h), kc.gamma2); }

__device__ __forceinline__ float dist2(float dx, float dy, float dz) {
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// the f64 position of particle j: one 16-byte and one 8-byte load from the same sector
__device__ __forceinline__ void load_pos(const PosD *__restrict__ pos, int64_t j, double &x, double &y, double &z) {
  const double2 xy = *reinterpret_cast<const double2 *>(&pos[j].x);
  x = xy.x;
  y = xy.y;
  z = pos[j].z;
}

// (float)(a - b) with the subtraction in f64
__device__ __forceinline__ float dsub_f(double a, double b) { return __double2float_rn(__dsub_rn(a, b)); }

// sort key: f64 dot product of the absolute position with the axis, rounded to f32
__device__ __forceinline__ float sort_key(double x, double y, double z, int sid) {
  const double s = __dadd_rn(__dadd_rn(__dmul_rn(x, kc.axis[sid][0]), __dmul_rn(y, kc.axis[sid][1])),
                             __dmul_rn(z, kc.axis[sid][2]));
  return __double2float_rn(s);
}

// kernel_deval, src/kernel_hydro.h:257-284 (cubic spline: 3 branches x 4 coefficients)
__device__ __forceinline__ void kernel_deval(float u, float &W, float &dW) {
  const float x = u * kc.gamma_inv;
  const int t = (int)(x * 2.f);
  float c0, c1, c2, c3;
  if (t <= 0) {
    c0 = 3.f; c1 = -3.f; c2 = 0.f; c3 = 0.5f;
  } else if (t == 1) {
    c0 = -1.f; c1 = 3.f; c2 = -3.f; c3 = 1.f;
  } else {
    c0 = 0.f; c1 = 0.f; c2 = 0.f; c3 = 0.f;
  }
  float w = c0 * x + c1;
  float dw = c0;
  dw = dw * x + w;
  w = x * w + c2;
  dw = dw * x + w;
  w = x * w + c3;
  w = fmaxf(w, 0.f);
  dw = fminf(dw, 0.f);
  W = w * kc.kconst * kc.ginv_dim;
  dW = dw * kc.kconst * kc.ginv_dim_p1;
}

struct DensAcc {
  float rho, rho_dh, wcount, wcount_dh, div_v, rx, ry, rz;
  __device__ __forceinline__ void zero() { rho = rho_dh = wcount = wcount_dh = div_v = rx = ry = rz = 0.f; }
};

// runner_iact_nonsym_density, src/hydro/Minimal/hydro_iact.h:134-184 (== Planetary :136-186)
__device__ __forceinline__ void iact_density(DensAcc &a, float r2, float dx, float dy, float dz, float hi_inv,
                                             float vix, float viy, float viz, float4 vmj) {
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float ui = r * hi_inv;
  float wi, wi_dx;
  kernel_deval(ui, wi, wi_dx);
  const float mj = vmj.w;
  a.rho += mj * wi;
  a.rho_dh -= mj * (3.f * wi + ui * wi_dx);
  a.wcount += wi;
  a.wcount_dh -= (3.f * wi + ui * wi_dx);
  const float faci = mj * wi_dx * r_inv;
  const float dvx = vix - vmj.x, dvy = viy - vmj.y, dvz = viz - vmj.z;
  const float dvdr = dvx * dx + dvy * dy + dvz * dz;
  a.div_v -= faci * dvdr;
  a.rx += faci * (dvy * dz - dvz * dy);
  a.ry += faci * (dvz * dx - dvx * dz);
  a.rz += faci * (dvx * dy - dvy * dx);
}

struct ForceAcc {
  float ax, ay, az, u_dt, h_dt, v_sig;
  int min_bin;
};

struct ForceI {  // per-i quantities hoisted out of the neighbour loop
  float h, hi_inv, hid_inv, m, rho, P, f, c, bal, vx, vy, vz;
  float Pr2;  // P / rho^2
  float m_inv;  // FAST form: 1 / m
  float alpha, adiff, u;  // SPHENIX: p->viscosity.alpha, p->diffusion.alpha, p->u
};
struct ForceJ {
  float h, m, rho, P, f, c, bal, vx, vy, vz;
  float Pr2;  // P / rho^2, formed once per particle by the ghost
  int bin;
  float alpha, adiff, u;  // SPHENIX
};

// runner_iact_nonsym_force: Minimal/hydro_iact.h:370-478, Planetary/hydro_iact.h:369-476
// + runner_iact_nonsym_timebin, src/timestep_limiter_iact.h:41-55
template <int SCHEME>
__device__ __forceinline__ void iact_force(ForceAcc &a, float r2, float dx, float dy, float dz, const ForceI &pi,
                                           const ForceJ &pj) {
  const float fac_mu = 1.f;  // pow_three_gamma_minus_five_over_two(a) for gamma = 5/3
  const float a2_Hubble = kc.a * kc.a * kc.H;
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float xi = r * pi.hi_inv;
  float wi, wi_dx;
  kernel_deval(xi, wi, wi_dx);
  const float wi_dr = pi.hid_inv * wi_dx;
  const float hj_inv = 1.0f / pj.h;
  const float hj2 = hj_inv * hj_inv;
  const float hjd_inv = hj2 * hj2;
  const float xj = r * hj_inv;
  float wj, wj_dx;
  kernel_deval(xj, wj, wj_dx);
  const float wj_dr = hjd_inv * wj_dx;
  const float f_ij = 1.f - pi.f / pj.m;
  const float f_ji = 1.f - pj.f / pi.m;
  const float P_over_rho2_i = pi.Pr2 * f_ij;  // pressure / (rho * rho) * f, :281-282
  const float P_over_rho2_j = pj.Pr2 * f_ji;
  float dvdr = (pi.vx - pj.vx) * dx + (pi.vy - pj.vy) * dy + (pi.vz - pj.vz) * dz;
  float dvdr_visc, visc, v_sig;
  const float rho_ij = 0.5f * (pi.rho + pj.rho);
  if (SCHEME == SWIFT_CUDA_SCHEME_PLANETARY) {
    dvdr = dvdr + a2_Hubble * r2;
    dvdr_visc = dvdr;
    const float mu_ij = fac_mu * r_inv * fminf(dvdr, 0.f);
    v_sig = pi.c + pj.c - kc.beta * mu_ij;
    visc = -0.25f * v_sig * mu_ij * (pi.bal + pj.bal) / rho_ij;
  } else {
    dvdr_visc = dvdr + a2_Hubble * r2;
    const float mu_ij = fac_mu * r_inv * fminf(dvdr_visc, 0.f);
    v_sig = pi.c + pj.c - kc.beta * mu_ij;
    visc = -0.25f * v_sig * (pi.bal + pj.bal) * mu_ij / rho_ij;
  }
  const float visc_acc_term = 0.5f * visc * (wi_dr * f_ij + wj_dr * f_ji) * r_inv;
  const float sph_acc_term = (P_over_rho2_i * wi_dr + P_over_rho2_j * wj_dr) * r_inv;
  const float acc = sph_acc_term + visc_acc_term;
  a.ax -= pj.m * acc * dx;
  a.ay -= pj.m * acc * dy;
  a.az -= pj.m * acc * dz;
  const float sph_du_term_i = P_over_rho2_i * dvdr * r_inv * wi_dr;
  const float visc_du_term = 0.5f * visc_acc_term * dvdr_visc;
  a.u_dt += (sph_du_term_i + visc_du_term) * pj.m;
  if (SCHEME == SWIFT_CUDA_SCHEME_PLANETARY)
    a.h_dt -= pj.m * dvdr * r_inv / pj.rho * wi_dr;
  else
    a.h_dt -= pj.m * dvdr * r_inv / pj.rho * wi_dr * f_ij;
  a.v_sig = fmaxf(a.v_sig, v_sig);
  if (pj.bin > 0) a.min_bin = min(a.min_bin, pj.bin);
}

// FAST form of runner_iact_nonsym_force (swift_cuda_params.fields_mode = SWIFT_CUDA_FIELDS_FAST): the same
// expressions with FMA contraction and the reciprocals / square root by MUFU (rcp.approx, rsqrt.approx:
// ~1 ulp) instead of IEEE division sequences.  Which pairs interact is decided elsewhere, exactly; the
// accelerations, du/dt, dh/dt and v_sig agree with the reference to ~1e-6 of the summed terms.
__device__ __forceinline__ float fast_rcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void kernel_deval_fast(float u, float &W, float &dW) {
  const float x = u * kc.gamma_inv;
  const int t = (int)(x * 2.f);
  float c0, c1, c2, c3;
  if (t <= 0) {
    c0 = 3.f; c1 = -3.f; c2 = 0.f; c3 = 0.5f;
  } else if (t == 1) {
    c0 = -1.f; c1 = 3.f; c2 = -3.f; c3 = 1.f;
  } else {
    c0 = 0.f; c1 = 0.f; c2 = 0.f; c3 = 0.f;
  }
  float w = __fmaf_rn(c0, x, c1);
  float dw = __fmaf_rn(c0, x, w);
  w = __fmaf_rn(x, w, c2);
  dw = __fmaf_rn(dw, x, w);
  w = __fmaf_rn(x, w, c3);
  w = fmaxf(w, 0.f);
  dw = fminf(dw, 0.f);
  W = w * (kc.kconst * kc.ginv_dim);
  dW = dw * (kc.kconst * kc.ginv_dim_p1);
}
template <int SCHEME>
__device__ __forceinline__ void iact_force_fast(ForceAcc &a, float r2, float dx, float dy, float dz, const ForceI &pi,
                                                const ForceJ &pj) {
  const float a2_Hubble = kc.a * kc.a * kc.H;
  const float r_inv = r2 > 0.f ? rsqrtf(r2) : 0.f;
  const float r = r2 * r_inv;
  float wi, wi_dx, wj, wj_dx;
  kernel_deval_fast(r * pi.hi_inv, wi, wi_dx);
  const float wi_dr = pi.hid_inv * wi_dx;
  const float hj_inv = fast_rcp(pj.h);
  const float hj2 = hj_inv * hj_inv;
  kernel_deval_fast(r * hj_inv, wj, wj_dx);
  const float wj_dr = hj2 * hj2 * wj_dx;
  const float f_ij = __fmaf_rn(-pi.f, fast_rcp(pj.m), 1.f);
  const float f_ji = __fmaf_rn(-pj.f, pi.m_inv, 1.f);
  const float P_over_rho2_i = pi.Pr2 * f_ij;
  const float P_over_rho2_j = pj.Pr2 * f_ji;
  float dvdr = __fmaf_rn(pi.vz - pj.vz, dz, __fmaf_rn(pi.vy - pj.vy, dy, (pi.vx - pj.vx) * dx));
  const float dvdr_H = __fmaf_rn(a2_Hubble, r2, dvdr);
  if (SCHEME == SWIFT_CUDA_SCHEME_PLANETARY) dvdr = dvdr_H;
  const float mu_ij = r_inv * fminf(dvdr_H, 0.f);
  const float v_sig = __fmaf_rn(-kc.beta, mu_ij, pi.c + pj.c);
  const float rho_ij_inv = fast_rcp(0.5f * (pi.rho + pj.rho));
  const float visc = -0.25f * v_sig * mu_ij * (pi.bal + pj.bal) * rho_ij_inv;
  const float visc_acc_term = 0.5f * visc * __fmaf_rn(wi_dr, f_ij, wj_dr * f_ji) * r_inv;
  const float sph_acc_term = __fmaf_rn(P_over_rho2_i, wi_dr, P_over_rho2_j * wj_dr) * r_inv;
  const float macc = pj.m * (sph_acc_term + visc_acc_term);
  a.ax = __fmaf_rn(-macc, dx, a.ax);
  a.ay = __fmaf_rn(-macc, dy, a.ay);
  a.az = __fmaf_rn(-macc, dz, a.az);
  const float sph_du_term_i = P_over_rho2_i * dvdr * r_inv * wi_dr;
  a.u_dt = __fmaf_rn(__fmaf_rn(0.5f * visc_acc_term, dvdr_H, sph_du_term_i), pj.m, a.u_dt);
  float hterm = pj.m * dvdr * r_inv * fast_rcp(pj.rho) * wi_dr;
  if (SCHEME != SWIFT_CUDA_SCHEME_PLANETARY) hterm *= f_ij;
  a.h_dt -= hterm;
  a.v_sig = fmaxf(a.v_sig, v_sig);
  if (pj.bin > 0) a.min_bin = min(a.min_bin, pj.bin);
}

// ---- SPHENIX: the gradient loop's interaction and the scheme's force interaction ----
struct GradAcc {
  float v_sig, lap, amax;  // p->viscosity.v_sig, p->diffusion.laplace_u, p->force.alpha_visc_max_ngb
  float div_v;             // carried through (the record sph_grad holds it next to the two sums)
};

// runner_iact_nonsym_gradient, src/hydro/SPHENIX/hydro_iact.h:272-329.  gi / gj: the gradient-loop
// record of a particle {u, rho, soundspeed, viscosity.alpha}.  max() is the reference's macro (a > b ? a : b).
__device__ __forceinline__ void iact_gradient(GradAcc &a, float r2, float dx, float dy, float dz, float hi,
                                              const float4 vmi, const float4 gi, const float4 vmj, const float4 gj) {
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float fac_mu = 1.f;
  const float a2_Hubble = kc.a * kc.a * kc.H;
  const float dvdr = (vmi.x - vmj.x) * dx + (vmi.y - vmj.y) * dy + (vmi.z - vmj.z) * dz;
  const float dvdr_Hubble = dvdr + a2_Hubble * r2;
  const float omega_ij = dvdr_Hubble < 0.f ? dvdr_Hubble : 0.f;
  const float mu_ij = fac_mu * r_inv * omega_ij;
  const float new_v_sig = gi.z + gj.z - kc.beta * mu_ij;  // hydro_signal_velocity, SPHENIX/hydro.h:490-498
  a.v_sig = a.v_sig > new_v_sig ? a.v_sig : new_v_sig;
  float wi, wi_dx;
  const float ui = r / hi;
  kernel_deval(ui, wi, wi_dx);
  const float delta_u_factor = (gi.x - gj.x) * r_inv;
  a.lap += vmj.w * delta_u_factor * wi_dx / gj.y;
  a.amax = a.amax > gj.w ? a.amax : gj.w;
}

// runner_iact_nonsym_force of SPHENIX, src/hydro/SPHENIX/hydro_iact.h:488-609 (+ runner_iact_nonsym_timebin).
// v_sig is not touched by this scheme's force loop.
__device__ __forceinline__ void iact_force_sphenix(ForceAcc &a, float r2, float dx, float dy, float dz,
                                                   const ForceI &pi, const ForceJ &pj) {
  const float fac_mu = 1.f;
  const float a2_Hubble = kc.a * kc.a * kc.H;
  const float r = sqrtf(r2);
  const float r_inv = r ? 1.0f / r : 0.0f;
  const float xi = r * pi.hi_inv;
  float wi, wi_dx;
  kernel_deval(xi, wi, wi_dx);
  const float wi_dr = pi.hid_inv * wi_dx;
  const float hj_inv = 1.0f / pj.h;
  const float hj2 = hj_inv * hj_inv;
  const float hjd_inv = hj2 * hj2;
  const float xj = r * hj_inv;
  float wj, wj_dx;
  kernel_deval(xj, wj, wj_dx);
  const float wj_dr = hjd_inv * wj_dx;
  const float dvdr = (pi.vx - pj.vx) * dx + (pi.vy - pj.vy) * dy + (pi.vz - pj.vz) * dz;
  const float dvdr_Hubble = dvdr + a2_Hubble * r2;
  const float omega_ij = dvdr_Hubble < 0.f ? dvdr_Hubble : 0.f;
  const float mu_ij = fac_mu * r_inv * omega_ij;
  const float v_sig = pi.c + pj.c - kc.beta * mu_ij;
  const float f_ij = 1.f - pi.f / pj.m;
  const float f_ji = 1.f - pj.f / pi.m;
  const float rho_ij = pi.rho + pj.rho;
  const float alpha = pi.alpha + pj.alpha;
  const float visc = -0.25f * alpha * v_sig * mu_ij * (pi.bal + pj.bal) / rho_ij;
  const float visc_acc_term = 0.5f * visc * (wi_dr * f_ij + wj_dr * f_ji) * r_inv;
  const float P_over_rho2_i = pi.Pr2 * f_ij;  // pressure / (rho * rho) * f
  const float P_over_rho2_j = pj.Pr2 * f_ji;
  const float sph_acc_term = (P_over_rho2_i * wi_dr + P_over_rho2_j * wj_dr) * r_inv;
  const float acc = sph_acc_term + visc_acc_term;
  a.ax -= pj.m * acc * dx;
  a.ay -= pj.m * acc * dy;
  a.az -= pj.m * acc * dz;
  const float sph_du_term_i = P_over_rho2_i * dvdr * r_inv * wi_dr;
  const float visc_du_term = 0.5f * visc_acc_term * dvdr_Hubble;
  const float alpha_diff = (pi.P * pi.adiff + pj.P * pj.adiff) / (pi.P + pj.P);
  const float v_diff =
      alpha_diff * 0.5f * (sqrtf(2.f * fabsf(pi.P - pj.P) / rho_ij) + fabsf(fac_mu * r_inv * dvdr_Hubble));
  const float diff_du_term = v_diff * (pi.u - pj.u) * (f_ij * wi_dr / pi.rho + f_ji * wj_dr / pj.rho);
  const float du_dt_i = sph_du_term_i + visc_du_term + diff_du_term;
  a.u_dt += du_dt_i * pj.m;
  a.h_dt -= pj.m * dvdr * r_inv / pj.rho * wi_dr;
  if (pj.bin > 0) a.min_bin = min(a.min_bin, pj.bin);
}

__device__ __forceinline__ void record_pair(const Dev &d, int loop, long long i, long long j) {
  const unsigned long long k = atomicAdd(&d.rec_count[loop], 1ULL);
  if ((long long)k < d.rec_capacity) {
    d.rec_pairs[loop][2 * k] = i;
    d.rec_pairs[loop][2 * k + 1] = j;
  }
}

// FAST form of the scheme's force interaction (fields_mode FAST): the same expressions with FMA contraction and
// MUFU reciprocals / rsqrt, as iact_force_fast; which pairs interact is decided elsewhere, exactly.
__device__ __forceinline__ void iact_force_sphenix_fast(ForceAcc &a, float r2, float dx, float dy, float dz,
                                                        const ForceI &pi, const ForceJ &pj) {
  const float a2_Hubble = kc.a * kc.a * kc.H;
  const float r_inv = r2 > 0.f ? rsqrtf(r2) : 0.f;
  const float r = r2 * r_inv;
  float wi, wi_dx, wj, wj_dx;
  kernel_deval_fast(r * pi.hi_inv, wi, wi_dx);
  const float wi_dr = pi.hid_inv * wi_dx;
  const float hj_inv = fast_rcp(pj.h);
  const float hj2 = hj_inv * hj_inv;
  kernel_deval_fast(r * hj_inv, wj, wj_dx);
  const float wj_dr = hj2 * hj2 * wj_dx;
  const float dvdr = __fmaf_rn(pi.vz - pj.vz, dz, __fmaf_rn(pi.vy - pj.vy, dy, (pi.vx - pj.vx) * dx));
  const float dvdr_H = __fmaf_rn(a2_Hubble, r2, dvdr);
  const float mu_ij = r_inv * fminf(dvdr_H, 0.f);
  const float v_sig = __fmaf_rn(-kc.beta, mu_ij, pi.c + pj.c);
  const float f_ij = __fmaf_rn(-pi.f, fast_rcp(pj.m), 1.f);
  const float f_ji = __fmaf_rn(-pj.f, pi.m_inv, 1.f);
  const float rho_ij_inv = fast_rcp(pi.rho + pj.rho);
  const float rhoi_inv = fast_rcp(pi.rho), rhoj_inv = fast_rcp(pj.rho);
  const float visc = -0.25f * (pi.alpha + pj.alpha) * v_sig * mu_ij * (pi.bal + pj.bal) * rho_ij_inv;
  const float visc_acc_term = 0.5f * visc * __fmaf_rn(wi_dr, f_ij, wj_dr * f_ji) * r_inv;
  const float P_over_rho2_i = pi.Pr2 * f_ij;
  const float P_over_rho2_j = pj.Pr2 * f_ji;
  const float sph_acc_term = __fmaf_rn(P_over_rho2_i, wi_dr, P_over_rho2_j * wj_dr) * r_inv;
  const float macc = pj.m * (sph_acc_term + visc_acc_term);
  a.ax = __fmaf_rn(-macc, dx, a.ax);
  a.ay = __fmaf_rn(-macc, dy, a.ay);
  a.az = __fmaf_rn(-macc, dz, a.az);
  const float sph_du_term_i = P_over_rho2_i * dvdr * r_inv * wi_dr;
  const float alpha_diff = __fmaf_rn(pi.P, pi.adiff, pj.P * pj.adiff) * fast_rcp(pi.P + pj.P);
  const float dP2 = 2.f * fabsf(pi.P - pj.P) * rho_ij_inv;
  const float v_diff = alpha_diff * 0.5f * ((dP2 > 0.f ? dP2 * rsqrtf(dP2) : 0.f) + fabsf(r_inv * dvdr_H));
  const float diff_du_term = v_diff * (pi.u - pj.u) * __fmaf_rn(f_ij * wi_dr, rhoi_inv, f_ji * wj_dr * rhoj_inv);
  const float du = __fmaf_rn(0.5f * visc_acc_term, dvdr_H, sph_du_term_i) + diff_du_term;
  a.u_dt = __fmaf_rn(du, pj.m, a.u_dt);
  a.h_dt -= pj.m * dvdr * r_inv * rhoj_inv * wi_dr;
  if (pj.bin > 0) a.min_bin = min(a.min_bin, pj.bin);
}

// the gradient loop's pairs (SPHENIX): slot 3 of the counters, its own buffer
__device__ __forceinline__ void record_pair_grad(const Dev &d, long long i, long long j) {
  const unsigned long long k = atomicAdd(&d.rec_count[3], 1ULL);
  if ((long long)k < d.rec_capacity) {
    d.rec_grad[2 * k] = i;
    d.rec_grad[2 * k + 1] = j;
  }
}

__device__ __forceinline__ unsigned long long block_sum(unsigned long long v) {
  __shared__ unsigned long long red[32];
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    v = (l < (blockDim.x >> 5)) ? red[l] : 0ULL;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  }
  return v;
}

// ---------------------------------------------------------------------------
// K2: 13-axis sorts (runner_do_hydro_sort, src/runner_sort.c:201-431)
// One CTA per (cell, axis) segment; bitonic sort of 64-bit keys
// (order-preserving key bits << 32 | local index) in shared memory, or in a
// global scratch area for cells above kSortSmem particles.
// ---------------------------------------------------------------------------
constexpr int kSortSmem = 1024;

__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// Which sorted lists does this step read?  The reference sorts a cell along an axis only when an
// ACTIVE pair task needs it (cell_activate_hydro_sorts, src/cell_unskip.c); with individual time steps
// most steps have few active particles.  One warp per leaf: if one of its particles is active, every
// (cell, axis) list its interactions walk is marked (both sides of a pair).
__global__ void k_mark_sorts(Dev d) {
  const int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (slot >= d.n_leaves) return;
  const CellDev c = d.cells[d.leaves[slot]];
  bool act = false;
  for (int k = lane; k < c.count; k += 32) act |= d.time_bin[c.first + k] <= kc.max_active_bin;
  if (!__any_sync(0xffffffffu, act)) return;
  for (int u = d.unit_begin[slot]; u < d.unit_begin[slot + 1]; u++) {
    const int q = d.lane_part[(int64_t)u * 32 + lane];
    const bool a = q >= 0 && d.time_bin[q] <= kc.max_active_bin;
    if (__any_sync(0xffffffffu, a) && lane == 0) d.unit_list[atomicAdd(d.unit_count, 1)] = u;
  }
  for (int pass = 0; pass < 2; pass++) {
    const int32_t *begin = pass == 0 ? d.item_begin : d.redo_begin;
    const Item *items = pass == 0 ? d.items : d.redo_items;
    for (int k = begin[slot] + lane; k < begin[slot + 1]; k += 32) {
      const Item it = items[k];
      if (it.kind == kSelf || it.kind == kSubSelf) continue;
      d.sort_need[(int64_t)it.partner * 13 + it.sid] = 1;
      d.sort_need[(int64_t)it.own * 13 + it.sid] = 1;
    }
  }
}

__device__ __forceinline__ void sort_segment(const Dev &d, int seg, const int32_t *seg_cell, int sid,
                                             const int64_t *seg_off, unsigned long long *scratch,
                                             const int64_t *scratch_off, unsigned long long *skeys) {
  const CellDev c = d.cells[seg_cell[seg]];
  const int n = c.count;
  int npad = 1;
  while (npad < n) npad <<= 1;
  if (npad <= 256) {
    // The common case (a leaf of <= 256 particles): rank sort.  Every thread counts the keys below
    // its own (four per 16-byte broadcast load) and writes its entry at that rank: ~3x fewer
    // instructions than a 36-stage bitonic network on 256 slots, one barrier.  Equal keys (rare)
    // are ordered by index, as sorting (key << 32 | index) does.
    uint32_t *sk = reinterpret_cast<uint32_t *>(skeys);
    const int t = threadIdx.x;
    uint32_t ku = 0xffffffffu;
    if (t < n) {
      double px, py, pz;
      load_pos(d.pos, c.first + t, px, py, pz);
      ku = float_to_ordered(sort_key(px, py, pz, sid));
    }
    sk[t] = ku;  // slots n .. 255: 0xffffffff, below nothing
    __syncthreads();
    if (t >= n) return;
    int lt = 0, eq = 0;
    const int n4 = (n + 3) & ~3;
    for (int k = 0; k < n4; k += 4) {
      const uint4 v = *reinterpret_cast<const uint4 *>(&sk[k]);
      lt += (v.x < ku) + (v.y < ku) + (v.z < ku) + (v.w < ku);
      eq += (v.x == ku) + (v.y == ku) + (v.z == ku) + (v.w == ku);
    }
    int rank = lt;
    if (eq > 1)
      for (int k = 0; k < t; k++) rank += (sk[k] == ku);
    SortEnt se;
    se.d = ordered_to_float(ku);
    se.i = t;
    d.sort[seg_off[seg] + rank] = se;
    return;
  }
  unsigned long long *keys = (npad <= kSortSmem) ? skeys : (scratch + scratch_off[seg]);
  for (int k = threadIdx.x; k < npad; k += blockDim.x) {
    unsigned long long key = ~0ULL;
    if (k < n) {
      const int64_t p = c.first + k;
      double px, py, pz;
      load_pos(d.pos, p, px, py, pz);
      const float dk = sort_key(px, py, pz, sid);
      key = ((unsigned long long)float_to_ordered(dk) << 32) | (unsigned)k;
    }
    keys[k] = key;
  }
  __syncthreads();
  for (int size = 2; size <= npad; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int k = threadIdx.x; k < (npad >> 1); k += blockDim.x) {
        const int lo = ((k / stride) * (stride << 1)) + (k % stride);
        const int hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const unsigned long long a = keys[lo], b = keys[hi];
        if ((a > b) == up) {
          keys[lo] = b;
          keys[hi] = a;
        }
      }
      __syncthreads();
    }
  const int64_t off = seg_off[seg];
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    const unsigned long long key = keys[k];
    SortEnt se;
    se.d = ordered_to_float((uint32_t)(key >> 32));
    se.i = (int32_t)(key & 0xffffffffu);
    d.sort[off + k] = se;
  }
}

// One CTA per (cell, axis) segment when every list is rebuilt; in time-integration mode (sort_need
// set) a small grid strides over the segments and sorts only the marked ones, so that a step with a
// handful of active particles does not pay for 13 launches' worth of empty CTAs per cell.
__global__ void __launch_bounds__(256) k_sort(Dev d, int n_seg, const int32_t *seg_cell, const int32_t *seg_sid,
                                              const int64_t *seg_off, unsigned long long *scratch,
                                              const int64_t *scratch_off) {
  __shared__ unsigned long long skeys[kSortSmem];
  if (!d.sort_need) {
    sort_segment(d, blockIdx.x, seg_cell, seg_sid[blockIdx.x], seg_off, scratch, scratch_off, skeys);
    return;
  }
  __shared__ int list[256];
  __shared__ int cnt;
  for (int base = blockIdx.x * 256; base < n_seg; base += gridDim.x * 256) {
    if (threadIdx.x == 0) cnt = 0;
    __syncthreads();
    const int mine = base + threadIdx.x;  // 256 flags checked at once, the marked ones sorted in turn
    if (mine < n_seg && d.sort_need[(int64_t)seg_cell[mine] * 13 + seg_sid[mine]]) list[atomicAdd(&cnt, 1)] = mine;
    __syncthreads();
    const int m = cnt;
    for (int k = 0; k < m; k++) {
      const int seg = list[k];
      sort_segment(d, seg, seg_cell, seg_sid[seg], seg_off, scratch, scratch_off, skeys);
      __syncthreads();  // skeys is reused by the next segment
    }
  }
}

// ---------------------------------------------------------------------------
// K1b: start of a step: zero sums (hydro_init_part), arm the ghost state
// ---------------------------------------------------------------------------
__global__ void k_init_step(Dev d, int64_t lo = 0, int64_t hi = INT64_MAX) {
  const int64_t p = lo + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n || p >= hi) return;
  const bool active = d.time_bin[p] <= kc.max_active_bin;
  d.gstate[p] = active ? 1 : 0;
  // time-integration mode (unit_list set): most steps have few active particles and nothing below is
  // read for the others, so their 45 bytes are not written
  if (!active && d.unit_list != nullptr) return;
  d.left[p] = 0.f;
  d.right[p] = kc.h_max;
  if (d.keep_sums) return;  // per-task shims: the records' running sums were unpacked into the caches
  d.dens0[p] = make_float4(0.f, 0.f, 0.f, 0.f);
  d.dens1[p] = make_float4(0.f, 0.f, 0.f, 0.f);
  d.min_bin[p] = kNumTimeBins + 1;
  if (active) {
    d.acc[p] = make_float4(0.f, 0.f, 0.f, 0.f);
    d.h_dt[p] = 0.f;
  }
}

// ---------------------------------------------------------------------------
// Neighbour-loop machinery shared by K3 (density), K5 (ghost redo) and K6 (force)
//
// Work decomposition.  The kernels are PERSISTENT: the grid is one resident
// wave of CTAs and every warp draws units from a device-side counter.  A unit
// is up to 32 particles of ONE leaf cell, one per lane (spatial neighbours:
// Morton order inside the leaf).  The warp walks that leaf's list of direct
// cell interactions in the order the reference's single-threaded runner
// executes them.  For every interaction:
//   1. skip it when the bounding box of the partner cell's particles cannot
//      reach the box of the unit's particles (each grown by its search radius);
//   2. stage the partner's particles 32 at a time in sorted-axis order, one per
//      lane, already converted to the pair's f32 frame, and keep only those
//      inside the unit's box and inside the union of the lanes' sort-key
//      windows; survivors are compacted, IN ORDER, into a warp-private
//      shared-memory ring;
//   3. every lane tests every surviving candidate against its own sorted-axis
//      cut and r2 < h^2 gamma^2 (FP32, one broadcast LDS.128 per candidate) and
//      appends its hits to a per-lane FIFO in global memory (L2 resident, kQD
//      entries deep: deep enough for ALL the hits of a particle in the
//      benchmark configurations).
// The interaction body (runner_iact_nonsym_*) runs when the walk is over (or
// when a FIFO is about to overflow), with every lane busy, consuming the hits
// in FIFO order into register accumulators: every particle sums its
// neighbours in exactly the order the reference does, so the results are
// bit-identical to the oracle.
// ---------------------------------------------------------------------------
constexpr int kQD = 128;    // deferred hits per lane
constexpr int kTile = 64;   // ring of staged candidates per warp (two 32-candidate tiles)
constexpr int kSide = 4;    // force: loop-2 hits of one pair kept aside per lane

struct WarpTile {
  float4 pos[kTile];  // frame-relative position of the candidate, sort key in w
  int idx[kTile];     // particle index of the candidate
  int sel[64];        // redo: compaction of the flagged particles of a leaf
};

__device__ __forceinline__ float next_down(float f) {
  // f finite; largest float < f
  if (f == 0.f) return -FLT_TRUE_MIN;
  int b = __float_as_int(f);
  return __int_as_float(f > 0.f ? b - 1 : b + 1);
}
__device__ __forceinline__ float next_up(float f) {
  if (f == 0.f) return FLT_TRUE_MIN;
  int b = __float_as_int(f);
  return __int_as_float(f > 0.f ? b + 1 : b - 1);
}
// largest float t with (double)t < y   ("sort.d < di" with a double di)
__device__ __forceinline__ float thr_lt(double y) {
  if (!(y > -(double)FLT_MAX)) return -INFINITY;
  if (y > (double)FLT_MAX) return FLT_MAX;
  float t = __double2float_rd(y);
  if ((double)t == y) t = next_down(t);
  return t;
}
// smallest float t with (double)t > y   ("sort.d > dj" with a double dj)
__device__ __forceinline__ float thr_gt(double y) {
  if (!(y < (double)FLT_MAX)) return INFINITY;
  if (y < -(double)FLT_MAX) return -FLT_MAX;
  float t = __double2float_ru(y);
  if ((double)t == y) t = next_up(t);
  return t;
}
// smallest float t with RN((double)t - rs) > y   ("sort.d - rshift > dj")
__device__ __forceinline__ float thr_gt_shift(double rs, double y) {
  if (rs == 0.0) return thr_gt(y);
  float t = __double2float_rn(__dadd_rn(y, rs));
  if (!(fabsf(t) < FLT_MAX)) return t > 0.f ? INFINITY : -FLT_MAX;
  for (int k = 0; k < 4; k++) {
    const float tm = next_down(t);
    if (__dsub_rn((double)tm, rs) > y) t = tm; else break;
  }
  for (int k = 0; k < 4; k++) {
    if (!(__dsub_rn((double)t, rs) > y)) t = next_up(t); else break;
  }
  return t;
}

// 16-byte store under a predicate: no branch, no reconvergence point in the candidate tests
__device__ __forceinline__ void st_pred(float4 *p, float x, float y, float z, float w, bool pred) {
  asm volatile("{ .reg .pred q; setp.ne.b32 q, %5, 0; @q st.global.cs.v4.f32 [%0], {%1, %2, %3, %4}; }" ::"l"(p), "f"(x),
               "f"(y), "f"(z), "f"(w), "r"((int)pred)
               : "memory");
}

// warp-wide min / max in one instruction (redux.sync f32: sm_100a)
__device__ __forceinline__ float warp_min(float v) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float warp_max(float v) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}

// periodic shift of an item along axis k, in box lengths {-1, 0, 1}
__device__ __forceinline__ int item_shift_i(int code, int k) {
  return (k == 0 ? code % 3 : (k == 1 ? (code / 3) % 3 : code / 9)) - 1;
}
__device__ __forceinline__ double item_shift(const Dev &d, int code, int k) {
  return (double)item_shift_i(code, k) * d.dim[k];
}
// rshift = sum_k shift[k] * runner_shift[sid][k], accumulated from 0 in k order as the reference does
__device__ __forceinline__ double item_rshift(const double sh[3], int sid) {
  double r = __dmul_rn(sh[0], kc.axis[sid][0]);
  r = __dadd_rn(r, __dmul_rn(sh[1], kc.axis[sid][1]));
  r = __dadd_rn(r, __dmul_rn(sh[2], kc.axis[sid][2]));
  return r;
}

// conservative FP32 cull threshold for pairs whose exact dx needs the f64 positions:
// every true hit (exact r2 < hg2) has an approximate r2 below this.
__device__ __forceinline__ float cull_thr(float hg2, float e) {
  return hg2 * 1.000002f + 3.5f * sqrtf(hg2) * e + 3.f * e * e;
}

// Box of a unit's particles (each grown by its search radius), given in the frame of the
// leaf (ulo/uhi, f32) -> the same box in the frame the candidates of an item are staged in.
// ol = (leaf origin) - (what is subtracted from a lane's position) - (lane frame origin).
// The margin covers every f32 rounding between the two frames (8 ulp of the magnitudes).
struct StageBox {
  float lo[3], hi[3];
};
__device__ __forceinline__ void stage_box(const float ulo[3], const float uhi[3], const double ol[3], float slack,
                                          StageBox &b) {
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float o = (float)ol[k];
    const float m = 2e-6f * (fabsf(o) + fabsf(ulo[k]) + fabsf(uhi[k])) + slack;
    b.lo[k] = ulo[k] + o - m;
    b.hi[k] = uhi[k] + o + m;
  }
}
// can a particle of cell cb (bounding box of its particles relative to its loc: bb[0..5]) lie in b?
// og = cb.loc - (frame origin of the staged candidates)
__device__ __forceinline__ bool box_reaches_cell(const StageBox &b, const float *__restrict__ bb, const double og[3]) {
  bool r = true;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float o = (float)og[k];
    const float lo = bb[k] + o, hi = bb[3 + k] + o;
    const float m = 2e-6f * (fabsf(o) + fabsf(lo) + fabsf(hi));
    r = r && (b.hi[k] >= lo - m) && (b.lo[k] <= hi + m);
  }
  return r;
}

// delta of the safe zone: a hit with r < H - delta passes every sorted-axis cut of the reference
// whatever the rounding of the keys (|key| <= sqrt(3) max|x|, f32)
__device__ __forceinline__ float cut_delta(const Dev &d) {
  return 6e-7f * (float)(fabs(d.dim[0]) + fabs(d.dim[1]) + fabs(d.dim[2]));
}
__device__ __forceinline__ float safe_r2(float H, float delta, bool force_slow) {
  const float rs = H - delta;
  return (force_slow || !(rs > 0.f)) ? -1.f : rs * rs * 0.999999f;
}

// Exact sorted-axis cuts of the density loops for one (lane, candidate): the slow path of hits
// within delta of the search radius.  key: sort key of the candidate.
// What the exact-cut slow paths read of Dev (a __noinline__ function taking Dev itself would force the
// kernel's copy of it onto the local-memory stack).
struct CutCtx {
  const CellDev *cells;
  const int64_t *sort_lookup;
  const SortEnt *sort;
  const float *cell_hmax;
  double dim[3];
};
__device__ __forceinline__ CutCtx cut_ctx(const Dev &d) {
  CutCtx c;
  c.cells = d.cells; c.sort_lookup = d.sort_lookup; c.sort = d.sort; c.cell_hmax = d.cell_hmax;
  c.dim[0] = d.dim[0]; c.dim[1] = d.dim[1]; c.dim[2] = d.dim[2];
  return c;
}
__device__ __noinline__ bool dens_cut_exact(const CutCtx d, const Item item, double xi, double yi, double zi,
                                            float hig, float key) {
  const int kind = item.kind, sid = item.sid;
  if (kind == kSelf || kind == kSubSelf) return true;
  const CellDev cB = d.cells[item.partner];
  double sh[3];
#pragma unroll
  for (int k = 0; k < 3; k++) sh[k] = (double)item_shift_i(item.shift, k) * d.dim[k];
  const int64_t segB = d.sort_lookup[(int64_t)item.partner * 13 + sid];
  if (kind == kPairLeft) {
    // DOPAIR1 loop 1 (:1122-1216)
    const double rshift = item_rshift(sh, sid);
    const float d_me = sort_key(xi, yi, zi, sid);
    const double dj_min = (double)d.sort[segB].d;
    const double hi_max = __dsub_rn((double)__fmul_rn(d.cell_hmax[item.own], kc.gamma), rshift);
    const double di = __dsub_rn((double)__fadd_rn(d_me, hig), rshift);
    if (!((__dadd_rn((double)d_me, hi_max) > dj_min) && !(di < dj_min))) return false;
    return key <= thr_lt(di);
  } else if (kind == kPairRight) {
    // DOPAIR1 loop 2 (:1222-1316)
    const double rshift = item_rshift(sh, sid);
    const float d_me = sort_key(xi, yi, zi, sid);
    const double di_max = __dsub_rn((double)d.sort[segB + cB.count - 1].d, rshift);
    const double hj_max = (double)__fmul_rn(d.cell_hmax[item.own], kc.gamma);
    const double dj = __dadd_rn((double)__fsub_rn(d_me, hig), rshift);
    if (!((__dsub_rn((double)d_me, hj_max) < di_max) && !(__dsub_rn(dj, rshift) > di_max))) return false;
    return key >= thr_gt(dj);
  }
  // DOPAIR_SUBSET (:745-770, 794-838)
  const double px = __dsub_rn(xi, sh[0]), py = __dsub_rn(yi, sh[1]), pz = __dsub_rn(zi, sh[2]);
  const float hg = (kind == kSubPair) ? __fadd_rn(hig, 0.f) : __fsub_rn(-hig, 0.f);
  const double cut = __dadd_rn(__dadd_rn(__dadd_rn((double)hg, __dmul_rn(px, kc.axis[sid][0])),
                                         __dmul_rn(py, kc.axis[sid][1])),
                               __dmul_rn(pz, kc.axis[sid][2]));
  return kind == kSubPair ? (key <= thr_lt(cut)) : (key >= thr_gt(cut));
}

// The same for a neighbour known only by its index (the FILT form of k_density decides when the hit is
// consumed, where the interaction it came from is no longer at hand): the interaction is found by the
// cell range j lies in, the neighbour's sort key is formed again from its position (as k_sort does).
__device__ __noinline__ bool dens_cut_exact_scan(const CutCtx d, const Item *__restrict__ items, int ib, int ie,
                                                 const PosD *__restrict__ pos, int j, double xi, double yi, double zi,
                                                 float hig) {
  for (int it = ib; it < ie; it++) {
    const Item item = items[it];
    if (item.kind == kSelf || item.kind == kSubSelf) continue;
    const CellDev cB = d.cells[item.partner];
    if (j < cB.first || j >= cB.first + cB.count) continue;
    double xj, yj, zj;
    load_pos(pos, j, xj, yj, zj);
    return dens_cut_exact(d, item, xi, yi, zi, hig, sort_key(xj, yj, zj, item.sid));
  }
  return true;  // own cell: no sorted-axis cut
}

// ---------------------------------------------------------------------------
// K3 / K5: density.
// MODE 0: first pass, DOSELF1 / DOPAIR1 arithmetic, every active particle;
//         a unit is 32 particles of a leaf.
// MODE 1: ghost redo sweep, DOSELF_SUBSET / DOPAIR_SUBSET arithmetic, only the
//         particles the ghost flagged (gstate == 2); a unit is a leaf, whose
//         flagged particles the warp first compacts into full groups of 32.
// LOOP 1 (with MODE 0): the GRADIENT loop of SPHENIX -- the reference pastes FUNCTION = gradient into
//         the same DOSELF1 / DOPAIR1 / DOSUB_*1 walks (src/runner_doiact_hydro.c:47-52), so this is the
//         first-pass walk with the converged h, runner_iact_nonsym_gradient as its body and one more
//         float4 gathered per hit (gin: u, rho, soundspeed, viscosity.alpha of the neighbour).
// LOOP 2 (with MODE 0): the time-step LIMITER loop -- FUNCTION = limiter in walks of the same shape
//         (src/runner_doiact_functions_limiter.h:233-450, 541-680; "starting" is the same comparison
//         time_bin <= max_active_bin, src/active.h:698-717, on the bins the time-step task has just
//         given).  runner_iact_nonsym_limiter (src/timestep_limiter_iact.h:107-117) writes the
//         NEIGHBOUR: wakeup[j] = max(wakeup[j], -time_bin_i) when time_bin_j > time_bin_i + 2.  No
//         sums, nothing of the lane's own particle is written.
// ---------------------------------------------------------------------------
template <int MODE, bool FILT = false, int LOOP = 0>
__global__ void __launch_bounds__(kBlock, SWIFT_MINB_DENSITY) k_density(Dev d, int sweep, int force_slow = 0) {
  if (MODE == 1 && d.walk_count[sweep] == 0) return;
  __shared__ WarpTile tiles[kWarps];
  WarpTile &ws = tiles[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  float4 *const gq = d.gq + (size_t)(blockIdx.x * kWarps + (threadIdx.x >> 5)) * (kQD * 32) + lane;
  // redo sweeps: while most particles are still flagged (the first sweeps) a warp takes a 32-particle
  // unit as the first pass does (5x more leaves in flight would not fit the L2); once the flagged
  // particles are few, a warp takes a whole leaf and compacts them into full groups of 32
  // (against the lanes that hold a particle of an active leaf, not d.n: a rank's halo copies are never flagged)
  const bool by_unit = MODE == 0 || (long long)d.walk_count[sweep] * 2 > (long long)d.n_units * 27;
  // few leaves with a flagged particle: walk the list k_ghost pushed; otherwise all leaves in tree order
  // (spatially coherent, the unflagged ones are skipped after one look)
  const bool by_list = MODE == 1 && !by_unit && (long long)d.leaf_count[sweep] * 4 < d.n_leaves;
  // first pass with few active units (time-integration mode): walk the list k_mark_sorts pushed
  const bool by_ulist = MODE == 0 && d.unit_list != nullptr && (long long)*d.unit_count * 4 < d.n_units;
  // first pass over a range of units only (swift_cuda_step_parts, records arriving in pieces)
  const bool ranged = MODE == 0 && !by_ulist && d.unit_hi > 0;
  const int total = by_ulist ? *d.unit_count
                             : (by_unit ? (ranged ? d.unit_hi : d.n_units) : (by_list ? d.leaf_count[sweep] : d.n_leaves));
  const int wid0 = ranged ? d.unit_lo : 0;
  int *const work = d.work_ctr + (LOOP >= 1 ? 2 + kc.max_iter : (MODE == 0 ? 0 : 1 + sweep));
  const Item *__restrict__ items = MODE == 0 ? d.items : d.redo_items;
  const bool rec = d.rec_capacity > 0;
  const float qnan = __int_as_float(0x7fc00000);
  unsigned long long nhits = 0;

  while (true) {
    int wid = 0;
    if (lane == 0) wid = atomicAdd(work, 1) + wid0;
    wid = __shfl_sync(kFull, wid, 0);
    if (wid >= total) break;
    if (by_ulist) wid = d.unit_list[wid];
    const int slot = by_unit ? d.unit_leaf[wid] : (by_list ? d.leaf_list[wid] : wid);
    const CellDev cL = d.cells[d.leaves[slot]];
    const int ib = (MODE == 0 ? d.item_begin : d.redo_begin)[slot];
    const int ie = (MODE == 0 ? d.item_begin : d.redo_begin)[slot + 1];
    int scan = 0, nsel = 0;
    bool first = true;

    while (true) {
      // ---- which particle does this lane own in this round?
      int p = -1;
      if (by_unit) {
        if (!first) break;
        first = false;
        const int q = d.lane_part[(int64_t)wid * 32 + lane];
        if (MODE == 0) {
          if (q >= 0 && d.time_bin[q] <= kc.max_active_bin) p = q;
        } else {
          if (q >= 0 && d.gstate[q] == 2) p = q;
        }
      } else {
        // the leaf's particles in lane (Morton) order
        const int64_t lp0 = (int64_t)d.unit_begin[slot] * 32;
        const int nlp = (d.unit_begin[slot + 1] - d.unit_begin[slot]) * 32;
        while (nsel < 32 && scan < nlp) {
          const int q = d.lane_part[lp0 + scan + lane];
          const bool f = q >= 0 && d.gstate[q] == 2;
          const unsigned b = __ballot_sync(kFull, f);
          if (f) ws.sel[nsel + __popc(b & lt_mask)] = q;
          nsel += __popc(b);
          scan += 32;
        }
        __syncwarp();
        if (nsel == 0) break;
        const int take = min(nsel, 32);
        if (lane < take) p = ws.sel[lane];
        const int carry = (lane + 32 < nsel) ? ws.sel[lane + 32] : 0;
        __syncwarp();
        ws.sel[lane] = carry;
        nsel -= take;
        __syncwarp();
      }
      if (!__any_sync(kFull, p >= 0)) continue;

      const bool mine = p >= 0;
      const int pp = mine ? p : cL.first;
      double xi, yi, zi;
      load_pos(d.pos, pp, xi, yi, zi);
      const float hi = d.h[pp];
      const float hig2 = sq_gamma2(hi);
      const float hi_inv = 1.f / hi;
      const float hig = __fmul_rn(hi, kc.gamma);
      const float4 vmi = d.vm[pp];
      const long long idi = rec ? d.id[pp] : 0;
      DensAcc acc;
      acc.zero();
      const int bin_i = LOOP == 2 ? (int)d.time_bin[pp] : 0;
      if (MODE == 0 && LOOP != 2 && d.keep_sums && mine) {
        const float4 s0 = d.dens0[p], s1 = d.dens1[p];
        acc.rho = s0.x; acc.rho_dh = s0.y; acc.wcount = s0.z; acc.wcount_dh = s0.w;
        acc.div_v = s1.x; acc.rx = s1.y; acc.ry = s1.z; acc.rz = s1.w;
      }
      GradAcc gacc;
      float4 gi = make_float4(0.f, 0.f, 0.f, 0.f);
      if (LOOP == 1) {
        // hydro_reset_gradient left v_sig = 2 c and alpha_visc_max_ngb = alpha; laplace_u is 0 since hydro_init_part
        gi = d.gin[pp];
        const float4 sg = d.sph_grad[pp];
        gacc.v_sig = d.v_sig[pp];
        gacc.div_v = sg.x; gacc.lap = sg.y; gacc.amax = sg.z;
      }
      int qcnt = 0;
      // list-recording sweep: the walk covers kGrowL x the search radius
      const bool recl = MODE == 1 && d.nl != nullptr;
      const float higR = recl ? __fmul_rn(hig, d.nl_grow + d.nl_slope * fabsf(hi / d.h_before[pp] - 1.f)) : hig;
      const float hig2R = recl ? __fmul_rn(__fmul_rn(higR, higR), 1.000001f) : hig2;
      const float r2safe = FILT ? safe_r2(hig, cut_delta(d), force_slow != 0) : 0.f;
      int ln = 0;
      float4 *nlp = nullptr;
      int ul = -1;
      if (recl && mine) {
        ul = d.part_lane[p];
        nlp = d.nl + ((size_t)(ul >> 5) * kLD) * 32 + (ul & 31);
      }

      // box of the unit's particles, each grown by its search radius, in the leaf's frame
      float ulo[3], uhi[3];
      {
        const float l[3] = {dsub_f(xi, cL.loc[0]), dsub_f(yi, cL.loc[1]), dsub_f(zi, cL.loc[2])};
        const float ru = higR * 1.0001f;
#pragma unroll
        for (int k = 0; k < 3; k++) {
          ulo[k] = warp_min(mine ? l[k] - ru : INFINITY);
          uhi[k] = warp_max(mine ? l[k] + ru : -INFINITY);
        }
      }

      // one deferred hit -> runner_iact_nonsym_density; vmj (and the f64 position of j when the
      // difference must be taken in f64) were gathered one round ahead
      auto consume = [&](const float4 e, const float4 vmj, const float4 gj, double xj, double yj, double zj) {
        const int w = __float_as_int(e.w);
        float dx, dy, dz;
        if (w & kDpFlag) {
          if (MODE == 0) {
            // own cell (DOSELF1): no periodic shift
            dx = dsub_f(xi, xj); dy = dsub_f(yi, yj); dz = dsub_f(zi, zj);
          } else {
            // e.xyz: periodic shift of the item in box lengths (-1, 0, 1)
            dx = dsub_f(__dsub_rn(xi, __dmul_rn((double)e.x, d.dim[0])), xj);
            dy = dsub_f(__dsub_rn(yi, __dmul_rn((double)e.y, d.dim[1])), yj);
            dz = dsub_f(__dsub_rn(zi, __dmul_rn((double)e.z, d.dim[2])), zj);
          }
        } else {
          dx = e.x; dy = e.y; dz = e.z;
        }
        const float r2 = dist2(dx, dy, dz);
        if (recl && r2 < hig2R) {
          if (ln < kLD) nlp[(size_t)ln * 32] = make_float4(dx, dy, dz, __int_as_float(w & kIdxMask));
          ln++;
        }
        bool hit = r2 < hig2 && (FILT || !recl || (w & kInWin));
        // FILT: the sorted-axis cut is implied by r2 < h^2 gamma^2 unless r is within delta of the radius
        if (FILT && hit && !(r2 < r2safe))
          hit = dens_cut_exact_scan(cut_ctx(d), items, ib, ie, d.pos, w & kIdxMask, xi, yi, zi, hig);
        if (hit && LOOP == 2) {
          const int j = w & kIdxMask;
          const int bin_j = (int)d.time_bin[j];
          bool wake = bin_j > bin_i + 2;  // time_bin_neighbour_max_delta_bin, src/timeline.h:51
          // DOSELF1 (:641-660): two starting particles in each other's range meet in the symmetric
          // runner_iact_limiter, which does nothing; the sorted pair walk knows no such case
          if (wake && (w & kDpFlag) && bin_j <= kc.max_active_bin && r2 < sq_gamma2(d.h[j])) wake = false;
          if (wake) atomicMax(&d.wakeup[j], -bin_i);
          nhits++;
        } else if (hit) {
          if (LOOP == 1) iact_gradient(gacc, r2, dx, dy, dz, hi, vmi, gi, vmj, gj);
          else iact_density(acc, r2, dx, dy, dz, hi_inv, vmi.x, vmi.y, vmi.z, vmj);
          nhits++;
          if (rec) {
            if (LOOP == 1) record_pair_grad(d, idi, d.id[w & kIdxMask]);
            else record_pair(d, MODE, idi, d.id[w & kIdxMask]);
          }
        }
      };
      // software pipeline over the FIFO: entry r+2 is being read, the gathers of entry r+1 are in
      // flight, entry r is consumed
      auto drain_all = [&]() {
        const int m = __reduce_max_sync(kFull, qcnt);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        auto gather = [&](const float4 e, bool ok, float4 &v, float4 &g, double &x, double &y, double &z) {
          if (ok) {
            const int w = __float_as_int(e.w);
            const int j = w & kIdxMask;
            if (LOOP != 2) v = d.vm[j];
            if (LOOP == 1) g = d.gin[j];
            if (w & kDpFlag) load_pos(d.pos, j, x, y, z);
          }
        };
        float4 e0 = qcnt > 0 ? __ldcs(&gq[0]) : z4;
        float4 e1 = qcnt > 1 ? __ldcs(&gq[32]) : z4;
        float4 v0 = z4, g0 = z4;
        double x0 = 0.0, y0 = 0.0, z0 = 0.0;
        gather(e0, qcnt > 0, v0, g0, x0, y0, z0);
        for (int r = 0; r < m; r++) {
          const float4 e2 = (r + 2 < qcnt) ? __ldcs(&gq[(r + 2) * 32]) : z4;
          float4 v1 = z4, g1 = z4;
          double x1 = 0.0, y1 = 0.0, z1 = 0.0;
          gather(e1, r + 1 < qcnt, v1, g1, x1, y1, z1);
          if (r < qcnt) consume(e0, v0, g0, x0, y0, z0);
          e0 = e1; e1 = e2; v0 = v1; g0 = g1; x0 = x1; y0 = y1; z0 = z1;
        }
        qcnt = 0;
      };

      for (int it = ib; it < ie; it++) {
        const Item item = items[it];
        const CellDev cA = d.cells[item.own];
        const CellDev cB = d.cells[item.partner];
        const int kind = item.kind;
        const int sid = item.sid;
        const bool sorted = !(kind == kSelf || kind == kSubSelf);
        const bool desc = (kind == kPairRight || kind == kSubPairFlip);
        const bool dp = (kind == kSelf || kind == kSubPair || kind == kSubPairFlip);
        double sh[3] = {0.0, 0.0, 0.0};
        float shf[3] = {0.f, 0.f, 0.f};
        if (sorted) {
#pragma unroll
          for (int k = 0; k < 3; k++) {
            const int s = item_shift_i(item.shift, k);
            shf[k] = (float)s;
            sh[k] = (double)s * d.dim[k];
          }
        }
        // frame origin of the staged candidates (fo) and of this lane's particle (fl);
        // sub: what DOPAIR_SUBSET takes off the lane's position first
        double fo[3], fl[3], sub[3] = {0.0, 0.0, 0.0};
        float slack = 0.f;
        if (kind == kSelf || kind == kSubSelf) {
#pragma unroll
          for (int k = 0; k < 3; k++) fo[k] = fl[k] = cA.loc[k];
          if (kind == kSelf) slack = 1e-6f * cA.width;
        } else if (kind == kPairLeft) {
#pragma unroll
          for (int k = 0; k < 3; k++) { fo[k] = cB.loc[k]; fl[k] = __dadd_rn(cB.loc[k], sh[k]); }
        } else if (kind == kPairRight) {
#pragma unroll
          for (int k = 0; k < 3; k++) { fo[k] = __dadd_rn(cA.loc[k], sh[k]); fl[k] = cA.loc[k]; }
        } else {
#pragma unroll
          for (int k = 0; k < 3; k++) { fo[k] = fl[k] = cB.loc[k]; sub[k] = sh[k]; }
          slack = 2e-6f * fmaxf(cA.width, cB.width);
        }
        StageBox sb;
        {
          double ol[3], og[3];
#pragma unroll
          for (int k = 0; k < 3; k++) {
            ol[k] = cL.loc[k] - sub[k] - fl[k];
            og[k] = cB.loc[k] - fo[k];
          }
          stage_box(ulo, uhi, ol, slack, sb);
          if (!box_reaches_cell(sb, d.cell_box + (int64_t)item.partner * 6, og)) continue;
        }

        const int64_t segB = sorted ? d.sort_lookup[(int64_t)item.partner * 13 + sid] : 0;
        double px = xi, py = yi, pz = zi;
        float lo = INFINITY, hi_cut = -INFINITY;  // lane's window on the candidates' sort keys (empty)
        float loR = INFINITY, hiR = -INFINITY;    // the same for the radius a list-recording sweep covers
        float thr = hig2R;
        if (kind == kSelf || kind == kSubSelf) {
          if (mine) { lo = loR = -INFINITY; hi_cut = hiR = INFINITY; }
          if (kind == kSelf) thr = cull_thr(hig2, 4e-7f * cA.width);
        } else if (FILT) {
          // the candidate test is a pure distance filter; what is needed here is only how far along the
          // sorted axis this lane can reach at all (|key_j - key_i'| <= r): it bounds the staging walk
          if (kind != kPairLeft && kind != kPairRight) {
            px = __dsub_rn(xi, sh[0]); py = __dsub_rn(yi, sh[1]); pz = __dsub_rn(zi, sh[2]);
            thr = cull_thr(hig2R, 8e-7f * fmaxf(cA.width, cB.width));
          }
          if (mine) {
            const double rshift = item_rshift(sh, sid);
            const float d_me = sort_key(xi, yi, zi, sid);
            const float kme = (float)(kind == kPairRight ? ((double)d_me + rshift) : ((double)d_me - rshift));
            const float m = 4e-7f * (fabsf(kme) + fabsf((float)rshift) + fabsf(d_me)) + 1e-30f;
            const float reach = higR * 1.000002f;
            if (!desc) { loR = -INFINITY; hiR = kme + reach + m; }
            else { loR = kme - reach - m; hiR = INFINITY; }
            lo = loR; hi_cut = hiR;
          }
        } else if (kind == kPairLeft) {
          // this leaf's particles are "pi" of DOPAIR1 loop 1 (:1122-1216): frame cj.loc + shift
          if (mine) {
            const double rshift = item_rshift(sh, sid);
            const float d_me = sort_key(xi, yi, zi, sid);
            const double dj_min = (double)d.sort[segB].d;
            const double hi_max = __dsub_rn((double)__fmul_rn(d.cell_hmax[item.own], kc.gamma), rshift);
            const double di = __dsub_rn((double)__fadd_rn(d_me, hig), rshift);
            if ((__dadd_rn((double)d_me, hi_max) > dj_min) && !(di < dj_min)) { lo = -INFINITY; hi_cut = thr_lt(di); }
          }
        } else if (kind == kPairRight) {
          // this leaf's particles are "pj" of DOPAIR1 loop 2 (:1222-1316): frame cj.loc (their own cell)
          if (mine) {
            const double rshift = item_rshift(sh, sid);
            const float d_me = sort_key(xi, yi, zi, sid);
            const double di_max = __dsub_rn((double)d.sort[segB + cB.count - 1].d, rshift);
            const double hj_max = (double)__fmul_rn(d.cell_hmax[item.own], kc.gamma);
            const double dj = __dadd_rn((double)__fsub_rn(d_me, hig), rshift);
            if ((__dsub_rn((double)d_me, hj_max) < di_max) && !(__dsub_rn(dj, rshift) > di_max)) {
              lo = thr_gt(dj);
              hi_cut = INFINITY;
            }
          }
        } else {
          // DOPAIR_SUBSET (:745-770, 794-838): pix = x - shift in f64, dx = (float)(pix - pj.x)
          px = __dsub_rn(xi, sh[0]); py = __dsub_rn(yi, sh[1]); pz = __dsub_rn(zi, sh[2]);
          thr = cull_thr(hig2R, 8e-7f * fmaxf(cA.width, cB.width));
          if (mine) {
            const float hg = (kind == kSubPair) ? __fadd_rn(hig, 0.f) : __fsub_rn(-hig, 0.f);
            const double cut = __dadd_rn(__dadd_rn(__dadd_rn((double)hg, __dmul_rn(px, kc.axis[sid][0])),
                                                   __dmul_rn(py, kc.axis[sid][1])),
                                         __dmul_rn(pz, kc.axis[sid][2]));
            if (kind == kSubPair) { lo = -INFINITY; hi_cut = thr_lt(cut); }
            else { lo = thr_gt(cut); hi_cut = INFINITY; }
            loR = lo; hiR = hi_cut;
            if (recl) {
              // the cut of the wider radius, generously: the list is a superset, the exact cut of the
              // current h travels with every entry (kInWin)
              const float m = (higR - hig) * 1.01f + 4.f * cut_delta(d);
              if (kind == kSubPair) hiR = hi_cut + m; else loR = lo - m;
            }
          }
        }
        if (!recl) { loR = lo; hiR = hi_cut; }
        // union of the lanes' windows: what is worth staging at all
        const float wlo = warp_min(loR), whi = warp_max(hiR);
        if (!(wlo <= whi)) continue;  // no lane takes part
        const float pfx = dsub_f(px, fl[0]), pfy = dsub_f(py, fl[1]), pfz = dsub_f(pz, fl[2]);

        // every lane against the staged candidates [st, st + n) of the ring (st is 0 or 32: no wrap);
        // n a multiple of 4.  The distance is formed with FMAs and compared with a threshold a few ulp
        // wider: the exact r2 < h^2 gamma^2 of the reference is redone when the hit is consumed.
        const float thr_c = thr * 1.000001f;
        const int flagbits = dp ? kDpFlag : 0;
        // SHIFTED (redo sweeps, pairs across the cell boundary): the entry carries the periodic shift
        // instead of dx, which the body takes again from the f64 positions
        auto test_impl = [&](auto shifted_c, int st, int n) {
          constexpr bool SHIFTED = decltype(shifted_c)::value;
          const float4 *__restrict__ tp = ws.pos + st;
          const int *__restrict__ ti = ws.idx + st;
          float4 *qp = gq + qcnt * 32;
          for (int k0 = 0; k0 < n; k0 += 4) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
              const float4 c = tp[k0 + u];
              const float dx = __fsub_rn(pfx, c.x), dy = __fsub_rn(pfy, c.y), dz = __fsub_rn(pfz, c.z);
              const float r2c = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
              const int j = ti[k0 + u];
              const bool push = FILT ? (mine && r2c < thr_c && j != pp)
                                     : (c.w >= loR && c.w <= hiR && r2c < thr_c && j != pp);
              const int inwin = (!FILT && MODE == 1 && c.w >= lo && c.w <= hi_cut) ? kInWin : 0;
              // straight-line: the store and the FIFO advance are predicated, not branched over
              if (SHIFTED) st_pred(qp, shf[0], shf[1], shf[2], __int_as_float(j | kDpFlag | inwin), push);
              else st_pred(qp, dx, dy, dz, __int_as_float(j | flagbits | inwin), push);
              qp += push ? 32 : 0;
            }
          }
          qcnt = (int)(qp - gq) >> 5;
        };
        auto test = [&](int st, int n) {
          if (__any_sync(kFull, qcnt > kQD - n)) drain_all();
          if (MODE == 1 && dp) test_impl(std::true_type{}, st, n);
          else test_impl(std::false_type{}, st, n);
        };

        int st = 0, ntile = 0, base = 0;
        bool done = false;
        // this lane's sort entry of a staging round, read one round ahead
        auto load_se = [&](int b) {
          SortEnt se;
          se.d = 0.f;
          se.i = b + lane;  // unsorted: index order
          if (sorted && b + lane < cB.count) se = d.sort[segB + (desc ? (cB.count - 1 - b - lane) : (b + lane))];
          return se;
        };
        SortEnt se_next = load_se(0);
        while (true) {
          if (base < cB.count) {
            const int nt = min(32, cB.count - base);
            const SortEnt se = se_next;
            if (base + 32 < cB.count) se_next = load_se(base + 32);
            const float dk = se.d;
            const int j = cB.first + se.i;
            bool keep = lane < nt;
            if (sorted) {
              const float k0 = __shfl_sync(kFull, dk, 0);
              if (desc ? !(k0 >= wlo) : !(k0 <= whi)) done = true;  // sorted axis: nothing further on can hit
              keep = keep && !done && dk >= wlo && dk <= whi;
            }
            float4 c = make_float4(qnan, qnan, qnan, qnan);
            if (keep) {
              double xj, yj, zj;
              load_pos(d.pos, j, xj, yj, zj);
              c = make_float4(dsub_f(xj, fo[0]), dsub_f(yj, fo[1]), dsub_f(zj, fo[2]), dk);
              keep = c.x >= sb.lo[0] && c.x <= sb.hi[0] && c.y >= sb.lo[1] && c.y <= sb.hi[1] && c.z >= sb.lo[2] &&
                     c.z <= sb.hi[2];
            }
            const unsigned b = __ballot_sync(kFull, keep);
            if (keep) {
              const int s = (st + ntile + __popc(b & lt_mask)) & (kTile - 1);
              ws.pos[s] = c;
              ws.idx[s] = j;
            }
            ntile += __popc(b);
            base += 32;
          } else {
            done = true;
          }
          int n = 0;
          if (ntile >= 32) {
            n = 32;
          } else if (done && ntile > 0) {
            n = (ntile + 3) & ~3;
            if (lane < n - ntile) ws.pos[(st + ntile + lane) & (kTile - 1)] = make_float4(qnan, qnan, qnan, qnan);
          }
          if (n > 0) {
            __syncwarp();
            test(st, n);
            st = (st + 32) & (kTile - 1);
            ntile = max(ntile - 32, 0);
          }
          if (done && ntile == 0) break;
        }
        __syncwarp();
      }
      drain_all();
      if (LOOP == 1) {
        if (mine) {
          d.v_sig[p] = gacc.v_sig;
          d.sph_grad[p] = make_float4(gacc.div_v, gacc.lap, gacc.amax, 0.f);
        }
      } else if (LOOP == 2) {
        // nothing of the lane's own particle changes
      } else if (mine) {
        d.dens0[p] = make_float4(acc.rho, acc.rho_dh, acc.wcount, acc.wcount_dh);
        d.dens1[p] = make_float4(acc.div_v, acc.rx, acc.ry, acc.rz);
        if (recl) {
          d.nl_cnt[ul] = ln <= kLD ? (unsigned short)ln : (unsigned short)0xffff;
          d.nl_R[ul] = hig2R;
        }
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) nhits += __shfl_down_sync(kFull, nhits, o);
  if (lane == 0 && nhits && LOOP != 2) atomicAdd(&d.hits[LOOP == 1 ? 3 : MODE], nhits);
}

// ---------------------------------------------------------------------------
// K4: ghost.  One thread per particle: hydro_end_density, Newton-Raphson /
// bisection step on h, convergence test, then either re-arm (zero the sums,
// flag for the next redo sweep) or hydro_prepare_force + hydro_reset_acceleration.
// runner_do_ghost, src/runner_ghost.c:1148-1483.
// ---------------------------------------------------------------------------
// expf as glibc >= 2.27 computes it (sysdeps/ieee754/flt-32/e_expf.c with the 32-entry table of
// exp2f_data.c: x * 32 / ln2 = k + r, 2^(k/32) from the table, a cubic in r, all in double, one
// rounding to float).  The reference calls libm's expf in the expanded / hybrid Tillotson states (tillotson.h) and when a
// drift changes h by more than ~20 % (hydro_predict_extra); CUDA's expf differs from it in the last bit
// for some arguments, which showed as 1-ulp pressure differences on low-density surface particles.
__device__ const unsigned long long kExp2fTab[32] = {
      0x3ff0000000000000ULL, 0x3fefd9b0d3158574ULL, 0x3fefb5586cf9890fULL, 0x3fef9301d0125b51ULL,
      0x3fef72b83c7d517bULL, 0x3fef54873168b9aaULL, 0x3fef387a6e756238ULL, 0x3fef1e9df51fdee1ULL,
      0x3fef06fe0a31b715ULL, 0x3feef1a7373aa9cbULL, 0x3feedea64c123422ULL, 0x3feece086061892dULL,
      0x3feebfdad5362a27ULL, 0x3feeb42b569d4f82ULL, 0x3feeab07dd485429ULL, 0x3feea47eb03a5585ULL,
      0x3feea09e667f3bcdULL, 0x3fee9f75e8ec5f74ULL, 0x3feea11473eb0187ULL, 0x3feea589994cce13ULL,
      0x3feeace5422aa0dbULL, 0x3feeb737b0cdc5e5ULL, 0x3feec49182a3f090ULL, 0x3feed503b23e255dULL,
      0x3feee89f995ad3adULL, 0x3feeff76f2fb5e47ULL, 0x3fef199bdd85529cULL, 0x3fef3720dcef9069ULL,
      0x3fef5818dcfba487ULL, 0x3fef7c97337b9b5fULL, 0x3fefa4afa2a490daULL, 0x3fefd0765b6e4540ULL};
__device__ float expf_glibc(float x) {
  if (x != x) return x;
  if (x > 0x1.62e42ep6f) return __int_as_float(0x7f800000);  // > log(2^128): overflow
  if (x < -0x1.9fe368p6f) return 0.f;                        // < log(2^-150): underflow
  const double inv_ln2_n = 0x1.71547652b82fep+0 * 32.0, shift = 0x1.8p+52;
  const double c0 = 0x1.c6af84b912394p-5 / 32.0 / 32.0 / 32.0, c1 = 0x1.ebfce50fac4f3p-3 / 32.0 / 32.0,
               c2 = 0x1.62e42ff0c52d6p-1 / 32.0;
  const double z = __dmul_rn(inv_ln2_n, (double)x);
  double kd = __dadd_rn(z, shift);
  const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
  kd = __dsub_rn(kd, shift);
  const double r = __dsub_rn(z, kd);
  const double s = __longlong_as_double((long long)(kExp2fTab[ki & 31] + (ki << 47)));
  const double zz = __dadd_rn(__dmul_rn(c0, r), c1);
  const double r2 = __dmul_rn(r, r);
  double y = __dadd_rn(__dmul_rn(c2, r), 1.0);
  y = __dadd_rn(__dmul_rn(zz, r2), y);
  return (float)__dmul_rn(y, s);
}

__device__ float til_pressure(float density, float u, const swift_cuda_til &m) {
  const float eta = density / m.rho_0;
  const float eta_sq = eta * eta;
  const float mu = eta - 1.f;
  const float nu = 1.f / eta - 1.f;
  const float w = u / (m.u_0 * eta_sq) + 1.f;
  const float w_inv = 1.f / w;
  float P_c = (m.a + m.b * w_inv) * density * u + m.A * mu + m.B * mu * mu;
  if (eta < m.eta_zero)
    P_c = 0.f;
  else if (eta < m.eta_min)
    P_c *= (eta - m.eta_zero) / (m.eta_min - m.eta_zero);
  const float P_e =
      m.a * density * u + (m.b * density * u * w_inv + m.A * mu * expf_glibc(-m.beta * nu)) * expf_glibc(-m.alpha * nu * nu);
  float P;
  if ((1.f < eta) || (u < m.u_iv))
    P = P_c;
  else if ((eta < 1.f) && (m.u_cv < u))
    P = P_e;
  else
    P = ((u - m.u_iv) * P_e + (m.u_cv - u) * P_c) / (m.u_cv - m.u_iv);
  if (P < m.P_min) P = m.P_min;
  return P;
}

__device__ float til_soundspeed(float density, float u, const swift_cuda_til &m) {
  const float rho_0_inv = 1.f / m.rho_0;
  const float eta = density * rho_0_inv;
  const float rho_inv = 1.f / density;
  const float eta_sq = eta * eta;
  const float mu = eta - 1.f;
  const float nu = 1.f / eta - 1.f;
  const float w = u / (m.u_0 * eta_sq) + 1.f;
  const float w_inv = 1.f / w;
  const float w_inv_sq = w_inv * w_inv;
  const float exp_beta = expf_glibc(-m.beta * nu);
  const float exp_alpha = expf_glibc(-m.alpha * nu * nu);
  float P_c = (m.a + m.b * w_inv) * density * u + m.A * mu + m.B * mu * mu;
  if (eta < m.eta_zero)
    P_c = 0.f;
  else if (eta < m.eta_min)
    P_c *= (eta - m.eta_zero) / (m.eta_min - m.eta_zero);
  const float c_sq_c = P_c * rho_inv * (1.f + m.a + m.b * w_inv) +
                       m.b * (w - 1.f) * w_inv_sq * (2.f * u - P_c * rho_inv) +
                       rho_inv * (m.A + m.B * (eta_sq - 1.f));
  const float P_e = m.a * density * u + (m.b * density * u * w_inv + m.A * mu * exp_beta) * exp_alpha;
  const float c_sq_e =
      P_e * rho_inv * (1.f + m.a + m.b * w_inv * exp_alpha) +
      (m.b * density * u * w_inv_sq / eta_sq *
           (rho_inv / m.u_0 * (2.f * u - P_e * rho_inv) + 2.f * m.alpha * nu * w * rho_0_inv) +
       m.A * rho_0_inv * (1.f + mu / eta_sq * (m.beta + 2.f * m.alpha * nu - eta)) * exp_beta) *
          exp_alpha;
  float c_sq;
  if ((1.f < eta) || (u < m.u_iv))
    c_sq = c_sq_c;
  else if ((eta < 1.f) && (m.u_cv < u))
    c_sq = c_sq_e;
  else
    c_sq = ((u - m.u_iv) * c_sq_e + (m.u_cv - u) * c_sq_c) / (m.u_cv - m.u_iv);
  c_sq = fmaxf(c_sq, m.A * rho_0_inv);
  return sqrtf(c_sq);
}

// gas_pressure_from_internal_energy / gas_soundspeed_*: ideal gas
// (src/equation_of_state/ideal_gas/equation_of_state.h) or the planetary
// dispatcher (src/equation_of_state/planetary/equation_of_state.h:910-1146, 1305-1457).
__device__ __forceinline__ void eos_eval(float rho, float u, int mat_id, float &P, float &cs) {
  if (kc.scheme != SWIFT_CUDA_SCHEME_PLANETARY) {
    P = 0.66666666666666667f * u * rho;
    cs = sqrtf(1.66666666666666667f * P / rho);
    return;
  }
  if (mat_id / 100 == 0) {
    const float g = kc.idg_gamma[mat_id % 100 & 3];
    P = (g - 1.f) * u * rho;
    cs = sqrtf(u * g * (g - 1.f));  // planetary/ideal_gas.h:161-165
    return;
  }
  if (mat_id >= 200) {
    // a material that lives in a table (HM80 200-202, SESAME 300-303, ANEOS 400-402, custom 900-909): eos_table.h
    for (int k = 0; k < kc.n_tab; k++)
      if (kc.tab[k].mat_id == mat_id) {
        et_eval(kc.tab[k], rho, u, P, cs);
        return;
      }
  }
  int m = 0;
  for (int k = 0; k < kc.n_til; k++)
    if (kc.til[k].mat_id == mat_id) m = k;
  P = til_pressure(rho, u, kc.til[m]);
  cs = til_soundspeed(rho, u, kc.til[m]);
}

// cbrtf as glibc 2.39 computes it (sysdeps/ieee754/flt-32/s_cbrtf.c: frexp, a quadratic
// seed, one Halley step in double, a table of 2^(k/3)), so that the rare bisection step of
// the ghost gives the same bits as the oracle on the host.  CUDA's cbrtf differs in the last ulp.
__device__ float cbrtf_glibc(float x) {
  const double factor[5] = {1.0 / 1.5874010519681994748, 1.0 / 1.2599210498948731648, 1.0, 1.2599210498948731648,
                            1.5874010519681994748};
  int xe;
  const float xm = frexpf(fabsf(x), &xe);
  if (xe == 0 && (x == 0.f || !isfinite(x))) return x + x;
  const float u = (float)__dadd_rn(
      0.492659620528969547,
      __dmul_rn(__dsub_rn(0.697570460207922770, __dmul_rn(0.191502161678719066, (double)xm)), (double)xm));
  const float t2 = __fmul_rn(__fmul_rn(u, u), u);
  const double num = __dmul_rn((double)u, __dadd_rn((double)t2, __dmul_rn(2.0, (double)xm)));
  const double den = __dadd_rn(__dmul_rn(2.0, (double)t2), (double)xm);
  const float ym = (float)__dmul_rn(__ddiv_rn(num, den), factor[2 + xe % 3]);
  return ldexpf(x > 0.f ? ym : -ym, xe / 3);
}

__device__ __forceinline__ void ghost_particle(const Dev &d, int sweep, int64_t p);
// SPHENIX: what runner_do_ghost does with a converged particle under EXTRA_HYDRO_LOOP (sphenix.cuh)
__device__ void sphenix_converged(const Dev &d, int64_t p, float h, float4 s0, float4 s1);

// Grid-stride: the launch sequence of a step always holds max_smoothing_iterations of these, and
// all but the first few find nothing pending; a small grid makes those launches free.
__global__ void k_ghost(Dev d, int sweep) {
  if (sweep > 0 && d.redo_count[sweep - 1] == 0) return;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < d.n; p += (int64_t)gridDim.x * blockDim.x)
    ghost_particle(d, sweep, p);
}

__device__ __forceinline__ void ghost_particle(const Dev &d, int sweep, int64_t p) {
  const uint8_t st = d.gstate[p];
  if (st != 1 && st != 2 && st != 4) return;
  const bool planetary = kc.scheme == SWIFT_CUDA_SCHEME_PLANETARY;
  float4 s0 = d.dens0[p];  // rho, rho_dh, wcount, wcount_dh
  float4 s1 = d.dens1[p];  // div_v, rot
  const float mass = d.vm[p].w;
  const float h_old = d.h[p];
  const float h_old_dim = h_old * h_old * h_old;
  const float h_old_dim_minus_one = h_old * h_old;
  float left = d.left[p], right = d.right[p];
  float h_new;
  float h = h_old;
  bool has_no_neighbours = false;
  bool finish = false;  // tidy up as converged
  if ((double)s0.z < 1.e-5 * (double)kc.kroot) {
    has_no_neighbours = true;
    h_new = 2.f * h_old;
  } else {
    // hydro_end_density, Minimal/hydro.h:543-575
    const float h_inv = 1.0f / h_old;
    const float h_inv_dim = h_inv * h_inv * h_inv;
    const float h_inv_dim_plus_one = h_inv_dim * h_inv;
    s0.x += mass * kc.kroot;
    s0.y -= 3.f * mass * kc.kroot;
    s0.z += kc.kroot;
    s0.w -= 3.f * kc.kroot;
    s0.x *= h_inv_dim;
    s0.y *= h_inv_dim_plus_one;
    s0.z *= h_inv_dim;
    s0.w *= h_inv_dim_plus_one;
    const float rho_inv = 1.f / s0.x;
    const float a_inv2 = 1.f / (kc.a * kc.a);
    s1.y *= h_inv_dim_plus_one * a_inv2 * rho_inv;
    s1.z *= h_inv_dim_plus_one * a_inv2 * rho_inv;
    s1.w *= h_inv_dim_plus_one * a_inv2 * rho_inv;
    s1.x *= h_inv_dim_plus_one * a_inv2 * rho_inv;
    const float n_sum = s0.z * h_old_dim;
    const float n_target = kc.eta_dim;
    const float f = n_sum - n_target;
    const float f_prime = s0.w * h_old_dim + 3.f * s0.z * h_old_dim_minus_one;
    if (n_sum < n_target)
      left = fmaxf(left, h_old);
    else if (n_sum > n_target)
      right = fminf(right, h_old);
    if (((h_old >= kc.h_max) && (f < 0.f)) || ((h_old <= kc.h_min) && (f > 0.f))) {
      finish = true;
      h_new = h_old;
    } else {
      h_new = h_old - f / (f_prime + FLT_MIN);
      h_new = fminf(h_new, 2.f * h_old);
      h_new = fmaxf(h_new, 0.5f * h_old);
      h_new = fmaxf(h_new, left);
      h_new = fminf(h_new, right);
    }
  }
  if (!finish && fabsf(h_new - h_old) > kc.h_tol * h_old) {
    if ((h_new == left && h_old == right) || (h_old == left && h_new == right))
      h = cbrtf_glibc(0.5f * (left * left * left + right * right * right));
    else
      h = h_new;
    if (h < kc.h_max && h > kc.h_min) {
      // another round: hydro_init_part
      d.h[p] = h;
      d.left[p] = left;
      d.right[p] = right;
      d.dens0[p] = make_float4(0.f, 0.f, 0.f, 0.f);
      d.dens1[p] = make_float4(0.f, 0.f, 0.f, 0.f);
      // 2: the next redo sweep walks the cells for this particle; 4: its recorded neighbour list covers the new h
      uint8_t ns = 2;
      if (d.nl != nullptr) {
        d.h_before[p] = h_old;
        const int ul = d.part_lane[p];
        if (ul >= 0 && d.nl_cnt[ul] != 0xffff && sq_gamma2(h) < d.nl_R[ul] * 0.999998f) ns = 4;
      }
      d.gstate[p] = ns;
      atomicAdd(&d.redo_count[sweep], 1);
      if (ns == 4) return;
      atomicAdd(&d.walk_count[sweep], 1);
      const int slot = d.part_slot[p];
      if (slot >= 0 && atomicExch(&d.leaf_mark[slot], sweep + 1) != sweep + 1)
        d.leaf_list[atomicAdd(&d.leaf_count[sweep], 1)] = slot;
      return;
    } else if (h <= kc.h_min) {
      h = kc.h_min;
    } else {
      h = kc.h_max;
      if (has_no_neighbours) {
        // hydro_part_has_no_neighbours, Minimal/hydro.h:626-650
        const float h_inv = 1.0f / h;
        const float h_inv_dim = h_inv * h_inv * h_inv;
        s0.x = mass * kc.kroot * h_inv_dim;
        s0.z = kc.kroot * h_inv_dim;
        s0.y = 0.f;
        s0.w = 0.f;
        s1 = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  if (kc.scheme == SWIFT_CUDA_SCHEME_SPHENIX) {
    // converged: hydro_prepare_gradient + hydro_reset_gradient (src/runner_ghost.c:1249-1264, 1429-1443)
    sphenix_converged(d, p, h, s0, s1);
    return;
  }
  // converged: hydro_prepare_force + hydro_reset_acceleration
  // (Minimal/hydro.h:669-761, Planetary/hydro.h:667-768)
  const float u = d.u[p];
  const float curl_v = sqrtf(s1.y * s1.y + s1.z * s1.z + s1.w * s1.w);
  const float div_physical_v = s1.x + 3.f * kc.H;
  const float abs_div = fabsf(div_physical_v);
  float pressure, soundspeed;
  eos_eval(s0.x, u, planetary ? d.mat_id[p] : 0, pressure, soundspeed);
  const float common_factor = h * 0.3333333333f / s0.z;
  float grad_h_term;
  if (h > (planetary ? 0.999f : 0.9999f) * kc.h_max) {
    grad_h_term = 0.f;
  } else {
    const float grad_W_term = common_factor * s0.w;
    grad_h_term = (grad_W_term < -0.9999f) ? 0.f : common_factor * s0.y / (1.f + grad_W_term);
  }
  const float fac_Balsara_eps = 1.f;  // a_factor_Balsara_eps at a = 1
  float balsara;
  if (planetary)
    balsara = kc.alpha * abs_div / (abs_div + curl_v + 0.0001f * fac_Balsara_eps * soundspeed / h);
  else
    balsara = kc.alpha * abs_div / (abs_div + curl_v + 0.0001f * fac_Balsara_eps * soundspeed * (1.f / h));
  d.h[p] = h;
  d.fin[p] = make_float4(s0.x, pressure, grad_h_term, soundspeed);
  // the rest of what a force neighbour needs, with P / rho^2 formed as runner_iact_force does (:281-282)
  d.hb[p] = make_float4(h, balsara, __int_as_float((int)d.time_bin[p]), pressure / (s0.x * s0.x));
  d.acc[p] = make_float4(0.f, 0.f, 0.f, 0.f);
  d.h_dt[p] = 0.f;
  d.v_sig[p] = planetary ? soundspeed : 2.f * soundspeed;
  d.gstate[p] = 3;
}

// c->hydro.h_max after the ghost (runner_ghost.c:1417-1420, 1565-1593): never
// below its value before, at least the largest h inside the cell.
__global__ void k_cell_hmax(Dev d) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= d.ncells) return;
  const CellDev cell = d.cells[c];
  float m = d.cell_hmax_old[c];
  for (int k = lane; k < cell.count; k += 32) m = fmaxf(m, d.h[cell.first + k]);
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) d.cell_hmax[c] = m;
}

// Bounding box of every cell's particles, f32 relative to the cell's loc (what the
// neighbour-loop kernels use to skip cell pairs and candidates out of reach; particles may
// have drifted out of their cell, so the cell geometry itself is not used).  One warp per
// LEAF folds its particles and pushes the box up through its ancestors with atomics on
// order-preserving integer images of the floats.  cell_box: [ncells][6] = lo xyz, hi xyz.
__global__ void k_cell_box_init(Dev d) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= d.ncells * 6) return;
  d.cell_box_u[i] = float_to_ordered((i % 6) < 3 ? INFINITY : -INFINITY);
}
// [first_lo, first_hi): only the leaves whose first particle lies in this range of cell-order slots
__global__ void k_cell_box(Dev d, int64_t first_lo = 0, int64_t first_hi = INT64_MAX) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= d.ncells || d.cell_split[c]) return;  // leaves only
  const CellDev cell = d.cells[c];
  if (cell.count == 0 || cell.first < first_lo || cell.first >= first_hi) return;
  double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int k = lane; k < cell.count; k += 32) {
    double p[3];
    load_pos(d.pos, cell.first + k, p[0], p[1], p[2]);
#pragma unroll
    for (int a = 0; a < 3; a++) { lo[a] = fmin(lo[a], p[a]); hi[a] = fmax(hi[a], p[a]); }
  }
#pragma unroll
  for (int a = 0; a < 3; a++)
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fmin(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmax(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
  if (lane >= 6) return;
  const int a = lane % 3;
  const bool is_hi = lane >= 3;
  const double v = is_hi ? (a == 0 ? hi[0] : a == 1 ? hi[1] : hi[2]) : (a == 0 ? lo[0] : a == 1 ? lo[1] : lo[2]);
  for (int q = c; q >= 0; q = d.cell_parent[q]) {
    const double rel = v - d.cells[q].loc[a];
    if (is_hi) atomicMax(&d.cell_box_u[(int64_t)q * 6 + lane], float_to_ordered(__double2float_ru(rel)));
    else atomicMin(&d.cell_box_u[(int64_t)q * 6 + lane], float_to_ordered(__double2float_rd(rel)));
  }
}
__global__ void k_cell_box_finish(Dev d) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= d.ncells * 6) return;
  d.cell_box[i] = ordered_to_float(d.cell_box_u[i]);
}

// ---------------------------------------------------------------------------
// K6: force.  Same persistent warp-per-unit walk as K3.  A pair (L in ci, R in
// cj) interacts when loop 1 of DOPAIR2 takes it (:1528-1737: r2 < hL^2 g^2
// with R inside L's sorted range) or loop 2 does (:1741-1952: r2 < hR^2 g^2
// and r2 >= hL^2 g^2 with L inside R's range).  The reference applies all
// loop-1 updates of a pair before any loop-2 update.  The candidates of a pair
// are walked ONCE: loop-1 hits go straight to the lane's FIFO, the (few)
// loop-2 hits wait in a small per-lane side list in shared memory and are
// appended when the pair is finished.  Should a side list overflow (kSide
// loop-2 hits of one lane in one pair: only with strongly varying h), the side
// lists are dropped and the pair is walked a second time for loop 2 alone.
//   tile.aux per candidate:  x = h^2 gamma^2 of the candidate (self: its cull threshold)
//                            y = candidate is cj: its "dj" of loop 2, +inf if loop 2 skips it
//                            z = candidate is ci: largest sort key of a cj particle inside its
//                                loop-1 range, -inf if loop 1 skips it
// ---------------------------------------------------------------------------
struct ForceTile {
  float4 pos[kTile];
  float4 aux[kTile];
  int idx[kTile];
  float4 side[kSide][32];
  float box[12];      // the unit's boxes (leaf frame): u0 lo/hi, uH lo/hi -- warp-uniform, kept out of the registers
};

template <int SCHEME, bool FAST = false>
__global__ void __launch_bounds__(kBlock, SWIFT_MINB_FORCE) k_force(Dev d) {
  __shared__ ForceTile tiles[kWarps];
  ForceTile &ws = tiles[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  float4 *const gq = d.gq + (size_t)(blockIdx.x * kWarps + (threadIdx.x >> 5)) * (kQD * 32) + lane;
  int *const work = d.work_ctr + 1 + kc.max_iter;
  const bool rec = d.rec_capacity > 0;
  const float qnan = __int_as_float(0x7fc00000);
  unsigned long long nhits = 0;
  const bool by_ulist = d.unit_list != nullptr && (long long)*d.unit_count * 4 < d.n_units;
  const int total = by_ulist ? *d.unit_count : (d.unit_hi > 0 ? d.unit_hi : d.n_units);
  const int wid0 = by_ulist ? 0 : d.unit_lo;

  while (true) {
    int wid = 0;
    if (lane == 0) wid = atomicAdd(work, 1) + wid0;
    wid = __shfl_sync(kFull, wid, 0);
    if (wid >= total) break;
    if (by_ulist) wid = d.unit_list[wid];
    const int slot = d.unit_leaf[wid];
    const CellDev cL = d.cells[d.leaves[slot]];
    const int ib = d.item_begin[slot], ie = d.item_begin[slot + 1];

    int p = -1;
    {
      const int q = d.lane_part[(int64_t)wid * 32 + lane];
      if (q >= 0 && d.time_bin[q] <= kc.max_active_bin) p = q;
    }
    if (!__any_sync(kFull, p >= 0)) continue;
    const bool mine = p >= 0;
    const int pp = mine ? p : cL.first;
    double xi, yi, zi;
    load_pos(d.pos, pp, xi, yi, zi);
    ForceI me;
    {
      const float4 vmi = d.vm[pp];
      const float4 fi = d.fin[pp];
      const float4 hbi = d.hb[pp];
      me.h = hbi.x;
      me.hi_inv = 1.0f / me.h;
      const float hi2 = me.hi_inv * me.hi_inv;
      me.hid_inv = hi2 * hi2;
      me.m = vmi.w; me.vx = vmi.x; me.vy = vmi.y; me.vz = vmi.z;
      me.rho = fi.x; me.P = fi.y; me.f = fi.z; me.c = fi.w;
      me.bal = hbi.y;
      me.Pr2 = hbi.w;
      me.m_inv = 1.0f / me.m;
      if (SCHEME == SWIFT_CUDA_SCHEME_SPHENIX) {
        const float4 sxi = d.sx[pp];
        me.alpha = sxi.x; me.adiff = sxi.y; me.u = sxi.z;
      }
    }
    const float hig2 = sq_gamma2(me.h);
    const float hig = __fmul_rn(me.h, kc.gamma);
    const long long idi = rec ? d.id[pp] : 0;
    ForceAcc acc;
    {
      const float4 a0 = d.acc[pp];
      acc.ax = a0.x; acc.ay = a0.y; acc.az = a0.z; acc.u_dt = a0.w;
      acc.h_dt = d.h_dt[pp];
      acc.v_sig = d.v_sig[pp];
      acc.min_bin = d.min_bin[pp];
    }
    int qcnt = 0;

    // boxes of the unit's particles in the leaf's frame: bare positions (u0), and grown by
    // each particle's own search radius (uH)
    float *const u0lo = ws.box, *const u0hi = ws.box + 3, *const uHlo = ws.box + 6, *const uHhi = ws.box + 9;
    __syncwarp();
    {
      const float l[3] = {dsub_f(xi, cL.loc[0]), dsub_f(yi, cL.loc[1]), dsub_f(zi, cL.loc[2])};
      const float ru = hig * 1.0001f;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const float a0 = warp_min(mine ? l[k] : INFINITY), a1 = warp_max(mine ? l[k] : -INFINITY);
        const float a2 = warp_min(mine ? l[k] - ru : INFINITY), a3 = warp_max(mine ? l[k] + ru : -INFINITY);
        if (lane == 0) { u0lo[k] = a0; u0hi[k] = a1; uHlo[k] = a2; uHhi[k] = a3; }
      }
    }
    __syncwarp();

    // one deferred hit -> runner_iact_nonsym_force; the three float4 of j (and its f64 position for
    // an own-cell pair) were gathered one round ahead
    auto consume = [&](const float4 e, const float4 hbj, const float4 vmj, const float4 fj, const float4 sxj,
                       double xj, double yj, double zj) {
      const int w = __float_as_int(e.w);
      float dx, dy, dz;
      bool hit = true;
      if (w & kDpFlag) {  // own cell: f64 position difference (:2349-2353, 2416-2420)
        dx = dsub_f(xi, xj); dy = dsub_f(yi, yj); dz = dsub_f(zi, zj);
        const float r2e = dist2(dx, dy, dz);
        hit = (r2e < hig2) || (r2e < sq_gamma2(hbj.x));
      } else {
        dx = e.x; dy = e.y; dz = e.z;
      }
      if (hit) {
        const float r2 = dist2(dx, dy, dz);
        ForceJ pj;
        pj.h = hbj.x; pj.bal = hbj.y; pj.bin = __float_as_int(hbj.z); pj.Pr2 = hbj.w;
        pj.m = vmj.w; pj.vx = vmj.x; pj.vy = vmj.y; pj.vz = vmj.z;
        pj.rho = fj.x; pj.P = fj.y; pj.f = fj.z; pj.c = fj.w;
        if (SCHEME == SWIFT_CUDA_SCHEME_SPHENIX) {
          pj.alpha = sxj.x; pj.adiff = sxj.y; pj.u = sxj.z;
          iact_force_sphenix(acc, r2, dx, dy, dz, me, pj);
        } else if (FAST) {
          iact_force_fast<SCHEME>(acc, r2, dx, dy, dz, me, pj);
        } else {
          iact_force<SCHEME>(acc, r2, dx, dy, dz, me, pj);
        }
        nhits++;
        if (rec) record_pair(d, 2, idi, d.id[w & (kDpFlag - 1)]);
      }
    };
    // software pipeline over the FIFO: entry r+2 is being read, the gathers of entry r+1 are in
    // flight, entry r is consumed
    auto drain_all = [&]() {
      const int m = __reduce_max_sync(kFull, qcnt);
      const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
      auto gather = [&](const float4 e, bool ok, float4 &hb, float4 &v, float4 &f, float4 &sxj, double &x, double &y,
                        double &z) {
        if (ok) {
          const int w = __float_as_int(e.w);
          const int j = w & (kDpFlag - 1);
          hb = d.hb[j];
          v = d.vm[j];
          f = d.fin[j];
          if (SCHEME == SWIFT_CUDA_SCHEME_SPHENIX) sxj = d.sx[j];
          if (w & kDpFlag) load_pos(d.pos, j, x, y, z);
        }
      };
      float4 e0 = qcnt > 0 ? __ldcs(&gq[0]) : z4;
      float4 e1 = qcnt > 1 ? __ldcs(&gq[32]) : z4;
      float4 hb0 = z4, v0 = z4, f0 = z4, s0 = z4;
      double x0 = 0.0, y0 = 0.0, z0 = 0.0;
      gather(e0, qcnt > 0, hb0, v0, f0, s0, x0, y0, z0);
      for (int r = 0; r < m; r++) {
        const float4 e2 = (r + 2 < qcnt) ? __ldcs(&gq[(r + 2) * 32]) : z4;
        float4 hb1 = z4, v1 = z4, f1 = z4, s1 = z4;
        double x1 = 0.0, y1 = 0.0, z1 = 0.0;
        gather(e1, r + 1 < qcnt, hb1, v1, f1, s1, x1, y1, z1);
        if (r < qcnt) consume(e0, hb0, v0, f0, s0, x0, y0, z0);
        e0 = e1; e1 = e2; hb0 = hb1; v0 = v1; f0 = f1; s0 = s1; x0 = x1; y0 = y1; z0 = z1;
      }
      qcnt = 0;
    };

    for (int it = ib; it < ie; it++) {
      const Item item = d.items[it];
      const CellDev cA = d.cells[item.own];
      const CellDev cB = d.cells[item.partner];
      const int kind = item.kind;
      const int sid = item.sid;
      const bool self = kind == kSelf;
      const bool left = kind == kPairLeft;  // this leaf's cell is the pair's ci
      const float hmaxA = d.cell_hmax[item.own], hmaxB = d.cell_hmax[item.partner];
      const float hmaxBg = __fmul_rn(hmaxB, kc.gamma);

      double sh[3] = {0.0, 0.0, 0.0};
      if (!self) {
#pragma unroll
        for (int k = 0; k < 3; k++) sh[k] = item_shift(d, item.shift, k);
      }
      // frame: ci particles relative to cj.loc + shift, cj particles relative to cj.loc (:1480-1482);
      // DOSELF2 (:2337-2471): relative to the cell
      double fo[3], fl[3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        if (self) { fo[k] = fl[k] = cA.loc[k]; }
        else if (left) { fo[k] = cB.loc[k]; fl[k] = __dadd_rn(cB.loc[k], sh[k]); }
        else { fo[k] = __dadd_rn(cA.loc[k], sh[k]); fl[k] = cA.loc[k]; }
      }
      // a hit needs r < max(H of the lane, H of the candidate): box of the lanes for this item
      StageBox sb;
      {
        float blo[3], bhi[3];
        const float rb = hmaxBg * 1.0001f;
        double ol[3], og[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          blo[k] = fminf(uHlo[k], u0lo[k] - rb);
          bhi[k] = fmaxf(uHhi[k], u0hi[k] + rb);
          ol[k] = cL.loc[k] - fl[k];
          og[k] = cB.loc[k] - fo[k];
        }
        stage_box(blo, bhi, ol, self ? 1e-6f * cA.width : 0.f, sb);
        if (!box_reaches_cell(sb, d.cell_box + (int64_t)item.partner * 6, og)) continue;
      }
      const float pfx = dsub_f(xi, fl[0]), pfy = dsub_f(yi, fl[1]), pfz = dsub_f(zi, fl[2]);

      // lane-side cuts of the two loops, and what bounds the walk along the sorted axis
      float cut1 = 0.f, cut2 = 0.f, bound1 = 0.f, bound2 = 0.f, thr_i = 0.f;
      int64_t segB = 0;
      double rshift = 0.0, dj_min = 0.0, hi_max_g = 0.0, hj_max_g = 0.0, rhs2 = 0.0;
      const float e_self = 4e-7f * cA.width;
      if (self) {
        thr_i = mine ? cull_thr(hig2, e_self) : -INFINITY;
      } else {
        segB = d.sort_lookup[(int64_t)item.partner * 13 + sid];
        const int64_t segA = d.sort_lookup[(int64_t)item.own * 13 + sid];
        rshift = item_rshift(sh, sid);
        const float d_me = sort_key(xi, yi, zi, sid);
        // constants of the pair, as the reference forms them (f64)
        dj_min = (double)d.sort[left ? segB : segA].d;                                            // first of cj
        const double di_max = (double)d.sort[left ? (segA + cA.count - 1) : (segB + cB.count - 1)].d;  // last of ci
        hi_max_g = __dmul_rn((double)(left ? hmaxA : hmaxB), (double)kc.gamma);
        hj_max_g = __dmul_rn((double)(left ? hmaxB : hmaxA), (double)kc.gamma);
        rhs2 = __dsub_rn(di_max, rshift);
        if (left) {
          // loop 1: I am pi; candidates R ascending while (double)dR < di
          const double di = __dsub_rn((double)__fadd_rn(d_me, hig), rshift);
          const bool ok1 = mine && (__dsub_rn(__dadd_rn((double)d_me, hi_max_g), rshift) > dj_min) && !(di < dj_min);
          cut1 = ok1 ? thr_lt(di) : -INFINITY;
          bound1 = cut1;
          // loop 2: I am a pi inside R's range: (double)d_me - rshift > dj(R)
          cut2 = mine ? thr_lt(__dsub_rn((double)d_me, rshift)) : -INFINITY;
          bound2 = -INFINITY;
          if (mine) {
            bound2 = cut2 + hmaxBg * 1.000001f + 1e-30f;
            bound2 = bound2 + fabsf(bound2) * 2e-7f;
          }
        } else {
          // loop 1: I am a pj inside L's range: (double)d_me < di(L); candidates L descending
          cut1 = mine ? d_me : INFINITY;
          float b = (float)((double)d_me + rshift) - hmaxBg * 1.000001f - 1e-30f;
          b = b - fabsf(b) * 2e-7f - fabsf((float)rshift) * 2e-7f;
          bound1 = mine ? b : INFINITY;
          // loop 2: I am pj; candidates L descending while (double)dL - rshift > dj
          const float djf = __fsub_rn(d_me, hig);
          const bool ok2 = mine && (__dsub_rn((double)d_me, hj_max_g) < rhs2) && !((double)djf > rhs2);
          cut2 = ok2 ? thr_gt_shift(rshift, (double)djf) : INFINITY;
          bound2 = cut2;
        }
      }

      // pass 0: both loops in one walk (loop-2 hits to the side list);
      // pass 1: only after a side-list overflow, loop 2 alone
      int scnt = 0;
      bool sovf = false;
      for (int pass = 0; pass < 2; pass++) {
        float wlim;  // how far along the sorted axis some lane can still hit
        if (self) wlim = 0.f;
        else if (left) wlim = warp_max(pass == 0 ? fmaxf(bound1, bound2) : bound2);
        else wlim = warp_min(pass == 0 ? fminf(bound1, bound2) : bound2);
        if (!self && (left ? (wlim == -INFINITY) : (wlim == INFINITY))) break;

        // KIND 0: own cell, 1: this leaf is the pair's ci, 2: the pair's cj.  Straight-line tests: the
        // FIFO store and its advance are predicated; only the rare loop-2 side push is a branch.
        auto test_impl = [&](auto kind_c, int st, int n) {
          constexpr int KIND = decltype(kind_c)::value;
          const float4 *__restrict__ tp = ws.pos + st;  // st is 0 or 32: no wrap inside a tile
          const float4 *__restrict__ ta = ws.aux + st;
          const int *__restrict__ ti = ws.idx + st;
          float4 *qp = gq + qcnt * 32;
          for (int k0 = 0; k0 < n; k0 += 4) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
              const float4 c = tp[k0 + u];
              const float4 a = ta[k0 + u];
              const int j = ti[k0 + u];
              const float dx = __fsub_rn(pfx, c.x), dy = __fsub_rn(pfy, c.y), dz = __fsub_rn(pfz, c.z);
              const float r2 = dist2(dx, dy, dz);
              if (KIND == 0) {
                const bool push = (r2 < thr_i || (mine && r2 < a.x)) && j != pp;
                st_pred(qp, 0.f, 0.f, 0.f, __int_as_float(j | kDpFlag), push);
                qp += push ? 32 : 0;
              } else {
                bool h1, h2;
                if (KIND == 1) {
                  h1 = c.w <= cut1 && r2 < hig2;
                  h2 = a.y <= cut2 && r2 < a.x && r2 >= hig2;
                } else {
                  h1 = cut1 <= a.z && r2 < a.x;
                  h2 = c.w >= cut2 && r2 < hig2 && r2 >= a.x;
                }
                const bool push = pass == 0 ? h1 : h2;
                st_pred(qp, dx, dy, dz, __int_as_float(j), push);
                qp += push ? 32 : 0;
                if (pass == 0 && h2) {
                  if (scnt < kSide) ws.side[scnt][lane] = make_float4(dx, dy, dz, __int_as_float(j));
                  else sovf = true;
                  scnt++;
                }
              }
            }
          }
          qcnt = (int)(qp - gq) >> 5;
        };
        auto test = [&](int st, int n) {
          if (__any_sync(kFull, qcnt > kQD - n - kSide)) drain_all();
          if (self) test_impl(std::integral_constant<int, 0>{}, st, n);
          else if (left) test_impl(std::integral_constant<int, 1>{}, st, n);
          else test_impl(std::integral_constant<int, 2>{}, st, n);
        };

        int st = 0, ntile = 0, base = 0;
        bool done = false;
        // this lane's sort entry of a staging round, read one round ahead
        auto load_se = [&](int b) {
          SortEnt se;
          se.d = 0.f;
          se.i = b + lane;  // own cell: index order
          if (!self && b + lane < cB.count) se = d.sort[segB + (left ? (b + lane) : (cB.count - 1 - b - lane))];
          return se;
        };
        SortEnt se_next = load_se(0);
        while (true) {
          if (base < cB.count) {
            const int nt = min(32, cB.count - base);
            const SortEnt se = se_next;
            if (base + 32 < cB.count) se_next = load_se(base + 32);
            const float dk = se.d;
            const int j = cB.first + se.i;
            bool keep = lane < nt;
            if (!self) {
              const float k0 = __shfl_sync(kFull, dk, 0);
              if (left ? !(k0 <= wlim) : !(k0 >= wlim)) done = true;  // sorted axis: nothing further on can hit
              keep = keep && !done && (left ? (dk <= wlim) : (dk >= wlim));
            }
            float4 c = make_float4(qnan, qnan, qnan, qnan);
            float4 a = make_float4(0.f, INFINITY, -INFINITY, 0.f);
            if (keep) {
              double xj, yj, zj;
              load_pos(d.pos, j, xj, yj, zj);
              c = make_float4(dsub_f(xj, fo[0]), dsub_f(yj, fo[1]), dsub_f(zj, fo[2]), dk);
              keep = c.x >= sb.lo[0] && c.x <= sb.hi[0] && c.y >= sb.lo[1] && c.y <= sb.hi[1] && c.z >= sb.lo[2] &&
                     c.z <= sb.hi[2];
            }
            if (keep) {
              const float hj = d.hb[j].x;
              const float hjg = __fmul_rn(hj, kc.gamma);
              a.x = sq_gamma2(hj);
              if (self) {
                a.x = cull_thr(a.x, e_self);
              } else if (left) {
                // candidate is R in cj: loop 2 outer conditions (:1741-1760)
                const float djf = __fsub_rn(dk, hjg);
                const bool ok = (__dsub_rn((double)dk, hj_max_g) < rhs2) && !((double)djf > rhs2);
                a.y = ok ? djf : INFINITY;
              } else {
                // candidate is L in ci: loop 1 outer conditions (:1528-1545)
                const double di = __dsub_rn((double)__fadd_rn(dk, hjg), rshift);
                const bool ok = (__dsub_rn(__dadd_rn((double)dk, hi_max_g), rshift) > dj_min) && !(di < dj_min);
                a.z = ok ? thr_lt(di) : -INFINITY;
              }
            }
            const unsigned b = __ballot_sync(kFull, keep);
            if (keep) {
              const int s = (st + ntile + __popc(b & lt_mask)) & (kTile - 1);
              ws.pos[s] = c;
              ws.aux[s] = a;
              ws.idx[s] = j;
            }
            ntile += __popc(b);
            base += 32;
          } else {
            done = true;
          }
          int n = 0;
          if (ntile >= 32) {
            n = 32;
          } else if (done && ntile > 0) {
            n = (ntile + 3) & ~3;
            if (lane < n - ntile) ws.pos[(st + ntile + lane) & (kTile - 1)] = make_float4(qnan, qnan, qnan, qnan);
          }
          if (n > 0) {
            __syncwarp();
            test(st, n);
            st = (st + 32) & (kTile - 1);
            ntile = max(ntile - 32, 0);
          }
          if (done && ntile == 0) break;
        }
        __syncwarp();
        if (self) break;
        if (pass == 0) {
          if (!__any_sync(kFull, sovf)) {
            // the pair's loop-2 hits follow its loop-1 hits
            for (int s = 0; s < scnt; s++) {
              gq[qcnt * 32] = ws.side[s][lane];
              qcnt++;
            }
            break;
          }
          scnt = 0;
        }
      }
    }
    drain_all();
    if (mine) {
      d.acc[p] = make_float4(acc.ax, acc.ay, acc.az, acc.u_dt);
      // hydro_end_force (Minimal/hydro.h:878-882), the whole of runner_do_end_hydro_force for these schemes
      d.h_dt[p] = d.no_end_force ? acc.h_dt : acc.h_dt * (me.h * 0.3333333333f);
      d.v_sig[p] = acc.v_sig;
      d.min_bin[p] = acc.min_bin;
    }
  }
  for (int o = 16; o > 0; o >>= 1) nhits += __shfl_down_sync(kFull, nhits, o);
  if (lane == 0 && nhits) atomicAdd(&d.hits[2], nhits);
}

// Are the flattened work lists still the reference's?  A sub-task recursed through a cell at the build
// because gamma * h_max + dx_max_part < dmin / 2 held (src/cell.h:761-782); the reference re-evaluates
// that every step with the current values (src/cell_unskip.c:808-815) and stops recursing when it fails,
// which changes the level -- and so the frame and the reach -- at which pairs are evaluated.
__global__ void k_check_lists(Dev d, float dx_max) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.ncells) return;
  const float lim = d.recurse_limit[c];
  if (!(kc.gamma * d.cell_hmax_old[c] + dx_max < lim)) atomicAdd(d.stale, 1);
}

// runner_do_end_hydro_force on its own (swift_cuda_step_parts without the force loop in the same call)
__global__ void k_end_force(Dev d) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n || d.time_bin[p] > kc.max_active_bin) return;
  d.h_dt[p] = d.h_dt[p] * (d.h[p] * 0.3333333333f);
}

}  // namespace swiftcuda
