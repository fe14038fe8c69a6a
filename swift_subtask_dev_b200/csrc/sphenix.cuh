// SPHENIX: the per-particle pieces the three-loop scheme adds around the neighbour
// loops.  The loops themselves are instantiations of the walks in kernels.cuh: the gradient loop is
// k_density<0, false, 1> (runner_iact_nonsym_gradient as its body), the force loop
// k_force<SWIFT_CUDA_SCHEME_SPHENIX> (the scheme's runner_iact_nonsym_force).
//
//   runner_do_ghost, converged particle under EXTRA_HYDRO_LOOP   src/runner_ghost.c:1249-1264, 1429-1443
//     hydro_end_density (div_v += 3 H)                           src/hydro/SPHENIX/hydro.h:599-641
//     hydro_prepare_gradient, hydro_reset_gradient               src/hydro/SPHENIX/hydro.h:654-733
//   runner_do_extra_ghost                                        src/runner_ghost.c:992-1075
//     hydro_end_gradient, hydro_prepare_force, hydro_reset_acceleration   src/hydro/SPHENIX/hydro.h:745-955
//
// Arithmetic in the reference's operand order, unfused (-fmad=false), IEEE division and square root: every
// field is bit-identical to the single-threaded reference (tests/test_gpu_sphenix.py).  The cosmological
// factors of struct cosmology (doubles) are all 1 here (swift_cuda_create refuses a != 1): a product of two
// floats formed in double and rounded once to float is the float product.
#pragma once
#include "kernels.cuh"
#include "pack.cuh"

namespace swiftcuda {

// s0 = rho, rho_dh, wcount, wcount_dh and s1 = div_v, rot_v after hydro_end_density (Minimal's form, which
// is SPHENIX's up to the Hubble term of div_v), h the converged smoothing length.
__device__ void sphenix_converged(const Dev &d, int64_t p, float h, float4 s0, float4 s1) {
  const float div_v = s1.x + kc.H * 3.f;  // p->viscosity.div_v += cosmo->H * hydro_dimension
  const float fac_B = 1.f;                // cosmo->a_factor_Balsara_eps at a = 1
  const float curl_v = sqrtf(s1.y * s1.y + s1.z * s1.z + s1.w * s1.w);
  const float abs_div_v = fabsf(div_v);
  const float u = d.u[p];
  float pressure, soundspeed;
  eos_eval(s0.x, u, 0, pressure, soundspeed);  // ideal gas: gas_pressure_from_internal_energy, gas_soundspeed_from_pressure
  const float balsara = abs_div_v / (abs_div_v + curl_v + 0.0001f * soundspeed * fac_B / h);
  const float common_factor = h * 0.3333333333f / s0.z;
  float grad_h_term;
  if (h > 0.9999f * kc.h_max) {
    grad_h_term = 0.f;
  } else {
    const float grad_W_term = common_factor * s0.w;
    grad_h_term = (grad_W_term < -0.9999f) ? 0.f : common_factor * s0.y / (1.f + grad_W_term);
  }
  const float alpha = d.sph_state[p].x;
  d.h[p] = h;
  d.fin[p] = make_float4(s0.x, pressure, grad_h_term, soundspeed);
  d.hb[p] = make_float4(h, balsara, __int_as_float((int)d.time_bin[p]), pressure / (s0.x * s0.x));
  // hydro_reset_gradient: v_sig = 2 c, alpha_visc_max_ngb = alpha; laplace_u is 0 since hydro_init_part
  d.v_sig[p] = 2.f * soundspeed;
  d.sph_grad[p] = make_float4(div_v, 0.f, alpha, 0.f);
  d.gin[p] = make_float4(u, s0.x, soundspeed, alpha);
  d.gstate[p] = 3;
}

// Start of a step: the neighbour records of the INACTIVE particles (their rho, c, alphas persist from their
// last active step; active particles get theirs from the ghost and the extra ghost).
// mode (swift_cuda_step_parts with phases in separate calls): bit 0 -- the ghost does not run in this call, the
// ACTIVE particles' gradient-loop record comes from the uploaded state too; bit 1 -- the same for the force-loop
// record when the extra ghost does not run.
__global__ void k_sphenix_prep(Dev d, int mode) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n) return;
  const bool active = d.time_bin[p] <= kc.max_active_bin;
  if (active && mode == 0) return;
  const float4 st = d.sph_state[p];
  const float4 f = d.fin[p];
  const float u = d.u[p];
  if (!active || (mode & 1)) d.gin[p] = make_float4(u, f.x, f.w, st.x);
  if (!active || (mode & 2)) d.sx[p] = make_float4(st.x, st.y, u, 0.f);
}

// runner_do_extra_ghost, one thread per particle
__global__ void k_extra_ghost(Dev d) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n) return;
  const int bin = d.time_bin[p];
  if (bin > kc.max_active_bin) return;
  const float h = d.h[p];
  const float u = d.u[p];
  const float rho = d.fin[p].x;
  float4 st = d.sph_state[p];   // alpha_visc, alpha_diff, div_v_prev, div_v_dt
  float4 sg = d.sph_grad[p];    // div_v, laplace_u (raw), alpha_visc_max_ngb
  // hydro_end_gradient
  {
    const float h_inv = 1.0f / h;
    const float h_inv_dim = h_inv * h_inv * h_inv;
    const float h_inv_dim_plus_one = h_inv_dim * h_inv;
    sg.y *= 2.f * h_inv_dim_plus_one;
  }
  // dt_alpha = get_timestep(p->time_bin, time_base) (src/timeline.h), a double handed over as a float
  const double dt_d = __dmul_rn(bin <= 0 ? 0.0 : (double)(1LL << (bin + 1)), kc.time_base);
  const float dt_alpha = __double2float_rn(dt_d);
  // hydro_prepare_force
  const float kernel_support_physical = h * kc.gamma;
  const float kernel_support_physical_inv = 1.f / kernel_support_physical;
  const float v_sig_physical = d.v_sig[p];
  float pressure, soundspeed_physical;
  eos_eval(rho, u, 0, pressure, soundspeed_physical);
  const float sound_crossing_time_inverse = soundspeed_physical * kernel_support_physical_inv;
  const float div_v = sg.x;
  const float div_v_dt = dt_alpha == 0.f ? 0.f : (div_v - st.z) / dt_alpha;
  const float mdt = -1.f * div_v_dt;
  const float S = div_v < 0.f ? kernel_support_physical * kernel_support_physical * (0.f > mdt ? 0.f : mdt) : 0.f;
  const float soundspeed_square = soundspeed_physical * soundspeed_physical;
  const float alpha_loc = kc.visc_alpha_max * S / (soundspeed_square + S);
  float alpha = st.x;
  if (alpha_loc > alpha) {
    alpha = alpha_loc;
  } else {
    const float timescale_ratio = dt_alpha * sound_crossing_time_inverse * kc.visc_length;
    alpha += alpha_loc * timescale_ratio;
    alpha /= (1.f + timescale_ratio);
  }
  alpha = alpha > kc.visc_alpha_min ? alpha : kc.visc_alpha_min;
  const float diffusion_timescale_physical_inverse = v_sig_physical * kernel_support_physical_inv;
  const float sqrt_u_inv = 1.f / sqrtf(u);
  float alpha_diff_dt = kc.diff_beta * kernel_support_physical * sg.y * sqrt_u_inv;
  alpha_diff_dt -= (st.y - kc.diff_alpha_min) * diffusion_timescale_physical_inverse;
  float new_diffusion_alpha = st.y;
  new_diffusion_alpha += alpha_diff_dt * dt_alpha;
  new_diffusion_alpha = new_diffusion_alpha > kc.diff_alpha_min ? new_diffusion_alpha : kc.diff_alpha_min;
  const float viscous_diffusion_limit = kc.diff_alpha_max * (1.f - sg.z / kc.visc_alpha_max);
  new_diffusion_alpha = new_diffusion_alpha < viscous_diffusion_limit ? new_diffusion_alpha : viscous_diffusion_limit;
  d.sph_state[p] = make_float4(alpha, new_diffusion_alpha, div_v, div_v_dt);
  d.sph_grad[p] = sg;
  d.sx[p] = make_float4(alpha, new_diffusion_alpha, u, 0.f);
  // hydro_reset_acceleration
  d.acc[p] = make_float4(0.f, 0.f, 0.f, 0.f);
  d.h_dt[p] = 0.f;
}

// caller order -> cell order (order[k] = caller index of the particle at slot k; null: the same)
__global__ void k_sphenix_set_state(Dev d, const int32_t *order, const float *alpha_visc, const float *alpha_diff,
                                    const float *div_v_prev) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= d.n) return;
  const int64_t s = order ? order[k] : k;
  float4 st = d.sph_state[k];
  if (alpha_visc) st.x = alpha_visc[s];
  if (alpha_diff) st.y = alpha_diff[s];
  if (div_v_prev) st.z = div_v_prev[s];
  d.sph_state[k] = st;
}

// cell order -> caller order, seven floats per particle: div_v, laplace_u, alpha_visc_max_ngb, alpha_visc,
// alpha_diff, div_v_previous_step, div_v_dt
__global__ void k_sphenix_unpack(Dev d, const int32_t *order, float *out) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= d.n) return;
  const int64_t s = order ? order[k] : k;
  const float4 st = d.sph_state[k];
  const float4 sg = d.sph_grad[k];
  out[s] = sg.x;
  out[d.n + s] = sg.y;
  out[2 * d.n + s] = sg.z;
  out[3 * d.n + s] = st.x;
  out[4 * d.n + s] = st.y;
  out[5 * d.n + s] = st.z;
  out[6 * d.n + s] = st.w;
}

// struct part write-back of a SPHENIX record (swift_cuda_step_parts): exactly the fields the reference's
// operators of the phases that ran write, for active particles.  ran: SWIFT_CUDA_PHASE_* bits.
__global__ void k_unpack_aos_sphenix(Dev d, unsigned char *aos, swift_cuda_part_layout L, const int32_t *order, int ran,
                                     int max_active_bin, int64_t lo, int64_t hi) {
  const int64_t k = lo + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= hi) return;
  if (d.time_bin[k] > max_active_bin) return;
  const int64_t s = order ? order[k] : k;
  unsigned char *r = aos + (size_t)s * L.stride;
  const float4 sg = d.sph_grad[k];
  if (ran & SWIFT_CUDA_PHASE_GHOST) {
    const float4 f = d.fin[k];
    aos_put<float>(r, L.h, d.h[k]);
    aos_put<float>(r, L.rho, f.x);
    aos_put<float>(r, L.pressure, f.y);
    aos_put<float>(r, L.f, f.z);
    aos_put<float>(r, L.soundspeed, f.w);
    aos_put<float>(r, L.balsara, d.hb[k].y);
    aos_put<float>(r, L.div_v, sg.x);
  }
  if (ran & (SWIFT_CUDA_PHASE_GHOST | SWIFT_CUDA_PHASE_GRADIENT)) {
    aos_put<float>(r, L.v_sig, d.v_sig[k]);
    aos_put<float>(r, L.alpha_visc_max_ngb, sg.z);
  }
  if (ran & (SWIFT_CUDA_PHASE_GHOST | SWIFT_CUDA_PHASE_GRADIENT | SWIFT_CUDA_PHASE_EXTRA_GHOST))
    aos_put<float>(r, L.laplace_u, sg.y);
  if (ran & SWIFT_CUDA_PHASE_EXTRA_GHOST) {
    const float4 st = d.sph_state[k];
    aos_put<float>(r, L.alpha_visc, st.x);
    aos_put<float>(r, L.alpha_diff, st.y);
    aos_put<float>(r, L.div_v_previous_step, st.z);
    aos_put<float>(r, L.div_v_dt, st.w);
  }
  if (ran & (SWIFT_CUDA_PHASE_EXTRA_GHOST | SWIFT_CUDA_PHASE_FORCE | SWIFT_CUDA_PHASE_END_FORCE)) {
    const float4 a = d.acc[k];
    aos_put<float>(r, L.a_hydro, a.x);
    aos_put<float>(r, L.a_hydro + 4, a.y);
    aos_put<float>(r, L.a_hydro + 8, a.z);
    aos_put<float>(r, L.u_dt, a.w);
    aos_put<float>(r, L.h_dt, d.h_dt[k]);
  }
  if ((ran & SWIFT_CUDA_PHASE_FORCE) && L.min_ngb_time_bin >= 0)
    aos_put<int8_t>(r, L.min_ngb_time_bin, (int8_t)d.min_bin[k]);
}

}  // namespace swiftcuda
