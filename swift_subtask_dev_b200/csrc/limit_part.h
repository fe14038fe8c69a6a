// timestep_limit_part (src/timestep_limiter.h:67-217) on one particle, written on scalars in plain C++ so that
// the SAME source is the body of k_limiter on the device (integrate.cuh; nvcc -fmad=false: no contraction) and,
// compiled by g++ with -ffp-contract=off, the subject of a host-side test against the reference's fixture
// (tests/test_limiter_oracle.py::test_kernel_body_on_the_host_*).  The host build exists in tests/ only: the
// library never runs it.
#pragma once

#ifdef __CUDACC__
#define SWIFT_LP_HD __host__ __device__ __forceinline__
#else
#define SWIFT_LP_HD inline
#endif

namespace swiftcuda {

constexpr int kLpNotAwake = -56;  // time_bin_not_awake = -num_time_bins, src/timeline.h:36, 48

struct LimitEnv {
  long long ti_current;  // e->ti_current
  double time_base;      // e->time_base
  int max_active_bin;    // e->max_active_bin
  float energy_min;      // hydro_kick_extra: max(minimal_internal_energy, entropy floor (none))
};

// src/timeline.h:59-63, 107-114, 126-138
SWIFT_LP_HD long long lp_timestep(int bin) { return bin <= 0 ? 0 : 1LL << (bin + 1); }
SWIFT_LP_HD long long lp_time_begin(long long ti_current, int bin) {
  const long long dti = lp_timestep(bin);
  return dti == 0 ? 0 : dti * ((ti_current - 1) / dti);
}
SWIFT_LP_HD long long lp_time_end(long long ti_current, int bin) {
  const long long dti = lp_timestep(bin);
  if (dti == 0) return 0;
  const long long mod = ti_current % dti;
  return mod == 0 ? ti_current : ti_current - mod + dti;
}

// kick_part (src/kick.h:244-278) + hydro_kick_extra (Minimal/hydro.h:901-930) over dt: v_full += a * dt with the
// product and the sum in double (float += float * double); u_full with dt as a float; u_dt = 0 at the floor.
SWIFT_LP_HD void lp_kick(float vf[4], const float a[3], float &u_dt, double dt, float energy_min, bool &dirty) {
  vf[0] = (float)((double)vf[0] + (double)a[0] * dt);
  vf[1] = (float)((double)vf[1] + (double)a[1] * dt);
  vf[2] = (float)((double)vf[2] + (double)a[2] * dt);
  const float dt_therm = (float)dt;
  const float delta_u = u_dt * dt_therm;
  const float ua = vf[3] + delta_u, ub = 0.5f * vf[3];
  vf[3] = ua > ub ? ua : ub;
  if (vf[3] < energy_min) {
    vf[3] = energy_min;
    u_dt = 0.f;
    dirty = true;
  }
}

// wake: p->limiter_data.wakeup (!= kLpNotAwake).  bin: p->time_bin in, the new bin out.  vf: xp->v_full[3],
// xp->u_full.  a: p->a_hydro.  u_dt: p->u_dt (dirty when it was zeroed).  Returns the particle's new end of step.
SWIFT_LP_HD long long limit_part(const LimitEnv &e, int wake, int &bin, float vf[4], const float a[3], float &u_dt,
                                 bool &dirty) {
  const int old_bin = bin;
  const int new_bin = (signed char)(-wake + 2);
  bin = new_bin;
  if (old_bin <= e.max_active_bin) {
    // active: only the length of the step that starts now changes (:83-95)
    return e.ti_current + lp_timestep(new_bin);
  }
  // inactive: in the middle of a step of its old bin (:97-215)
  const long long ti_beg_old = lp_time_begin(e.ti_current, old_bin);
  const long long ti_end_old = lp_time_end(e.ti_current, old_bin);
  const long long dti_old = ti_end_old - ti_beg_old;
  const long long dti_new = lp_timestep(new_bin);
  // the reference: k = 0; while (ti_beg_old + k * dti_new <= ti_current) k++; ti_beg_new = ti_beg_old + (k - 1) * dti_new
  const long long ti_beg_new = ti_beg_old + ((e.ti_current - ti_beg_old) / dti_new) * dti_new;
  const long long ti_end_half_old = ti_beg_old + dti_old / 2;
  const long long ti_end_half_new = ti_beg_new + dti_new / 2;
  // undo the kick1 of the old step (the minus sign is applied to the double), then kick to the new beginning
  lp_kick(vf, a, u_dt, -((double)(ti_end_half_old - ti_beg_old) * e.time_base), e.energy_min, dirty);
  lp_kick(vf, a, u_dt, (double)(ti_beg_new - ti_beg_old) * e.time_base, e.energy_min, dirty);
  if (new_bin > e.max_active_bin) {
    // the new step is under way: its kick1 is due here, the kick task will not see the particle
    lp_kick(vf, a, u_dt, (double)(ti_end_half_new - ti_beg_new) * e.time_base, e.energy_min, dirty);
    return ti_beg_new + dti_new;
  }
  return e.ti_current + lp_timestep(new_bin);
}

}  // namespace swiftcuda
