// Time integration either side of the neighbour loops: drift, kick and time-step
// as per-particle kernels on the SoA state the loops already keep in HBM, so that the particles stay
// resident on the device across steps.  Restates, per particle,
//   drift_part                      src/drift.h:143-233   (+ hydro_predict_extra, Minimal/hydro.h:815-866,
//                                                           Planetary/hydro.h:823-870; tail of the loop of
//                                                           cell_drift_part, src/cell_drift.c:303-318)
//   kick_part                       src/kick.h:215-279    (+ hydro_kick_extra, Minimal/hydro.h:901-930,
//                                                           Planetary/hydro.h:904-925) as driven by
//   runner_do_kick1 / runner_do_kick2   src/runner_time_integration.c:141-180, 412-457
//                                   (kick2 ends with hydro_reset_predicted_values, Minimal/hydro.h:771-796)
//   runner_do_timestep              src/runner_time_integration.c:696-745, 790-800 with get_part_timestep and
//                                   make_integer_timestep (src/timestep.h:45-82, 141-222)
//   runner_do_limiter               src/runner_time_integration.c:1324-1467 with timestep_limit_part
//                                   (src/timestep_limiter.h:67-217): k_limiter, after the limiter LOOP
//                                   (k_density<0, false, 2>, kernels.cuh) has filled Dev::wakeup
// without cosmology (a = 1), gravity or cooling.  All of them are HBM-bound streaming
// kernels (one thread per particle, 16-byte vector accesses); arithmetic is in the reference's types
// and operand order, products with the double time steps in f64, so the results are bit-identical.
#pragma once
#include "kernels.cuh"
#include "limit_part.h"

namespace swiftcuda {

constexpr int kMaxDeltaBin = 2;  // time_bin_neighbour_max_delta_bin, src/timeline.h:51
constexpr long long kMaxNrTimesteps = 1LL << (kNumTimeBins + 1);

struct IntegPrm {
  float cfl, log_max_h_change, min_u;
  double dt_min, dt_max, time_base, time_base_inv;
  long long ti_current;
};

struct IntegDev {
  float4 *vu_full;        // xp->v_full[3], xp->u_full
  float4 *xdiff;          // xp->x_diff[3]
  float4 *xdiff_sort;     // xp->x_diff_sort[3]
  unsigned *dx2_max;      // [2] bit images of the (non-negative) maxima of |x_diff|^2, |x_diff_sort|^2
  unsigned long long *ti_end_min;
  unsigned long long *updated;
  int *too_small;         // particles that want a step below dt_min
};

// src/timeline.h:59-63, 78-83, 107-114, 126-138
__device__ __forceinline__ long long int_timestep(int bin) { return bin <= 0 ? 0 : 1LL << (bin + 1); }
__device__ __forceinline__ int time_bin_of(long long dti) { return 62 - __clzll(dti); }
__device__ __forceinline__ long long int_time_begin(long long ti_current, int bin) {
  const long long dti = int_timestep(bin);
  return dti == 0 ? 0 : dti * ((ti_current - 1) / dti);
}
__device__ __forceinline__ long long int_time_end(long long ti_current, int bin) {
  const long long dti = int_timestep(bin);
  if (dti == 0) return 0;
  const long long mod = ti_current % dti;
  return mod == 0 ? ti_current : ti_current - mod + dti;
}

// src/approx_math.h:35-37
__device__ __forceinline__ float approx_expf(float x) {
  const float c = __fadd_rn(1.f / 6.f, __fmul_rn(1.f / 24.f, x));
  const float b = __fadd_rn(0.5f, __fmul_rn(x, c));
  const float a = __fadd_rn(1.f, __fmul_rn(x, b));
  return __fadd_rn(1.f, __fmul_rn(x, a));
}

// v += a * dt with a float accumulator and a double dt: (float)((double)v + (double)a * dt)
__device__ __forceinline__ float kick_f(float v, float a, double dt) {
  return (float)__dadd_rn((double)v, __dmul_rn((double)a, dt));
}

// hydro_first_init_part (Minimal/hydro.h:983-995): v_full = v, u_full = u; offsets since the last rebuild = 0
__global__ void k_integ_init(Dev d, IntegDev g) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n) return;
  const float4 vm = d.vm[p];
  g.vu_full[p] = make_float4(vm.x, vm.y, vm.z, d.u[p]);
  g.xdiff[p] = make_float4(0.f, 0.f, 0.f, 0.f);
  g.xdiff_sort[p] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// Every particle from ti_old to ti_current (engine_policy_drift_all).  dt = (ti_current - ti_old) *
// time_base is at once dt_drift, dt_kick_hydro and dt_therm (cell_drift.c:232-236, no cosmology).
__global__ void __launch_bounds__(256, 4) k_drift(Dev d, IntegDev g, IntegPrm ip, double dt) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  float dx2 = 0.f, dx2s = 0.f;
  if (p < d.n) {
    const bool planetary = kc.scheme == SWIFT_CUDA_SCHEME_PLANETARY;
    const float4 vf = g.vu_full[p];
    PosD *pos = const_cast<PosD *>(d.pos);
    PosD q = pos[p];
    q.x = __dadd_rn(q.x, __dmul_rn((double)vf.x, dt));
    q.y = __dadd_rn(q.y, __dmul_rn((double)vf.y, dt));
    q.z = __dadd_rn(q.z, __dmul_rn((double)vf.z, dt));
    pos[p] = q;
    const float4 acc = d.acc[p];  // a_hydro, u_dt
    float4 vm = d.vm[p];
    vm.x = kick_f(vm.x, acc.x, dt);
    vm.y = kick_f(vm.y, acc.y, dt);
    vm.z = kick_f(vm.z, acc.z, dt);
    d.vm[p] = vm;
    // hydro_predict_extra
    const float dtf = (float)dt;
    float *up = const_cast<float *>(d.u);
    float u = __fadd_rn(up[p], __fmul_rn(acc.w, dtf));
    if (planetary) u = u > ip.min_u ? u : ip.min_u;
    float h = d.h[p];
    const float h_inv = 1.f / h;
    const float w1 = __fmul_rn(__fmul_rn(d.h_dt[p], h_inv), dtf);
    h = __fmul_rn(h, fabsf(w1) < 0.2f ? approx_expf(w1) : expf_glibc(w1));
    float4 fin = d.fin[p];  // rho, P, f, c
    const float w2 = __fmul_rn(-3.f, w1);
    fin.x = __fmul_rn(fin.x, fabsf(w2) < 0.2f ? approx_expf(w2) : expf_glibc(w2));
    if (!planetary) {
      u = u > 0.f ? u : 0.f;  // entropy floor "none": floor_u = 0
      u = u > ip.min_u ? u : ip.min_u;
    }
    up[p] = u;
    float P, cs;
    eos_eval(fin.x, u, planetary ? d.mat_id[p] : 0, P, cs);
    fin.y = P;
    fin.w = cs;
    d.fin[p] = fin;
    const float two_cs = __fmul_rn(2.f, cs);
    const float vs = d.v_sig[p];
    d.v_sig[p] = vs > two_cs ? vs : two_cs;
    // offsets since the last rebuild / sort
    const float ox = (float)__dmul_rn((double)vf.x, dt), oy = (float)__dmul_rn((double)vf.y, dt),
                oz = (float)__dmul_rn((double)vf.z, dt);
    float4 xd = g.xdiff[p], xs = g.xdiff_sort[p];
    xd.x = __fsub_rn(xd.x, ox); xd.y = __fsub_rn(xd.y, oy); xd.z = __fsub_rn(xd.z, oz);
    xs.x = __fsub_rn(xs.x, ox); xs.y = __fsub_rn(xs.y, oy); xs.z = __fsub_rn(xs.z, oz);
    g.xdiff[p] = xd;
    g.xdiff_sort[p] = xs;
    // cell_drift.c:303-304
    h = h < kc.h_max ? h : kc.h_max;
    h = h > kc.h_min ? h : kc.h_min;
    d.h[p] = h;
    float4 hb = d.hb[p];  // the force neighbour record follows: h, balsara, time_bin, P / rho^2
    hb.x = h;
    hb.w = P / (fin.x * fin.x);
    d.hb[p] = hb;
    dx2 = __fadd_rn(__fadd_rn(__fmul_rn(xd.x, xd.x), __fmul_rn(xd.y, xd.y)), __fmul_rn(xd.z, xd.z));
    dx2s = __fadd_rn(__fadd_rn(__fmul_rn(xs.x, xs.x), __fmul_rn(xs.y, xs.y)), __fmul_rn(xs.z, xs.z));
  }
  // max over the block, then one atomic each (non-negative floats order like their bit images)
  for (int o = 16; o > 0; o >>= 1) {
    dx2 = fmaxf(dx2, __shfl_xor_sync(0xffffffffu, dx2, o));
    dx2s = fmaxf(dx2s, __shfl_xor_sync(0xffffffffu, dx2s, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (dx2 > 0.f) atomicMax(&g.dx2_max[0], __float_as_uint(dx2));
    if (dx2s > 0.f) atomicMax(&g.dx2_max[1], __float_as_uint(dx2s));
  }
}

// which = 1: first half-kick of the step that starts now (runner_do_kick1, starting particles);
// which = 2: second half-kick of the step that ends now (runner_do_kick2, active particles), then
// hydro_reset_predicted_values.
__global__ void k_kick(Dev d, IntegDev g, IntegPrm ip, int which) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n) return;
  const int bin = d.time_bin[p];
  if (bin > kc.max_active_bin) return;
  const bool planetary = kc.scheme == SWIFT_CUDA_SCHEME_PLANETARY;
  const long long ti_step = int_timestep(bin);
  const long long ti_begin =
      which == 1 ? int_time_begin(ip.ti_current + 1, bin) : int_time_begin(ip.ti_current, bin) + ti_step / 2;
  const long long ti_end = ti_begin + ti_step / 2;
  const double dt = __dmul_rn((double)(ti_end - ti_begin), ip.time_base);
  float4 acc = d.acc[p];
  float4 vf = g.vu_full[p];
  vf.x = kick_f(vf.x, acc.x, dt);
  vf.y = kick_f(vf.y, acc.y, dt);
  vf.z = kick_f(vf.z, acc.z, dt);
  // hydro_kick_extra
  const float delta_u = __fmul_rn(acc.w, (float)dt);
  const float ua = __fadd_rn(vf.w, delta_u), ub = __fmul_rn(0.5f, vf.w);
  vf.w = ua > ub ? ua : ub;
  const float energy_min = planetary ? ip.min_u : (ip.min_u > 0.f ? ip.min_u : 0.f);
  if (vf.w < energy_min) {
    vf.w = energy_min;
    acc.w = 0.f;
    d.acc[p] = acc;
  }
  g.vu_full[p] = vf;
  if (which == 2) {
    float4 vm = d.vm[p];
    vm.x = vf.x; vm.y = vf.y; vm.z = vf.z;
    d.vm[p] = vm;
    const_cast<float *>(d.u)[p] = vf.w;
    float4 fin = d.fin[p];
    float P, cs;
    eos_eval(fin.x, vf.w, planetary ? d.mat_id[p] : 0, P, cs);
    fin.y = P;
    fin.w = cs;
    d.fin[p] = fin;
    float4 hb = d.hb[p];
    hb.w = P / (fin.x * fin.x);
    d.hb[p] = hb;
    const float two_cs = __fmul_rn(2.f, cs);
    const float vs = d.v_sig[p];
    d.v_sig[p] = vs > two_cs ? vs : two_cs;
  }
}

// New time bin of every active particle; the next sync point of the run.
__global__ void k_timestep(Dev d, IntegDev g, IntegPrm ip) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  long long t_end = kMaxNrTimesteps;
  int upd = 0;
  if (p < d.n) {
    const int old_bin = d.time_bin[p];
    if (old_bin > kc.max_active_bin) {
      t_end = int_time_end(ip.ti_current, old_bin);
    } else {
      const float h = d.h[p];
      // hydro_compute_timestep: cosmo->a and a_factor_sound_speed are doubles (= 1)
      const float front = __fmul_rn(__fmul_rn(2.f, kc.gamma), ip.cfl);
      const float dt_cfl = (float)__ddiv_rn(__dmul_rn((double)front, (double)h), (double)d.v_sig[p]);
      const float h_dt = d.h_dt[p];
      const float dt_h_change = (h_dt != 0.0f) ? fabsf(__fmul_rn(ip.log_max_h_change, h) / h_dt) : 3.402823466e+38f;
      float new_dt = dt_cfl < dt_h_change ? dt_cfl : dt_h_change;
      new_dt = (double)new_dt < ip.dt_max ? new_dt : (float)ip.dt_max;
      if ((double)new_dt < ip.dt_min) atomicAdd(g.too_small, 1);
      // make_integer_timestep
      long long new_dti = (long long)__dmul_rn((double)new_dt, ip.time_base_inv);
      int new_bin = (signed char)time_bin_of(new_dti);
      const int ngb = d.min_bin[p] + kMaxDeltaBin;
      new_bin = (signed char)(new_bin < ngb ? new_bin : ngb);
      new_dti = int_timestep(new_bin);
      const long long current_dti = int_timestep(old_bin);
      const long long ti_end = int_time_end(ip.ti_current, old_bin);
      if (old_bin > 0) new_dti = new_dti < 2 * current_dti ? new_dti : 2 * current_dti;
      long long dti_timeline = kMaxNrTimesteps;
      while (new_dti < dti_timeline) dti_timeline /= 2LL;
      new_dti = dti_timeline;
      if (new_dti > current_dti) {
        if ((kMaxNrTimesteps - ti_end) % new_dti > 0) new_dti = current_dti;
      }
      const int bin = time_bin_of(new_dti);
      const_cast<int8_t *>(d.time_bin)[p] = (int8_t)bin;
      float4 hb = d.hb[p];
      hb.z = __int_as_float(bin);
      d.hb[p] = hb;
      upd = 1;
      t_end = ip.ti_current + new_dti;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const long long other = __shfl_xor_sync(0xffffffffu, t_end, o);
    t_end = other < t_end ? other : t_end;
  }
  const unsigned b = __ballot_sync(0xffffffffu, upd);
  if ((threadIdx.x & 31) == 0) {
    // most warps cannot lower the minimum any more: look before the atomic (one address for 5e5 warps)
    if ((unsigned long long)t_end < *(volatile unsigned long long *)g.ti_end_min)
      atomicMin(g.ti_end_min, (unsigned long long)t_end);
    if (b) atomicAdd(g.updated, (unsigned long long)__popc(b));
  }
}

// runner_do_limiter on every particle the limiter loop woke (Dev::wakeup != time_bin_not_awake):
// timestep_limit_part, src/timestep_limiter.h:67-217, whose arithmetic is limit_part() of limit_part.h -- plain
// C++ on scalars, so that the same source also compiles for the host, where tests/test_limiter_oracle.py holds
// it against the reference's fixture.  An ACTIVE particle only takes the bin -wakeup + 2 for the step that
// starts now.  An INACTIVE one is in the middle of a step of its old bin: its kick1 is undone (negative dt), it
// is kicked from the old beginning to the beginning of the new, shorter step that contains ti_current, and --
// when that bin does not start now, so that the kick task will not see the particle -- given the first
// half-kick of the new step on the spot.  One thread per particle, HBM-bound like k_kick.
__global__ void k_limiter(Dev d, IntegDev g, IntegPrm ip) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  long long t_end = kMaxNrTimesteps;
  int lim = 0;
  if (p < d.n) {
    const int wake = d.wakeup[p];
    if (wake != kLpNotAwake) {
      lim = 1;
      LimitEnv e;
      e.ti_current = ip.ti_current;
      e.time_base = ip.time_base;
      e.max_active_bin = kc.max_active_bin;
      const bool planetary = kc.scheme == SWIFT_CUDA_SCHEME_PLANETARY;
      e.energy_min = planetary ? ip.min_u : (ip.min_u > 0.f ? ip.min_u : 0.f);
      int bin = d.time_bin[p];
      const bool inactive = bin > kc.max_active_bin;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f), vf4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (inactive) {
        acc = d.acc[p];
        vf4 = g.vu_full[p];
      }
      float vf[4] = {vf4.x, vf4.y, vf4.z, vf4.w};
      const float a[3] = {acc.x, acc.y, acc.z};
      float u_dt = acc.w;
      bool dirty = false;
      t_end = limit_part(e, wake, bin, vf, a, u_dt, dirty);
      if (inactive) {
        g.vu_full[p] = make_float4(vf[0], vf[1], vf[2], vf[3]);
        if (dirty) {
          acc.w = u_dt;
          d.acc[p] = acc;
        }
      }
      const_cast<int8_t *>(d.time_bin)[p] = (int8_t)bin;
      float4 hb = d.hb[p];
      hb.z = __int_as_float(bin);
      d.hb[p] = hb;
      d.wakeup[p] = kLpNotAwake;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const long long other = __shfl_xor_sync(0xffffffffu, t_end, o);
    t_end = other < t_end ? other : t_end;
  }
  const unsigned b = __ballot_sync(0xffffffffu, lim);
  if ((threadIdx.x & 31) == 0 && b) {
    atomicMin(g.ti_end_min, (unsigned long long)t_end);
    atomicAdd(g.updated, (unsigned long long)__popc(b));
  }
}

// The limiter loop's per-particle slate: nobody is awake (time_bin_not_awake = -num_time_bins).
__global__ void k_wakeup_reset(int64_t n, int *wakeup) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p < n) wakeup[p] = -kNumTimeBins;
}

// Time bins handed over by a caller that runs the time-step task itself (caller order -> device slots;
// the force-loop record of a neighbour carries the bin too).
__global__ void k_set_bins(int64_t n, const int32_t *order, const int8_t *src, int8_t *time_bin, float4 *hb) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int bin = src[order ? order[k] : k];
  time_bin[k] = (int8_t)bin;
  float4 r = hb[k];
  r.z = __int_as_float(bin);
  hb[k] = r;
}

// p->limiter_data.wakeup in caller order
__global__ void k_scatter_wakeup(int64_t n, const int32_t *order, const int *wakeup, int8_t *dst) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  dst[order ? order[k] : k] = (int8_t)wakeup[k];
}

// State inspection: scatter `width` floats per particle from a strided source to caller order.
__global__ void k_scatter_f32(int64_t n, const int32_t *order, const float *src, int src_stride, int width,
                              float *dst) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t s = order ? order[k] : k;
  for (int c = 0; c < width; c++) dst[s * width + c] = src[k * src_stride + c];
}
__global__ void k_scatter_pos(int64_t n, const int32_t *order, const PosD *src, double *dst) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t s = order ? order[k] : k;
  const PosD q = src[k];
  dst[3 * s] = q.x; dst[3 * s + 1] = q.y; dst[3 * s + 2] = q.z;
}
__global__ void k_scatter_bins(int64_t n, const int32_t *order, const int8_t *tb, const int *min_bin, int8_t *dst_tb,
                               int8_t *dst_min) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t s = order ? order[k] : k;
  dst_tb[s] = tb[k];
  dst_min[s] = (int8_t)min_bin[k];
}

}  // namespace swiftcuda
