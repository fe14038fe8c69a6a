// Ghost redo sweeps from recorded neighbour lists (kernels.cuh, kLD, Dev::nl_grow).
//
// The first redo sweep of a step walks the cells (k_density<1>) and, while it is at it, records for
// every lane the candidates within kGrowL x its search radius as (dx, dy, dz, j): dx exactly as
// DOSELF_SUBSET / DOPAIR_SUBSET form it (src/runner_doiact_functions_hydro.h:745-770, 974-1003),
// which does not depend on h.  Newton-Raphson then changes h by a few per cent at most, so the
// following sweeps need no walk: k_density_list re-tests the ~57 recorded candidates of a particle
// against the new h, in the recorded (= the reference's) order, with the same unfused arithmetic:
// bit-identical sums.  A particle whose new h outgrows its list, whose list overflowed, or that has
// a hit within rounding of the sorted-axis cut (the cut is implied by r2 < h^2 gamma^2 only up to
// `cut_delta`) goes back to the walking kernel for that sweep, which records a fresh list.
#pragma once
#include "kernels.cuh"

namespace swiftcuda {

__global__ void __launch_bounds__(kBlock) k_density_list(Dev d, int sweep, int force_slow) {
  if (d.redo_count[sweep] == 0) return;
  const int lane = threadIdx.x & 31;
  const int u = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (u >= d.n_units) return;
  const int ul = u * 32 + lane;
  const int q = d.lane_part[ul];
  const bool on = q >= 0 && d.gstate[q] == 4;
  if (!__any_sync(kFull, on)) return;
  const bool rec = d.rec_capacity > 0;
  const int pp = on ? q : 0;
  const float hi = d.h[pp];
  const float hig2 = sq_gamma2(hi);
  const float hi_inv = 1.f / hi;
  const float hig = __fmul_rn(hi, kc.gamma);
  // force_slow (test hook): every hit counts as unsafe, i.e. every listed particle goes back to the walk
  const float r2safe = safe_r2(hig, cut_delta(d), force_slow != 0);
  const float4 vmi = d.vm[pp];
  const int cnt = on ? (int)d.nl_cnt[ul] : 0;
  const float4 *__restrict__ nlp = d.nl + ((size_t)u * kLD) * 32 + lane;
  const int m = __reduce_max_sync(kFull, cnt);
  bool abort = false;
  if (rec) {
    // test hook: pairs are recorded as they are consumed, so find out first whether the lane goes back
    for (int r = 0; r < m; r++) {
      if (r < cnt) {
        const float4 e = nlp[(size_t)r * 32];
        const float r2 = dist2(e.x, e.y, e.z);
        if (r2 < hig2 && !(r2 < r2safe)) abort = true;
      }
    }
  }
  DensAcc acc;
  acc.zero();
  unsigned long long nhits = 0;
  const long long idi = rec ? d.id[pp] : 0;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  // software pipeline: entry r + 2 being read, v/m of entry r + 1 in flight, entry r computed
  float4 e0 = cnt > 0 ? __ldcs(&nlp[0]) : z4;
  float4 e1 = cnt > 1 ? __ldcs(&nlp[32]) : z4;
  float4 v0 = cnt > 0 ? d.vm[__float_as_int(e0.w)] : z4;
  for (int r = 0; r < m; r++) {
    const float4 e2 = (r + 2 < cnt) ? __ldcs(&nlp[(size_t)(r + 2) * 32]) : z4;
    const float4 v1 = (r + 1 < cnt) ? d.vm[__float_as_int(e1.w)] : z4;
    if (r < cnt && !abort) {
      const float r2 = dist2(e0.x, e0.y, e0.z);
      if (r2 < hig2) {
        if (!(r2 < r2safe)) {
          abort = true;
        } else {
          iact_density(acc, r2, e0.x, e0.y, e0.z, hi_inv, vmi.x, vmi.y, vmi.z, v0);
          nhits++;
          if (rec) record_pair(d, 1, idi, d.id[__float_as_int(e0.w)]);
        }
      }
    }
    e0 = e1; e1 = e2; v0 = v1;
  }
  if (on) {
    if (abort) {
      // back to the walking kernel for this sweep (it runs after this one and records a new list)
      d.gstate[q] = 2;
      d.nl_cnt[ul] = 0xffff;
      nhits = 0;
      atomicAdd(&d.walk_count[sweep], 1);
      const int slot = d.part_slot[q];
      if (slot >= 0 && atomicExch(&d.leaf_mark[slot], sweep + 1) != sweep + 1)
        d.leaf_list[atomicAdd(&d.leaf_count[sweep], 1)] = slot;
    } else {
      d.dens0[q] = make_float4(acc.rho, acc.rho_dh, acc.wcount, acc.wcount_dh);
      d.dens1[q] = make_float4(acc.div_v, acc.rx, acc.ry, acc.rz);
    }
  }
  for (int o = 16; o > 0; o >>= 1) nhits += __shfl_down_sync(kFull, nhits, o);
  if (lane == 0 && nhits) atomicAdd(&d.hits[1], nhits);
}

}  // namespace swiftcuda
