// K6 in its FAST form (swift_cuda_params.fields_mode = SWIFT_CUDA_FIELDS_FAST): the same persistent
// warp-per-unit walk as k_force, restructured around what the FAST mode no longer has to keep:
//
//   * the pair SET is what must be exact, not the order the hits are summed in: loop-1 and loop-2
//     hits of DOPAIR2 (src/runner_doiact_functions_hydro.h:1528-1737, 1741-1952) are consumed as they
//     are found, one pass per pair, no side lists;
//   * the candidate test is a pure distance filter: r2 (FMA form) against max(H_i^2, H_j^2) widened by
//     1e-6 -- 12 instructions per candidate instead of 22, and no per-candidate cut constants to
//     stage (the ring keeps the candidate's (h gamma)^2 in the w slot of its position);
//   * the reference's decision is made when the hit is consumed, from the stored dx exactly as the
//     reference forms it: r2 < H_i^2 (loop 1) or else r2 < H_j^2 (loop 2), unfused arithmetic.  The
//     sorted-axis cuts of the reference are implied by that decision except when r is within
//     `cut_delta` of the deciding radius; only then the cuts are evaluated, exactly, by
//     force_cut_exact (which finds the interaction the neighbour belongs to by its cell range);
//   * the body is iact_force_fast (FMA contraction, MUFU reciprocals).
//
// Same pair set as k_force, bit for bit; a, u_dt, h_dt, v_sig within the FAST tolerance.
#pragma once
#include "kernels.cuh"

namespace swiftcuda {

// What force_cut_exact reads of Dev (see CutCtx in kernels.cuh: a __noinline__ function taking Dev
// itself would force the kernel's copy of it onto the local-memory stack).
struct ForceCutCtx {
  const CellDev *cells;
  const int64_t *sort_lookup;
  const SortEnt *sort;
  const float *cell_hmax;
  const Item *items;
  const PosD *pos;
  double dim[3];
};

// The exact DOPAIR2 decision for one (lane particle i, neighbour j) pair inside the unsafe band of one of
// the two radii: which loop of the pair's task visits it (i in ci: loop 1 by H_i, else loop 2 by H_j;
// i in cj: loop 1 by H_j, else loop 2 by H_i), and does that loop's sorted-axis cut let it through?
// Mirrors the per-lane / per-candidate constants of k_force (cut1, cut2, aux.y, aux.z).
// hig2 / hjg2: (h gamma)^2 of the lane / of the neighbour as the reference forms them.
__device__ __noinline__ bool force_cut_exact(const ForceCutCtx d, int ib, int ie, int j, double xi, double yi, double zi,
                                             float hi, float hj, float r2) {
  const float hig = __fmul_rn(hi, kc.gamma), hjg = __fmul_rn(hj, kc.gamma);
  const float hig2 = sq_gamma2(hi), hjg2 = sq_gamma2(hj);
  for (int it = ib; it < ie; it++) {
    const Item item = d.items[it];
    if (item.kind == kSelf) continue;
    const CellDev cB = d.cells[item.partner];
    if (j < cB.first || j >= cB.first + cB.count) continue;
    const CellDev cA = d.cells[item.own];
    const int sid = item.sid;
    const bool left = item.kind == kPairLeft;
    double sh[3];
#pragma unroll
    for (int k = 0; k < 3; k++) sh[k] = (double)item_shift_i(item.shift, k) * d.dim[k];
    const double rshift = item_rshift(sh, sid);
    const int64_t segB = d.sort_lookup[(int64_t)item.partner * 13 + sid];
    const int64_t segA = d.sort_lookup[(int64_t)item.own * 13 + sid];
    const float hmaxA = d.cell_hmax[item.own], hmaxB = d.cell_hmax[item.partner];
    const double dj_min = (double)d.sort[left ? segB : segA].d;                                         // first of cj
    const double di_max = (double)d.sort[left ? (segA + cA.count - 1) : (segB + cB.count - 1)].d;       // last of ci
    const double hi_max_g = __dmul_rn((double)(left ? hmaxA : hmaxB), (double)kc.gamma);
    const double hj_max_g = __dmul_rn((double)(left ? hmaxB : hmaxA), (double)kc.gamma);
    const double rhs2 = __dsub_rn(di_max, rshift);
    const float d_me = sort_key(xi, yi, zi, sid);
    double xj, yj, zj;
    load_pos(d.pos, j, xj, yj, zj);
    const float dk = sort_key(xj, yj, zj, sid);
    if (left) {
      // loop 1: I am pi; R in cj while (double)dR < di
      const double di = __dsub_rn((double)__fadd_rn(d_me, hig), rshift);
      const bool ok1 = (__dsub_rn(__dadd_rn((double)d_me, hi_max_g), rshift) > dj_min) && !(di < dj_min);
      const bool h1 = ok1 && dk <= thr_lt(di) && r2 < hig2;
      // loop 2: the candidate is pj; I am inside its range
      const float djf = __fsub_rn(dk, hjg);
      const bool ok2 = (__dsub_rn((double)dk, hj_max_g) < rhs2) && !((double)djf > rhs2);
      const bool h2 = ok2 && djf <= thr_lt(__dsub_rn((double)d_me, rshift)) && r2 < hjg2 && r2 >= hig2;
      return h1 || h2;
    }
    // loop 1 of the task: the candidate is pi; I am inside its range
    const double di = __dsub_rn((double)__fadd_rn(dk, hjg), rshift);
    const bool ok1 = (__dsub_rn(__dadd_rn((double)dk, hi_max_g), rshift) > dj_min) && !(di < dj_min);
    const bool h1 = ok1 && d_me <= thr_lt(di) && r2 < hjg2;
    // loop 2 of the task: I am pj; L in ci descending while (double)dL - rshift > dj
    const float djf = __fsub_rn(d_me, hig);
    const bool ok2 = (__dsub_rn((double)d_me, hj_max_g) < rhs2) && !((double)djf > rhs2);
    const bool h2 = ok2 && dk >= thr_gt_shift(rshift, (double)djf) && r2 < hig2 && r2 >= hjg2;
    return h1 || h2;
  }
  return true;  // not found among the pair interactions: own cell, DOSELF2 has no sorted-axis cut
}

// FIFO traffic with L2 eviction priorities: a stored hit should stay in the L2 until it is consumed
// (evict_last), a consumed one should be the first to go (evict_first).  SWIFT_FIFO_POLICY=0: the
// streaming st.global.cs / ld.global.cs of the first-generation kernels.
#ifndef SWIFT_FIFO_POLICY
#define SWIFT_FIFO_POLICY 1
#endif
__device__ __forceinline__ unsigned long long l2_policy_evict_last() {
  unsigned long long p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ unsigned long long l2_policy_evict_first() {
  unsigned long long p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void st_pred_pol(float4 *p, float x, float y, float z, float w, bool pred,
                                            unsigned long long pol) {
  asm volatile(
      "{ .reg .pred q; setp.ne.b32 q, %5, 0; @q st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %6; }" ::"l"(p),
      "f"(x), "f"(y), "f"(z), "f"(w), "r"((int)pred), "l"(pol)
      : "memory");
}
__device__ __forceinline__ float4 ld_pol(const float4 *p, unsigned long long pol) {
  float4 r;
  asm volatile("ld.global.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p), "l"(pol)
               : "memory");
  return r;
}

struct FastTile {
  float4 pos[kTile];  // frame position of the candidate; w: its (h gamma)^2 (own cell: its cull threshold)
  int idx[kTile];
  float box[12];
};

template <int SCHEME>
__global__ void __launch_bounds__(kBlock, SWIFT_MINB_FORCE) k_force_fast(Dev d, int force_slow, int qd) {
  __shared__ FastTile tiles[kWarps];
  FastTile &ws = tiles[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  // qd <= kQD rows of the per-warp FIFO are used: the FAST mode may consume its hits whenever the FIFO
  // fills up, and a shorter FIFO keeps the hit traffic of all resident warps inside the L2
  float4 *const gq = d.gq + (size_t)(blockIdx.x * kWarps + (threadIdx.x >> 5)) * (qd * 32) + lane;
  int *const work = d.work_ctr + 1 + kc.max_iter;
  const bool rec = d.rec_capacity > 0;
  const float qnan = __int_as_float(0x7fc00000);
  const float delta = cut_delta(d);
  const unsigned long long pol_keep = l2_policy_evict_last(), pol_done = l2_policy_evict_first();
  (void)pol_keep; (void)pol_done;
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
    const float r2safe_i = safe_r2(hig, delta, force_slow != 0);
    const float hig2_t = mine ? hig2 * 1.000001f : -INFINITY;  // the filter's threshold of this lane
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

    // one deferred candidate: the reference's decision, then runner_iact_nonsym_force (FAST arithmetic)
    auto consume = [&](const float4 e, const float4 hbj, const float4 vmj, const float4 fj, const float4 sxj,
                       double xj, double yj, double zj) {
      const int w = __float_as_int(e.w);
      const int j = w & (kDpFlag - 1);
      float dx, dy, dz;
      if (w & kDpFlag) {  // own cell: f64 position difference (:2349-2353, 2416-2420)
        dx = dsub_f(xi, xj); dy = dsub_f(yi, yj); dz = dsub_f(zi, zj);
      } else {
        dx = e.x; dy = e.y; dz = e.z;
      }
      const float r2 = dist2(dx, dy, dz);
      const float hjg2 = sq_gamma2(hbj.x);
      const bool in_i = r2 < hig2, in_j = r2 < hjg2;
      bool hit = in_i || in_j;
      if (hit && !(w & kDpFlag)) {
        // inside a radius but within delta of it: that radius' sorted-axis cut is not implied
        const float safe_j = safe_r2(__fmul_rn(hbj.x, kc.gamma), delta, force_slow != 0);
        if ((in_i && !(r2 < r2safe_i)) || (in_j && !(r2 < safe_j))) {
          ForceCutCtx cc;
          cc.cells = d.cells; cc.sort_lookup = d.sort_lookup; cc.sort = d.sort; cc.cell_hmax = d.cell_hmax;
          cc.items = d.items; cc.pos = d.pos;
          cc.dim[0] = d.dim[0]; cc.dim[1] = d.dim[1]; cc.dim[2] = d.dim[2];
          hit = force_cut_exact(cc, ib, ie, j, xi, yi, zi, me.h, hbj.x, r2);
        }
      }
      if (hit) {
        ForceJ pj;
        pj.h = hbj.x; pj.bal = hbj.y; pj.bin = __float_as_int(hbj.z); pj.Pr2 = hbj.w;
        pj.m = vmj.w; pj.vx = vmj.x; pj.vy = vmj.y; pj.vz = vmj.z;
        pj.rho = fj.x; pj.P = fj.y; pj.f = fj.z; pj.c = fj.w;
        if (SCHEME == SWIFT_CUDA_SCHEME_SPHENIX) {
          pj.alpha = sxj.x; pj.adiff = sxj.y; pj.u = sxj.z;
          iact_force_sphenix_fast(acc, r2, dx, dy, dz, me, pj);
        } else {
          iact_force_fast<SCHEME>(acc, r2, dx, dy, dz, me, pj);
        }
        nhits++;
        if (rec) record_pair(d, 2, idi, d.id[j]);
      }
    };
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
#if SWIFT_FIFO_POLICY
#define FIFO_LD(ptr) ld_pol(ptr, pol_done)
#else
#define FIFO_LD(ptr) __ldcs(ptr)
#endif
      float4 e0 = qcnt > 0 ? FIFO_LD(&gq[0]) : z4;
      float4 e1 = qcnt > 1 ? FIFO_LD(&gq[32]) : z4;
      float4 hb0 = z4, v0 = z4, f0 = z4, s0 = z4;
      double x0 = 0.0, y0 = 0.0, z0 = 0.0;
      gather(e0, qcnt > 0, hb0, v0, f0, s0, x0, y0, z0);
      for (int r = 0; r < m; r++) {
        const float4 e2 = (r + 2 < qcnt) ? FIFO_LD(&gq[(r + 2) * 32]) : z4;
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
      const bool left = kind == kPairLeft;
      const float hmaxBg = __fmul_rn(d.cell_hmax[item.partner], kc.gamma);

      double sh[3] = {0.0, 0.0, 0.0};
      if (!self) {
#pragma unroll
        for (int k = 0; k < 3; k++) sh[k] = item_shift(d, item.shift, k);
      }
      double fo[3], fl[3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        if (self) { fo[k] = fl[k] = cA.loc[k]; }
        else if (left) { fo[k] = cB.loc[k]; fl[k] = __dadd_rn(cB.loc[k], sh[k]); }
        else { fo[k] = __dadd_rn(cA.loc[k], sh[k]); fl[k] = cA.loc[k]; }
      }
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
      const float e_self = 4e-7f * cA.width;
      // the lane's threshold for this interaction (own cell: the cull threshold of the f64 difference)
      const float thr_me = self ? (mine ? cull_thr(hig2, e_self) : -INFINITY) : hig2_t;

      // how far along the sorted axis some lane can still interact: a pair needs
      // |key_j - (key_i -+ rshift)| <= r < max(H_i, H_j) <= max(H_i, gamma h_max of the partner)
      float wlim = 0.f;
      int64_t segB = 0;
      if (!self) {
        segB = d.sort_lookup[(int64_t)item.partner * 13 + sid];
        const double rshift = item_rshift(sh, sid);
        const float d_me = sort_key(xi, yi, zi, sid);
        const float reach = fmaxf(hig, hmaxBg) * 1.000002f;
        const float kme = (float)(left ? ((double)d_me - rshift) : ((double)d_me + rshift));
        const float m = 4e-7f * (fabsf(kme) + fabsf((float)rshift) + fabsf(d_me)) + 1e-30f;
        if (left) wlim = warp_max(mine ? kme + reach + m : -INFINITY);
        else wlim = warp_min(mine ? kme - reach - m : INFINITY);
      }

      auto test = [&](int st, int n) {
        if (__any_sync(kFull, qcnt > qd - n)) drain_all();
        const float4 *__restrict__ tp = ws.pos + st;
        const int *__restrict__ ti = ws.idx + st;
        float4 *qp = gq + qcnt * 32;
        const int flag = self ? kDpFlag : 0;
        for (int k0 = 0; k0 < n; k0 += 4) {
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const float4 c = tp[k0 + u];
            const int j = ti[k0 + u];
            const float dx = __fsub_rn(pfx, c.x), dy = __fsub_rn(pfy, c.y), dz = __fsub_rn(pfz, c.z);
            const float r2c = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
            // candidate's own radius counts only for lanes that take part (mine) and never for oneself
            const bool push = mine && r2c < fmaxf(thr_me, c.w) && j != pp;
#if SWIFT_FIFO_POLICY
            st_pred_pol(qp, dx, dy, dz, __int_as_float(j | flag), push, pol_keep);
#else
            st_pred(qp, dx, dy, dz, __int_as_float(j | flag), push);
#endif
            qp += push ? 32 : 0;
          }
        }
        qcnt = (int)(qp - gq) >> 5;
      };

      int st = 0, ntile = 0, base = 0;
      bool done = false;
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
            if (left ? !(k0 <= wlim) : !(k0 >= wlim)) done = true;  // sorted axis: nothing further on can interact
            keep = keep && !done && (left ? (dk <= wlim) : (dk >= wlim));
          }
          float4 c = make_float4(qnan, qnan, qnan, -INFINITY);
          if (keep) {
            double xj, yj, zj;
            load_pos(d.pos, j, xj, yj, zj);
            c = make_float4(dsub_f(xj, fo[0]), dsub_f(yj, fo[1]), dsub_f(zj, fo[2]), 0.f);
            keep = c.x >= sb.lo[0] && c.x <= sb.hi[0] && c.y >= sb.lo[1] && c.y <= sb.hi[1] && c.z >= sb.lo[2] &&
                   c.z <= sb.hi[2];
          }
          if (keep) {
            const float hjg2 = sq_gamma2(d.hb[j].x);
            c.w = self ? cull_thr(hjg2, e_self) : hjg2 * 1.000001f;
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
          if (lane < n - ntile) {
            ws.pos[(st + ntile + lane) & (kTile - 1)] = make_float4(qnan, qnan, qnan, -INFINITY);
            ws.idx[(st + ntile + lane) & (kTile - 1)] = -1;
          }
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
    if (mine) {
      d.acc[p] = make_float4(acc.ax, acc.ay, acc.az, acc.u_dt);
      d.h_dt[p] = d.no_end_force ? acc.h_dt : acc.h_dt * (me.h * 0.3333333333f);
      d.v_sig[p] = acc.v_sig;
      d.min_bin[p] = acc.min_bin;
    }
  }
  for (int o = 16; o > 0; o >>= 1) nhits += __shfl_down_sync(kFull, nhits, o);
  if (lane == 0 && nhits) atomicAdd(&d.hits[2], nhits);
}

}  // namespace swiftcuda
