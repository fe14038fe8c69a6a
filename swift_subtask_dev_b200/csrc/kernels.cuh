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

#include "device_types.cuh"

namespace swiftcuda {

__constant__ Consts kc;

constexpr int kBlock = 128;     // threads per CTA of the pair kernels
constexpr int kTile = 128;      // partner particles staged per shared-memory tile
constexpr int kNumTimeBins = 56;

// ---------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float sq_gamma2(float h) { return __fmul_rn(__fmul_rn(h, h), kc.gamma2); }

__device__ __forceinline__ float dist2(float dx, float dy, float dz) {
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
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
};
struct ForceJ {
  float h, m, rho, P, f, c, bal, vx, vy, vz;
  int bin;
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
  const float P_over_rho2_i = pi.P / (pi.rho * pi.rho) * f_ij;
  const float P_over_rho2_j = pj.P / (pj.rho * pj.rho) * f_ji;
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

__device__ __forceinline__ void record_pair(const Dev &d, int loop, long long i, long long j) {
  const unsigned long long k = atomicAdd(&d.rec_count[loop], 1ULL);
  if ((long long)k < d.rec_capacity) {
    d.rec_pairs[loop][2 * k] = i;
    d.rec_pairs[loop][2 * k + 1] = j;
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
constexpr int kSortSmem = 4096;

__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void __launch_bounds__(256) k_sort(Dev d, const int32_t *seg_cell, const int32_t *seg_sid,
                                              const int64_t *seg_off, unsigned long long *scratch,
                                              const int64_t *scratch_off) {
  __shared__ unsigned long long skeys[kSortSmem];
  const int seg = blockIdx.x;
  const CellDev c = d.cells[seg_cell[seg]];
  const int sid = seg_sid[seg];
  const int n = c.count;
  int npad = 1;
  while (npad < n) npad <<= 1;
  unsigned long long *keys = (npad <= kSortSmem) ? skeys : (scratch + scratch_off[seg]);
  for (int k = threadIdx.x; k < npad; k += blockDim.x) {
    unsigned long long key = ~0ULL;
    if (k < n) {
      const int64_t p = c.first + k;
      const float dk = sort_key(d.x[p], d.y[p], d.z[p], sid);
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
    d.sort_d[off + k] = ordered_to_float((uint32_t)(key >> 32));
    d.sort_i[off + k] = (int32_t)(key & 0xffffffffu);
  }
}

// ---------------------------------------------------------------------------
// K1b: start of a step: zero sums (hydro_init_part), arm the ghost state
// ---------------------------------------------------------------------------
__global__ void k_init_step(Dev d) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n) return;
  d.dens0[p] = make_float4(0.f, 0.f, 0.f, 0.f);
  d.dens1[p] = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool active = d.time_bin[p] <= kc.max_active_bin;
  d.gstate[p] = active ? 1 : 0;
  d.left[p] = 0.f;
  d.right[p] = kc.h_max;
  d.min_bin[p] = kNumTimeBins + 1;
}

// ---------------------------------------------------------------------------
// K3/K5: density.  One CTA per updated cell; a thread owns one particle i of
// the cell and keeps its eight neighbour sums in registers across all of the
// cell's interactions; partner particles are staged through shared memory in
// sorted-axis order and walked with the reference's early exit.
// MODE 0: first pass (DOSELF1/DOPAIR1 arithmetic, every active particle).
// MODE 1: ghost redo (DOSELF_SUBSET/DOPAIR_SUBSET arithmetic, gstate == 2 only).
// ---------------------------------------------------------------------------
struct TileJ {
  double x[kTile], y[kTile], z[kTile];  // absolute positions (self, subset pair)
  float fx[kTile], fy[kTile], fz[kTile];  // frame-relative positions (pair, subset self)
  float d[kTile];                       // sort key
  float4 vm[kTile];
  long long id[kTile];
  int idx[kTile];
};

template <int MODE>
__global__ void __launch_bounds__(kBlock) k_density(Dev d, int sweep) {
  if (MODE == 1 && d.redo_count[sweep] == 0) return;
  __shared__ TileJ t;
  __shared__ int s_list[kBlock];
  __shared__ int s_nlist;
  const int A = d.icells[blockIdx.x];
  const CellDev cA = d.cells[A];
  const int ib = d.item_begin[blockIdx.x], ie = d.item_begin[blockIdx.x + 1];
  const bool rec = d.rec_capacity > 0;
  const int loop = MODE;
  unsigned long long nhits = 0;

  for (int chunk = 0; chunk < cA.count; chunk += kBlock) {
    // which particle does this thread own in this chunk?
    int li = chunk + threadIdx.x;
    bool mine = li < cA.count;
    if (MODE == 0) {
      if (mine) mine = d.time_bin[cA.first + li] <= kc.max_active_bin;
    } else {
      // compact the flagged particles of this chunk so that the CTA can leave early
      if (threadIdx.x == 0) s_nlist = 0;
      __syncthreads();
      if (mine && d.gstate[cA.first + li] == 2) s_list[atomicAdd(&s_nlist, 1)] = li;
      __syncthreads();
      const int nl = s_nlist;
      if (nl == 0) {
        __syncthreads();
        continue;
      }
      mine = threadIdx.x < nl;
      li = mine ? s_list[threadIdx.x] : 0;
      __syncthreads();
    }
    const int64_t pi = cA.first + (mine ? li : 0);
    const double xi = d.x[pi], yi = d.y[pi], zi = d.z[pi];
    const float hi = d.h[pi];
    const float hig2 = sq_gamma2(hi);
    const float hi_inv = 1.f / hi;
    const float4 vmi = d.vm[pi];
    const long long idi = rec ? d.id[pi] : 0;
    DensAcc acc;
    acc.zero();

    for (int it = ib; it < ie; it++) {
      const Item item = d.items[it];
      const CellDev cB = d.cells[item.partner];
      if (item.role == kRoleSelf) {
        // ---- own cell ----
        float pfx = 0.f, pfy = 0.f, pfz = 0.f;
        if (MODE == 1) {
          pfx = dsub_f(xi, cA.loc[0]);
          pfy = dsub_f(yi, cA.loc[1]);
          pfz = dsub_f(zi, cA.loc[2]);
        }
        for (int base = 0; base < cB.count; base += kTile) {
          const int nt = min(kTile, cB.count - base);
          __syncthreads();
          for (int k = threadIdx.x; k < nt; k += kBlock) {
            const int64_t pj = cB.first + base + k;
            const double xj = d.x[pj], yj = d.y[pj], zj = d.z[pj];
            if (MODE == 0) {
              t.x[k] = xj; t.y[k] = yj; t.z[k] = zj;
            } else {
              t.fx[k] = dsub_f(xj, cA.loc[0]);
              t.fy[k] = dsub_f(yj, cA.loc[1]);
              t.fz[k] = dsub_f(zj, cA.loc[2]);
            }
            t.vm[k] = d.vm[pj];
            if (rec) t.id[k] = d.id[pj];
          }
          __syncthreads();
          if (mine) {
            for (int k = 0; k < nt; k++) {
              if (base + k == li) continue;
              float dx, dy, dz;
              if (MODE == 0) {
                dx = dsub_f(xi, t.x[k]); dy = dsub_f(yi, t.y[k]); dz = dsub_f(zi, t.z[k]);
              } else {
                dx = __fsub_rn(pfx, t.fx[k]); dy = __fsub_rn(pfy, t.fy[k]); dz = __fsub_rn(pfz, t.fz[k]);
              }
              const float r2 = dist2(dx, dy, dz);
              if (r2 < hig2) {
                iact_density(acc, r2, dx, dy, dz, hi_inv, vmi.x, vmi.y, vmi.z, t.vm[k]);
                nhits++;
                if (rec) record_pair(d, loop, idi, t.id[k]);
              }
            }
          }
        }
      } else {
        // ---- pair ----
        const int sid = item.sid;
        const int64_t segB = d.sort_lookup[(int64_t)item.partner * 13 + sid];
        const int64_t segA = d.sort_lookup[(int64_t)A * 13 + sid];
        const bool left = item.role == kRoleLeft;
        const float d_me = sort_key(xi, yi, zi, sid);
        // frame origins: particles of ci are taken relative to cj.loc + shift, those of cj relative to cj.loc
        const double oix = __dadd_rn(item.org[0], item.shift[0]), oiy = __dadd_rn(item.org[1], item.shift[1]),
                     oiz = __dadd_rn(item.org[2], item.shift[2]);
        float pfx, pfy, pfz;
        double cut;           // di (left) / dj (right) / subset di
        double px = 0, py = 0, pz = 0;
        bool skip = !mine;
        if (MODE == 0) {
          if (left) {
            pfx = dsub_f(xi, oix); pfy = dsub_f(yi, oiy); pfz = dsub_f(zi, oiz);
            const double dj_min = (double)d.sort_d[segB];
            const double hi_max = __dsub_rn((double)__fmul_rn(d.cell_hmax[A], kc.gamma), item.rshift);
            if (!(__dadd_rn((double)d_me, hi_max) > dj_min)) skip = true;
            cut = __dsub_rn((double)__fadd_rn(d_me, __fmul_rn(hi, kc.gamma)), item.rshift);
            if (cut < dj_min) skip = true;
          } else {
            pfx = dsub_f(xi, item.org[0]); pfy = dsub_f(yi, item.org[1]); pfz = dsub_f(zi, item.org[2]);
            const double di_max = __dsub_rn((double)d.sort_d[segB + cB.count - 1], item.rshift);
            const double hj_max = (double)__fmul_rn(d.cell_hmax[A], kc.gamma);
            if (!(__dsub_rn((double)d_me, hj_max) < di_max)) skip = true;
            cut = __dadd_rn((double)__fsub_rn(d_me, __fmul_rn(hi, kc.gamma)), item.rshift);
            if (__dsub_rn(cut, item.rshift) > di_max) skip = true;
          }
        } else {
          // DOPAIR_SUBSET: shift seen from this cell; "flipped" == this cell is the right one
          const double sgn = left ? 1.0 : -1.0;
          px = __dsub_rn(xi, sgn * item.shift[0]);
          py = __dsub_rn(yi, sgn * item.shift[1]);
          pz = __dsub_rn(zi, sgn * item.shift[2]);
          const double proj0 = __dmul_rn(px, kc.axis[sid][0]), proj1 = __dmul_rn(py, kc.axis[sid][1]),
                       proj2 = __dmul_rn(pz, kc.axis[sid][2]);
          const float hg = left ? __fmul_rn(hi, kc.gamma) : __fmul_rn(-hi, kc.gamma);
          cut = __dadd_rn(__dadd_rn(__dadd_rn((double)hg, proj0), proj1), proj2);
          pfx = pfy = pfz = 0.f;
        }
        (void)segA;
        // walk the partner's sorted list from the interface outwards
        int done = skip;
        for (int base = 0; base < cB.count; base += kTile) {
          const int nt = min(kTile, cB.count - base);
          if (__syncthreads_and(done)) break;
          for (int k = threadIdx.x; k < nt; k += kBlock) {
            const int e = left ? (base + k) : (cB.count - 1 - base - k);
            const int lj = d.sort_i[segB + e];
            const int64_t pj = cB.first + lj;
            t.d[k] = d.sort_d[segB + e];
            const double xj = d.x[pj], yj = d.y[pj], zj = d.z[pj];
            if (MODE == 0) {
              if (left) {  // partner is cj
                t.fx[k] = dsub_f(xj, item.org[0]); t.fy[k] = dsub_f(yj, item.org[1]); t.fz[k] = dsub_f(zj, item.org[2]);
              } else {     // partner is ci
                t.fx[k] = dsub_f(xj, oix); t.fy[k] = dsub_f(yj, oiy); t.fz[k] = dsub_f(zj, oiz);
              }
            } else {
              t.x[k] = xj; t.y[k] = yj; t.z[k] = zj;
            }
            t.vm[k] = d.vm[pj];
            if (rec) t.id[k] = d.id[pj];
          }
          __syncthreads();
          if (!done) {
            for (int k = 0; k < nt; k++) {
              const double dk = (double)t.d[k];
              if (left ? !(dk < cut) : !(dk > cut)) {
                done = 1;
                break;
              }
              float dx, dy, dz;
              if (MODE == 0) {
                dx = __fsub_rn(pfx, t.fx[k]); dy = __fsub_rn(pfy, t.fy[k]); dz = __fsub_rn(pfz, t.fz[k]);
              } else {
                dx = dsub_f(px, t.x[k]); dy = dsub_f(py, t.y[k]); dz = dsub_f(pz, t.z[k]);
              }
              const float r2 = dist2(dx, dy, dz);
              if (r2 < hig2) {
                iact_density(acc, r2, dx, dy, dz, hi_inv, vmi.x, vmi.y, vmi.z, t.vm[k]);
                nhits++;
                if (rec) record_pair(d, loop, idi, t.id[k]);
              }
            }
          }
        }
        __syncthreads();
      }
    }
    if (mine) {
      float *o0 = reinterpret_cast<float *>(&d.dens0[pi]);
      float *o1 = reinterpret_cast<float *>(&d.dens1[pi]);
      atomicAdd(o0 + 0, acc.rho);
      atomicAdd(o0 + 1, acc.rho_dh);
      atomicAdd(o0 + 2, acc.wcount);
      atomicAdd(o0 + 3, acc.wcount_dh);
      atomicAdd(o1 + 0, acc.div_v);
      atomicAdd(o1 + 1, acc.rx);
      atomicAdd(o1 + 2, acc.ry);
      atomicAdd(o1 + 3, acc.rz);
    }
    __syncthreads();
  }
  const unsigned long long tot = block_sum(nhits);
  if (threadIdx.x == 0 && tot) atomicAdd(&d.hits[MODE], tot);
}

// ---------------------------------------------------------------------------
// K4: ghost.  One thread per particle: hydro_end_density, Newton-Raphson /
// bisection step on h, convergence test, then either re-arm (zero the sums,
// flag for the next redo sweep) or hydro_prepare_force + hydro_reset_acceleration.
// runner_do_ghost, src/runner_ghost.c:1148-1483.
// ---------------------------------------------------------------------------
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
      m.a * density * u + (m.b * density * u * w_inv + m.A * mu * expf(-m.beta * nu)) * expf(-m.alpha * nu * nu);
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
  const float exp_beta = expf(-m.beta * nu);
  const float exp_alpha = expf(-m.alpha * nu * nu);
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
    cs = sqrtf(g * (g - 1.f) * u);
    return;
  }
  int m = 0;
  for (int k = 0; k < kc.n_til; k++)
    if (kc.til[k].mat_id == mat_id) m = k;
  P = til_pressure(rho, u, kc.til[m]);
  cs = til_soundspeed(rho, u, kc.til[m]);
}

__global__ void k_ghost(Dev d, int sweep) {
  if (sweep > 0 && d.redo_count[sweep - 1] == 0) return;
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n) return;
  const uint8_t st = d.gstate[p];
  if (st != 1 && st != 2) return;
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
      h = cbrtf(0.5f * (left * left * left + right * right * right));
    else
      h = h_new;
    if (h < kc.h_max && h > kc.h_min) {
      // another round: hydro_init_part
      d.h[p] = h;
      d.left[p] = left;
      d.right[p] = right;
      d.dens0[p] = make_float4(0.f, 0.f, 0.f, 0.f);
      d.dens1[p] = make_float4(0.f, 0.f, 0.f, 0.f);
      d.gstate[p] = 2;
      atomicAdd(&d.redo_count[sweep], 1);
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
  d.balsara[p] = balsara;
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

// ---------------------------------------------------------------------------
// K6: force.  Same decomposition as K3.  A pair (L in ci, R in cj) interacts
// when loop 1 of DOPAIR2 takes it (r2 < hL^2 g^2 inside L's sorted range) or
// loop 2 does (r2 < hR^2 g^2, r2 >= hL^2 g^2, inside R's sorted range).
// ---------------------------------------------------------------------------
struct TileF {
  double x[kTile], y[kTile], z[kTile];
  float fx[kTile], fy[kTile], fz[kTile];
  float d[kTile];
  float4 vm[kTile];
  float4 fin[kTile];
  float h[kTile], bal[kTile];
  int bin[kTile];
  long long id[kTile];
};

struct PairCuts {  // per (pair, direction) constants of DOPAIR2, all f64 as in the reference
  double rshift, dj_min, di_max, hi_max_g, hj_max_g;
};

// loop 1 (:1528-1737): L is the "pi" of the loop, R a particle of cj in L's range
__device__ __forceinline__ bool force_loop1(const PairCuts &c, float dL, float hL, float dR, float r2, float hLg2) {
  if (!(__dsub_rn(__dadd_rn((double)dL, c.hi_max_g), c.rshift) > c.dj_min)) return false;
  const double di = __dsub_rn((double)__fadd_rn(dL, __fmul_rn(hL, kc.gamma)), c.rshift);
  if (di < c.dj_min) return false;
  if (!((double)dR < di)) return false;
  return r2 < hLg2;
}
// loop 2 (:1741-1952): R is the "pj" of the loop, L a particle of ci in R's range
__device__ __forceinline__ bool force_loop2(const PairCuts &c, float dL, float dR, float hR, float r2, float hLg2,
                                            float hRg2) {
  const double rhs = __dsub_rn(c.di_max, c.rshift);
  if (!(__dsub_rn((double)dR, c.hj_max_g) < rhs)) return false;
  const double dj = (double)__fsub_rn(dR, __fmul_rn(hR, kc.gamma));
  if (dj > rhs) return false;
  if (!(__dsub_rn((double)dL, c.rshift) > dj)) return false;
  return (r2 < hRg2) && (r2 >= hLg2);
}

template <int SCHEME>
__global__ void __launch_bounds__(kBlock) k_force(Dev d) {
  __shared__ TileF t;
  const int A = d.icells[blockIdx.x];
  const CellDev cA = d.cells[A];
  const int ib = d.item_begin[blockIdx.x], ie = d.item_begin[blockIdx.x + 1];
  const bool rec = d.rec_capacity > 0;
  unsigned long long nhits = 0;

  for (int chunk = 0; chunk < cA.count; chunk += kBlock) {
    const int li = chunk + threadIdx.x;
    bool mine = li < cA.count;
    if (mine) mine = d.time_bin[cA.first + li] <= kc.max_active_bin;
    const int64_t pi = cA.first + (mine ? li : 0);
    const double xi = d.x[pi], yi = d.y[pi], zi = d.z[pi];
    ForceI me;
    {
      const float4 vmi = d.vm[pi];
      const float4 fi = d.fin[pi];
      me.h = d.h[pi];
      me.hi_inv = 1.0f / me.h;
      const float hi2 = me.hi_inv * me.hi_inv;
      me.hid_inv = hi2 * hi2;
      me.m = vmi.w; me.vx = vmi.x; me.vy = vmi.y; me.vz = vmi.z;
      me.rho = fi.x; me.P = fi.y; me.f = fi.z; me.c = fi.w;
      me.bal = d.balsara[pi];
    }
    const float hig2 = sq_gamma2(me.h);
    const long long idi = rec ? d.id[pi] : 0;
    ForceAcc acc;
    acc.ax = acc.ay = acc.az = acc.u_dt = acc.h_dt = 0.f;
    acc.v_sig = 0.f;  // combined with the stored value by max at the end
    acc.min_bin = kNumTimeBins + 1;

    for (int it = ib; it < ie; it++) {
      const Item item = d.items[it];
      const CellDev cB = d.cells[item.partner];
      const bool self = item.role == kRoleSelf;
      const bool left = item.role == kRoleLeft;
      const int sid = self ? 0 : item.sid;
      const int64_t segB = self ? 0 : d.sort_lookup[(int64_t)item.partner * 13 + sid];
      const int64_t segA = self ? 0 : d.sort_lookup[(int64_t)A * 13 + sid];
      float d_me = 0.f, pfx = 0.f, pfy = 0.f, pfz = 0.f;
      double oix = 0, oiy = 0, oiz = 0;
      PairCuts pc;
      double bound = 0;  // conservative end of the walk along the partner's sorted list
      if (!self) {
        d_me = sort_key(xi, yi, zi, sid);
        oix = __dadd_rn(item.org[0], item.shift[0]);
        oiy = __dadd_rn(item.org[1], item.shift[1]);
        oiz = __dadd_rn(item.org[2], item.shift[2]);
        pc.rshift = item.rshift;
        const float hmaxA = d.cell_hmax[A], hmaxB = d.cell_hmax[item.partner];
        if (left) {  // A = ci, B = cj
          pfx = dsub_f(xi, oix); pfy = dsub_f(yi, oiy); pfz = dsub_f(zi, oiz);
          pc.dj_min = (double)d.sort_d[segB];
          pc.di_max = (double)d.sort_d[segA + cA.count - 1];
          pc.hi_max_g = __dmul_rn((double)hmaxA, (double)kc.gamma);
          pc.hj_max_g = __dmul_rn((double)hmaxB, (double)kc.gamma);
          const double hmx = (double)fmaxf(me.h, hmaxB) * (double)kc.gamma * 1.000001 + 1e-30;
          bound = (double)d_me - pc.rshift + hmx + 4.0 * fabs((double)d_me) * 1.2e-7;
        } else {     // A = cj, B = ci
          pfx = dsub_f(xi, item.org[0]); pfy = dsub_f(yi, item.org[1]); pfz = dsub_f(zi, item.org[2]);
          pc.dj_min = (double)d.sort_d[segA];
          pc.di_max = (double)d.sort_d[segB + cB.count - 1];
          pc.hi_max_g = __dmul_rn((double)hmaxB, (double)kc.gamma);
          pc.hj_max_g = __dmul_rn((double)hmaxA, (double)kc.gamma);
          const double hmx = (double)fmaxf(me.h, hmaxB) * (double)kc.gamma * 1.000001 + 1e-30;
          bound = (double)d_me + pc.rshift - hmx - 4.0 * fabs((double)d_me) * 1.2e-7;
        }
      }
      int done = !mine;
      for (int base = 0; base < cB.count; base += kTile) {
        const int nt = min(kTile, cB.count - base);
        if (__syncthreads_and(done)) break;
        for (int k = threadIdx.x; k < nt; k += kBlock) {
          int lj;
          if (self) {
            lj = base + k;
          } else {
            const int e = left ? (base + k) : (cB.count - 1 - base - k);
            lj = d.sort_i[segB + e];
            t.d[k] = d.sort_d[segB + e];
          }
          const int64_t pj = cB.first + lj;
          const double xj = d.x[pj], yj = d.y[pj], zj = d.z[pj];
          if (self) {
            t.x[k] = xj; t.y[k] = yj; t.z[k] = zj;
            t.d[k] = __int_as_float(lj);
          } else if (left) {
            t.fx[k] = dsub_f(xj, item.org[0]); t.fy[k] = dsub_f(yj, item.org[1]); t.fz[k] = dsub_f(zj, item.org[2]);
          } else {
            t.fx[k] = dsub_f(xj, oix); t.fy[k] = dsub_f(yj, oiy); t.fz[k] = dsub_f(zj, oiz);
          }
          t.vm[k] = d.vm[pj];
          t.fin[k] = d.fin[pj];
          t.h[k] = d.h[pj];
          t.bal[k] = d.balsara[pj];
          t.bin[k] = d.time_bin[pj];
          if (rec) t.id[k] = d.id[pj];
        }
        __syncthreads();
        if (!done) {
          for (int k = 0; k < nt; k++) {
            float dx, dy, dz, r2;
            bool hit;
            const float hj = t.h[k];
            if (self) {
              if (base + k == li) continue;
              dx = dsub_f(xi, t.x[k]); dy = dsub_f(yi, t.y[k]); dz = dsub_f(zi, t.z[k]);
              r2 = dist2(dx, dy, dz);
              hit = (r2 < hig2) || (r2 < sq_gamma2(hj));
            } else {
              const float dk = t.d[k];
              if (left ? !((double)dk < bound) : !((double)dk > bound)) {
                done = 1;
                break;
              }
              dx = __fsub_rn(pfx, t.fx[k]); dy = __fsub_rn(pfy, t.fy[k]); dz = __fsub_rn(pfz, t.fz[k]);
              r2 = dist2(dx, dy, dz);
              const float hjg2 = sq_gamma2(hj);
              if (left)
                hit = force_loop1(pc, d_me, me.h, dk, r2, hig2) || force_loop2(pc, d_me, dk, hj, r2, hig2, hjg2);
              else
                hit = force_loop1(pc, dk, hj, d_me, r2, hjg2) || force_loop2(pc, dk, d_me, me.h, r2, hjg2, hig2);
            }
            if (hit) {
              ForceJ pj;
              const float4 vmj = t.vm[k];
              const float4 fj = t.fin[k];
              pj.h = hj; pj.m = vmj.w; pj.vx = vmj.x; pj.vy = vmj.y; pj.vz = vmj.z;
              pj.rho = fj.x; pj.P = fj.y; pj.f = fj.z; pj.c = fj.w;
              pj.bal = t.bal[k]; pj.bin = t.bin[k];
              iact_force<SCHEME>(acc, r2, dx, dy, dz, me, pj);
              nhits++;
              if (rec) record_pair(d, 2, idi, t.id[k]);
            }
          }
        }
      }
      __syncthreads();
    }
    if (mine) {
      float *o = reinterpret_cast<float *>(&d.acc[pi]);
      atomicAdd(o + 0, acc.ax);
      atomicAdd(o + 1, acc.ay);
      atomicAdd(o + 2, acc.az);
      atomicAdd(o + 3, acc.u_dt);
      atomicAdd(&d.h_dt[pi], acc.h_dt);
      if (acc.v_sig > 0.f) atomicMax(reinterpret_cast<int *>(&d.v_sig[pi]), __float_as_int(acc.v_sig));
      atomicMin(&d.min_bin[pi], acc.min_bin);
    }
    __syncthreads();
  }
  const unsigned long long tot = block_sum(nhits);
  if (threadIdx.x == 0 && tot) atomicAdd(&d.hits[2], tot);
}

// runner_do_end_hydro_force (src/runner_others.c:618-690): hydro_end_force, Minimal/hydro.h:878-882
__global__ void k_end_force(Dev d) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= d.n) return;
  if (d.time_bin[p] <= kc.max_active_bin) d.h_dt[p] *= d.h[p] * 0.3333333333f;
}

}  // namespace swiftcuda
