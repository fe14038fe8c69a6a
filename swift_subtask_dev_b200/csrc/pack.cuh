// K1 / K7: particle pack and unpack on the device (role of cache_read_particles,
// src/cache.h:183, and of the write-back into struct part).  The caller's
// arrays are copied to the device as they are (caller order, one cudaMemcpy
// each, straight from the caller's buffers) and permuted into the cell-ordered
// SoA layout by k_pack; k_unpack writes the results back in caller order into
// flat device arrays that are then copied out, one cudaMemcpy per field.  No
// per-particle work is left on the host.  Both kernels are HBM-bound gathers.
#pragma once
#include "device_types.cuh"

namespace swiftcuda {

// Raw caller-order inputs on the device.
struct RawIn {
  const double *x;   // [n][3]
  const float *v;    // [n][3]
  const float *mass, *h, *u;
  const int8_t *time_bin;                                   // may be null
  const int32_t *mat_id;                                    // may be null
  const int64_t *id;                                        // may be null
  const float *rho, *pressure, *f, *soundspeed, *balsara;   // carried state, each may be null
  const int32_t *order;                                     // order[k] = caller index of cell-order slot k; null = identity
};

// Caller-order outputs on the device (null = not requested).
struct RawOut {
  float *h, *rho, *pressure, *soundspeed, *f, *balsara, *a, *u_dt, *h_dt, *v_sig;
  int8_t *min_bin;
  float *d_rho, *d_rho_dh, *d_wcount, *d_wcount_dh, *d_div_v, *d_rot_v;
  const int32_t *order;
};

struct PackDst {  // mutable views of the Dev arrays k_pack fills
  PosD *pos;
  float4 *vm, *fin, *hb;
  float *h, *h0, *u;
  int8_t *time_bin;
  int32_t *mat_id;
  int64_t *id;
};

// what: 1 = positions, 2 = everything else, 3 = both (the overlapped step uploads and packs the
// positions first so that the sorts can start while the other fields are still on the bus)
__global__ void k_pack(int64_t n, RawIn in, PackDst o, int with_static, int what) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t s = in.order ? in.order[k] : k;
  if (what & 1) {
    PosD q;
    q.x = in.x[3 * s]; q.y = in.x[3 * s + 1]; q.z = in.x[3 * s + 2]; q.pad = 0.0;
    o.pos[k] = q;
  }
  if (!(what & 2)) return;
  o.vm[k] = make_float4(in.v[3 * s], in.v[3 * s + 1], in.v[3 * s + 2], in.mass[s]);
  const float h = in.h[s];
  o.h[k] = h;
  o.h0[k] = h;
  o.u[k] = in.u[s];
  const int8_t tb = in.time_bin ? in.time_bin[s] : (int8_t)1;
  o.time_bin[k] = tb;
  // state carried by inactive particles: p->rho, p->force.{pressure, f, soundspeed, balsara}
  const float rho = in.rho ? in.rho[s] : 0.f;
  const float P = in.pressure ? in.pressure[s] : 0.f;
  o.fin[k] = make_float4(rho, P, in.f ? in.f[s] : 0.f, in.soundspeed ? in.soundspeed[s] : 0.f);
  o.hb[k] = make_float4(h, in.balsara ? in.balsara[s] : 0.f, __int_as_float((int)tb), P / (rho * rho));
  if (with_static) {
    o.mat_id[k] = in.mat_id ? in.mat_id[s] : 0;
    o.id[k] = in.id ? in.id[s] : s;
  }
}

__global__ void k_unpack(Dev d, RawOut o, const float4 *dens0_first, const float4 *dens1_first) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= d.n) return;
  const int64_t s = o.order ? o.order[k] : k;
  if (o.h) o.h[s] = d.h[k];
  if (o.rho || o.pressure || o.f || o.soundspeed) {
    const float4 f = d.fin[k];
    if (o.rho) o.rho[s] = f.x;
    if (o.pressure) o.pressure[s] = f.y;
    if (o.f) o.f[s] = f.z;
    if (o.soundspeed) o.soundspeed[s] = f.w;
  }
  if (o.balsara) o.balsara[s] = d.hb[k].y;
  if (o.a || o.u_dt) {
    const float4 a = d.acc[k];
    if (o.a) {
      o.a[3 * s] = a.x;
      o.a[3 * s + 1] = a.y;
      o.a[3 * s + 2] = a.z;
    }
    if (o.u_dt) o.u_dt[s] = a.w;
  }
  if (o.h_dt) o.h_dt[s] = d.h_dt[k];
  if (o.v_sig) o.v_sig[s] = d.v_sig[k];
  if (o.min_bin) o.min_bin[s] = (int8_t)d.min_bin[k];
  if (o.d_rho || o.d_rho_dh || o.d_wcount || o.d_wcount_dh) {
    const float4 f = dens0_first[k];
    if (o.d_rho) o.d_rho[s] = f.x;
    if (o.d_rho_dh) o.d_rho_dh[s] = f.y;
    if (o.d_wcount) o.d_wcount[s] = f.z;
    if (o.d_wcount_dh) o.d_wcount_dh[s] = f.w;
  }
  if (o.d_div_v || o.d_rot_v) {
    const float4 f = dens1_first[k];
    if (o.d_div_v) o.d_div_v[s] = f.x;
    if (o.d_rot_v) {
      o.d_rot_v[3 * s] = f.y;
      o.d_rot_v[3 * s + 1] = f.z;
      o.d_rot_v[3 * s + 2] = f.w;
    }
  }
}

// c->hydro.h_max of every cell from the particles' h (after a pack): one warp per cell
__global__ void k_cell_hmax_init(Dev d, float *hmax, float *hmax_old) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= d.ncells) return;
  const CellDev cell = d.cells[c];
  float m = 0.f;
  for (int k = lane; k < cell.count; k += 32) m = fmaxf(m, d.h[cell.first + k]);
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) {
    hmax[c] = m;
    hmax_old[c] = m;
  }
}

}  // namespace swiftcuda
