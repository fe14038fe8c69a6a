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

// A time bin with SWIFT_CUDA_FOREIGN_BIN added marks a copy of a particle another rank owns: never
// active here (every real bin is <= 56 < 64) while its owner's bin still reaches the neighbour-bin
// limit of the particles that see it (runner_iact_nonsym_timebin reads bin & kBinMask).
constexpr int kBinMask = 63;

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
// m: only caller indices below m were refreshed on the device (swift_cuda_set_transfer_count): the
// others keep their device state
__global__ void k_pack(int64_t n, int64_t m, RawIn in, PackDst o, int with_static, int what) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t s = in.order ? in.order[k] : k;
  if (s >= m) return;
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
  o.hb[k] = make_float4(h, in.balsara ? in.balsara[s] : 0.f, __int_as_float((int)tb & kBinMask), P / (rho * rho));
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

// ---------------------------------------------------------------------------
// K1' / K7': the caller's AoS records (struct part, 128 B) <-> the SoA caches, on the device.
// One thread per particle; a warp reads 32 consecutive records (4 KB, fully coalesced: every byte
// of every sector is a record byte).  order: cell-order slot -> record index (null: identity).
// ---------------------------------------------------------------------------
struct AosDst {
  float4 *dens0, *dens1, *acc;
  float *h_dt, *v_sig;
  int *min_bin;
  float4 *sph_state, *sph_grad;  // SPHENIX records only (null otherwise)
};
template <typename T>
__device__ __forceinline__ T aos_get(const unsigned char *rec, int off) {
  return *reinterpret_cast<const T *>(rec + off);
}
template <typename T>
__device__ __forceinline__ void aos_put(unsigned char *rec, int off, T v) {
  *reinterpret_cast<T *>(rec + off) = v;
}
// sums: 1 = also the running density sums (KEEP); force: 1 = also a_hydro, u_dt, h_dt, v_sig, min_bin
__global__ void k_pack_aos(int64_t n, const unsigned char *aos, swift_cuda_part_layout L, const int32_t *order,
                           PackDst o, AosDst a, int with_static, int sums, int force, int64_t lo = 0) {
  const int64_t k = lo + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;  // n: end of the range of slots
  if (k >= n) return;
  const int64_t s = order ? order[k] : k;
  const unsigned char *r = aos + (size_t)s * L.stride;
  PosD q;
  q.x = aos_get<double>(r, L.x); q.y = aos_get<double>(r, L.x + 8); q.z = aos_get<double>(r, L.x + 16); q.pad = 0.0;
  o.pos[k] = q;
  o.vm[k] = make_float4(aos_get<float>(r, L.v), aos_get<float>(r, L.v + 4), aos_get<float>(r, L.v + 8),
                        aos_get<float>(r, L.mass));
  const float h = aos_get<float>(r, L.h);
  o.h[k] = h;
  o.h0[k] = h;
  o.u[k] = aos_get<float>(r, L.u);
  const int8_t tb = aos_get<int8_t>(r, L.time_bin);
  o.time_bin[k] = tb;
  const float rho = aos_get<float>(r, L.rho);
  if (sums) {
    a.dens0[k] = make_float4(rho, aos_get<float>(r, L.rho_dh), aos_get<float>(r, L.wcount), aos_get<float>(r, L.wcount_dh));
    a.dens1[k] = make_float4(aos_get<float>(r, L.div_v), aos_get<float>(r, L.rot_v), aos_get<float>(r, L.rot_v + 4),
                             aos_get<float>(r, L.rot_v + 8));
  } else {
    // state the force loop reads of particles the ghost does not touch in this call
    const float P = aos_get<float>(r, L.pressure);
    o.fin[k] = make_float4(rho, P, aos_get<float>(r, L.f), aos_get<float>(r, L.soundspeed));
    o.hb[k] = make_float4(h, aos_get<float>(r, L.balsara), __int_as_float((int)tb & kBinMask), P / (rho * rho));
  }
  if (force) {
    a.acc[k] = make_float4(aos_get<float>(r, L.a_hydro), aos_get<float>(r, L.a_hydro + 4),
                           aos_get<float>(r, L.a_hydro + 8), aos_get<float>(r, L.u_dt));
    a.h_dt[k] = aos_get<float>(r, L.h_dt);
    a.v_sig[k] = aos_get<float>(r, L.v_sig);
    a.min_bin[k] = L.min_ngb_time_bin >= 0 ? (int)aos_get<int8_t>(r, L.min_ngb_time_bin) : 57;
  }
  if (a.sph_state != nullptr) {
    // SPHENIX: what the scheme carries in the record from step to step, and the gradient loop's running
    // values (read by active lanes only when the ghost did not run in this call)
    a.sph_state[k] = make_float4(aos_get<float>(r, L.alpha_visc), aos_get<float>(r, L.alpha_diff),
                                 aos_get<float>(r, L.div_v_previous_step), aos_get<float>(r, L.div_v_dt));
    a.sph_grad[k] = make_float4(aos_get<float>(r, L.div_v), aos_get<float>(r, L.laplace_u),
                                aos_get<float>(r, L.alpha_visc_max_ngb), 0.f);
  }
  if (with_static) {
    o.mat_id[k] = L.mat_id >= 0 ? aos_get<int32_t>(r, L.mat_id) : 0;
    o.id[k] = L.id >= 0 ? aos_get<long long>(r, L.id) : (long long)s;
  }
}

// what: last phase that ran -- 2 density (raw sums), 4 ghost, 8 force (16 with the ghost's fields too)
// [lo, hi): the cell-order slots written by this launch (the force loop of swift_cuda_step_parts hands its
// results back in pieces so that the copy of one piece overlaps the loop over the next)
__global__ void k_unpack_aos(Dev d, unsigned char *aos, swift_cuda_part_layout L, const int32_t *order, int what,
                             int max_active_bin, int64_t lo, int64_t hi) {
  const int64_t k = lo + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= hi) return;
  if (d.time_bin[k] > max_active_bin) return;  // the reference's operators update active particles only
  const int64_t s = order ? order[k] : k;
  unsigned char *r = aos + (size_t)s * L.stride;
  if (what == 2) {
    const float4 s0 = d.dens0[k], s1 = d.dens1[k];
    aos_put<float>(r, L.rho, s0.x);
    aos_put<float>(r, L.rho_dh, s0.y);
    aos_put<float>(r, L.wcount, s0.z);
    aos_put<float>(r, L.wcount_dh, s0.w);
    aos_put<float>(r, L.div_v, s1.x);
    aos_put<float>(r, L.rot_v, s1.y);
    aos_put<float>(r, L.rot_v + 4, s1.z);
    aos_put<float>(r, L.rot_v + 8, s1.w);
    return;
  }
  if (what == 4 || what == 16) {
    const float4 f = d.fin[k];
    aos_put<float>(r, L.h, d.h[k]);
    aos_put<float>(r, L.rho, f.x);
    aos_put<float>(r, L.pressure, f.y);
    aos_put<float>(r, L.f, f.z);
    aos_put<float>(r, L.soundspeed, f.w);
    aos_put<float>(r, L.balsara, d.hb[k].y);
  }
  const float4 a = d.acc[k];
  aos_put<float>(r, L.a_hydro, a.x);
  aos_put<float>(r, L.a_hydro + 4, a.y);
  aos_put<float>(r, L.a_hydro + 8, a.z);
  aos_put<float>(r, L.u_dt, a.w);
  aos_put<float>(r, L.h_dt, d.h_dt[k]);
  aos_put<float>(r, L.v_sig, d.v_sig[k]);
  if (what != 4 && L.min_ngb_time_bin >= 0) aos_put<int8_t>(r, L.min_ngb_time_bin, (int8_t)d.min_bin[k]);
}

// c->hydro.h_max of every cell from the particles' h (after a pack): one warp per cell
// (end_lo, end_hi]: only the cells whose last particle lies in this range of cell-order slots
__global__ void k_cell_hmax_init(Dev d, float *hmax, float *hmax_old, int64_t end_lo = -1, int64_t end_hi = INT64_MAX) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= d.ncells) return;
  const CellDev cell = d.cells[c];
  const int64_t end = (int64_t)cell.first + cell.count;
  if (end <= end_lo || end > end_hi) return;
  float m = 0.f;
  for (int k = lane; k < cell.count; k += 32) m = fmaxf(m, d.h[cell.first + k]);
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) {
    hmax[c] = m;
    hmax_old[c] = m;
  }
}

}  // namespace swiftcuda
