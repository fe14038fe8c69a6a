// Device-side data layout of the SPH backend (SoA particle caches in HBM, the
// role of src/cache.h in the reference) and the numeric constants of
// src/kernel_hydro.h / src/dimension.h / src/adiabatic_index.h.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/swift_cuda.h"
#include "eos_table.h"
#include "host_tree.hpp"

namespace swiftcuda {

// Numeric constants, evaluated ON THE HOST with the reference's own expressions
// (src/kernel_hydro.h:52-53, 205-241; cubic spline, 3D) and uploaded, so that
// device and CPU use the same bits.
struct Consts {
  float gamma;          // kernel_gamma
  float gamma2;         // kernel_gamma2
  float gamma_inv;      // kernel_gamma_inv
  float kconst;         // kernel_constant
  float ginv_dim;       // kernel_gamma_inv_dim
  float ginv_dim_p1;    // kernel_gamma_inv_dim_plus_one
  float kroot;          // kernel_root
  float eta_dim;        // pow_dimension(eta_neighbours)
  float h_tol, h_max, h_min;
  float alpha, beta;    // viscosity alpha, const_viscosity_beta
  float a, H;           // cosmology
  int max_active_bin;
  int scheme;
  int max_iter;
  int n_til;
  swift_cuda_til til[8];
  float idg_gamma[4];
  double axis[13][3];   // runner_shift
  // SPHENIX (appended: the offsets above are what the two-loop kernels were compiled against):
  // hydro_props->viscosity.{alpha_max, alpha_min, length}, ->diffusion.{beta, alpha_max, alpha_min}, e->time_base
  float visc_alpha_max, visc_alpha_min, visc_length;
  float diff_beta, diff_alpha_max, diff_alpha_min;
  double time_base;
  // table equations of state of the Planetary scheme (appended; eos_table.h): device arrays
  int n_tab;
  int pad_tab_;
  EosTable tab[8];
};

struct CellDev {
  double loc[3];
  int first;
  int count;
  float width;   // largest edge (bounds the f32 error of the conservative candidate cull)
  int pad;
};

struct __align__(16) PosD {
  double x, y, z, pad;
};

struct __align__(8) SortEnt {
  float d;
  int32_t i;
};

// Everything a kernel needs, passed by value.
struct Dev {
  int64_t n;
  int ncells;
  int n_leaves;
  int n_units;
  int unit_lo, unit_hi;   // force loop: walk the units [unit_lo, unit_hi) only (unit_hi == 0: all of them)
  double dim[3];
  // particles, cell order
  const PosD *pos;     // x, y, z f64 in one 32-byte record: a gather touches ONE sector
  float4 *vm;          // vx, vy, vz, mass
  float *h;
  const float *u;
  const int8_t *time_bin;
  const int32_t *mat_id;
  const int64_t *id;
  // density sums
  float4 *dens0;       // rho, rho_dh, wcount, wcount_dh
  float4 *dens1;       // div_v, rot_v[3]
  // force inputs (after ghost)
  float4 *fin;         // rho, pressure, grad-h f, soundspeed
  float4 *hb;          // h, balsara, time_bin (int bits), unused: the rest of a force neighbour
  // force outputs
  float4 *acc;         // a_hydro[3], u_dt
  float *h_dt;
  float *v_sig;
  int *min_bin;
  // ghost state
  float *left, *right;
  uint8_t *gstate;     // 0 idle/inactive, 1 pending, 2 redo this sweep, 3 done
  // tree
  const CellDev *cells;
  float *cell_hmax;        // current c->hydro.h_max
  const int32_t *cell_parent;  // -1 for a top-level cell
  const uint8_t *cell_split;   // 0: leaf
  float *cell_box;         // [ncells][6] bounding box of the cell's particles relative to loc (lo xyz, hi xyz)
  uint32_t *cell_box_u;    // the same while it is being reduced (order-preserving integer images)
  const float *cell_hmax_old;  // value at build (h_max_old)
  const float *recurse_limit;  // [ncells] 0.5 * dmin of the cells the work lists recursed through (+inf: none)
  int *stale;                  // [1] cells whose h_max (+ dx_max) outgrew recurse_limit: the lists need a rebuild
  const int32_t *leaves;       // [n_leaves] cell index
  const int32_t *item_begin;   // [n_leaves+1]
  const Item *items;           // first density pass + force
  const int32_t *redo_begin;   // [n_leaves+1]
  const Item *redo_items;      // ghost redo passes
  const int32_t *unit_leaf;    // [n_units] warp unit -> leaf slot
  const int32_t *unit_begin;   // [n_leaves+1] first unit of a leaf
  const int32_t *lane_part;    // [n_units*32] particle of each lane, or -1
  // leaf-resident kernels (leaf_kernels.cuh): groups of up to kLW consecutive units of one leaf
  const int4 *groups;          // [n_groups] {leaf slot, first unit, units, 0}
  int n_groups;
  // sparse redo sweeps: the leaves that hold a particle flagged in this sweep, pushed by k_ghost
  const int32_t *part_slot;    // [n] leaf slot of a particle (-1: its leaf has no work)
  int *leaf_mark;              // [n_leaves] sweep + 1 of the last push
  int *leaf_list;              // [n_leaves]
  int *leaf_count;             // [max_iter + 2] entries of leaf_list per sweep
  // sorts
  const int64_t *sort_lookup;  // [ncells*13]
  SortEnt *sort;           // struct sort_entry (src/sort_part.h:34-41): key d, index i inside the cell
  uint8_t *sort_need;      // [ncells*13] or null: only these (cell, axis) lists are rebuilt (k_mark_sorts)
  int *unit_list;          // [n_units] or null: the units with an active lane, pushed by k_mark_sorts
  int *unit_count;         // [1]
  // persistent kernels: per-warp FIFO of deferred hits [n_qwarps][kQD][32], unit counters
  // [0] first density pass, [1 + sweep] redo sweeps, [1 + max_iter] force
  float4 *gq;
  int *work_ctr;
  // counters: [0] density, [1] redo, [2] force
  unsigned long long *hits;
  int *redo_count;         // [max_iter+1]
  int *walk_count;         // [max_iter+2] of those, the particles the walking kernel redoes (the others: k_density_list)
  // neighbour lists of the ghost iteration (null: not recorded): [n_units][kLD][32] (dx, dy, dz, j)
  float4 *nl;
  unsigned short *nl_cnt;      // [n_units*32] entries of a lane's list, 0xffff: none / overflowed
  float *nl_R;                 // [n_units*32] (h gamma)^2 the list covers
  float nl_grow;               // the recording sweep covers (nl_grow + nl_slope |h / h_before - 1|) x the search radius
  float nl_slope;
  float *h_before;             // [n] h before the ghost's last update (how far Newton-Raphson just moved it)
  const int32_t *part_lane;    // [n] unit * 32 + lane of a particle, -1: none
  // swift_cuda_step_parts: continue the running density sums / leave hydro_end_force to its own phase
  int keep_sums;
  int no_end_force;
  // pair recording (test hook)
  long long *rec_pairs[3];
  unsigned long long *rec_count;  // [4]
  long long rec_capacity;
  // ---- SPHENIX only (null otherwise; appended, see Consts) ----
  float4 *sph_state;   // p->viscosity.alpha, p->diffusion.alpha, p->viscosity.div_v_previous_step, p->viscosity.div_v_dt
  float4 *sph_grad;    // p->viscosity.div_v, p->diffusion.laplace_u, p->force.alpha_visc_max_ngb, unused
  float4 *gin;         // what a gradient-loop neighbour is read for: u, rho, soundspeed, viscosity.alpha
  float4 *sx;          // the rest of a force-loop neighbour: viscosity.alpha, diffusion.alpha, u, unused
  long long *rec_grad; // pair recording of the gradient loop (counter: rec_count[3])
  int *wakeup;         // time-step limiter loop: p->limiter_data.wakeup per particle (null until the loop is used);
                       // takes the place of the padding that kept sizeof(Dev) a multiple of 16
};

}  // namespace swiftcuda
