/* swift_cuda.h -- C ABI of the B200-native SPH neighbour-loop backend.
 *
 * This is the whole drop-in boundary: plain pointers and sizes, no C++ or
 * torch types.  It replaces, for one (sub-)step of a two-loop SPH scheme
 * (Minimal or Planetary), the reference's task subgraph
 *
 *   sort -> density(self/pair/sub) -> ghost -> force(self/pair/sub) -> end_force
 *
 * i.e. the operators declared in
 *   src/runner_doiact_hydro.h:168-191   runner_do{self,pair}{1,2}_branch_{density,force},
 *                                       runner_dosub_{self,pair}{1,2}_{density,force},
 *                                       runner_do{self,pair}_subset_branch_density,
 *                                       runner_dosub_subset_density
 *   src/runner.h:94-119                 runner_do_ghost, runner_do_hydro_sort,
 *                                       runner_do_end_hydro_force
 * and called from src/runner_main.c:196-204, 246-254, 294-302, 340-348,
 * 386-394, 417-418, 455-456.  INTEGRATION.md shows the C shim a SWIFT
 * maintainer adds in src/cuda/ to route those tasks here.
 *
 * Error convention: every function returns 0 on success, non-zero on failure;
 * swift_cuda_last_error() gives the message.  (The reference aborts through
 * error(), src/error.h:42-78; the shim turns a non-zero status into error().)
 * There is no CPU fallback: without a CUDA device every compute entry fails. */
#ifndef SWIFT_CUDA_H
#define SWIFT_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SWIFT_CUDA_ABI_VERSION 4

/* SPHENIX (the reference's configure default, configure.ac:1958-1959) is the three-loop scheme: see the
 * "SPHENIX" section below for what it adds and which entries support it. */
enum { SWIFT_CUDA_SCHEME_MINIMAL = 0, SWIFT_CUDA_SCHEME_PLANETARY = 1, SWIFT_CUDA_SCHEME_SPHENIX = 2 };

/* How the per-particle FIELDS are computed.  Which pairs interact -- the neighbour lists and
 * counts of every loop, and therefore h -- is bit-exact in both modes.
 *   EXACT  every field bit-identical to the single-threaded reference built without FMA
 *          contraction: reference operand order, IEEE division, hits consumed in task order;
 *   FAST   the force loop's arithmetic uses FMA contraction and MUFU reciprocals / rsqrt:
 *          a, u_dt, h_dt, v_sig agree with the reference within the tolerances of the project
 *          (<= 1e-4 relative, max over particles); h, rho, P, c, f, balsara stay bit-identical. */
enum { SWIFT_CUDA_FIELDS_EXACT = 0, SWIFT_CUDA_FIELDS_FAST = 1 };

/* Tillotson material, INTERNAL units -- a copy of the reference's
 * struct Til_params after convert_units_Til
 * (src/equation_of_state/planetary/tillotson.h:43-143). */
typedef struct {
  float rho_0, a, b, A, B, u_0, u_iv, u_cv, alpha, beta, eta_min, eta_zero, P_min;
  int mat_id;
} swift_cuda_til;

/* Run-time scalars the operators read through r->e:
 * e->s->dim/periodic, e->hydro_properties->{eta_neighbours, h_tolerance, h_max,
 * h_min, max_smoothing_iterations, viscosity.alpha}, e->max_active_bin,
 * e->cosmology->{a, H}. */
typedef struct {
  double dim[3];
  int periodic;
  float eta_neighbours;
  float h_tolerance;
  float h_max;
  float h_min;
  int max_smoothing_iterations;
  float viscosity_alpha;
  int max_active_bin;
  int scheme;      /* SWIFT_CUDA_SCHEME_* */
  float cosmo_a;   /* 1 without cosmology */
  float cosmo_H;   /* 0 without cosmology */
  int n_til;
  swift_cuda_til til[8];
  float idg_gamma[4]; /* planetary ideal-gas gammas for mat_id 0..3 */
  int fields_mode;    /* SWIFT_CUDA_FIELDS_* */
} swift_cuda_params;

/* One cell of the caller's tree: struct cell {loc, width, hydro.parts - parts,
 * hydro.count, parent, progeny, depth, split} (src/cell.h:354). */
typedef struct {
  double loc[3];
  double width[3];
  int first;
  int count;
  int parent;
  int progeny[8];
  int depth;
  int split;
} swift_cuda_cell;

/* One hydro task as left by scheduler_splittask_hydro (src/scheduler.c:1154):
 * type 0 self, 1 pair, 2 sub_self, 3 sub_pair. */
typedef struct {
  int type;
  int ci;
  int cj;
} swift_cuda_task;

/* Particle arrays, host memory, n entries each ([n][3] row-major for x, v).
 * Field meaning as struct part (src/hydro/Minimal/hydro_part.h:98-189). */
typedef struct {
  int64_t n;
  const double *x;
  const float *v;
  const float *mass;
  const float *h;
  const float *u;
  const int32_t *mat_id;  /* Planetary only; may be NULL */
  const int8_t *time_bin; /* may be NULL: all particles in bin 1 */
  const int64_t *id;      /* may be NULL: ids are the input positions */
  /* State that INACTIVE particles (time_bin > max_active_bin) carry over from
   * their last active step and that the force loop of their active neighbours
   * reads: p->rho, p->force.{pressure, f, soundspeed, balsara}.  Each may be
   * NULL (zeros); ignored for active particles, which get theirs from the ghost. */
  const float *rho, *pressure, *f, *soundspeed, *balsara;
} swift_cuda_parts_in;

/* Results, host memory, written in the caller's particle order; any pointer
 * may be NULL.  a[n][3]. */
typedef struct {
  float *h;          /* p->h after the ghost */
  float *rho;        /* p->rho */
  float *pressure;   /* p->force.pressure */
  float *soundspeed; /* p->force.soundspeed */
  float *f;          /* p->force.f (grad-h term) */
  float *balsara;    /* p->force.balsara */
  float *a;          /* p->a_hydro */
  float *u_dt;       /* p->u_dt */
  float *h_dt;       /* p->force.h_dt after hydro_end_force */
  float *v_sig;      /* p->force.v_sig */
  int8_t *min_ngb_time_bin; /* p->limiter_data.min_ngb_time_bin */
  /* raw neighbour sums of the FIRST density pass (before hydro_end_density) */
  float *d_rho, *d_rho_dh, *d_wcount, *d_wcount_dh, *d_div_v, *d_rot_v;
} swift_cuda_parts_out;

/* Integer work counters of the last step (BASELINE.md "Metric"). */
typedef struct {
  long long n_density;      /* ordered hits of the first density pass */
  long long n_density_redo; /* ordered hits of all ghost redo sweeps */
  long long n_force;        /* ordered hits of the force loop */
  int ghost_sweeps;         /* redo sweeps that had work */
  int ghost_unconverged;    /* particles left unconverged (error if > 0) */
  long long n_leaf_interactions; /* direct self + pair cell interactions */
  long long n_sort_entries;      /* sort entries built */
  int n_kernel_launches;         /* kernels launched by the last step */
} swift_cuda_counters;

/* Device timings of the last step, milliseconds (CUDA events). */
typedef struct {
  float sort_ms, density_ms, ghost_ms, force_ms, total_ms;
} swift_cuda_timings;

typedef struct swift_cuda_ctx swift_cuda_ctx;

int swift_cuda_abi_version(void);
/* sizeof of the ABI structs as compiled into the library, for binding checks:
 * 0 params, 1 cell, 2 task, 3 parts_in, 4 parts_out, 5 counters, 6 timings, 7 til, 8 integrator,
 * 9 state_out, 10 part_layout, 11 sphenix_props, 12 sphenix_in, 13 sphenix_out. */
size_t swift_cuda_sizeof(int which);
const char *swift_cuda_last_error(void);

/* Context on CUDA device `device`. */
int swift_cuda_create(const swift_cuda_params *params, int device, swift_cuda_ctx **out);
void swift_cuda_destroy(swift_cuda_ctx *ctx);

/* (A) Library-built tree: bins the particles into the top-level grid and
 * octree exactly as space_regrid / space_split do (src/space_regrid.c:120-301,
 * src/space_split.c:199-262), creates and splits the hydro tasks
 * (src/engine_maketasks.c:4060-4182, src/scheduler.c:1154-1378), packs the
 * particles to the device in cell order.  Role of cache.h / space_rebuild. */
int swift_cuda_build_from_particles(swift_cuda_ctx *ctx, const swift_cuda_parts_in *parts);

/* (B) Caller-built tree (the drop-in path): particles are already in cell
 * order (a cell's particles are [first, first+count)), `cells` and `tasks`
 * describe the caller's struct cell tree and density task list. */
int swift_cuda_build_from_cells(swift_cuda_ctx *ctx, const swift_cuda_parts_in *parts, int ncells,
                                const swift_cuda_cell *cells, int ntasks, const swift_cuda_task *tasks);

/* Refresh particle state (x, v, h, u, time_bin) after a drift with the tree
 * unchanged; same order as the build call. */
int swift_cuda_update_particles(swift_cuda_ctx *ctx, const swift_cuda_parts_in *parts);

/* Multi-GPU hosts: only the first `count` particles (caller order) are moved by the host <->
 * device copies of swift_cuda_update_particles / swift_cuda_get_results / swift_cuda_step_io; the
 * others (halo copies, refreshed over NVLink by the halo entries below) keep their device state.
 * count < 0: all particles again.  Reset by a build call. */
int swift_cuda_set_transfer_count(swift_cuda_ctx *ctx, int64_t count);

/* Puts the device copy of the particle state (h and the cells' h_max) back to
 * what the last build / update call uploaded, without touching host memory:
 * repeats a step from the same initial smoothing lengths. */
int swift_cuda_restore_particles(swift_cuda_ctx *ctx);

/* One density -> ghost -> force -> end_force pass on the device.
 * flags: SWIFT_CUDA_STEP_SORT rebuilds the 13-axis sorts first
 * (runner_do_hydro_sort); SWIFT_CUDA_STEP_RESTORE first puts h back to the
 * values of the last build / update call (device-side, inside the timed
 * region), so that a repeated step does the work of the first one. */
enum { SWIFT_CUDA_STEP_SORT = 1, SWIFT_CUDA_STEP_RESTORE = 2 };
int swift_cuda_step(swift_cuda_ctx *ctx, int flags);

/* swift_cuda_update_particles + swift_cuda_step + swift_cuda_get_results as ONE call, with the
 * transfers overlapped with the kernels (host buffers should be page-locked for that): the
 * positions are uploaded first and the sorts start while v, m, h, u are still on the bus; what
 * the ghost leaves (h, rho, P, c, f, balsara) is copied out while the force loop runs.  Same
 * results, bit for bit, as the three separate calls. */
int swift_cuda_step_io(swift_cuda_ctx *ctx, const swift_cuda_parts_in *parts, int flags,
                       const swift_cuda_parts_out *out);

/* ---- The caller's AoS particle records (struct part) in, the same records out ------------------
 * The reference keeps its particles as an array of 128-byte `struct part`
 * (src/hydro/Minimal/hydro_part.h:98-189, src/hydro/Planetary/hydro_part.h:100-192; the density-loop
 * fields share a union with the force-loop fields) and its operators update those records IN PLACE.
 * These entries take the array as it is: the records are copied to the device whole, unpacked there
 * into the SoA caches (one coalesced read of the record per particle; role of cache_read_particles,
 * src/cache.h:183) and the results are written back into the device copy of the records and copied
 * out whole.  The library never sees the struct: the shim compiled against the reference's headers
 * passes the byte offsets (offsetof).  An offset < 0 means the record has no such field. */
typedef struct {
  int stride;                                          /* sizeof(struct part) */
  int x, v, a_hydro, mass, h, u, u_dt, rho;            /* p->x[3] (double), p->v[3], p->a_hydro[3], ... */
  int rho_dh, wcount, wcount_dh, div_v, rot_v;         /* p->density.* (union with p->force.*) */
  int f, pressure, soundspeed, v_sig, h_dt, balsara;   /* p->force.* */
  int time_bin, min_ngb_time_bin, id, mat_id;          /* p->time_bin, p->limiter_data.min_ngb_time_bin, p->id, p->mat_id */
  /* SPHENIX records (src/hydro/SPHENIX/hydro_part.h, 160 bytes; div_v and v_sig above are then
   * p->viscosity.div_v / p->viscosity.v_sig); -1 in the two-loop schemes */
  int alpha_visc, div_v_previous_step, div_v_dt;       /* p->viscosity.{alpha, div_v_previous_step, div_v_dt} */
  int alpha_diff, laplace_u;                           /* p->diffusion.{alpha, laplace_u} */
  int alpha_visc_max_ngb;                              /* p->force.alpha_visc_max_ngb */
} swift_cuda_part_layout;

/* Phases of swift_cuda_step_parts.  With SWIFT_CUDA_PHASE_KEEP the density loop continues the running
 * sums found in the records (p->rho, p->density.*) instead of starting from hydro_init_part: what a
 * per-task shim of runner_do{self,pair}1_branch_density needs, which the reference calls once per cell
 * pair on records that already hold the contributions of the pairs before.  The force loop always
 * continues p->a_hydro, p->u_dt, p->force.h_dt, p->force.v_sig, p->limiter_data.min_ngb_time_bin as found;
 * without SWIFT_CUDA_PHASE_GHOST in the same call it takes rho, P, f, c, balsara of every particle from
 * the records.  SWIFT_CUDA_PHASE_END_FORCE is runner_do_end_hydro_force (hydro_end_force). */
enum {
  SWIFT_CUDA_PHASE_SORT = 1,
  SWIFT_CUDA_PHASE_DENSITY = 2,
  SWIFT_CUDA_PHASE_GHOST = 4,
  SWIFT_CUDA_PHASE_FORCE = 8,
  SWIFT_CUDA_PHASE_END_FORCE = 16,
  SWIFT_CUDA_PHASE_KEEP = 32,
  SWIFT_CUDA_PHASE_ALL = 1 | 2 | 4 | 8 | 16,
  /* SPHENIX contexts only: the gradient loop (continues p->viscosity.v_sig, p->diffusion.laplace_u,
   * p->force.alpha_visc_max_ngb as found in the records unless SWIFT_CUDA_PHASE_GHOST runs in the same call)
   * and runner_do_extra_ghost */
  SWIFT_CUDA_PHASE_GRADIENT = 64,
  SWIFT_CUDA_PHASE_EXTRA_GHOST = 128,
  SWIFT_CUDA_PHASE_ALL_SPHENIX = 1 | 2 | 4 | 8 | 16 | 64 | 128
};

/* As swift_cuda_build_from_cells, from `n` records in cell order. */
int swift_cuda_build_from_cells_parts(swift_cuda_ctx *ctx, const swift_cuda_part_layout *layout, const void *parts,
                                      int64_t n, int ncells, const swift_cuda_cell *cells, int ntasks,
                                      const swift_cuda_task *tasks);
/* As swift_cuda_build_from_particles (library-built tree), from `n` records in any order. */
int swift_cuda_build_from_parts(swift_cuda_ctx *ctx, const swift_cuda_part_layout *layout, const void *parts,
                                int64_t n);
/* Uploads the records as they are now (drifted x, v, h, u, time_bin and the running sums), runs the
 * phases, writes what they produce back into the records (active particles only, as the reference):
 *   DENSITY            p->rho, p->density.{rho_dh, wcount, wcount_dh, div_v, rot_v}   (raw sums)
 *   GHOST              p->h, p->rho, p->force.{f, pressure, soundspeed, balsara, v_sig, h_dt}, p->a_hydro, p->u_dt
 *   FORCE / END_FORCE  p->a_hydro, p->u_dt, p->force.{h_dt, v_sig}, p->limiter_data.min_ngb_time_bin
 * SPHENIX records (what the reference's operators of that scheme write, no more):
 *   GHOST              p->h, p->rho, p->force.{f, pressure, soundspeed, balsara, alpha_visc_max_ngb},
 *                      p->viscosity.{div_v, v_sig}, p->diffusion.laplace_u
 *   GRADIENT           p->viscosity.v_sig, p->diffusion.laplace_u, p->force.alpha_visc_max_ngb
 *   EXTRA_GHOST        p->diffusion.{laplace_u, alpha}, p->viscosity.{alpha, div_v_previous_step, div_v_dt},
 *                      p->a_hydro, p->u_dt, p->force.h_dt
 *   FORCE / END_FORCE  p->a_hydro, p->u_dt, p->force.h_dt, p->limiter_data.min_ngb_time_bin
 * `parts` should be page-locked for the copies to run at full PCIe rate. */
int swift_cuda_step_parts(swift_cuda_ctx *ctx, void *parts, int phases);
/* e->max_active_bin of the step that starts (outside the time-integration mode, where
 * swift_cuda_set_time does it): which particles the operators update. */
int swift_cuda_set_max_active_bin(swift_cuda_ctx *ctx, int max_active_bin);

/* Operator-level entry points (same phases, separately callable; used by the
 * per-task shims and the parity tests). */
int swift_cuda_do_sort(swift_cuda_ctx *ctx);
int swift_cuda_do_density(swift_cuda_ctx *ctx);
int swift_cuda_do_ghost(swift_cuda_ctx *ctx);
int swift_cuda_do_force(swift_cuda_ctx *ctx);

int swift_cuda_get_results(swift_cuda_ctx *ctx, const swift_cuda_parts_out *out);
int swift_cuda_get_counters(swift_cuda_ctx *ctx, swift_cuda_counters *out);
int swift_cuda_get_timings(swift_cuda_ctx *ctx, swift_cuda_timings *out);

/* Sustained FP32 FMA rate of the context's GPU in TFLOP/s (microbenchmark;
 * the roofline denominator of the pair kernels).  < 0 on error. */
double swift_cuda_measure_fp32_peak(swift_cuda_ctx *ctx);

/* Tree introspection (library-built tree): sizes, then copies. */
int swift_cuda_get_tree_size(swift_cuda_ctx *ctx, int *ncells, int *ntasks, int cdim[3]);
int swift_cuda_get_tree(swift_cuda_ctx *ctx, swift_cuda_cell *cells, swift_cuda_task *tasks, int32_t *order);

/* Host-only planning (no GPU needed): the tree, task list and flattened
 * interaction count the library would build for these particles.  Arrays are
 * malloc'ed by the library; release with swift_cuda_host_free.  `order` has n
 * entries (caller-allocated). */
int swift_cuda_host_plan(const swift_cuda_params *params, const swift_cuda_parts_in *parts,
                         swift_cuda_cell **cells, int *ncells, swift_cuda_task **tasks, int *ntasks,
                         int32_t *order, int cdim[3], long long *n_interactions, long long *n_items);
void swift_cuda_host_free(void *p);
/* A 64-bit digest of everything the host builder would hand to the device for these particles (tree, tasks,
 * cell order, both item lists, units, lanes, sort segments).  The builder runs on several host threads
 * (SWIFT_CUDA_HOST_THREADS in the environment, default min(16, cores)); with 1 it is the single-threaded form the
 * threaded one must reproduce byte for byte. */
int swift_cuda_host_plan_digest(const swift_cuda_params *params, const swift_cuda_parts_in *parts,
                                unsigned long long *digest);

/* Test hook: record ordered (updated id, neighbour id) pairs of the next
 * phases into device buffers of `capacity` pairs per loop
 * (0 first density pass, 1 redo sweeps, 2 force); fetch copies them out. */
int swift_cuda_record_pairs(swift_cuda_ctx *ctx, long long capacity);
int swift_cuda_fetch_pairs(swift_cuda_ctx *ctx, int loop, long long *pairs, long long capacity, long long *count);

/* Multi-GPU: device pointers of the packed halo buffers so a
 * host can move them with NCCL send/recv (phase 0: xv before density, phase 1:
 * rho after the ghost).  See swift_subtask_dev_b200/multi.py.
 * A SPHENIX context has three exchanges, as the reference under EXTRA_HYDRO_LOOP (send_xv, send_rho,
 * send_gradient, src/engine_maketasks.c:141-324): phase 0 xv; phase 3 after the ghost (h, rho, P, f, c, balsara
 * and what a gradient-loop neighbour is read for: u, the viscosity alpha of before the extra ghost); phase 2
 * after the extra ghost (the evolved viscosity / diffusion alphas the force loop reads). */
int swift_cuda_halo_pack(swift_cuda_ctx *ctx, int phase, const int32_t *d_index, int64_t count, void *d_buffer,
                         void *stream);
int swift_cuda_halo_unpack(swift_cuda_ctx *ctx, int phase, const int32_t *d_index, int64_t count,
                           const void *d_buffer, void *stream);
size_t swift_cuda_halo_bytes_per_particle(int phase);
/* d_index entries are positions in the library's cell order: cell_order_slot =
 * inverse of the `order` table swift_cuda_get_tree returns. */

/* mask[k] = 1 (caller order, n bytes) when particle k lies in a cell that an
 * interaction of an active leaf reads: what a rank must keep fresh in its halo. */
int swift_cuda_get_touched(swift_cuda_ctx *ctx, uint8_t *mask);

/* Recomputes every cell's h_max from the device copy of h.  after_ghost = 0:
 * from scratch, also h_max_old (after an xv refresh); 1: raise to cover the
 * current h (after a rho refresh), as runner_do_ghost leaves it. */
int swift_cuda_refresh_cell_hmax(swift_cuda_ctx *ctx, int after_ghost);

/* ---- Time integration on the device -------------------
 * The step either side of the neighbour loops, so that the particles stay resident in HBM across
 * steps and only snapshots cross the bus.  Replaces, for gas particles without gravity, cosmology or
 * cooling (the time-step limiter has a section of its own below):
 *   runner_do_drift_part -> cell_drift_part   src/cell_drift.c:145-380  (drift_part, src/drift.h:143-233,
 *                                             hydro_predict_extra, src/hydro/<scheme>/hydro.h)
 *   runner_do_kick1 / runner_do_kick2         src/runner_time_integration.c:87-220, 359-500 (kick_part,
 *                                             src/kick.h:215-279, hydro_kick_extra, hydro_reset_predicted_values)
 *   runner_do_timestep                        src/runner_time_integration.c:637-1000 (get_part_timestep,
 *                                             make_integer_timestep, src/timestep.h:45-82, 141-222)
 * Engine-step order (src/engine.c, engine_step), with every particle drifted at every step
 * (engine_policy_drift_all):
 *   set_time(ti_current, get_max_active_bin(ti_current)); drift(ti_old); step(SORT); kick(2); timestep; kick(1)
 * The context must have been built with every cell in the work lists (max_active_bin = 56 at the build:
 * which particles are active then changes per step on the device only).  In this mode the raw first-pass
 * sums (swift_cuda_parts_out.d_*) are not kept. */
typedef struct {
  float cfl_condition;           /* hydro_props->CFL_condition */
  float log_max_h_change;        /* hydro_props->log_max_h_change */
  float minimal_internal_energy; /* hydro_props->minimal_internal_energy */
  double dt_min, dt_max;         /* e->dt_min, e->dt_max */
  double time_base;              /* e->time_base */
} swift_cuda_integrator;

/* The whole per-particle state, host memory, caller order; any pointer may be NULL. */
typedef struct {
  double *x;                     /* p->x */
  float *v, *v_full;             /* p->v (predicted), xp->v_full */
  float *u, *u_full;             /* p->u (predicted), xp->u_full */
  float *h, *rho, *pressure, *soundspeed, *v_sig;
  float *a, *u_dt, *h_dt;
  float *x_diff, *x_diff_sort;   /* xp->x_diff, xp->x_diff_sort */
  int8_t *time_bin, *min_ngb_time_bin;
} swift_cuda_state_out;

/* hydro_first_init_part: v_full = v, u_full = u, x_diff = x_diff_sort = 0 from the uploaded particles. */
int swift_cuda_integrator_init(swift_cuda_ctx *ctx, const swift_cuda_integrator *p);
/* e->ti_current and e->max_active_bin of the step that starts. */
int swift_cuda_set_time(swift_cuda_ctx *ctx, long long ti_current, int max_active_bin);
/* Drifts every particle from ti_old to ti_current, refreshes the cells' h_max; dx_max[0] = largest
 * |x_diff| (c->hydro.dx_max_part over all cells), dx_max[1] = largest |x_diff_sort|: what the host's
 * rebuild / re-sort criterion reads.  dx_max may be NULL. */
int swift_cuda_drift(swift_cuda_ctx *ctx, long long ti_old, float dx_max[2]);
/* which = 1: runner_do_kick1 (particles starting a step); 2: runner_do_kick2 (particles ending one). */
int swift_cuda_kick(swift_cuda_ctx *ctx, int which);
/* New time bins of the active particles.  ti_end_min: the next ti_current; updated: particles that got a
 * new step.  Fails ("wants a time-step below dt_min", as the reference's error()) when one does. */
int swift_cuda_timestep(swift_cuda_ctx *ctx, long long *ti_end_min, long long *updated);
/* Zeroes x_diff_sort (after a re-sort) and/or x_diff (after a rebuild): what = 1 sort, 2 both. */
int swift_cuda_reset_offsets(swift_cuda_ctx *ctx, int what);
int swift_cuda_get_state(swift_cuda_ctx *ctx, const swift_cuda_state_out *out);
/* Device time (CUDA events on the context's stream) of the last drift, kick1, kick2, timestep, ms. */
int swift_cuda_get_integrator_timings(swift_cuda_ctx *ctx, float ms[4]);

/* ---- Time-step limiter (swift --limiter; the remainder of SURVEY.md 8(f) N3) --------------------
 * Replaces
 *   runner_do{self,pair}1_branch_limiter,     src/runner_doiact_limiter.h:94-102 = the DOSELF1 / DOPAIR1 /
 *   runner_dosub_{self,pair}1_limiter         DOSUB_*1 walks of src/runner_doiact_functions_limiter.h:233-828
 *                                             with runner_iact_nonsym_limiter (src/timestep_limiter_iact.h:
 *                                             107-117); call sites src/runner_main.c:205-206, 255-256,
 *                                             303-304, 349-350
 *   runner_do_limiter                         src/runner_time_integration.c:1324-1467 with
 *                                             timestep_limit_part (src/timestep_limiter.h:67-217); call
 *                                             site src/runner_main.c:467-468
 * Order in a step (src/engine_maketasks.c: kick2 -> timestep -> limiter loop -> timestep_limiter -> kick1):
 *   ... step; kick(2); timestep; do_limiter_loop; do_limiter; kick(1)
 * the next ti_current being the smaller of what swift_cuda_timestep and swift_cuda_do_limiter return.
 *
 * swift_cuda_do_limiter_loop: for every particle that STARTS a step now (time_bin <= max_active_bin, on the
 * bins the time-step task has just given) and every neighbour j within its kernel support whose bin is more
 * than time_bin_neighbour_max_delta_bin (2) above: wakeup[j] = max(wakeup[j], -time_bin_i).  Two starting
 * particles of the same leaf cell that lie in each other's support do not wake one another (DOSELF1 hands them
 * to the symmetric runner_iact_limiter, which does nothing); in a pair of cells they do -- as the reference.
 * Positions, smoothing lengths, sorts and work lists are those of the step just done (kick2 and the time-step
 * task move neither x nor h).
 *   time_bin: NULL in the time-integration mode (the bins are on the device).  A caller that runs
 *             runner_do_timestep itself passes p->time_bin of all n particles (caller order).
 *   wakeup:   NULL, or n bytes (caller order) that receive p->limiter_data.wakeup (time_bin_not_awake = -56
 *             where nobody asked) for the caller's own runner_do_limiter.
 * Single device: a woken halo copy is not sent back to its owner (the reference's limiter pack / unpack
 * tasks, src/runner_main.c:539-543, have no counterpart in swift_subtask_dev_b200/multi.py). */
int swift_cuda_do_limiter_loop(swift_cuda_ctx *ctx, const signed char *time_bin, signed char *wakeup);
/* runner_do_limiter on the device state (time-integration mode): every woken particle takes the bin
 * -wakeup + 2; an inactive one has the kick1 of its running step undone, is kicked to the beginning of the
 * new step that contains ti_current and, unless that step starts now, given its first half-kick.
 * ti_end_min: earliest new end of step (2^57 if nobody was woken); limited: how many were. */
int swift_cuda_do_limiter(swift_cuda_ctx *ctx, long long *ti_end_min, long long *limited);
/* Device time of the last limiter loop and the last swift_cuda_do_limiter, ms. */
int swift_cuda_get_limiter_timings(swift_cuda_ctx *ctx, float ms[2]);

/* ---- Table equations of state of the Planetary scheme -----------------------
 * Replaces, for materials that live in tables,
 *   load_table_SESAME / prepare_table_SESAME / convert_units_SESAME   src/equation_of_state/planetary/sesame.h:149-352
 *   load_table_HM80 / prepare_table_HM80 / convert_units_HM80         planetary/hm80.h:76-178
 *   SESAME_{pressure,soundspeed}_from_internal_energy                 sesame.h:480-594, 605-721
 *   HM80_{pressure,soundspeed}_from_internal_energy                   hm80.h:224-300
 * as called by gas_pressure_from_internal_energy / gas_soundspeed_from_internal_energy of the planetary
 * dispatcher (planetary/equation_of_state.h:910-1146, 1305-1457) from hydro_prepare_force, the drift and kick2.
 * SESAME-style tables serve SESAME iron / basalt / water (mat_id 300-302), SS08 water (303), ANEOS forsterite /
 * iron / Fe85Si15 (400-402) and custom materials (900-909); HM80 tables HHe / ice / rock (200-202).
 * The reference's table FILES are not part of its tree (examples/Planetary/EoSTables/get_eos_tables.sh downloads
 * them): parity is pinned on synthetic tables in the reference's file formats, read by the reference's own
 * loaders and evaluated by its own functions (tests/test_eos_tables.py). */
typedef struct {
  int kind;              /* 1 SESAME-style, 2 HM80 */
  int mat_id;            /* enum eos_planetary_material_id */
  int num_rho, num_T;    /* SESAME: without the rho = 0 / T = 0 entries of the file; HM80: num_rho, num_u */
  float *log_rho;        /* SESAME [num_rho]: log(rho), internal units */
  float *log_u;          /* SESAME [num_rho][num_T]: log(u) */
  float *P;              /* SESAME [num_rho][num_T]: P; HM80 [num_rho][num_u]: log(P) */
  float *c;              /* SESAME [num_rho][num_T]: sound speed */
  float P_tiny, c_tiny;  /* SESAME */
  float log_rho_min, log_rho_step, inv_log_rho_step; /* HM80 */
  float log_u_min, log_u_step, inv_log_u_step;       /* HM80 */
  float bulk_mod, P_min_for_c_min;                   /* HM80 */
} swift_cuda_eos_table;

/* units_cgs_conversion_factor(&si, X) / units_cgs_conversion_factor(us, X) for X = UNIT_CONV_DENSITY,
 * _ENERGY_PER_UNIT_MASS, _PRESSURE, _SPEED, as the caller's unit system gives them (doubles: the tables are
 * scaled in the reference's operand order, float *= double and float += logf(double)). */
typedef struct {
  double density, energy_per_unit_mass, pressure, speed;
} swift_cuda_eos_units;

/* Host only (no GPU needed).  Read a table file in the reference's text format and make it what the reference
 * holds after load + prepare + convert_units.  version_date: what set_<material> expects (20220714 for the
 * SESAME / ANEOS materials, 0 = any for custom ones; 20201003 for HM80, whose bulk modulus and P_min_for_c_min
 * follow from mat_id as in set_HM80_*).  The arrays belong to the table: swift_cuda_eos_table_free. */
int swift_cuda_eos_table_load_sesame(const char *path, int mat_id, int version_date, const swift_cuda_eos_units *units,
                                     swift_cuda_eos_table *out);
int swift_cuda_eos_table_load_hm80(const char *path, int mat_id, const swift_cuda_eos_units *units,
                                   swift_cuda_eos_table *out);
/* The same table as one flat little-endian binary file (64-byte header, then the float arrays): what a
 * production run reads instead of parsing 1e6 numbers of text at every start. */
int swift_cuda_eos_table_write_bin(const char *path, const swift_cuda_eos_table *table);
int swift_cuda_eos_table_read_bin(const char *path, swift_cuda_eos_table *out);
void swift_cuda_eos_table_free(swift_cuda_eos_table *table);
/* Copies the table to the context's device (at most 8 tables per context, Planetary scheme only); particles
 * whose mat_id equals the table's are evaluated from it from then on; a second table for the same mat_id replaces
 * the first.  Tables outlive rebuilds of the context.  The caller must set a table for every table material its
 * particles carry: a mat_id >= 200 without one is NOT detected (it is evaluated as the first Tillotson material,
 * as any unknown material id has been since round 1). */
int swift_cuda_set_eos_table(swift_cuda_ctx *ctx, const swift_cuda_eos_table *table);

/* ---- Flat binary particle files (the other half of SURVEY.md 8(f) N4) ---------------------------
 * The field set, names and types of the reference's HDF5 initial conditions and snapshots for gas
 * (hydro_read_particles / hydro_write_particles, src/hydro/Planetary/hydro_io.h:55-84, 190-228; the other
 * schemes lack MaterialIDs), as one little-endian file: a 128-byte header (magic "SWCUPRT1", n, the box, a mask
 * of the fields present), then the arrays in the order below, each padded to 16 bytes.  HDF5 is not needed
 * (nor present here): a converter from the reference's files is a dozen lines of h5py on a machine that has it.
 * Values are in internal units; nothing is converted.  Host only. */
#define SWIFT_CUDA_FIELD_COORDINATES 1       /* double [n][3]   p->x          compulsory */
#define SWIFT_CUDA_FIELD_VELOCITIES 2        /* float  [n][3]   p->v          compulsory */
#define SWIFT_CUDA_FIELD_MASSES 4            /* float  [n]      p->mass       compulsory */
#define SWIFT_CUDA_FIELD_SMOOTHING_LENGTHS 8 /* float  [n]      p->h          compulsory */
#define SWIFT_CUDA_FIELD_INTERNAL_ENERGIES 16 /* float [n]      p->u          compulsory */
#define SWIFT_CUDA_FIELD_PARTICLE_IDS 32     /* int64  [n]      p->id         compulsory */
#define SWIFT_CUDA_FIELD_MATERIAL_IDS 64     /* int32  [n]      p->mat_id     Planetary */
#define SWIFT_CUDA_FIELD_DENSITIES 128       /* float  [n]      p->rho        optional (snapshots) */
#define SWIFT_CUDA_FIELD_ACCELERATIONS 256   /* float  [n][3]   p->a_hydro    optional */
#define SWIFT_CUDA_FIELD_PRESSURES 512       /* float  [n]      convert_P     optional (snapshots) */
#define SWIFT_CUDA_FIELDS_COMPULSORY 63
typedef struct {
  int64_t n;
  double dim[3];          /* box size (Header/BoxSize) */
  double time;            /* Header/Time */
  int fields;             /* SWIFT_CUDA_FIELD_* present; absent arrays are NULL */
  double *coordinates;
  float *velocities, *masses, *smoothing_lengths, *internal_energies;
  long long *particle_ids;
  int *material_ids;
  float *densities, *accelerations, *pressures;
} swift_cuda_particle_file;
/* Writes the arrays whose bit is set in `fields` (the compulsory ones must be). */
int swift_cuda_particles_write_bin(const char *path, const swift_cuda_particle_file *p);
/* Reads a file written by the above; the arrays belong to the struct (swift_cuda_particles_free).  The arrays
 * of a read file can go straight into swift_cuda_parts_in. */
int swift_cuda_particles_read_bin(const char *path, swift_cuda_particle_file *out);
void swift_cuda_particles_free(swift_cuda_particle_file *p);

/* ---- SPHENIX: the gradient loop and the extra ghost -------------------------
 * params.scheme = SWIFT_CUDA_SCHEME_SPHENIX (+ ideal gas) turns the step into the reference's
 * EXTRA_HYDRO_LOOP subgraph (src/engine_maketasks.c:2306-2317, src/runner_main.c:199-200, 249-250,
 * 297-298, 343-344, 420-421):
 *
 *   sort -> density -> ghost -> GRADIENT (self/pair/sub) -> extra_ghost -> force -> end_force
 *
 *   runner_do{self,pair}1_branch_gradient, runner_dosub_{self,pair}1_gradient   src/runner_doiact_hydro.h:168-191
 *     with runner_iact_nonsym_gradient                                          src/hydro/SPHENIX/hydro_iact.h:272-329
 *   runner_do_ghost: hydro_prepare_gradient + hydro_reset_gradient at convergence   src/runner_ghost.c:1249-1264
 *   runner_do_extra_ghost: hydro_end_gradient, hydro_prepare_force (the viscosity and diffusion
 *     alphas evolve over dt_alpha = get_timestep(p->time_bin, e->time_base)), hydro_reset_acceleration
 *                                                                               src/runner_ghost.c:992-1075
 *   runner_iact_nonsym_force of the scheme                                      src/hydro/SPHENIX/hydro_iact.h:488-609
 *
 * Supported entries: swift_cuda_build_from_particles / _from_cells, swift_cuda_update_particles,
 * swift_cuda_step, swift_cuda_do_*, swift_cuda_get_results (v_sig = p->viscosity.v_sig, d_div_v the raw
 * first-pass sum), the pair recorder (loop 3 = gradient), the time-integration entries (the scheme's
 * hydro_predict_extra / hydro_kick_extra / hydro_compute_timestep are Minimal's arithmetic;
 * swift_cuda_integrator_init makes its time_base the one of the extra ghost's dt_alpha) and the entries
 * below; both fields modes (FAST: k_force_fast with the scheme's interaction; the gradient loop and the extra
 * ghost stay exact, they decide the alphas); the struct part entries with the scheme's 160-byte record
 * (swift_cuda_part_layout's SPHENIX offsets, SWIFT_CUDA_PHASE_GRADIENT / _EXTRA_GHOST); the halo entries with the
 * scheme's three exchanges; swift_cuda_step_io (the scheme's own state goes through swift_cuda_sphenix_set_state /
 * _get_results around it). */
typedef struct {
  float viscosity_alpha_max;  /* hydro_props->viscosity.alpha_max, SPH:viscosity_alpha_max (2.0) */
  float viscosity_alpha_min;  /* ->viscosity.alpha_min (0.0) */
  float viscosity_length;     /* ->viscosity.length (0.05) */
  float diffusion_alpha;      /* ->diffusion.alpha: initial value of p->diffusion.alpha (0.0) */
  float diffusion_beta;       /* ->diffusion.beta (1.0) */
  float diffusion_alpha_max;  /* ->diffusion.alpha_max (1.0) */
  float diffusion_alpha_min;  /* ->diffusion.alpha_min (0.0) */
  double time_base;           /* e->time_base */
} swift_cuda_sphenix_props;
/* Before the build call.  Without it a SPHENIX context uses the defaults of
 * src/hydro/SPHENIX/hydro_parameters.h:53-101 and time_base = 1e-3; params.viscosity_alpha is
 * hydro_props->viscosity.alpha, the initial p->viscosity.alpha. */
int swift_cuda_sphenix_set_props(swift_cuda_ctx *ctx, const swift_cuda_sphenix_props *props);

/* The state the scheme carries in struct part from step to step, host arrays in the caller's particle
 * order; NULL: left as it is on the device (after a build: the initial values of the props). */
typedef struct {
  const float *alpha_visc;          /* p->viscosity.alpha */
  const float *alpha_diff;          /* p->diffusion.alpha */
  const float *div_v_previous_step; /* p->viscosity.div_v_previous_step */
} swift_cuda_sphenix_in;
int swift_cuda_sphenix_set_state(swift_cuda_ctx *ctx, const swift_cuda_sphenix_in *in);

/* Results, host arrays in the caller's particle order, any pointer may be NULL. */
typedef struct {
  float *div_v;               /* p->viscosity.div_v (after the ghost) */
  float *laplace_u;           /* p->diffusion.laplace_u: the raw sum after the gradient loop, x 2 / h^(d+1) after the extra ghost */
  float *alpha_visc_max_ngb;  /* p->force.alpha_visc_max_ngb */
  float *alpha_visc;          /* p->viscosity.alpha */
  float *alpha_diff;          /* p->diffusion.alpha */
  float *div_v_previous_step; /* p->viscosity.div_v_previous_step */
  float *div_v_dt;            /* p->viscosity.div_v_dt */
} swift_cuda_sphenix_out;
int swift_cuda_sphenix_get_results(swift_cuda_ctx *ctx, const swift_cuda_sphenix_out *out);

/* Operator-level entries of the two phases the scheme adds (swift_cuda_step runs them between
 * swift_cuda_do_ghost and swift_cuda_do_force). */
int swift_cuda_do_gradient(swift_cuda_ctx *ctx);
int swift_cuda_do_extra_ghost(swift_cuda_ctx *ctx);
/* ordered hits of the gradient loop in the last step (swift_cuda_counters keeps its layout) */
int swift_cuda_sphenix_get_counter(swift_cuda_ctx *ctx, long long *n_gradient);

/* ---- Allocation-free time-step entry ----------------------------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * runner_do_timestep (the arithmetic of swift_cuda_timestep / k_timestep) on arrays the CALLER owns,
 * without a context: no cudaMalloc, cudaFree or allocating library routine is called.  The kernel's
 * reduction scratch (one 16-byte partial per 256-particle block + a 256-byte header: 1 048 832 bytes)
 * is a file-scope `static __device__` array of the library; every GPU that loads the library holds it
 * until exit.  Hence the limit: count <= SWIFT_CUDA_TIMESTEP_STATIC_MAX (65 536 blocks x 256), more is
 * rejected with an error.  Every call on a device shares that array: calls that overlap on the same
 * GPU (two streams, two host threads) race; serialise them.
 *
 * device: the GPU the arrays live on; it is made current (cudaSetDevice).  All array pointers are
 * DEVICE pointers of that GPU (checked with cudaPointerGetAttributes).  time_bin is read and written
 * in place: bins of particles with time_bin <= max_active_bin are replaced by their new bin, the
 * others are left alone (and count towards ti_end_min with the end of their current step).
 * stream: a cudaStream_t of that device, NULL = the CUDA default stream.  The call returns after the
 * stream has been synchronised, with *ti_end_min / *updated as swift_cuda_timestep gives them
 * (either may be NULL); it fails, like swift_cuda_timestep, when a particle wants a step below dt_min. */
#define SWIFT_CUDA_TIMESTEP_STATIC_MAX 16777216LL
typedef struct {
  int64_t count;
  const float *h;                   /* [count] p->h */
  const float *v_sig;               /* [count] p->force.v_sig */
  const float *h_dt;                /* [count] p->force.h_dt */
  const int8_t *min_ngb_time_bin;   /* [count] p->limiter_data.min_ngb_time_bin */
  int8_t *time_bin;                 /* [count] p->time_bin, in and out */
  long long ti_current;             /* e->ti_current */
  int max_active_bin;               /* e->max_active_bin */
  swift_cuda_integrator integ;      /* CFL, log_max_h_change, dt_min, dt_max, time_base */
} swift_cuda_timestep_static_args;
int swift_cuda_timestep_static(int device, const swift_cuda_timestep_static_args *args, long long *ti_end_min,
                               long long *updated, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SWIFT_CUDA_H */
