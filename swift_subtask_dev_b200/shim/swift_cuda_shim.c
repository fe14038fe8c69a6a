/* swift_cuda_shim.c -- what a SWIFT maintainer adds as src/cuda/swift_cuda_shim.c so that the
 * reference's hydro operators run on a B200 through libswiftcuda.so (include/swift_cuda.h).
 *
 * Plain C, compiled against the REFERENCE's own headers and config.h: struct part, struct cell,
 * struct engine, struct runner are the reference's, never redefined.  It exports
 *
 *  (1) the reference's operator symbols, same names and signatures
 *        src/runner_doiact_hydro.h:168-191   runner_do{self,pair}{1,2}_branch_{density,force},
 *                                            runner_dosub_{self,pair}{1,2}_{density,force}
 *        src/runner.h:94-119                 runner_do_ghost, runner_do_hydro_sort,
 *                                            runner_do_end_hydro_force
 *      so that src/runner_main.c:196-204, 246-254, 386-394, 417-418, 455-456 and the reference's
 *      unit tests (tests/test27cells.c:358-369, tests/test125cells.c:453-465) link unchanged against
 *      this file + libswiftcuda.so INSTEAD OF runner_doiact_hydro.o / runner_ghost.o / runner_sort.o.
 *      Each call packs the one or two cells it is given (struct part records as they are, the
 *      running sums included), runs the one phase on the device and writes the records back in
 *      place: the reference's in-place `+=` semantics.  One launch per cell pair cannot feed a
 *      B200; these per-task entries exist for parity with the reference's tests and call sites;
 *
 *  (2) the batch entries a production integration uses instead: swift_cuda_shim_rebuild(e) after
 *      engine_rebuild (flattens the cell tree and the density tasks of e->sched, hands s->parts over)
 *      and swift_cuda_shim_step(e) in place of the sort / density / ghost / force / end_force task
 *      subgraph of a step (one H2D copy of s->parts, the device step, one D2H copy).
 *
 * Errors: a non-zero status of the library becomes the reference's error() (src/error.h:42-78). */
#include <config.h>

#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#include "active.h"
#include "cell.h"
#include "engine.h"
#include "error.h"
#include "hydro.h"
#include "hydro_properties.h"
#include "runner.h"
#include "scheduler.h"
#include "space.h"
#include "swift_cuda.h"
#include "task.h"
#ifdef PLANETARY_SPH
#include "equation_of_state.h"
#endif

#define CUDA_OK(call)                                   \
  do {                                                  \
    if ((call) != 0) error("%s", swift_cuda_last_error()); \
  } while (0)

/* ---- struct part as the library sees it: byte offsets only ------------------------------------ */
static swift_cuda_part_layout shim_layout(void) {
  swift_cuda_part_layout L;
  L.stride = (int)sizeof(struct part);
  L.x = (int)offsetof(struct part, x);
  L.v = (int)offsetof(struct part, v);
  L.a_hydro = (int)offsetof(struct part, a_hydro);
  L.mass = (int)offsetof(struct part, mass);
  L.h = (int)offsetof(struct part, h);
  L.u = (int)offsetof(struct part, u);
  L.u_dt = (int)offsetof(struct part, u_dt);
  L.rho = (int)offsetof(struct part, rho);
  L.rho_dh = (int)offsetof(struct part, density.rho_dh);
  L.wcount = (int)offsetof(struct part, density.wcount);
  L.wcount_dh = (int)offsetof(struct part, density.wcount_dh);
#ifdef SPHENIX_SPH
  L.div_v = (int)offsetof(struct part, viscosity.div_v);
#else
  L.div_v = (int)offsetof(struct part, density.div_v);
#endif
  L.rot_v = (int)offsetof(struct part, density.rot_v);
  L.f = (int)offsetof(struct part, force.f);
  L.pressure = (int)offsetof(struct part, force.pressure);
  L.soundspeed = (int)offsetof(struct part, force.soundspeed);
#ifdef SPHENIX_SPH
  L.v_sig = (int)offsetof(struct part, viscosity.v_sig);
#else
  L.v_sig = (int)offsetof(struct part, force.v_sig);
#endif
  L.h_dt = (int)offsetof(struct part, force.h_dt);
  L.balsara = (int)offsetof(struct part, force.balsara);
  L.time_bin = (int)offsetof(struct part, time_bin);
  L.min_ngb_time_bin = (int)offsetof(struct part, limiter_data.min_ngb_time_bin);
  L.id = (int)offsetof(struct part, id);
#ifdef PLANETARY_SPH
  L.mat_id = (int)offsetof(struct part, mat_id);
#else
  L.mat_id = -1;
#endif
#ifdef SPHENIX_SPH
  L.alpha_visc = (int)offsetof(struct part, viscosity.alpha);
  L.div_v_previous_step = (int)offsetof(struct part, viscosity.div_v_previous_step);
  L.div_v_dt = (int)offsetof(struct part, viscosity.div_v_dt);
  L.alpha_diff = (int)offsetof(struct part, diffusion.alpha);
  L.laplace_u = (int)offsetof(struct part, diffusion.laplace_u);
  L.alpha_visc_max_ngb = (int)offsetof(struct part, force.alpha_visc_max_ngb);
#else
  L.alpha_visc = L.div_v_previous_step = L.div_v_dt = L.alpha_diff = L.laplace_u = L.alpha_visc_max_ngb = -1;
#endif
  return L;
}

/* ---- what the operators read through r->e ------------------------------------ */
static void shim_params(const struct engine *e, swift_cuda_params *p) {
  memset(p, 0, sizeof(*p));
  const struct space *s = e->s;
  for (int k = 0; k < 3; k++) p->dim[k] = s->dim[k];
  p->periodic = s->periodic;
  const struct hydro_props *hp = e->hydro_properties;
  p->eta_neighbours = hp->eta_neighbours;
  p->h_tolerance = hp->h_tolerance;
  p->h_max = hp->h_max;
  p->h_min = hp->h_min;
  p->max_smoothing_iterations = hp->max_smoothing_iterations;
  p->viscosity_alpha = hp->viscosity.alpha;
  p->max_active_bin = e->max_active_bin;
  p->cosmo_a = e->cosmology ? (float)e->cosmology->a : 1.f;
  p->cosmo_H = e->cosmology ? (float)e->cosmology->H : 0.f;
  p->fields_mode = SWIFT_CUDA_FIELDS_EXACT;
#ifdef PLANETARY_SPH
  p->scheme = SWIFT_CUDA_SCHEME_PLANETARY;
  /* Tillotson materials in internal units, as convert_units_Til leaves them
     (src/equation_of_state/planetary/tillotson.h:43-143) */
  const struct Til_params *til[4] = {&eos.Til_iron, &eos.Til_granite, &eos.Til_water, &eos.Til_basalt};
  p->n_til = 4;
  for (int k = 0; k < 4; k++) {
    swift_cuda_til *t = &p->til[k];
    t->rho_0 = til[k]->rho_0; t->a = til[k]->a; t->b = til[k]->b; t->A = til[k]->A; t->B = til[k]->B;
    t->u_0 = til[k]->u_0; t->u_iv = til[k]->u_iv; t->u_cv = til[k]->u_cv; t->alpha = til[k]->alpha;
    t->beta = til[k]->beta; t->eta_min = til[k]->eta_min; t->eta_zero = til[k]->eta_zero;
    t->P_min = til[k]->P_min; t->mat_id = til[k]->mat_id;
  }
  for (int k = 0; k < 4; k++) p->idg_gamma[k] = 5.f / 3.f;
#elif defined(SPHENIX_SPH)
  p->scheme = SWIFT_CUDA_SCHEME_SPHENIX;
  for (int k = 0; k < 4; k++) p->idg_gamma[k] = 5.f / 3.f;
#else
  p->scheme = SWIFT_CUDA_SCHEME_MINIMAL;
  for (int k = 0; k < 4; k++) p->idg_gamma[k] = 5.f / 3.f;
#endif
}

#ifdef PLANETARY_SPH
/* Materials that live in tables: eos_init has read the files into the reference's global `eos`
 * (src/equation_of_state/planetary/equation_of_state.h:1629-1745: load_table_*, prepare_table_*, convert_units_*);
 * the prepared arrays are handed to the library as they are, which copies them to the device.  A material whose
 * table was not loaded (its arrays are NULL) is skipped. */
static void shim_set_sesame(swift_cuda_ctx *ctx, const struct SESAME_params *m) {
  if (m->table_log_rho == NULL || m->table_P_rho_T == NULL) return;
  swift_cuda_eos_table t;
  memset(&t, 0, sizeof(t));
  t.kind = 1;
  t.mat_id = (int)m->mat_id;
  t.num_rho = m->num_rho;
  t.num_T = m->num_T;
  t.log_rho = m->table_log_rho;
  t.log_u = m->table_log_u_rho_T;
  t.P = m->table_P_rho_T;
  t.c = m->table_c_rho_T;
  t.P_tiny = m->P_tiny;
  t.c_tiny = m->c_tiny;
  CUDA_OK(swift_cuda_set_eos_table(ctx, &t));
}
static void shim_set_hm80(swift_cuda_ctx *ctx, const struct HM80_params *m) {
  if (m->table_log_P_rho_u == NULL) return;
  swift_cuda_eos_table t;
  memset(&t, 0, sizeof(t));
  t.kind = 2;
  t.mat_id = (int)m->mat_id;
  t.num_rho = m->num_rho;
  t.num_T = m->num_u;
  t.P = m->table_log_P_rho_u;
  t.log_rho_min = m->log_rho_min;
  t.log_rho_step = m->log_rho_step;
  t.inv_log_rho_step = m->inv_log_rho_step;
  t.log_u_min = m->log_u_min;
  t.log_u_step = m->log_u_step;
  t.inv_log_u_step = m->inv_log_u_step;
  t.bulk_mod = m->bulk_mod;
  t.P_min_for_c_min = m->P_min_for_c_min;
  CUDA_OK(swift_cuda_set_eos_table(ctx, &t));
}
static void shim_set_tables(swift_cuda_ctx *ctx) {
  /* at most 8 tables per context: the materials of one run */
  shim_set_hm80(ctx, &eos.HM80_HHe);
  shim_set_hm80(ctx, &eos.HM80_ice);
  shim_set_hm80(ctx, &eos.HM80_rock);
  shim_set_sesame(ctx, &eos.SESAME_iron);
  shim_set_sesame(ctx, &eos.SESAME_basalt);
  shim_set_sesame(ctx, &eos.SESAME_water);
  shim_set_sesame(ctx, &eos.SS08_water);
  shim_set_sesame(ctx, &eos.ANEOS_forsterite);
  shim_set_sesame(ctx, &eos.ANEOS_iron);
  shim_set_sesame(ctx, &eos.ANEOS_Fe85Si15);
  for (int k = 0; k < 10; k++) shim_set_sesame(ctx, &eos.custom[k]);
}
#endif

/* A context for this engine: swift_cuda_create, and for SPHENIX the scheme's hydro_props and e->time_base
 * (the extra ghost's dt_alpha is get_timestep(p->time_bin, e->time_base), src/runner_ghost.c:1045-1046). */
static void shim_create(const struct engine *e, swift_cuda_ctx **ctx) {
  swift_cuda_params prm;
  shim_params(e, &prm);
  CUDA_OK(swift_cuda_create(&prm, 0, ctx));
#ifdef PLANETARY_SPH
  shim_set_tables(*ctx);
#endif
#ifdef SPHENIX_SPH
  const struct hydro_props *hp = e->hydro_properties;
  swift_cuda_sphenix_props sp;
  sp.viscosity_alpha_max = hp->viscosity.alpha_max;
  sp.viscosity_alpha_min = hp->viscosity.alpha_min;
  sp.viscosity_length = hp->viscosity.length;
  sp.diffusion_alpha = hp->diffusion.alpha;
  sp.diffusion_beta = hp->diffusion.beta;
  sp.diffusion_alpha_max = hp->diffusion.alpha_max;
  sp.diffusion_alpha_min = hp->diffusion.alpha_min;
  sp.time_base = e->time_base;
  CUDA_OK(swift_cuda_sphenix_set_props(*ctx, &sp));
#endif
}

/* ---- flattening a struct cell tree ------------------------------------------------------------ */
struct flat {
  swift_cuda_cell *cells;
  struct cell **src; /* the struct cell of every flattened cell */
  int n, cap;
};

static int flat_push(struct flat *f) {
  if (f->n == f->cap) {
    f->cap = f->cap ? 2 * f->cap : 64;
    f->cells = (swift_cuda_cell *)realloc(f->cells, sizeof(swift_cuda_cell) * f->cap);
    f->src = (struct cell **)realloc(f->src, sizeof(struct cell *) * f->cap);
    if (f->cells == NULL || f->src == NULL) error("Can't allocate the flattened cell tree.");
  }
  return f->n++;
}

/* Depth-first, parents before progeny; `first` is the offset of the cell's particles from
 * `base` plus `shift` records (the per-task shims put ci and cj one after the other). */
static int flatten_cell(struct flat *f, struct cell *c, int parent, const struct part *base, long shift) {
  const int me = flat_push(f);
  swift_cuda_cell *o = &f->cells[me];
  f->src[me] = c;
  for (int k = 0; k < 3; k++) {
    o->loc[k] = c->loc[k];
    o->width[k] = c->width[k];
  }
  o->first = (int)((c->hydro.parts - base) + shift);
  o->count = c->hydro.count;
  o->parent = parent;
  o->depth = c->depth;
  o->split = c->split;
  for (int k = 0; k < 8; k++) f->cells[me].progeny[k] = -1;
  if (c->split)
    for (int k = 0; k < 8; k++)
      if (c->progeny[k] != NULL && c->progeny[k]->hydro.count > 0) {
        const int child = flatten_cell(f, c->progeny[k], me, base, shift);
        f->cells[me].progeny[k] = child; /* f->cells may have moved: index again */
      }
  return me;
}

static void flat_free(struct flat *f) {
  free(f->cells);
  free(f->src);
  memset(f, 0, sizeof(*f));
}

/* c->hydro.h_max / h_max_active from the records, leaves first (runner_ghost.c:1565-1593) */
static void refresh_hmax(const struct flat *f, const struct engine *e) {
  for (int k = f->n - 1; k >= 0; k--) {
    struct cell *c = f->src[k];
    float h_max = 0.f, h_max_active = 0.f;
    if (c->split) {
      for (int j = 0; j < 8; j++)
        if (c->progeny[j] != NULL && c->progeny[j]->hydro.count > 0) {
          h_max = max(h_max, c->progeny[j]->hydro.h_max);
          h_max_active = max(h_max_active, c->progeny[j]->hydro.h_max_active);
        }
    } else {
      for (int j = 0; j < c->hydro.count; j++) {
        const struct part *p = &c->hydro.parts[j];
        h_max = max(h_max, p->h);
        if (part_is_active(p, e)) h_max_active = max(h_max_active, p->h);
      }
    }
    c->hydro.h_max = h_max;
    c->hydro.h_max_active = h_max_active;
  }
}

/* ---- (1) per-task entries ---------------------------------------------------------------------- */
/* One task on one or two cells: records of ci then cj in a scratch array, one context, one phase. */
/* subset != NULL: only the `count` particles ci->hydro.parts[subset[k]] are updated (the *_subset_* entries below):
 * in the scratch copy they alone carry an active time bin; the bins of the records that go back are the caller's. */
static void shim_run_task_subset(struct runner *r, struct cell *ci, struct cell *cj, int task_type, int phases,
                                 const int *subset, int count) {
  const struct engine *e = r->e;
  if (ci->hydro.count == 0 || (cj != NULL && cj->hydro.count == 0)) return;
  const int ni = ci->hydro.count, nj = cj ? cj->hydro.count : 0;
  struct part *buf = NULL;
  if (posix_memalign((void **)&buf, part_align, sizeof(struct part) * (size_t)(ni + nj)) != 0)
    error("Can't allocate the scratch particle array.");
  memcpy(buf, ci->hydro.parts, sizeof(struct part) * ni);
  if (cj) memcpy(buf + ni, cj->hydro.parts, sizeof(struct part) * nj);
  timebin_t *saved_bins = NULL;
  if (subset != NULL) {
    if ((saved_bins = (timebin_t *)malloc(sizeof(timebin_t) * (size_t)(ni + nj))) == NULL)
      error("Can't allocate the time-bin scratch array.");
    for (int k = 0; k < ni + nj; k++) {
      saved_bins[k] = buf[k].time_bin;
      buf[k].time_bin = (timebin_t)(e->max_active_bin + 1); /* inactive */
    }
    for (int k = 0; k < count; k++) buf[subset[k]].time_bin = (timebin_t)e->max_active_bin;
  }

  struct flat f = {0};
  flatten_cell(&f, ci, -1, ci->hydro.parts, 0);
  const int root_j = cj ? flatten_cell(&f, cj, -1, cj->hydro.parts, ni) : -1;
  swift_cuda_task t = {task_type, 0, root_j};

  const swift_cuda_part_layout L = shim_layout();
  swift_cuda_ctx *ctx = NULL;
  shim_create(e, &ctx);
  CUDA_OK(swift_cuda_build_from_cells_parts(ctx, &L, buf, ni + nj, f.n, f.cells, 1, &t));
  CUDA_OK(swift_cuda_step_parts(ctx, buf, phases));
  swift_cuda_destroy(ctx);

  if (saved_bins != NULL) {
    for (int k = 0; k < ni + nj; k++) buf[k].time_bin = saved_bins[k];
    free(saved_bins);
  }
  memcpy(ci->hydro.parts, buf, sizeof(struct part) * ni);
  if (cj) memcpy(cj->hydro.parts, buf + ni, sizeof(struct part) * nj);
  flat_free(&f);
  free(buf);
}

static void shim_run_task(struct runner *r, struct cell *ci, struct cell *cj, int task_type, int phases) {
  shim_run_task_subset(r, ci, cj, task_type, phases, NULL, 0);
}

#define DENSITY_PHASES (SWIFT_CUDA_PHASE_SORT | SWIFT_CUDA_PHASE_DENSITY | SWIFT_CUDA_PHASE_KEEP)
#define FORCE_PHASES (SWIFT_CUDA_PHASE_SORT | SWIFT_CUDA_PHASE_FORCE)

/* src/runner_doiact_functions_hydro.h:2271-2296, 1331-1413: nothing to do on cells without an active
 * particle; the library decides particle by particle (time_bin <= e->max_active_bin). */
void runner_doself1_branch_density(struct runner *r, struct cell *c) {
  if (!cell_is_active_hydro(c, r->e)) return;
  shim_run_task(r, c, NULL, 0, DENSITY_PHASES);
}
void runner_dopair1_branch_density(struct runner *r, struct cell *ci, struct cell *cj) {
  if (!cell_is_active_hydro(ci, r->e) && !cell_is_active_hydro(cj, r->e)) return;
  shim_run_task(r, ci, cj, 1, DENSITY_PHASES);
}
/* runner_do{self,pair}_subset_branch_density (src/runner_doiact_functions_hydro.h:884-937, 1047-1057; the entries
 * runner_do_ghost calls for the particles whose h has not converged, and tests/test27cells.c built as
 * test27cells_subset calls directly): only parts[ind[0..count)] of ci are updated, from ALL particles of ci (self)
 * or of cj (pair).  That is the density task with nobody but the subset active -- DOSELF1 / DOPAIR1 update an active
 * particle from every neighbour, active or not -- so the scratch copy carries an active bin for the subset alone.
 * `parts` must be ci->hydro.parts, as at every call site of the reference. */
void runner_doself_subset_branch_density(struct runner *r, struct cell *restrict ci, struct part *restrict parts,
                                         int *restrict ind, int count) {
  if (parts != ci->hydro.parts) error("runner_doself_subset_branch_density: parts is not ci->hydro.parts");
  if (count == 0) return;
  shim_run_task_subset(r, ci, NULL, 0, DENSITY_PHASES, ind, count);
}
void runner_dopair_subset_branch_density(struct runner *r, struct cell *restrict ci, struct part *restrict parts_i,
                                         int *restrict ind, int count, struct cell *restrict cj) {
  if (parts_i != ci->hydro.parts) error("runner_dopair_subset_branch_density: parts_i is not ci->hydro.parts");
  if (count == 0) return;
  shim_run_task_subset(r, ci, cj, 1, DENSITY_PHASES, ind, count);
}
void runner_doself2_branch_force(struct runner *r, struct cell *c) {
  if (!cell_is_active_hydro(c, r->e)) return;
  shim_run_task(r, c, NULL, 0, FORCE_PHASES);
}
void runner_dopair2_branch_force(struct runner *r, struct cell *ci, struct cell *cj) {
  if (!cell_is_active_hydro(ci, r->e) && !cell_is_active_hydro(cj, r->e)) return;
  shim_run_task(r, ci, cj, 1, FORCE_PHASES);
}
/* :2524-2719: the run-time recursion is replayed by the library's work-list builder from the
 * same criteria (cell_can_recurse_in_{self,pair}_hydro_task on the cells' h_max). */
void runner_dosub_self1_density(struct runner *r, struct cell *ci, int gettimer) {
  if (!cell_is_active_hydro(ci, r->e)) return;
  shim_run_task(r, ci, NULL, 2, DENSITY_PHASES);
}
void runner_dosub_pair1_density(struct runner *r, struct cell *ci, struct cell *cj, int gettimer) {
  if (!cell_is_active_hydro(ci, r->e) && !cell_is_active_hydro(cj, r->e)) return;
  shim_run_task(r, ci, cj, 3, DENSITY_PHASES);
}
void runner_dosub_self2_force(struct runner *r, struct cell *ci, int gettimer) {
  if (!cell_is_active_hydro(ci, r->e)) return;
  shim_run_task(r, ci, NULL, 2, FORCE_PHASES);
}
void runner_dosub_pair2_force(struct runner *r, struct cell *ci, struct cell *cj, int gettimer) {
  if (!cell_is_active_hydro(ci, r->e) && !cell_is_active_hydro(cj, r->e)) return;
  shim_run_task(r, ci, cj, 3, FORCE_PHASES);
}

/* src/runner_ghost.c:1085-1596 for one (super-)cell whose density sums are complete.  The redo
 * passes of the reference walk c->hydro.density (the task links); a cell handed over on its own
 * has none -- as in tests/test125cells.c -- and then neither side re-runs the density loop.  With
 * links the batch entry below is the one to use: it has every cell of the rank on the device. */
void runner_do_ghost(struct runner *r, struct cell *c, int timer) {
  if (c->hydro.count == 0 || !cell_is_active_hydro(c, r->e)) return;
  if (c->hydro.density != NULL)
    error("runner_do_ghost on a cell with density task links: use swift_cuda_shim_step()");
  const struct engine *e = r->e;
  struct flat f = {0};
  flatten_cell(&f, c, -1, c->hydro.parts, 0);
  const swift_cuda_part_layout L = shim_layout();
  swift_cuda_ctx *ctx = NULL;
  shim_create(e, &ctx);
  CUDA_OK(swift_cuda_build_from_cells_parts(ctx, &L, c->hydro.parts, c->hydro.count, f.n, f.cells, 0, NULL));
  CUDA_OK(swift_cuda_step_parts(ctx, c->hydro.parts, SWIFT_CUDA_PHASE_GHOST | SWIFT_CUDA_PHASE_KEEP));
  swift_cuda_destroy(ctx);
  refresh_hmax(&f, e);
  flat_free(&f);
}

#ifdef EXTRA_HYDRO_LOOP
/* ---- SPHENIX: the gradient loop and the extra ghost (src/runner_doiact_hydro.h:168-191 with FUNCTION =
 * gradient, src/runner_main.c:199-200, 249-250, 297-298, 343-344; src/runner_ghost.c:992-1075) --------- */
#define GRADIENT_PHASES (SWIFT_CUDA_PHASE_SORT | SWIFT_CUDA_PHASE_GRADIENT)
void runner_doself1_branch_gradient(struct runner *r, struct cell *c) {
  if (!cell_is_active_hydro(c, r->e)) return;
  shim_run_task(r, c, NULL, 0, GRADIENT_PHASES);
}
void runner_dopair1_branch_gradient(struct runner *r, struct cell *ci, struct cell *cj) {
  if (!cell_is_active_hydro(ci, r->e) && !cell_is_active_hydro(cj, r->e)) return;
  shim_run_task(r, ci, cj, 1, GRADIENT_PHASES);
}
void runner_dosub_self1_gradient(struct runner *r, struct cell *ci, int gettimer) {
  if (!cell_is_active_hydro(ci, r->e)) return;
  shim_run_task(r, ci, NULL, 2, GRADIENT_PHASES);
}
void runner_dosub_pair1_gradient(struct runner *r, struct cell *ci, struct cell *cj, int gettimer) {
  if (!cell_is_active_hydro(ci, r->e) && !cell_is_active_hydro(cj, r->e)) return;
  shim_run_task(r, ci, cj, 3, GRADIENT_PHASES);
}
/* hydro_end_gradient, hydro_prepare_force, hydro_reset_acceleration for the active particles of one cell */
void runner_do_extra_ghost(struct runner *r, struct cell *c, int timer) {
  if (c->hydro.count == 0 || !cell_is_active_hydro(c, r->e)) return;
  const struct engine *e = r->e;
  struct flat f = {0};
  flatten_cell(&f, c, -1, c->hydro.parts, 0);
  const swift_cuda_part_layout L = shim_layout();
  swift_cuda_ctx *ctx = NULL;
  shim_create(e, &ctx);
  CUDA_OK(swift_cuda_build_from_cells_parts(ctx, &L, c->hydro.parts, c->hydro.count, f.n, f.cells, 0, NULL));
  CUDA_OK(swift_cuda_step_parts(ctx, c->hydro.parts, SWIFT_CUDA_PHASE_EXTRA_GHOST));
  swift_cuda_destroy(ctx);
  flat_free(&f);
}
#endif /* EXTRA_HYDRO_LOOP */

/* src/runner_sort.c:201-431.  The sorted lists live on the device and are rebuilt there by every
 * call that walks a pair; what is left for the host is the bookkeeping the callers look at. */
void runner_do_hydro_sort(struct runner *r, struct cell *c, int flags, int cleanup, int rt_requests_sort,
                          int clock) {
  if (c->split)
    for (int k = 0; k < 8; k++)
      if (c->progeny[k] != NULL) runner_do_hydro_sort(r, c->progeny[k], flags, cleanup, rt_requests_sort, 0);
  c->hydro.sorted |= flags;
  c->hydro.do_sort = 0;
  c->hydro.dx_max_sort = 0.f;
  c->hydro.dx_max_sort_old = 0.f;
  if (c->hydro.xparts != NULL)
    for (int k = 0; k < c->hydro.count; k++)
      c->hydro.xparts[k].x_diff_sort[0] = c->hydro.xparts[k].x_diff_sort[1] = c->hydro.xparts[k].x_diff_sort[2] = 0.f;
}

/* src/runner_others.c:618-690: hydro_end_force is one multiplication per particle and stays on the
 * host for a cell handed over on its own; the batch step does it in the force kernel's epilogue. */
void runner_do_end_hydro_force(struct runner *r, struct cell *c, int timer) {
  const struct engine *e = r->e;
  if (!cell_is_active_hydro(c, e)) return;
  if (c->split) {
    for (int k = 0; k < 8; k++)
      if (c->progeny[k] != NULL) runner_do_end_hydro_force(r, c->progeny[k], 0);
  } else {
    for (int k = 0; k < c->hydro.count; k++) {
      struct part *p = &c->hydro.parts[k];
      if (part_is_active(p, e)) hydro_end_force(p, e->cosmology);
    }
  }
}

/* ---- (2) batch entries ------------------------------------------------------------------------- */
static swift_cuda_ctx *shim_ctx = NULL;
static struct flat shim_flat = {0};

static int find_flat(const struct flat *f, const struct cell *c) {
  /* cells are flattened depth first from s->cells_top: small trees, linear search per task end */
  for (int k = 0; k < f->n; k++)
    if (f->src[k] == c) return k;
  return -1;
}

/* After space_rebuild + engine_maketasks (src/engine.c, engine_rebuild). */
void swift_cuda_shim_rebuild(struct engine *e) {
  struct space *s = e->s;
  flat_free(&shim_flat);
  for (int k = 0; k < s->nr_cells; k++)
    if (s->cells_top[k].hydro.count > 0) flatten_cell(&shim_flat, &s->cells_top[k], -1, s->parts, 0);

  /* the density tasks as scheduler_splittask_hydro left them (src/scheduler.c:1154-1378) */
  const struct scheduler *sched = &e->sched;
  swift_cuda_task *tasks = (swift_cuda_task *)malloc(sizeof(swift_cuda_task) * (sched->nr_tasks + 1));
  if (tasks == NULL) error("Can't allocate the task list.");
  int ntasks = 0;
  for (int k = 0; k < sched->nr_tasks; k++) {
    const struct task *t = &sched->tasks[k];
    if (t->subtype != task_subtype_density) continue;
    int type;
    if (t->type == task_type_self) type = 0;
    else if (t->type == task_type_pair) type = 1;
    else if (t->type == task_type_sub_self) type = 2;
    else if (t->type == task_type_sub_pair) type = 3;
    else continue;
    if (t->ci->hydro.count == 0 || (t->cj != NULL && t->cj->hydro.count == 0)) continue;
    tasks[ntasks].type = type;
    tasks[ntasks].ci = find_flat(&shim_flat, t->ci);
    tasks[ntasks].cj = t->cj ? find_flat(&shim_flat, t->cj) : -1;
    ntasks++;
  }
  const swift_cuda_part_layout L = shim_layout();
  if (shim_ctx == NULL) shim_create(e, &shim_ctx);
  CUDA_OK(swift_cuda_build_from_cells_parts(shim_ctx, &L, s->parts, (int64_t)s->nr_parts, shim_flat.n,
                                            shim_flat.cells, ntasks, tasks));
  free(tasks);
}

/* Once per step, in place of the sort / density / ghost / force / end_force tasks (the scheduler
 * leaves them skipped), after the drift tasks have run. */
void swift_cuda_shim_step(struct engine *e, int resort) {
  if (shim_ctx == NULL) error("swift_cuda_shim_step before swift_cuda_shim_rebuild");
  CUDA_OK(swift_cuda_set_max_active_bin(shim_ctx, e->max_active_bin));
  int phases = (resort ? SWIFT_CUDA_PHASE_SORT : 0) | SWIFT_CUDA_PHASE_DENSITY | SWIFT_CUDA_PHASE_GHOST |
               SWIFT_CUDA_PHASE_FORCE | SWIFT_CUDA_PHASE_END_FORCE;
#ifdef EXTRA_HYDRO_LOOP
  phases |= SWIFT_CUDA_PHASE_GRADIENT | SWIFT_CUDA_PHASE_EXTRA_GHOST;
#endif
  CUDA_OK(swift_cuda_step_parts(shim_ctx, e->s->parts, phases));
  refresh_hmax(&shim_flat, e);
}

/* With engine_policy_timestep_limiter (swift --limiter): once per step AFTER the time-step tasks, in place of the
 * tasks of subtype limiter (runner_do{self,pair}1_branch_limiter, runner_dosub_{self,pair}1_limiter,
 * src/runner_main.c:205-206, 255-256, 303-304, 349-350), which the scheduler leaves skipped like the other
 * loops.  The new time bins go up (1 byte per particle), the limiter loop runs on the positions, smoothing
 * lengths and sorts swift_cuda_shim_step left on the device, and p->limiter_data.wakeup comes back (1 byte per
 * particle) for the reference's own runner_do_limiter tasks (src/runner_time_integration.c:1324), which stay
 * where they are. */
void swift_cuda_shim_limiter(struct engine *e) {
  if (shim_ctx == NULL) error("swift_cuda_shim_limiter before swift_cuda_shim_rebuild");
  struct space *s = e->s;
  const size_t n = s->nr_parts;
  signed char *bins = (signed char *)malloc(2 * n + 2);
  if (bins == NULL) error("Can't allocate the time-bin buffer.");
  signed char *wake = bins + n;
  for (size_t k = 0; k < n; k++) bins[k] = s->parts[k].time_bin;
  CUDA_OK(swift_cuda_do_limiter_loop(shim_ctx, bins, wake));
  for (size_t k = 0; k < n; k++)
    if (wake[k] > s->parts[k].limiter_data.wakeup) s->parts[k].limiter_data.wakeup = wake[k]; /* accumulate_max_c */
  free(bins);
}

void swift_cuda_shim_clean(void) {
  if (shim_ctx != NULL) swift_cuda_destroy(shim_ctx);
  shim_ctx = NULL;
  flat_free(&shim_flat);
}
