/* TEST INFRASTRUCTURE ONLY -- never linked into or called by the product.
 *
 * Harness around the REFERENCE's own compiled hot-path operators (built from
 * /root/reference/src by oracle/Makefile.ref into oracle/_ref/).  It receives
 * particles, a cell tree and a task list as flat arrays, builds the reference's
 * `struct cell` / `struct part` / `struct task` objects from them and runs
 *   hydro_init_part -> runner_do_hydro_sort -> density tasks -> runner_do_ghost
 *   -> force tasks -> runner_do_end_hydro_force
 * calling the operators directly, single threaded, in the order
 * tests/test125cells.c:660-822 does.  The ghost's redo
 * passes see the density tasks through hand-made `struct link` lists, as
 * engine_maketasks.c would have linked them (runner_ghost.c:1503-1546).
 *
 * This file contains no physics: every number it returns was computed by the
 * reference's code. */
#include <config.h>

#include <float.h>
#include <stdlib.h>
#include <string.h>

#include "active.h"
#include "cell.h"
#include "cosmology.h"
#include "engine.h"
#include "hydro.h"
#include "hydro_properties.h"
#include "physical_constants.h"
#include "runner.h"
#include "runner_doiact_hydro.h"
#include "space.h"
#include "task.h"
#include "timeline.h"

#include "oracle_api.h"

/* ---- pair log ----------------------------------------------------------- */
static oracle_pairlog *g_log = NULL;
static int g_store = 0;
static int g_stage = 0; /* 0 density, 1 ghost redo, 2 force */

static oracle_sphenix *g_sx = NULL; /* SPHENIX: extra inputs / outputs, the gradient loop's pair log */

/* the fields that moved between the schemes' struct part */
#ifdef SPHENIX_SPH
#define PART_DIV_V(p) ((p)->viscosity.div_v)
#define PART_V_SIG(p) ((p)->viscosity.v_sig)
#else
#define PART_DIV_V(p) ((p)->density.div_v)
#define PART_V_SIG(p) ((p)->force.v_sig)
#endif

void ref_record_pair(int loop, long long id_updated, long long id_other) {
  if (loop == 2) { /* gradient loop */
    if (g_sx == NULL) return;
    const long long k = g_sx->grad_count++;
    if (g_store && g_sx->grad_pairs != NULL && k < g_sx->grad_capacity) {
      g_sx->grad_pairs[2 * k] = id_updated;
      g_sx->grad_pairs[2 * k + 1] = id_other;
    }
    return;
  }
  if (g_log == NULL) return;
  const int slot = (loop == 1) ? 2 : g_stage;
  const long long k = g_log->count[slot]++;
  if (g_store && g_log->pairs[slot] != NULL && k < g_log->capacity[slot]) {
    g_log->pairs[slot][2 * k] = id_updated;
    g_log->pairs[slot][2 * k + 1] = id_other;
  }
}

int ref_sizeof_part(void) { return (int)sizeof(struct part); }

/* ---- phases, serial or on a pool of pthreads ---------------------------------------------------
 * ref_set_threads(1) (the default): every phase runs on the calling thread in creation order -- what the
 * golden fixtures were made with.  ref_set_threads(T > 1): the TIMED CPU baseline of bench.py -- the
 * reference's operators on ONE box shared by T threads, the role of its threaded runner
 * (src/runner_main.c, src/engine.c:1692-1736): each thread has its own struct runner, draws tasks from a
 * shared counter and takes the locks of the top-level cells a task updates (both cells of a pair; the
 * reference's cell_locktree, src/cell_lock.c, at the granularity of top-level cells); a task whose cells
 * are busy is put aside and done when the list is through.  Sums then arrive in another order than in the
 * serial run (rounding differs), which is why nothing is CHECKED in this mode. */
#include <pthread.h>
#include <time.h>
enum { PH_SORT, PH_DENSITY, PH_GHOST, PH_GRADIENT, PH_EXTRA_GHOST, PH_FORCE, PH_END_FORCE, PH_LIMITER };
/* src/runner_doiact_limiter.h:94-102 (FUNCTION = limiter, compiled in src/runner_doiact_limiter.c) */
void runner_doself1_branch_limiter(struct runner *r, struct cell *c);
void runner_dopair1_branch_limiter(struct runner *r, struct cell *ci, struct cell *cj);
void runner_dosub_self1_limiter(struct runner *r, struct cell *ci, int gettimer);
void runner_dosub_pair1_limiter(struct runner *r, struct cell *ci, struct cell *cj, int gettimer);
static int g_threads = 1;
static double g_last_seconds = 0.0;
void ref_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
double ref_last_step_seconds(void) { return g_last_seconds; }

struct phase_ctx {
  int phase;
  struct engine *e;
  struct cell *cells;
  int ncells;
  struct task *tasks;
  int ntasks;
  pthread_mutex_t *locks; /* one per cell slot; only those of top-level cells are used */
  int next;               /* shared work counter */
};

static void phase_task(struct runner *r, struct task *t, int phase) {
  if (phase == PH_DENSITY) {
    switch (t->type) {
      case task_type_self: runner_doself1_branch_density(r, t->ci); break;
      case task_type_pair: runner_dopair1_branch_density(r, t->ci, t->cj); break;
      case task_type_sub_self: runner_dosub_self1_density(r, t->ci, 0); break;
      default: runner_dosub_pair1_density(r, t->ci, t->cj, 0); break;
    }
#ifdef EXTRA_HYDRO_LOOP
  } else if (phase == PH_GRADIENT) {
    switch (t->type) {
      case task_type_self: runner_doself1_branch_gradient(r, t->ci); break;
      case task_type_pair: runner_dopair1_branch_gradient(r, t->ci, t->cj); break;
      case task_type_sub_self: runner_dosub_self1_gradient(r, t->ci, 0); break;
      default: runner_dosub_pair1_gradient(r, t->ci, t->cj, 0); break;
    }
#endif
  } else if (phase == PH_LIMITER) {
    switch (t->type) {
      case task_type_self: runner_doself1_branch_limiter(r, t->ci); break;
      case task_type_pair: runner_dopair1_branch_limiter(r, t->ci, t->cj); break;
      case task_type_sub_self: runner_dosub_self1_limiter(r, t->ci, 0); break;
      default: runner_dosub_pair1_limiter(r, t->ci, t->cj, 0); break;
    }
  } else {
    switch (t->type) {
      case task_type_self: runner_doself2_branch_force(r, t->ci); break;
      case task_type_pair: runner_dopair2_branch_force(r, t->ci, t->cj); break;
      case task_type_sub_self: runner_dosub_self2_force(r, t->ci, 0); break;
      default: runner_dosub_pair2_force(r, t->ci, t->cj, 0); break;
    }
  }
}

static void phase_cell(struct runner *r, struct cell *c, int phase) {
  if (phase == PH_SORT) {
    if (c->hydro.count > 0) runner_do_hydro_sort(r, c, 0x1FFF, 0, 0, 0);
  } else if (phase == PH_GHOST) {
    runner_do_ghost(r, c, 0);
#ifdef EXTRA_HYDRO_LOOP
  } else if (phase == PH_EXTRA_GHOST) {
    runner_do_extra_ghost(r, c, 0);
#endif
  } else {
    runner_do_end_hydro_force(r, c, 0);
  }
}

static int phase_is_tasks(int phase) {
  return phase == PH_DENSITY || phase == PH_GRADIENT || phase == PH_FORCE || phase == PH_LIMITER;
}

static void *phase_worker(void *arg) {
  struct phase_ctx *x = (struct phase_ctx *)arg;
  struct runner r;
  bzero(&r, sizeof(r));
  r.e = x->e;
  if (!phase_is_tasks(x->phase)) {
    for (;;) {
      const int k = __atomic_fetch_add(&x->next, 1, __ATOMIC_RELAXED);
      if (k >= x->ncells) break;
      if (x->cells[k].parent == NULL) phase_cell(&r, &x->cells[k], x->phase);
    }
    return NULL;
  }
  int *later = (int *)malloc(sizeof(int) * (size_t)(x->ntasks + 1));
  int nlater = 0;
  for (;;) {
    const int k = __atomic_fetch_add(&x->next, 1, __ATOMIC_RELAXED);
    if (k >= x->ntasks) break;
    struct task *t = &x->tasks[k];
    pthread_mutex_t *la = &x->locks[t->ci->top - x->cells];
    pthread_mutex_t *lb = (t->cj != NULL && t->cj->top != t->ci->top) ? &x->locks[t->cj->top - x->cells] : NULL;
    if (pthread_mutex_trylock(la) != 0) {
      later[nlater++] = k;
      continue;
    }
    if (lb != NULL && pthread_mutex_trylock(lb) != 0) {
      pthread_mutex_unlock(la);
      later[nlater++] = k;
      continue;
    }
    phase_task(&r, t, x->phase);
    if (lb != NULL) pthread_mutex_unlock(lb);
    pthread_mutex_unlock(la);
  }
  /* what was busy: now with blocking locks, always in address order */
  for (int i = 0; i < nlater; i++) {
    struct task *t = &x->tasks[later[i]];
    pthread_mutex_t *la = &x->locks[t->ci->top - x->cells];
    pthread_mutex_t *lb = (t->cj != NULL && t->cj->top != t->ci->top) ? &x->locks[t->cj->top - x->cells] : NULL;
    if (lb != NULL && lb < la) {
      pthread_mutex_t *tmp = la;
      la = lb;
      lb = tmp;
    }
    pthread_mutex_lock(la);
    if (lb != NULL) pthread_mutex_lock(lb);
    phase_task(&r, t, x->phase);
    if (lb != NULL) pthread_mutex_unlock(lb);
    pthread_mutex_unlock(la);
  }
  free(later);
  return NULL;
}

static void run_phase(struct phase_ctx *x, struct runner *serial_runner, int phase) {
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  x->phase = phase;
  x->next = 0;
  if (g_threads <= 1) {
    if (phase_is_tasks(phase)) {
      for (int k = 0; k < x->ntasks; k++) phase_task(serial_runner, &x->tasks[k], phase);
    } else {
      for (int k = 0; k < x->ncells; k++)
        if (x->cells[k].parent == NULL) phase_cell(serial_runner, &x->cells[k], phase);
    }
  } else {
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)g_threads);
    for (int k = 0; k < g_threads; k++) pthread_create(&th[k], NULL, phase_worker, x);
    for (int k = 0; k < g_threads; k++) pthread_join(th[k], NULL);
    free(th);
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  g_last_seconds += (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
#ifdef PLANETARY_SPH
static void ref_planetary_eos_init(void);
#endif

/* ---- helpers ------------------------------------------------------------ */
static void add_link(struct cell *c, struct task *t, struct link *pool,
                     int *used) {
  struct link *l = &pool[(*used)++];
  l->t = t;
  l->next = c->hydro.density;
  c->hydro.density = l;
}

static int cell_has_active(const struct cell *c, const struct engine *e) {
  for (int k = 0; k < c->hydro.count; k++)
    if (c->hydro.parts[k].time_bin <= e->max_active_bin) return 1;
  return 0;
}

static int run_step(const oracle_params *prm, oracle_parts *io, int ncells, const oracle_cell *cd, int ntasks,
                    const oracle_task *td, oracle_pairlog *log, oracle_sphenix *sx);

int ref_run_step(const oracle_params *prm, oracle_parts *io, int ncells,
                 const oracle_cell *cd, int ntasks, const oracle_task *td,
                 oracle_pairlog *log) {
#ifdef SPHENIX_SPH
  return 1; /* this build runs the three-loop scheme: ref_run_step_sphenix */
#else
  return run_step(prm, io, ncells, cd, ntasks, td, log, NULL);
#endif
}

int ref_run_step_sphenix(const oracle_params *prm, oracle_parts *io, int ncells, const oracle_cell *cd, int ntasks,
                         const oracle_task *td, oracle_pairlog *log, oracle_sphenix *sx) {
#ifdef SPHENIX_SPH
  if (sx == NULL) return 1;
  return run_step(prm, io, ncells, cd, ntasks, td, log, sx);
#else
  return 1;
#endif
}

static int run_step(const oracle_params *prm, oracle_parts *io, int ncells, const oracle_cell *cd, int ntasks,
                    const oracle_task *td, oracle_pairlog *log, oracle_sphenix *sx) {
  const int n = io->n;
  const integertime_t ti_current = 8;

  /* ---- particles ---- */
  struct part *parts = NULL;
  struct xpart *xparts = NULL;
  if (posix_memalign((void **)&parts, part_align, (size_t)(n + 1) * sizeof(struct part)) != 0) return 1;
  if (posix_memalign((void **)&xparts, xpart_align, (size_t)(n + 1) * sizeof(struct xpart)) != 0) return 1;
  bzero(parts, (size_t)(n + 1) * sizeof(struct part));
  bzero(xparts, (size_t)(n + 1) * sizeof(struct xpart));
  for (int k = 0; k < n; k++) {
    struct part *p = &parts[k];
    for (int d = 0; d < 3; d++) {
      p->x[d] = io->x[3 * k + d];
      p->v[d] = io->v[3 * k + d];
    }
    p->mass = io->mass[k];
    p->h = io->h[k];
    p->u = io->u[k];
#ifdef PLANETARY_SPH
    p->mat_id = io->mat_id ? io->mat_id[k] : 0;
#endif
    hydro_first_init_part(p, &xparts[k]);
    p->id = io->id[k];
    p->time_bin = io->time_bin ? io->time_bin[k] : 1;
    p->limiter_data.min_ngb_time_bin = num_time_bins + 1;
#ifdef SPHENIX_SPH
    /* what hydro_convert_quantities / the previous step left in the particle */
    p->viscosity.alpha = sx->in_alpha_visc ? sx->in_alpha_visc[k] : prm->viscosity_alpha;
    p->diffusion.alpha = sx->in_alpha_diff ? sx->in_alpha_diff[k] : 0.f;
    p->viscosity.div_v_previous_step = sx->in_div_v_prev ? sx->in_div_v_prev[k] : 0.f;
#endif
  }

  /* ---- engine & friends (as tests/test125cells.c:566-619) ---- */
  struct space space;
  bzero(&space, sizeof(space));
  space.periodic = prm->periodic;
  for (int d = 0; d < 3; d++) space.dim[d] = prm->dim[d];
  hydro_space_init(&space.hs, &space);

  struct phys_const prog_const;
  bzero(&prog_const, sizeof(prog_const));
  prog_const.const_vacuum_permeability = 1.0;

  struct hydro_props hp;
  bzero(&hp, sizeof(hp));
  hydro_props_init_no_hydro(&hp);
  hp.eta_neighbours = prm->eta_neighbours;
  hp.h_tolerance = prm->h_tolerance;
  hp.h_max = prm->h_max;
  hp.h_min = prm->h_min;
  hp.h_min_ratio = 0.f;
  hp.max_smoothing_iterations = prm->max_smoothing_iterations;
  hp.CFL_condition = 0.1f;
  hp.viscosity.alpha = prm->viscosity_alpha;
  hp.use_mass_weighted_num_ngb = 0;
#ifdef SPHENIX_SPH
  hp.viscosity.alpha_max = sx->viscosity_alpha_max;
  hp.viscosity.alpha_min = sx->viscosity_alpha_min;
  hp.viscosity.length = sx->viscosity_length;
  hp.diffusion.beta = sx->diffusion_beta;
  hp.diffusion.alpha_max = sx->diffusion_alpha_max;
  hp.diffusion.alpha_min = sx->diffusion_alpha_min;
#endif

  struct cosmology cosmo;
  cosmology_init_no_cosmo(&cosmo);

  struct engine engine;
  bzero(&engine, sizeof(engine));
  engine.hydro_properties = &hp;
  engine.physical_constants = &prog_const;
  engine.s = &space;
  engine.time = 0.1f;
  engine.ti_current = ti_current;
  engine.max_active_bin = prm->max_active_bin;
  engine.nodeID = 0;
  engine.cosmology = &cosmo;
  engine.time_base = sx ? sx->time_base : 1e-3;
  engine.policy = engine_policy_hydro;
  struct lightcone_array_props lcp;
  bzero(&lcp, sizeof(lcp));
  engine.lightcone_array_properties = &lcp;
  struct pressure_floor_props pfloor;
  bzero(&pfloor, sizeof(pfloor));
  engine.pressure_floor_props = &pfloor;
  space.e = &engine;

  struct runner runner;
  bzero(&runner, sizeof(runner));
  runner.e = &engine;

  /* ---- cells ---- */
  struct cell *cells = NULL;
  if (posix_memalign((void **)&cells, cell_align, (size_t)ncells * sizeof(struct cell)) != 0) return 1;
  bzero(cells, (size_t)ncells * sizeof(struct cell));
  for (int k = 0; k < ncells; k++) {
    struct cell *c = &cells[k];
    for (int d = 0; d < 3; d++) {
      c->loc[d] = cd[k].loc[d];
      c->width[d] = cd[k].width[d];
    }
    c->dmin = fmin(c->width[0], fmin(c->width[1], c->width[2]));
    c->depth = cd[k].depth;
    c->split = cd[k].split;
    c->parent = cd[k].parent >= 0 ? &cells[cd[k].parent] : NULL;
    for (int j = 0; j < 8; j++)
      c->progeny[j] = cd[k].progeny[j] >= 0 ? &cells[cd[k].progeny[j]] : NULL;
    c->hydro.parts = &parts[cd[k].first];
    c->hydro.xparts = &xparts[cd[k].first];
    c->hydro.count = cd[k].count;
    float h_max = 0.f, h_max_old = 0.f;
    for (int j = 0; j < c->hydro.count; j++) {
      h_max = max(h_max, c->hydro.parts[j].h);
      h_max_old = max(h_max_old, io->in_h_build ? io->in_h_build[cd[k].first + j] : c->hydro.parts[j].h);
    }
    c->hydro.h_max = h_max;
    c->hydro.h_max_old = h_max_old;
    c->hydro.h_max_active = h_max;
    c->hydro.ti_old_part = ti_current;
    c->hydro.ti_end_min = cell_has_active(c, &engine) ? ti_current : 2 * ti_current;
    c->nodeID = 0;
  }
  for (int k = 0; k < ncells; k++) {
    struct cell *top = &cells[k];
    while (top->parent != NULL) top = top->parent;
    cells[k].top = top;
    cells[k].hydro.super = top;
    cells[k].super = top;
  }

  /* ---- tasks and their links ---- */
  struct task *tasks = (struct task *)calloc((size_t)ntasks + 1, sizeof(struct task));
  struct link *links = (struct link *)calloc(2 * (size_t)ntasks + 2, sizeof(struct link));
  int nlinks = 0;
  for (int k = 0; k < ntasks; k++) {
    struct task *t = &tasks[k];
    t->ci = &cells[td[k].ci];
    t->cj = td[k].cj >= 0 ? &cells[td[k].cj] : NULL;
    t->subtype = task_subtype_density;
    switch (td[k].type) {
      case ORACLE_TASK_SELF: t->type = task_type_self; break;
      case ORACLE_TASK_PAIR: t->type = task_type_pair; break;
      case ORACLE_TASK_SUB_SELF: t->type = task_type_sub_self; break;
      default: t->type = task_type_sub_pair; break;
    }
    add_link(t->ci, t, links, &nlinks);
    if (t->cj != NULL) add_link(t->cj, t, links, &nlinks);
  }

#ifdef PLANETARY_SPH
  ref_planetary_eos_init();
#endif
  g_log = log;
  g_sx = sx;
  if (sx) sx->grad_count = 0;
  g_store = prm->record_pairs;
  if (log) {
    log->count[0] = log->count[1] = log->count[2] = 0;
    log->ghost_sweeps = 0;
  }

  /* ---- init + sort (all 13 axes on every top-level cell) ---- */
  /* only active particles are re-initialised (src/cell_drift.c:342); inactive ones keep
   * the state of their last active step, given by the caller */
  for (int k = 0; k < n; k++) {
    struct part *p = &parts[k];
    if (part_is_active(p, &engine)) {
      hydro_init_part(p, &space.hs);
    } else {
      p->rho = io->in_rho ? io->in_rho[k] : 0.f;
      p->force.pressure = io->in_pressure ? io->in_pressure[k] : 0.f;
      p->force.f = io->in_f ? io->in_f[k] : 0.f;
      p->force.soundspeed = io->in_soundspeed ? io->in_soundspeed[k] : 0.f;
      p->force.balsara = io->in_balsara ? io->in_balsara[k] : 0.f;
    }
  }
  struct phase_ctx px;
  bzero(&px, sizeof(px));
  px.e = &engine;
  px.cells = cells;
  px.ncells = ncells;
  px.tasks = tasks;
  px.ntasks = ntasks;
  px.locks = (pthread_mutex_t *)malloc(sizeof(pthread_mutex_t) * (size_t)(ncells + 1));
  for (int k = 0; k < ncells; k++) pthread_mutex_init(&px.locks[k], NULL);
  g_last_seconds = 0.0;
  if (g_threads > 1) g_log = NULL, g_sx = NULL; /* the pair log is not thread safe: nothing is recorded or counted */
  run_phase(&px, &runner, PH_SORT);
  if (io->l_only) goto limiter_loop;

  /* ---- density ---- */
  g_stage = 0;
  run_phase(&px, &runner, PH_DENSITY);
  for (int k = 0; k < n; k++) {
    const struct part *p = &parts[k];
    if (io->d_rho) io->d_rho[k] = p->rho;
    if (io->d_rho_dh) io->d_rho_dh[k] = p->density.rho_dh;
    if (io->d_wcount) io->d_wcount[k] = p->density.wcount;
    if (io->d_wcount_dh) io->d_wcount_dh[k] = p->density.wcount_dh;
    if (io->d_div_v) io->d_div_v[k] = PART_DIV_V(p);
    if (io->d_rot_v)
      for (int d = 0; d < 3; d++) io->d_rot_v[3 * k + d] = p->density.rot_v[d];
  }

  /* ---- ghost ---- */
  g_stage = 1;
  run_phase(&px, &runner, PH_GHOST);
  for (int k = 0; k < n; k++) {
    const struct part *p = &parts[k];
    if (io->g_h) io->g_h[k] = p->h;
    if (io->g_rho) io->g_rho[k] = p->rho;
    if (io->g_f) io->g_f[k] = p->force.f;
    if (io->g_pressure) io->g_pressure[k] = p->force.pressure;
    if (io->g_soundspeed) io->g_soundspeed[k] = p->force.soundspeed;
    if (io->g_balsara) io->g_balsara[k] = p->force.balsara;
#ifdef SPHENIX_SPH
    if (sx->g_div_v) sx->g_div_v[k] = p->viscosity.div_v;
#endif
  }

#ifdef EXTRA_HYDRO_LOOP
  /* ---- gradient loop + extra ghost (src/runner_main.c:199-200, 249-250, 297-298, 343-344, 420-421) ---- */
  run_phase(&px, &runner, PH_GRADIENT);
  for (int k = 0; k < n; k++) {
    const struct part *p = &parts[k];
    if (sx->r_v_sig) sx->r_v_sig[k] = p->viscosity.v_sig;
    if (sx->r_laplace_u) sx->r_laplace_u[k] = p->diffusion.laplace_u;
    if (sx->r_alpha_visc_max_ngb) sx->r_alpha_visc_max_ngb[k] = p->force.alpha_visc_max_ngb;
  }
  run_phase(&px, &runner, PH_EXTRA_GHOST);
  for (int k = 0; k < n; k++) {
    const struct part *p = &parts[k];
    if (sx->e_laplace_u) sx->e_laplace_u[k] = p->diffusion.laplace_u;
    if (sx->e_alpha_visc) sx->e_alpha_visc[k] = p->viscosity.alpha;
    if (sx->e_alpha_diff) sx->e_alpha_diff[k] = p->diffusion.alpha;
    if (sx->e_div_v_prev) sx->e_div_v_prev[k] = p->viscosity.div_v_previous_step;
    if (sx->e_div_v_dt) sx->e_div_v_dt[k] = p->viscosity.div_v_dt;
  }
#endif

  /* ---- force ---- */
  g_stage = 2;
  run_phase(&px, &runner, PH_FORCE);
  run_phase(&px, &runner, PH_END_FORCE);
  for (int k = 0; k < n; k++) {
    const struct part *p = &parts[k];
    if (io->f_a)
      for (int d = 0; d < 3; d++) io->f_a[3 * k + d] = p->a_hydro[d];
    if (io->f_u_dt) io->f_u_dt[k] = p->u_dt;
    if (io->f_h_dt) io->f_h_dt[k] = p->force.h_dt;
    if (io->f_v_sig) io->f_v_sig[k] = PART_V_SIG(p);
    if (io->f_min_ngb_time_bin) io->f_min_ngb_time_bin[k] = p->limiter_data.min_ngb_time_bin;
  }

limiter_loop:
  /* ---- time-step limiter loop (src/runner_main.c:205-206, 255-256, 303-304, 349-350): the tasks of
   * subtype limiter sit on the same cells as the density tasks (src/engine_maketasks.c: the limiter
   * loop is linked next to density / force for every hydro self / pair) and run after the time-step
   * task, i.e. with the NEW bins; kick2 and the time-step task move neither x nor h ---- */
  if (io->l_time_bin != NULL && io->l_wakeup != NULL) {
    for (int k = 0; k < n; k++) {
      parts[k].time_bin = io->l_time_bin[k];
      parts[k].limiter_data.wakeup = time_bin_not_awake;
    }
    for (int k = 0; k < ncells; k++) {
      struct cell *c = &cells[k];
      /* cell_is_starting_hydro (src/active.h:588-601); the flags are what cell_activate_subcell_hydro_tasks
       * (with_timestep_limiter) and cell_activate_limiter leave on the cells of an active task */
      c->hydro.ti_beg_max = cell_has_active(c, &engine) ? ti_current : 0;
      cell_set_flag(c, cell_flag_do_hydro_limiter | cell_flag_do_hydro_sub_limiter);
    }
    run_phase(&px, &runner, PH_LIMITER);
    for (int k = 0; k < n; k++) io->l_wakeup[k] = parts[k].limiter_data.wakeup;
  }
  for (int k = 0; k < ncells; k++) pthread_mutex_destroy(&px.locks[k]);
  free(px.locks);

  g_log = NULL;
  g_sx = NULL;
  for (int k = 0; k < ncells; k++) cell_free_hydro_sorts(&cells[k]);
  free(links);
  free(tasks);
  free(cells);
  free(parts);
  free(xparts);
  return 0;
}

/* ---- time integration ---------------------------------- */
/* The reference's own cell_drift_part / runner_do_kick1 / runner_do_kick2 / runner_do_timestep run
 * on ONE fake leaf cell holding all particles (these tasks are per particle: the cell only
 * carries the loop and the min/max reductions). */
#include "cell.h"
#include "entropy_floor.h"

typedef struct {
  struct space space;
  struct hydro_props hp;
  struct cosmology cosmo;
  struct phys_const pc;
  struct engine engine;
  struct lightcone_array_props lcp;
  struct pressure_floor_props pfloor;
  struct entropy_floor_properties efloor;
  struct runner runner;
  struct cell cell;
  struct part *parts;
  struct xpart *xparts;
} ienv;

static ienv *ienv_make(const oracle_iparams *ip, const oracle_istate *s) {
  ienv *v = NULL;
  if (posix_memalign((void **)&v, 128, sizeof(ienv)) != 0) return NULL;
  bzero(v, sizeof(*v));
  const int n = s->n;
  if (posix_memalign((void **)&v->parts, part_align, (size_t)(n + 1) * sizeof(struct part)) != 0) return NULL;
  if (posix_memalign((void **)&v->xparts, xpart_align, (size_t)(n + 1) * sizeof(struct xpart)) != 0) return NULL;
  bzero(v->parts, (size_t)(n + 1) * sizeof(struct part));
  bzero(v->xparts, (size_t)(n + 1) * sizeof(struct xpart));
  v->space.periodic = 1; /* nobody is removed for leaving the box (cell_drift.c:274) */
  for (int d = 0; d < 3; d++) v->space.dim[d] = 1.0;
  hydro_space_init(&v->space.hs, &v->space);
  v->pc.const_vacuum_permeability = 1.0;
  hydro_props_init_no_hydro(&v->hp);
  v->hp.CFL_condition = ip->cfl_condition;
  v->hp.log_max_h_change = ip->log_max_h_change;
  v->hp.minimal_internal_energy = ip->minimal_internal_energy;
  v->hp.h_max = ip->h_max;
  v->hp.h_min = ip->h_min;
  cosmology_init_no_cosmo(&v->cosmo);
  struct engine *e = &v->engine;
  e->hydro_properties = &v->hp;
  e->physical_constants = &v->pc;
  e->s = &v->space;
  e->cosmology = &v->cosmo;
  e->lightcone_array_properties = &v->lcp;
  e->pressure_floor_props = &v->pfloor;
  e->entropy_floor = &v->efloor;
  e->policy = engine_policy_hydro;
  e->time_base = ip->time_base;
  e->time_base_inv = 1.0 / ip->time_base;
  e->dt_min = ip->dt_min;
  e->dt_max = ip->dt_max;
  e->dt_max_RMS_displacement = FLT_MAX;
  e->ti_current = ip->ti_current;
  e->max_active_bin = ip->max_active_bin;
  e->nodeID = 0;
  v->space.e = e;
  v->runner.e = e;
#ifdef PLANETARY_SPH
  ref_planetary_eos_init();
#endif
  for (int k = 0; k < n; k++) {
    struct part *p = &v->parts[k];
    struct xpart *xp = &v->xparts[k];
    for (int d = 0; d < 3; d++) {
      p->x[d] = s->x[3 * k + d];
      p->v[d] = s->v[3 * k + d];
      xp->v_full[d] = s->v_full[3 * k + d];
      p->a_hydro[d] = s->a[3 * k + d];
      xp->x_diff[d] = s->x_diff[3 * k + d];
      xp->x_diff_sort[d] = s->x_diff_sort[3 * k + d];
    }
    p->u = s->u[k];
    xp->u_full = s->u_full[k];
    p->h = s->h[k];
    p->rho = s->rho[k];
    p->force.pressure = s->pressure[k];
    p->force.soundspeed = s->soundspeed[k];
    PART_V_SIG(p) = s->v_sig[k];
    p->u_dt = s->u_dt[k];
    p->force.h_dt = s->h_dt[k];
    p->time_bin = s->time_bin[k];
    p->limiter_data.min_ngb_time_bin = s->min_ngb_time_bin ? s->min_ngb_time_bin[k] : num_time_bins + 1;
    p->limiter_data.wakeup = time_bin_not_awake;
    p->mass = 1.f;
    p->id = k;
#ifdef PLANETARY_SPH
    p->mat_id = s->mat_id ? s->mat_id[k] : 0;
#endif
  }
  struct cell *c = &v->cell;
  c->hydro.parts = v->parts;
  c->hydro.xparts = v->xparts;
  c->hydro.count = n;
  c->split = 0;
  c->nodeID = 0;
  c->top = c;
  c->super = c;
  c->hydro.super = c;
  c->width[0] = c->width[1] = c->width[2] = 1.0;
  c->dmin = 1.0;
  return v;
}

static void ienv_store(ienv *v, oracle_istate *s) {
  for (int k = 0; k < s->n; k++) {
    const struct part *p = &v->parts[k];
    const struct xpart *xp = &v->xparts[k];
    for (int d = 0; d < 3; d++) {
      s->x[3 * k + d] = p->x[d];
      s->v[3 * k + d] = p->v[d];
      s->v_full[3 * k + d] = xp->v_full[d];
      s->a[3 * k + d] = p->a_hydro[d];
      s->x_diff[3 * k + d] = xp->x_diff[d];
      s->x_diff_sort[3 * k + d] = xp->x_diff_sort[d];
    }
    s->u[k] = p->u;
    s->u_full[k] = xp->u_full;
    s->h[k] = p->h;
    s->rho[k] = p->rho;
    s->pressure[k] = p->force.pressure;
    s->soundspeed[k] = p->force.soundspeed;
    s->v_sig[k] = PART_V_SIG(p);
    s->u_dt[k] = p->u_dt;
    s->h_dt[k] = p->force.h_dt;
    s->time_bin[k] = p->time_bin;
  }
  free(v->parts);
  free(v->xparts);
  free(v);
}

int ref_drift(const oracle_iparams *ip, oracle_istate *s, long long ti_old, float dx_max[2]) {
  ienv *v = ienv_make(ip, s);
  if (v == NULL) return 1;
  /* nobody is "active" during the drift itself: cell_drift.c:342 would otherwise zero the density
   * sums of the active particles (hydro_init_part), which the caller does at the start of the step */
  v->engine.max_active_bin = -1;
  v->cell.hydro.ti_old_part = ti_old;
  cell_drift_part(&v->cell, &v->engine, /*force=*/1, NULL);
  dx_max[0] = v->cell.hydro.dx_max_part;
  dx_max[1] = v->cell.hydro.dx_max_sort;
  ienv_store(v, s);
  return 0;
}

int ref_kick(const oracle_iparams *ip, oracle_istate *s, int which) {
  ienv *v = ienv_make(ip, s);
  if (v == NULL) return 1;
  /* make the cell pass cell_is_starting_hydro / cell_is_active_hydro (src/active.h) */
  v->cell.hydro.ti_beg_max = ip->ti_current;
  v->cell.hydro.ti_end_min = ip->ti_current;
  v->space.periodic = 0; /* no mesh-gravity kick (runner_time_integration.c:125: e->mesh is not set up) */
  if (which == 1)
    runner_do_kick1(&v->runner, &v->cell, 0);
  else
    runner_do_kick2(&v->runner, &v->cell, 0);
  ienv_store(v, s);
  return 0;
}

int ref_timestep(const oracle_iparams *ip, oracle_istate *s, long long *ti_end_min, long long *updated) {
  ienv *v = ienv_make(ip, s);
  if (v == NULL) return 1;
  v->cell.hydro.ti_end_min = ip->ti_current;
  runner_do_timestep(&v->runner, &v->cell, 0);
  *ti_end_min = v->cell.hydro.ti_end_min;
  *updated = v->cell.hydro.updated;
  ienv_store(v, s);
  return 0;
}

int ref_limiter(const oracle_iparams *ip, oracle_istate *s, signed char *wakeup, long long *ti_end_min,
                long long *limited) {
  ienv *v = ienv_make(ip, s);
  if (v == NULL) return 1;
  long long count = 0;
  for (int k = 0; k < s->n; k++) {
    v->parts[k].limiter_data.wakeup = wakeup[k];
    v->parts[k].limiter_data.to_be_synchronized = 0;
    if (wakeup[k] != time_bin_not_awake) count++;
  }
  v->cell.hydro.ti_end_min = max_nr_timesteps;
  v->cell.hydro.ti_beg_max = 0;
  v->space.periodic = 0; /* as in ref_kick: no mesh */
  runner_do_limiter(&v->runner, &v->cell, /*force=*/1, 0);
  *ti_end_min = v->cell.hydro.ti_end_min;
  *limited = count;
  for (int k = 0; k < s->n; k++) wakeup[k] = v->parts[k].limiter_data.wakeup;
  ienv_store(v, s);
  return 0;
}

/* ---- planetary equation of state ------------------------------------------ */
#ifdef PLANETARY_SPH
#include "equation_of_state.h"
#include "units.h"

/* Fills the reference's global `eos` with its own setters and unit conversion
 * (set_idg_def, set_Til_*, convert_units_Til: what eos_init does for
 * EoS:planetary_use_idg_def / _Til_iron / _Til_granite / _Til_water /
 * _Til_basalt = 1, src/equation_of_state/planetary/equation_of_state.h:1599-1627)
 * in the internal units of examples/Planetary/EarthImpact/earth_impact.yml:2-7. */
static int g_eos_ready = 0;
static void ref_planetary_eos_init(void) {
  if (g_eos_ready) return;
  struct unit_system us;
  units_init(&us, 1e27, 1e8, 1.0, 1.0, 1.0); /* mass g, length cm, time s (velocity 1e8 cm/s) */
  bzero(&eos, sizeof(eos));
  set_idg_def(&eos.idg_def, eos_planetary_id_idg_def);
  set_Til_iron(&eos.Til_iron, eos_planetary_id_Til_iron);
  convert_units_Til(&eos.Til_iron, &us);
  set_Til_granite(&eos.Til_granite, eos_planetary_id_Til_granite);
  convert_units_Til(&eos.Til_granite, &us);
  set_Til_water(&eos.Til_water, eos_planetary_id_Til_water);
  convert_units_Til(&eos.Til_water, &us);
  set_Til_basalt(&eos.Til_basalt, eos_planetary_id_Til_basalt);
  convert_units_Til(&eos.Til_basalt, &us);
  g_eos_ready = 1;
}

static void copy_til(oracle_til *o, const struct Til_params *m) {
  o->rho_0 = m->rho_0; o->a = m->a; o->b = m->b; o->A = m->A; o->B = m->B;
  o->u_0 = m->u_0; o->u_iv = m->u_iv; o->u_cv = m->u_cv; o->alpha = m->alpha; o->beta = m->beta;
  o->eta_min = m->eta_min; o->eta_zero = m->eta_zero; o->P_min = m->P_min;
  o->mat_id = (int)m->mat_id;
}

int ref_get_eos(oracle_eos *out) {
  ref_planetary_eos_init();
  bzero(out, sizeof(*out));
  out->n_til = 4;
  copy_til(&out->til[0], &eos.Til_iron);
  copy_til(&out->til[1], &eos.Til_granite);
  copy_til(&out->til[2], &eos.Til_water);
  copy_til(&out->til[3], &eos.Til_basalt);
  for (int k = 0; k < 4; k++) out->idg_gamma[k] = eos.idg_def.gamma;
  return 0;
}

/* ---- table materials: the reference's own loaders and interpolation ---- */
static struct SESAME_params *sesame_slot(int mat_id) {
  switch (mat_id) {
    case eos_planetary_id_SESAME_iron: return &eos.SESAME_iron;
    case eos_planetary_id_SESAME_basalt: return &eos.SESAME_basalt;
    case eos_planetary_id_SESAME_water: return &eos.SESAME_water;
    case eos_planetary_id_SS08_water: return &eos.SS08_water;
    case eos_planetary_id_ANEOS_forsterite: return &eos.ANEOS_forsterite;
    case eos_planetary_id_ANEOS_iron: return &eos.ANEOS_iron;
    case eos_planetary_id_ANEOS_Fe85Si15: return &eos.ANEOS_Fe85Si15;
    default:
      if (mat_id >= 900 && mat_id < 910) return &eos.custom[mat_id - 900];
      return NULL;
  }
}
static struct HM80_params *hm80_slot(int mat_id) {
  switch (mat_id) {
    case eos_planetary_id_HM80_HHe: return &eos.HM80_HHe;
    case eos_planetary_id_HM80_ice: return &eos.HM80_ice;
    case eos_planetary_id_HM80_rock: return &eos.HM80_rock;
    default: return NULL;
  }
}

int ref_eos_table_load(int mat_id, const char *path) {
  ref_planetary_eos_init();
  struct unit_system us;
  units_init(&us, 1e27, 1e8, 1.0, 1.0, 1.0);
  char file[1024];
  strncpy(file, path, sizeof(file) - 1);
  file[sizeof(file) - 1] = 0;
  struct HM80_params *h = hm80_slot(mat_id);
  if (h != NULL) {
    if (mat_id == eos_planetary_id_HM80_HHe) set_HM80_HHe(h, mat_id);
    else if (mat_id == eos_planetary_id_HM80_ice) set_HM80_ice(h, mat_id);
    else set_HM80_rock(h, mat_id);
    load_table_HM80(h, file);
    prepare_table_HM80(h);
    convert_units_HM80(h, &us);
    return 0;
  }
  struct SESAME_params *m = sesame_slot(mat_id);
  if (m == NULL) return 1;
  switch (mat_id) {
    case eos_planetary_id_SESAME_iron: set_SESAME_iron(m, mat_id); break;
    case eos_planetary_id_SESAME_basalt: set_SESAME_basalt(m, mat_id); break;
    case eos_planetary_id_SESAME_water: set_SESAME_water(m, mat_id); break;
    case eos_planetary_id_SS08_water: set_SS08_water(m, mat_id); break;
    case eos_planetary_id_ANEOS_forsterite: set_ANEOS_forsterite(m, mat_id); break;
    case eos_planetary_id_ANEOS_iron: set_ANEOS_iron(m, mat_id); break;
    case eos_planetary_id_ANEOS_Fe85Si15: set_ANEOS_Fe85Si15(m, mat_id); break;
    default: set_custom(m, mat_id); break;
  }
  load_table_SESAME(m, file);
  prepare_table_SESAME(m);
  convert_units_SESAME(m, &us);
  return 0;
}

int ref_eos_unit_ratios(double out[4]) {
  struct unit_system us, si;
  units_init(&us, 1e27, 1e8, 1.0, 1.0, 1.0);
  units_init_si(&si);
  out[0] = units_cgs_conversion_factor(&si, UNIT_CONV_DENSITY) / units_cgs_conversion_factor(&us, UNIT_CONV_DENSITY);
  out[1] = units_cgs_conversion_factor(&si, UNIT_CONV_ENERGY_PER_UNIT_MASS) /
           units_cgs_conversion_factor(&us, UNIT_CONV_ENERGY_PER_UNIT_MASS);
  out[2] = units_cgs_conversion_factor(&si, UNIT_CONV_PRESSURE) / units_cgs_conversion_factor(&us, UNIT_CONV_PRESSURE);
  out[3] = units_cgs_conversion_factor(&si, UNIT_CONV_SPEED) / units_cgs_conversion_factor(&us, UNIT_CONV_SPEED);
  return 0;
}

int ref_eos_table_get(int mat_id, oracle_eos_table *o) {
  bzero(o, sizeof(*o));
  o->mat_id = mat_id;
  const struct HM80_params *h = hm80_slot(mat_id);
  if (h != NULL) {
    if (h->table_log_P_rho_u == NULL) return 2;
    o->kind = 2;
    o->num_rho = h->num_rho;
    o->num_T = h->num_u;
    o->P = h->table_log_P_rho_u;
    o->log_rho_min = h->log_rho_min; o->log_rho_step = h->log_rho_step; o->inv_log_rho_step = h->inv_log_rho_step;
    o->log_u_min = h->log_u_min; o->log_u_step = h->log_u_step; o->inv_log_u_step = h->inv_log_u_step;
    o->bulk_mod = h->bulk_mod; o->P_min_for_c_min = h->P_min_for_c_min;
    return 0;
  }
  const struct SESAME_params *m = sesame_slot(mat_id);
  if (m == NULL) return 1;
  if (m->table_log_rho == NULL) return 2;
  o->kind = 1;
  o->num_rho = m->num_rho;
  o->num_T = m->num_T;
  o->log_rho = m->table_log_rho;
  o->log_u = m->table_log_u_rho_T;
  o->P = m->table_P_rho_T;
  o->c = m->table_c_rho_T;
  o->P_tiny = m->P_tiny;
  o->c_tiny = m->c_tiny;
  return 0;
}

int ref_eos_table_eval(int mat_id, int n, const float *rho, const float *u, float *P, float *c) {
  ref_planetary_eos_init();
  for (int k = 0; k < n; k++) {
    P[k] = gas_pressure_from_internal_energy(rho[k], u[k], (enum eos_planetary_material_id)mat_id);
    c[k] = gas_soundspeed_from_internal_energy(rho[k], u[k], (enum eos_planetary_material_id)mat_id);
  }
  return 0;
}
#else
int ref_get_eos(oracle_eos *out) {
  (void)out;
  return 1;
}
int ref_eos_table_load(int mat_id, const char *path) { (void)mat_id; (void)path; return 1; }
int ref_eos_unit_ratios(double out[4]) { (void)out; return 1; }
int ref_eos_table_get(int mat_id, oracle_eos_table *o) { (void)mat_id; (void)o; return 1; }
int ref_eos_table_eval(int mat_id, int n, const float *rho, const float *u, float *P, float *c) {
  (void)mat_id; (void)n; (void)rho; (void)u; (void)P; (void)c;
  return 1;
}
#endif
