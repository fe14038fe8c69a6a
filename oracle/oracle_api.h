/* TEST INFRASTRUCTURE ONLY.  Plain-C interface of oracle/_ref/libswiftref_*.so:
 * the harness that drives the reference's own operators
 * (runner_do_hydro_sort, runner_dosub_{self,pair}1_density, runner_do_ghost,
 * runner_dosub_{self,pair}2_force, runner_do_end_hydro_force, ...) on a cell
 * tree described by flat arrays, in the order tests/test125cells.c:660-822
 * calls them.  Shared by ref_harness.c, sph_oracle.c and the ctypes tests. */
#ifndef ORACLE_API_H
#define ORACLE_API_H

#ifdef __cplusplus
extern "C" {
#endif

/* One cell of the tree (top-level or progeny).  Particles of a cell are the
 * contiguous range [first, first+count) of the particle arrays, and a
 * progeny's range lies inside its parent's (src/space_split.c). */
typedef struct {
  double loc[3];
  double width[3];
  int first;
  int count;
  int parent;     /* -1 at top level */
  int progeny[8]; /* -1 where absent */
  int depth;
  int split;
} oracle_cell;

/* A hydro task after scheduler_splittask_hydro (src/scheduler.c:1154). */
enum { ORACLE_TASK_SELF = 0, ORACLE_TASK_PAIR = 1, ORACLE_TASK_SUB_SELF = 2, ORACLE_TASK_SUB_PAIR = 3 };
typedef struct {
  int type;
  int ci;
  int cj; /* -1 for self tasks */
} oracle_task;

typedef struct {
  double dim[3];
  int periodic;
  float eta_neighbours;  /* SPH:resolution_eta */
  float h_tolerance;     /* SPH:h_tolerance */
  float h_max;           /* SPH:h_max */
  float h_min;
  int max_smoothing_iterations;
  float viscosity_alpha;
  int max_active_bin;    /* particles with time_bin <= this are active */
  int record_pairs;      /* log ordered pairs into the buffers below */
  int scheme;            /* ORACLE_SCHEME_*; the ref harness ignores it (compile-time there) */
} oracle_params;

enum { ORACLE_SCHEME_MINIMAL = 0, ORACLE_SCHEME_PLANETARY = 1, ORACLE_SCHEME_SPHENIX = 2 };

/* Tillotson material in INTERNAL units (src/equation_of_state/planetary/tillotson.h:43-143) */
typedef struct {
  float rho_0, a, b, A, B, u_0, u_iv, u_cv, alpha, beta, eta_min, eta_zero, P_min;
  int mat_id;
} oracle_til;

/* ---- table equations of state of the Planetary scheme ----
 * One material's table as the reference holds it after load_table_* + prepare_table_* + convert_units_*
 * (src/equation_of_state/planetary/sesame.h:149-352, hm80.h:76-178), internal units of ref_get_eos.  Pointers
 * are into the reference's own arrays. */
typedef struct {
  int kind;                 /* 1 SESAME-style (SESAME, SS08, ANEOS, custom), 2 HM80 */
  int mat_id;
  int num_rho, num_T;       /* HM80: num_rho, num_u */
  const float *log_rho;     /* SESAME [num_rho] */
  const float *log_u;       /* SESAME [num_rho][num_T] */
  const float *P;           /* SESAME P [num_rho][num_T]; HM80 log P [num_rho][num_u] */
  const float *c;           /* SESAME [num_rho][num_T] */
  float P_tiny, c_tiny;
  float log_rho_min, log_rho_step, inv_log_rho_step, log_u_min, log_u_step, inv_log_u_step;
  float bulk_mod, P_min_for_c_min;
} oracle_eos_table;

typedef struct {
  int n_til;
  oracle_til til[8];
  float idg_gamma[4]; /* planetary ideal gas gammas, mat_id 0..2 (planetary/ideal_gas.h) */
  /* materials that live in tables (mat_id >= 200): prepared tables, e.g. those of ref_eos_table_get or of
   * the fixture tests/golden/eos_tables/reference_results.npz.  n_tab = 0: none. */
  int n_tab;
  const oracle_eos_table *tab;
} oracle_eos;

/* Particle arrays (SoA, n entries; vectors are [n][3] row-major). */
typedef struct {
  int n;
  /* inputs */
  const double *x;
  const float *v;
  const float *mass;
  const float *h;
  const float *u;
  const int *mat_id;            /* Planetary only, may be NULL */
  const long long *id;
  const signed char *time_bin;  /* may be NULL: all 1 */
  /* stage 0: raw neighbour sums after the first density pass */
  float *d_rho, *d_rho_dh, *d_wcount, *d_wcount_dh, *d_div_v, *d_rot_v;
  /* stage 1: after the ghost */
  float *g_h, *g_rho, *g_f, *g_pressure, *g_soundspeed, *g_balsara;
  /* stage 2: after force + end_force */
  float *f_a, *f_u_dt, *f_h_dt, *f_v_sig;
  signed char *f_min_ngb_time_bin;
  /* optional inputs: the state INACTIVE particles carry over from their last
   * active step (p->rho, p->force.{pressure, f, soundspeed, balsara}); the
   * force loop of their active neighbours reads it.  NULL: zeros. */
  const float *in_rho, *in_pressure, *in_f, *in_soundspeed, *in_balsara;
  /* optional: the smoothing lengths at the last rebuild, from which c->hydro.h_max_old is taken (it
   * decides how deep the sub-tasks recurse and stays fixed between rebuilds, src/space_split.c).
   * NULL: the current h (a step right after a rebuild). */
  const float *in_h_build;
  /* optional, the time-step LIMITER loop (task_subtype_limiter, src/runner_doiact_limiter.c with
   * runner_iact_nonsym_limiter, src/timestep_limiter_iact.h:107-117): run after end_force on the same
   * positions and smoothing lengths, with the time bins runner_do_timestep has just given
   * (l_time_bin, every particle); l_wakeup = p->limiter_data.wakeup afterwards (time_bin_not_awake =
   * -56 where nobody asked).  Both NULL: the loop is not run. */
  const signed char *l_time_bin;
  signed char *l_wakeup;
  /* != 0: sort and the limiter loop ONLY, on x, h as given (the engine loop of tests/integ_lib.py calls
   * the step first, then the time-step task, then this: the new bins are not known earlier) */
  int l_only;
} oracle_parts;

/* Ordered (updated id, neighbour id) pairs.  loop 0: first density pass,
 * loop 1: ghost redo passes (all sweeps concatenated), loop 2: force.
 * count[] is always the true number, even beyond capacity. */
typedef struct {
  long long *pairs[3]; /* [capacity][2] each, may be NULL */
  long long capacity[3];
  long long count[3];
  int ghost_sweeps;    /* number of redo density sweeps that had work */
} oracle_pairlog;

/* Runs sort -> density -> ghost -> force -> end_force with the REFERENCE's
 * compiled operators.  Returns 0 on success. */
int ref_run_step(const oracle_params *params, oracle_parts *parts,
                 int ncells, const oracle_cell *cells, int ntasks,
                 const oracle_task *tasks, oracle_pairlog *log);

/* sizeof(struct part), offsetof checks for the INTEGRATION shim */
int ref_sizeof_part(void);
/* Planetary build only: the reference's own eos_init() result (internal units
 * UnitMass 1e27 g, UnitLength 1e8 cm, UnitVelocity 1e8 cm/s as
 * examples/Planetary/EarthImpact/earth_impact.yml:2-7). Returns 0 when compiled in. */
int ref_get_eos(oracle_eos *out);

/* ---- table equations of state: the reference's own loaders and functions (libswiftref_planetary.so only) ---- */
/* what eos_init does for EoS:planetary_use_<material> = 1 with EoS:planetary_<material>_table_file = path
 * (src/equation_of_state/planetary/equation_of_state.h:1629-1745); mat_id: 200-202, 300-303, 400-402, 900-909 */
int ref_eos_table_load(int mat_id, const char *path);
/* units_cgs_conversion_factor(&si, X) / units_cgs_conversion_factor(us, X) for X = density, energy per unit mass,
 * pressure, speed in the harness' unit system: what convert_units_* scale the tables with */
int ref_eos_unit_ratios(double out[4]);
int ref_eos_table_get(int mat_id, oracle_eos_table *out);
/* gas_pressure_from_internal_energy / gas_soundspeed_from_internal_energy (rho[k], u[k], mat_id) */
int ref_eos_table_eval(int mat_id, int n, const float *rho, const float *u, float *P, float *c);

/* The cell tree and the hydro task list of a particle set by the reference's OWN space_regrid,
 * space_split_recursive, engine_make_hydroloop_tasks_mapper and scheduler_splittasks (oracle/ref_tree.c).
 * order[k] = input index of the particle at cell-order slot k; arrays are malloc'ed (free()). */
int ref_build_tree_tasks(const oracle_params *prm, int n, const double *x, const float *h, int *order,
                         oracle_cell **cells_out, int *ncells_out, int cdim_out[3], oracle_task **tasks_out,
                         int *ntasks_out);

/* ---- SPHENIX: the three-loop scheme, src/hydro/SPHENIX -------------------
 * sort -> density -> ghost (hydro_prepare_gradient + hydro_reset_gradient at convergence) ->
 * GRADIENT loop (runner_do{self,pair}1_branch_gradient, runner_dosub_{self,pair}1_gradient: the
 * DOSELF1 / DOPAIR1 walks with runner_iact_nonsym_gradient) -> runner_do_extra_ghost
 * (hydro_end_gradient, hydro_prepare_force, hydro_reset_acceleration) -> force -> end_force.
 * Everything the two-loop interface above does not carry travels in this struct; the oracle_parts
 * fields keep their meaning (d_div_v / f_v_sig are p->viscosity.div_v / p->viscosity.v_sig). */
typedef struct {
  /* hydro_props->viscosity / ->diffusion (src/hydro/SPHENIX/hydro_parameters.h:105-140) and
   * e->time_base (dt_alpha = get_timestep(p->time_bin, time_base), src/runner_ghost.c:1045-1046) */
  float viscosity_alpha_max, viscosity_alpha_min, viscosity_length;
  float diffusion_beta, diffusion_alpha_max, diffusion_alpha_min;
  double time_base;
  /* inputs, every particle: p->viscosity.alpha, p->diffusion.alpha, p->viscosity.div_v_previous_step */
  const float *in_alpha_visc, *in_alpha_diff, *in_div_v_prev;
  /* after the ghost: p->viscosity.div_v (hydro_end_density) */
  float *g_div_v;
  /* after the gradient loop, before the extra ghost: p->viscosity.v_sig, p->diffusion.laplace_u (raw
   * sum), p->force.alpha_visc_max_ngb */
  float *r_v_sig, *r_laplace_u, *r_alpha_visc_max_ngb;
  /* after the extra ghost: p->diffusion.laplace_u, p->viscosity.{alpha, div_v_previous_step, div_v_dt},
   * p->diffusion.alpha */
  float *e_laplace_u, *e_alpha_visc, *e_alpha_diff, *e_div_v_prev, *e_div_v_dt;
  /* ordered (updated id, neighbour id) pairs of the gradient loop */
  long long *grad_pairs; /* [grad_capacity][2], may be NULL */
  long long grad_capacity;
  long long grad_count;
} oracle_sphenix;

/* The reference's own operators (libswiftref_sphenix.so only; returns 1 elsewhere). */
int ref_run_step_sphenix(const oracle_params *params, oracle_parts *parts, int ncells, const oracle_cell *cells,
                         int ntasks, const oracle_task *tasks, oracle_pairlog *log, oracle_sphenix *sx);

/* ---- the plain-C restatement (oracle/sph_oracle.c) ---- */
/* params->scheme must be ORACLE_SCHEME_SPHENIX */
int oracle_run_step_sphenix(const oracle_params *prm, oracle_parts *parts, int ncells, const oracle_cell *cells,
                            int ntasks, const oracle_task *tasks, oracle_pairlog *log, oracle_sphenix *sx);
int oracle_build_tree(const oracle_params *prm, int n, const double *x, const float *h, int *order,
                      oracle_cell **cells_out, int *ncells_out, int cdim_out[3]);
int oracle_build_tasks(const oracle_params *prm, int ncells, const oracle_cell *cells, const int cdim[3],
                       const float *h_sorted, oracle_task **tasks_out, int *ntasks_out);
int oracle_run_step(const oracle_params *prm, const oracle_eos *eos, oracle_parts *parts, int ncells,
                    const oracle_cell *cells, int ntasks, const oracle_task *tasks, oracle_pairlog *log);
void oracle_free(void *p);

/* ---- time integration either side of the path --------
 * drift (src/drift.h:143-233 inside the particle loop of src/cell_drift.c:233-340),
 * kick (src/kick.h:215-279 as called by src/runner_time_integration.c:141-180, 412-457),
 * time-step (src/timestep.h:45-82, 141-222; src/runner_time_integration.c:696-745),
 * without cosmology, gravity, cooling; the time-step limiter is ref_limiter / oracle_limiter below and the
 * l_* fields of oracle_parts.  All arrays have n entries ([n][3] vectors). */
typedef struct {
  int scheme;                    /* ORACLE_SCHEME_* (the ref harness ignores it) */
  float cfl_condition;           /* hydro_props->CFL_condition */
  float log_max_h_change;        /* hydro_props->log_max_h_change */
  float minimal_internal_energy; /* hydro_props->minimal_internal_energy */
  float h_max, h_min;            /* hydro_props->h_max / h_min (cell_drift.c:303-304) */
  double dt_min, dt_max;         /* e->dt_min, e->dt_max */
  double time_base;              /* e->time_base; time_base_inv = 1 / time_base */
  int max_active_bin;            /* e->max_active_bin */
  long long ti_current;          /* e->ti_current */
} oracle_iparams;

typedef struct {
  int n;
  double *x;
  float *v, *v_full;             /* p->v (predicted), xp->v_full */
  float *u, *u_full;             /* p->u (predicted), xp->u_full */
  float *h, *rho, *pressure, *soundspeed, *v_sig;
  float *a, *u_dt, *h_dt;        /* p->a_hydro, p->u_dt, p->force.h_dt */
  float *x_diff, *x_diff_sort;   /* xp->x_diff, xp->x_diff_sort */
  signed char *time_bin;
  const signed char *min_ngb_time_bin;
  const int *mat_id;             /* Planetary only, may be NULL */
} oracle_istate;

/* every particle drifts from ti_old to ip->ti_current (dt = (ti_current - ti_old) * time_base, the
 * non-cosmological branch of cell_drift.c:232-236): drift_part + the h clamp; dx_max[0] =
 * c->hydro.dx_max_part, dx_max[1] = c->hydro.dx_max_sort of a cell holding all particles */
int ref_drift(const oracle_iparams *ip, oracle_istate *s, long long ti_old, float dx_max[2]);
int oracle_drift(const oracle_iparams *ip, const oracle_eos *eos, oracle_istate *s, long long ti_old,
                 float dx_max[2]);
/* which = 1: runner_do_kick1 (starting particles, first half step of their bin);
 * which = 2: runner_do_kick2 (active particles, second half) + hydro_reset_predicted_values */
int ref_kick(const oracle_iparams *ip, oracle_istate *s, int which);
int oracle_kick(const oracle_iparams *ip, const oracle_eos *eos, oracle_istate *s, int which);
/* runner_do_timestep on the active particles: new time_bin; ti_end_min = min(ti_current + new step),
 * updated = particles that got a new step.  Returns 2 when a particle wants a step below dt_min. */
int ref_timestep(const oracle_iparams *ip, oracle_istate *s, long long *ti_end_min, long long *updated);
int oracle_timestep(const oracle_iparams *ip, oracle_istate *s, long long *ti_end_min, long long *updated);
/* runner_do_limiter (src/runner_time_integration.c:1324-1467, force = 1) on a cell holding all particles:
 * timestep_limit_part (src/timestep_limiter.h:67-217) for every particle with wakeup[k] != -56
 * (time_bin_not_awake) -- an active one only takes the bin -wakeup + 2; an inactive one has its kick1
 * undone, is kicked to the start of its new, shorter step and, if that bin does not start now, given
 * the first half-kick of it.  Updates v_full, u_full, u_dt, time_bin; wakeup[k] is reset.  ti_end_min =
 * the earliest new end of step (max_nr_timesteps if nobody was woken); limited = their number. */
int ref_limiter(const oracle_iparams *ip, oracle_istate *s, signed char *wakeup, long long *ti_end_min,
                long long *limited);
int oracle_limiter(const oracle_iparams *ip, const oracle_eos *eos, oracle_istate *s, signed char *wakeup,
                   long long *ti_end_min, long long *limited);

#ifdef __cplusplus
}
#endif
#endif
