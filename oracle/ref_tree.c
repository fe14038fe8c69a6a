/* TEST INFRASTRUCTURE ONLY.  Builds the cell tree and the hydro task list of a particle set with the
 * REFERENCE'S OWN code -- space_regrid (src/space_regrid.c:36), space_split_recursive
 * (src/space_split.c:53), engine_make_hydroloop_tasks_mapper (src/engine_maketasks.c:4060) and
 * scheduler_splittasks (src/scheduler.c:1664, which runs scheduler_splittask_hydro, :1154-1378) --
 * and flattens the result into the oracle's cell / task arrays, so that the restatements
 * (oracle_build_tree / oracle_build_tasks, and the product's csrc/host_tree.cpp) can be pinned
 * against the reference itself rather than against each other.
 * What is restated here is only the glue space_rebuild does around those calls: binning the
 * particles into the top-level cells with cell_getid (src/cell.h) and keeping them contiguous. */
#include <config.h>

#include <float.h>
#include <stdlib.h>
#include <string.h>

#include "cell.h"
#include "engine.h"
#include "error.h"
#include "hydro_properties.h"
#include "scheduler.h"
#include "space.h"
#include "space_getsid.h"
#include "task.h"
#include "threadpool.h"

#include "oracle_api.h"

void space_split_recursive(struct space *s, struct cell *c, struct cell_buff *restrict buff,
                           struct cell_buff *restrict sbuff, struct cell_buff *restrict bbuff,
                           struct cell_buff *restrict gbuff, struct cell_buff *restrict sink_buff,
                           const short int tpid);
void engine_make_hydroloop_tasks_mapper(void *map_data, int num_elements, void *extra_data);

struct flat_out {
  oracle_cell *cells;
  struct cell **src;
  int n, cap;
};

static int flat_add(struct flat_out *f, struct cell *c, int parent, const struct part *base) {
  if (f->n == f->cap) {
    f->cap = f->cap ? 2 * f->cap : 1024;
    f->cells = (oracle_cell *)realloc(f->cells, sizeof(oracle_cell) * f->cap);
    f->src = (struct cell **)realloc(f->src, sizeof(struct cell *) * f->cap);
  }
  const int me = f->n++;
  oracle_cell *o = &f->cells[me];
  memset(o, 0, sizeof(*o));
  f->src[me] = c;
  for (int k = 0; k < 3; k++) {
    o->loc[k] = c->loc[k];
    o->width[k] = c->width[k];
  }
  o->first = c->hydro.count > 0 ? (int)(c->hydro.parts - base) : 0;
  o->count = c->hydro.count;
  o->parent = parent;
  o->depth = c->depth;
  o->split = c->split;
  for (int k = 0; k < 8; k++) o->progeny[k] = -1;
  return me;
}

static void flat_progeny(struct flat_out *f, int me, const struct part *base) {
  struct cell *c = f->src[me];
  if (!c->split) return;
  for (int k = 0; k < 8; k++)
    if (c->progeny[k] != NULL && c->progeny[k]->hydro.count > 0) {
      const int child = flat_add(f, c->progeny[k], me, base);
      f->cells[me].progeny[k] = child;
      flat_progeny(f, child, base);
    }
}

static int flat_find(const struct flat_out *f, const struct cell *c) {
  for (int k = 0; k < f->n; k++)
    if (f->src[k] == c) return k;
  return -1;
}

/* order[k] = input index of the particle at cell-order slot k.  Top-level cells come first
 * (indices 0..ntop-1, as s->cells_top), then every top-level cell's progeny depth first. */
int ref_build_tree_tasks(const oracle_params *prm, int n, const double *x, const float *h, int *order,
                         oracle_cell **cells_out, int *ncells_out, int cdim_out[3], oracle_task **tasks_out,
                         int *ntasks_out) {
  static struct space s;
  static struct engine e;
  static struct hydro_props hp;
  static struct threadpool tp;
  static int tp_ready = 0;
  memset(&s, 0, sizeof(s));
  memset(&e, 0, sizeof(e));
  memset(&hp, 0, sizeof(hp));
  hp.h_max = prm->h_max;
  hp.h_min = prm->h_min;
  hp.eta_neighbours = prm->eta_neighbours;
  e.hydro_properties = &hp;
  e.policy = engine_policy_hydro;
  e.nodeID = 0;
  e.nr_nodes = 1;
  e.ti_current = 0;
  e.max_active_bin = prm->max_active_bin;
  if (!tp_ready) {
    threadpool_init(&tp, 1);
    tp_ready = 1;
  }
  e.threadpool = tp;

  for (int k = 0; k < 3; k++) s.dim[k] = prm->dim[k];
  s.periodic = prm->periodic;
  s.with_hydro = 1;
  s.nr_queues = 1;
  const double dmax = fmax(s.dim[0], fmax(s.dim[1], s.dim[2]));
  s.cell_min = 0.99 * dmax / space_max_top_level_cells_default; /* src/space.c:1160 */
  s.nr_parts = n;
  s.size_parts = n;
  if (posix_memalign((void **)&s.parts, part_align, sizeof(struct part) * (size_t)(n > 0 ? n : 1)) != 0 ||
      posix_memalign((void **)&s.xparts, xpart_align, sizeof(struct xpart) * (size_t)(n > 0 ? n : 1)) != 0)
    return 1;
  memset(s.parts, 0, sizeof(struct part) * (size_t)n);
  memset(s.xparts, 0, sizeof(struct xpart) * (size_t)n);
  s.cells_sub = (struct cell **)calloc(4, sizeof(struct cell *));
  s.multipoles_sub = (struct gravity_tensors **)calloc(4, sizeof(struct gravity_tensors *));

  /* the particles, h clamped as space_init / the drift would (h <= h_max) */
  for (int k = 0; k < n; k++) {
    struct part *p = &s.parts[k];
    p->x[0] = x[3 * k]; p->x[1] = x[3 * k + 1]; p->x[2] = x[3 * k + 2];
    p->h = h[k] < prm->h_max ? h[k] : prm->h_max;
    p->id = k;
    p->time_bin = 1;
    p->mass = 1.f;
  }

  /* 1. the top-level grid: the reference's space_regrid (s->e == NULL is a supported state there) */
  s.e = NULL;
  space_regrid(&s, 0);
  s.e = &e;
  e.s = &s;
  for (int k = 0; k < 3; k++) cdim_out[k] = s.cdim[k];
  const int ntop = s.nr_cells;

  /* 2. what space_rebuild does between regrid and split: bin by cell_getid, keep cells contiguous
   *    (a counting sort: the particles of a cell keep their input order) */
  int *top_of = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
  long *start = (long *)calloc((size_t)ntop + 1, sizeof(long));
  for (int k = 0; k < n; k++) {
    const struct part *p = &s.parts[k];
    const int cid = cell_getid(s.cdim, p->x[0] * s.iwidth[0], p->x[1] * s.iwidth[1], p->x[2] * s.iwidth[2]);
    top_of[k] = cid;
    start[cid + 1]++;
  }
  for (int c = 0; c < ntop; c++) start[c + 1] += start[c];
  struct part *sorted = NULL;
  if (posix_memalign((void **)&sorted, part_align, sizeof(struct part) * (size_t)(n > 0 ? n : 1)) != 0) return 1;
  {
    long *fill = (long *)malloc(sizeof(long) * (size_t)(ntop + 1));
    memcpy(fill, start, sizeof(long) * (size_t)ntop);
    for (int k = 0; k < n; k++) sorted[fill[top_of[k]]++] = s.parts[k];
    free(fill);
  }
  free(s.parts);
  s.parts = sorted;
  for (int c = 0; c < ntop; c++) {
    struct cell *ct = &s.cells_top[c];
    ct->hydro.parts = &s.parts[start[c]];
    ct->hydro.xparts = &s.xparts[start[c]];
    ct->hydro.count = (int)(start[c + 1] - start[c]);
    ct->nodeID = 0;
  }

  /* 3. the reference's own split of every top-level cell */
  for (int c = 0; c < ntop; c++)
    if (s.cells_top[c].hydro.count > 0) space_split_recursive(&s, &s.cells_top[c], NULL, NULL, NULL, NULL, NULL, 0);

  /* flatten: top-level cells 0..ntop-1, then their progeny */
  struct flat_out f = {0};
  for (int c = 0; c < ntop; c++) flat_add(&f, &s.cells_top[c], -1, s.parts);
  for (int c = 0; c < ntop; c++) flat_progeny(&f, c, s.parts);
  for (int k = 0; k < n; k++) order[k] = (int)s.parts[k].id;

  /* 4. the reference's own task creation and splitting */
  *tasks_out = NULL;
  *ntasks_out = 0;
  if (tasks_out != NULL) {
    scheduler_init(&e.sched, &s, 64 * ntop + 4096, 1, 0, 0, &e.threadpool);
    scheduler_reset(&e.sched, 64 * ntop + 4096);
    engine_make_hydroloop_tasks_mapper((void *)0, ntop, &e);
    scheduler_splittasks(&e.sched, 0, 0);
    oracle_task *out = (oracle_task *)malloc(sizeof(oracle_task) * (size_t)(e.sched.nr_tasks + 1));
    int nt = 0;
    for (int k = 0; k < e.sched.nr_tasks; k++) {
      const struct task *t = &e.sched.tasks[k];
      if (t->subtype != task_subtype_density) continue;
      int type;
      if (t->type == task_type_self) type = 0;
      else if (t->type == task_type_pair) type = 1;
      else if (t->type == task_type_sub_self) type = 2;
      else if (t->type == task_type_sub_pair) type = 3;
      else continue;
      out[nt].type = type;
      out[nt].ci = flat_find(&f, t->ci);
      out[nt].cj = t->cj ? flat_find(&f, t->cj) : -1;
      nt++;
    }
    *tasks_out = out;
    *ntasks_out = nt;
  }
  *cells_out = f.cells;
  *ncells_out = f.n;
  free(f.src);
  free(top_of);
  free(start);
  return 0;
}
