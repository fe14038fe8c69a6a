/* TEST INFRASTRUCTURE ONLY.  Force-included (gcc -include) in front of the
 * reference's src/runner_doiact_hydro.c by oracle/Makefile.ref so that every
 * particle interaction the reference's own loops apply is also logged as an
 * ordered (updated particle id, neighbour id) pair.  The reference source is
 * compiled where it lies; nothing is copied.  The IACT/IACT_NONSYM macros of
 * src/runner_doiact_hydro.h:98-102 paste the names runner_iact[_nonsym]_<loop>,
 * which are then macro-expanded again into the wrappers below. */
#ifndef ORACLE_REF_RECORD_H
#define ORACLE_REF_RECORD_H

#include <config.h>

#include "hydro.h"

/* implemented in ref_harness.c */
void ref_record_pair(int loop, long long id_updated, long long id_other);

__attribute__((always_inline)) static inline void rec_iact_nonsym_density(
    const float r2, const float dx[3], const float hi, const float hj,
    struct part *restrict pi, const struct part *restrict pj, const float a,
    const float H) {
  ref_record_pair(0, pi->id, pj->id);
  runner_iact_nonsym_density(r2, dx, hi, hj, pi, pj, a, H);
}
__attribute__((always_inline)) static inline void rec_iact_density(
    const float r2, const float dx[3], const float hi, const float hj,
    struct part *restrict pi, struct part *restrict pj, const float a,
    const float H) {
  ref_record_pair(0, pi->id, pj->id);
  ref_record_pair(0, pj->id, pi->id);
  runner_iact_density(r2, dx, hi, hj, pi, pj, a, H);
}
__attribute__((always_inline)) static inline void rec_iact_nonsym_force(
    const float r2, const float dx[3], const float hi, const float hj,
    struct part *restrict pi, const struct part *restrict pj, const float a,
    const float H) {
  ref_record_pair(1, pi->id, pj->id);
  runner_iact_nonsym_force(r2, dx, hi, hj, pi, pj, a, H);
}
__attribute__((always_inline)) static inline void rec_iact_force(
    const float r2, const float dx[3], const float hi, const float hj,
    struct part *restrict pi, struct part *restrict pj, const float a,
    const float H) {
  ref_record_pair(1, pi->id, pj->id);
  ref_record_pair(1, pj->id, pi->id);
  runner_iact_force(r2, dx, hi, hj, pi, pj, a, H);
}

#ifdef EXTRA_HYDRO_LOOP
/* SPHENIX: the gradient loop (loop 2 of the recorder) */
__attribute__((always_inline)) static inline void rec_iact_nonsym_gradient(
    const float r2, const float dx[3], const float hi, const float hj,
    struct part *restrict pi, struct part *restrict pj, const float a,
    const float H) {
  ref_record_pair(2, pi->id, pj->id);
  runner_iact_nonsym_gradient(r2, dx, hi, hj, pi, pj, a, H);
}
__attribute__((always_inline)) static inline void rec_iact_gradient(
    const float r2, const float dx[3], const float hi, const float hj,
    struct part *restrict pi, struct part *restrict pj, const float a,
    const float H) {
  ref_record_pair(2, pi->id, pj->id);
  ref_record_pair(2, pj->id, pi->id);
  runner_iact_gradient(r2, dx, hi, hj, pi, pj, a, H);
}
#define runner_iact_nonsym_gradient rec_iact_nonsym_gradient
#define runner_iact_gradient rec_iact_gradient
#endif

#define runner_iact_nonsym_density rec_iact_nonsym_density
#define runner_iact_density rec_iact_density
#define runner_iact_nonsym_force rec_iact_nonsym_force
#define runner_iact_force rec_iact_force

#endif
