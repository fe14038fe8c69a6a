// TEST INFRASTRUCTURE ONLY.  swift_subtask_dev_b200/csrc/eos_table.h -- the very source nvcc compiles into the
// device's eos_eval -- built for the host (g++ -O2 -ffp-contract=off, the counterpart of nvcc -fmad=false)
// behind loops over arrays, so that tests/test_eos_tables.py can hold it against the reference's own
// SESAME_* / HM80_* functions on the same tables.  Never linked into libswiftcuda.so, never used by the product.
#include "../../swift_subtask_dev_b200/csrc/eos_table.h"

#include "../../include/swift_cuda.h"

extern "C" int eos_table_eval_host(const swift_cuda_eos_table *t, int n, const float *rho, const float *u, float *P,
                                   float *c) {
  swiftcuda::EosTable e;
  e.kind = t->kind;
  e.mat_id = t->mat_id;
  e.num_rho = t->num_rho;
  e.num_T = t->num_T;
  e.log_rho = t->log_rho;
  e.log_u = t->log_u;
  e.P = t->P;
  e.c = t->c;
  e.P_tiny = t->P_tiny;
  e.c_tiny = t->c_tiny;
  e.log_rho_min = t->log_rho_min;
  e.log_rho_step = t->log_rho_step;
  e.inv_log_rho_step = t->inv_log_rho_step;
  e.log_u_min = t->log_u_min;
  e.log_u_step = t->log_u_step;
  e.inv_log_u_step = t->inv_log_u_step;
  e.bulk_mod = t->bulk_mod;
  e.P_min_for_c_min = t->P_min_for_c_min;
  for (int k = 0; k < n; k++) swiftcuda::et_eval(e, rho[k], u[k], P[k], c[k]);
  return 0;
}

extern "C" void eos_table_logf_expf_host(int n, const float *x, float *log_out, float *exp_out) {
  for (int k = 0; k < n; k++) {
    log_out[k] = swiftcuda::et_logf(x[k]);
    exp_out[k] = swiftcuda::et_expf(x[k]);
  }
}
