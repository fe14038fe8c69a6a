// TEST INFRASTRUCTURE ONLY.  The body of the device kernel k_limiter -- limit_part() of
// swift_subtask_dev_b200/csrc/limit_part.h, the very source nvcc compiles -- built for the host (g++ -O2
// -ffp-contract=off, the counterpart of nvcc -fmad=false) behind a loop over arrays, so that
// tests/test_limiter_oracle.py can hold it against the fixture the compiled reference wrote and against the
// oracle.  Never linked into libswiftcuda.so, never used by the product.
#include "../../swift_subtask_dev_b200/csrc/limit_part.h"

extern "C" int limit_parts_host(int n, long long ti_current, double time_base, int max_active_bin, float energy_min,
                                signed char *wakeup, signed char *time_bin, float *v_full /* [n][3] */,
                                float *u_full, const float *a /* [n][3] */, float *u_dt, long long *ti_end_min,
                                long long *limited) {
  swiftcuda::LimitEnv e;
  e.ti_current = ti_current;
  e.time_base = time_base;
  e.max_active_bin = max_active_bin;
  e.energy_min = energy_min;
  long long t_min = 1LL << 57, count = 0;
  for (int k = 0; k < n; k++) {
    if (wakeup[k] == swiftcuda::kLpNotAwake) continue;
    int bin = time_bin[k];
    float vf[4] = {v_full[3 * k], v_full[3 * k + 1], v_full[3 * k + 2], u_full[k]};
    bool dirty = false;
    const bool inactive = bin > max_active_bin;
    const long long t = swiftcuda::limit_part(e, wakeup[k], bin, vf, &a[3 * k], u_dt[k], dirty);
    if (inactive) {  // as the kernel: an active particle's xpart is not written
      v_full[3 * k] = vf[0];
      v_full[3 * k + 1] = vf[1];
      v_full[3 * k + 2] = vf[2];
      u_full[k] = vf[3];
    }
    time_bin[k] = (signed char)bin;
    wakeup[k] = swiftcuda::kLpNotAwake;
    t_min = t < t_min ? t : t_min;
    count++;
  }
  *ti_end_min = t_min;
  *limited = count;
  return 0;
}
