// Table equations of state of the Planetary scheme: P(rho, u) and c(rho, u) by bilinear
// interpolation in (log rho, log u) --
//   SESAME-style tables (SESAME iron / basalt / water, SS08 water, ANEOS forsterite / iron / Fe85Si15, custom):
//       SESAME_pressure_from_internal_energy, SESAME_soundspeed_from_internal_energy,
//       src/equation_of_state/planetary/sesame.h:480-594, 605-721, with find_value_in_monot_incr_array,
//       src/utilities.h:33-57;
//   Hubbard & MacFarlane (1980) tables (HHe, ice, rock):
//       HM80_pressure_from_internal_energy, HM80_soundspeed_from_internal_energy, planetary/hm80.h:224-300 --
// in the reference's operand order, with logf and expf as glibc computes them (the reference calls libm's;
// CUDA's differ in the last bit for some arguments).  Plain C++ on scalars and raw pointers: the SAME source is
// compiled by nvcc into the device's eos_eval (-fmad=false: no contraction) and, by g++ -ffp-contract=off, into
// the host-side test that holds it against the reference's own functions on the same tables
// (tests/test_eos_tables.py).  The host build exists in tests/ only: the library never evaluates an EoS on the host.
#pragma once
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define SWIFT_ET_HD __host__ __device__ __forceinline__
#else
#define SWIFT_ET_HD inline
#endif

namespace swiftcuda {

enum { kEosTableNone = 0, kEosTableSesame = 1, kEosTableHM80 = 2 };

// One material's table, arrays in the memory of whoever evaluates it (device arrays in the library's constants).
struct EosTable {
  int kind;                 // kEosTable*
  int mat_id;               // enum eos_planetary_material_id (200.., 300.., 400.., 900..)
  int num_rho, num_T;       // SESAME: after dropping the rho = 0 / T = 0 entries; HM80: num_rho, num_u
  const float *log_rho;     // SESAME [num_rho]: log(rho), internal units
  const float *log_u;       // SESAME [num_rho][num_T]: log(u)
  const float *P;           // SESAME [num_rho][num_T]: P;  HM80 [num_rho][num_u]: log(P)
  const float *c;           // SESAME [num_rho][num_T]: sound speed
  float P_tiny, c_tiny;     // SESAME
  float log_rho_min, log_rho_step, inv_log_rho_step;  // HM80
  float log_u_min, log_u_step, inv_log_u_step;        // HM80
  float bulk_mod, P_min_for_c_min;                    // HM80
};

SWIFT_ET_HD uint32_t et_asuint(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  return u;
}
SWIFT_ET_HD float et_asfloat(uint32_t u) {
  float x;
  memcpy(&x, &u, 4);
  return x;
}
SWIFT_ET_HD uint64_t et_asuint64(double x) {
  uint64_t u;
  memcpy(&u, &x, 8);
  return u;
}
SWIFT_ET_HD double et_asdouble(uint64_t u) {
  double x;
  memcpy(&x, &u, 8);
  return x;
}

// logf as glibc >= 2.27 computes it (sysdeps/ieee754/flt-32/e_logf.c with the 16-entry table of logf_data.c):
// x = 2^k z with z in [OFF, 2 OFF), r = z / c - 1 with c near the centre of the i-th subinterval, a cubic in r
// plus log(c) + k ln2, all in double, one rounding to float.  Verified against this machine's libm on every
// positive float (2 139 095 039 arguments, 0 differences; tests/test_eos_tables.py repeats a sample).
SWIFT_ET_HD float et_logf(float x) {
  static const double T[16][2] = {
      {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
      {0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2},  {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
      {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3},
      {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
      {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
      {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4},
      {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3},
      {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};
  const double Ln2 = 0x1.62e42fefa39efp-1;
  const double A0 = -0x1.00ea348b88334p-2, A1 = 0x1.5575b0be00b6ap-2, A2 = -0x1.ffffef20a4123p-2;
  uint32_t ix = et_asuint(x);
  if (ix == 0x3f800000u) return 0.f;
  if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
    // x < 2^-126, inf or nan
    if (ix * 2u == 0u) return et_asfloat(0xff800000u);  // log(+-0) = -inf
    if (ix == 0x7f800000u) return x;                    // log(inf) = inf
    if ((ix & 0x80000000u) || ix * 2u >= 0xff000000u) return et_asfloat(0x7fc00000u);  // log(< 0), log(nan)
    ix = et_asuint(x * 0x1p23f);  // subnormal: normalise
    ix -= 23u << 23;
  }
  const uint32_t tmp = ix - 0x3f330000u;
  const int i = (int)((tmp >> 19) % 16u);
  const int k = (int32_t)tmp >> 23;
  const uint32_t iz = ix - (tmp & (0x1ffu << 23));
  const double z = (double)et_asfloat(iz);
  const double r = z * T[i][0] - 1;
  const double y0 = T[i][1] + (double)k * Ln2;
  const double r2 = r * r;
  double y = A1 * r + A2;
  y = A0 * r2 + y;
  y = y * r2 + (y0 + r);
  return (float)y;
}

// expf as glibc >= 2.27 computes it (sysdeps/ieee754/flt-32/e_expf.c with the 32-entry table of exp2f_data.c);
// the arithmetic of expf_glibc in kernels.cuh, on plain operators.
SWIFT_ET_HD float et_expf(float x) {
  static const uint64_t tab[32] = {
      0x3ff0000000000000ULL, 0x3fefd9b0d3158574ULL, 0x3fefb5586cf9890fULL, 0x3fef9301d0125b51ULL,
      0x3fef72b83c7d517bULL, 0x3fef54873168b9aaULL, 0x3fef387a6e756238ULL, 0x3fef1e9df51fdee1ULL,
      0x3fef06fe0a31b715ULL, 0x3feef1a7373aa9cbULL, 0x3feedea64c123422ULL, 0x3feece086061892dULL,
      0x3feebfdad5362a27ULL, 0x3feeb42b569d4f82ULL, 0x3feeab07dd485429ULL, 0x3feea47eb03a5585ULL,
      0x3feea09e667f3bcdULL, 0x3fee9f75e8ec5f74ULL, 0x3feea11473eb0187ULL, 0x3feea589994cce13ULL,
      0x3feeace5422aa0dbULL, 0x3feeb737b0cdc5e5ULL, 0x3feec49182a3f090ULL, 0x3feed503b23e255dULL,
      0x3feee89f995ad3adULL, 0x3feeff76f2fb5e47ULL, 0x3fef199bdd85529cULL, 0x3fef3720dcef9069ULL,
      0x3fef5818dcfba487ULL, 0x3fef7c97337b9b5fULL, 0x3fefa4afa2a490daULL, 0x3fefd0765b6e4540ULL};
  if (x != x) return x;
  if (x > 0x1.62e42ep6f) return et_asfloat(0x7f800000u);  // > log(2^128): overflow
  if (x < -0x1.9fe368p6f) return 0.f;                     // < log(2^-150): underflow
  const double inv_ln2_n = 0x1.71547652b82fep+0 * 32.0, shift = 0x1.8p+52;
  const double c0 = 0x1.c6af84b912394p-5 / 32.0 / 32.0 / 32.0, c1 = 0x1.ebfce50fac4f3p-3 / 32.0 / 32.0,
               c2 = 0x1.62e42ff0c52d6p-1 / 32.0;
  const double z = inv_ln2_n * (double)x;
  double kd = z + shift;
  const uint64_t ki = et_asuint64(kd);
  kd = kd - shift;
  const double r = z - kd;
  const double s = et_asdouble(tab[ki & 31] + (ki << 47));
  const double zz = c0 * r + c1;
  const double r2 = r * r;
  double y = c2 * r + 1.0;
  y = zz * r2 + y;
  return (float)(y * s);
}

// src/utilities.h:33-57: -1 and n for x below and above the array's edge values
SWIFT_ET_HD int et_find(float x, const float *array, int n) {
  int index_low = 0, index_high = n;
  while (index_high - index_low > 1) {
    const int index_mid = (index_high + index_low) / 2;
    if (array[index_mid] <= x)
      index_low = index_mid;
    else
      index_high = index_mid;
  }
  if (x < array[0]) return -1;
  if (array[n - 1] <= x) return n;
  return index_low;
}

// The common front of SESAME_pressure_from_internal_energy and SESAME_soundspeed_from_internal_energy
// (sesame.h:488-544 = 613-669): indices and interpolation weights in (log rho, log u).
// For a density outside the table the reference searches the u slice of row idx_rho = -1 or num_rho (+1) BEFORE
// it clamps idx_rho, i.e. it reads beyond the array (undefined); here those slice rows are clamped into the
// table, so results agree with the reference's for table_rho[0] <= rho < table_rho[num_rho - 1] and are merely
// defined outside.
struct EtSesameIdx {
  int idx_rho, idx_u_1, idx_u_2;
  float intp_rho, intp_u_1, intp_u_2;
};
SWIFT_ET_HD EtSesameIdx et_sesame_index(float density, float u, const EosTable &m) {
  EtSesameIdx s;
  const float log_rho = et_logf(density);
  const float log_u = et_logf(u);
  const int nT = m.num_T;
  int idx_rho = et_find(log_rho, m.log_rho, m.num_rho);
  const int row1 = idx_rho < 0 ? 0 : (idx_rho > m.num_rho - 1 ? m.num_rho - 1 : idx_rho);
  const int row2 = idx_rho + 1 < 0 ? 0 : (idx_rho + 1 > m.num_rho - 1 ? m.num_rho - 1 : idx_rho + 1);
  int idx_u_1 = et_find(log_u, m.log_u + (size_t)row1 * nT, nT);
  int idx_u_2 = et_find(log_u, m.log_u + (size_t)row2 * nT, nT);
  // outside the table: extrapolate from the edge and edge-but-one values
  if (idx_rho <= -1) idx_rho = 0;
  else if (idx_rho >= m.num_rho) idx_rho = m.num_rho - 2;
  if (idx_u_1 <= -1) idx_u_1 = 0;
  else if (idx_u_1 >= nT) idx_u_1 = nT - 2;
  if (idx_u_2 <= -1) idx_u_2 = 0;
  else if (idx_u_2 >= nT) idx_u_2 = nT - 2;
  // duplicates in the tables
  if (m.log_rho[idx_rho + 1] != m.log_rho[idx_rho])
    s.intp_rho = (log_rho - m.log_rho[idx_rho]) / (m.log_rho[idx_rho + 1] - m.log_rho[idx_rho]);
  else
    s.intp_rho = 1.f;
  const float *u1 = m.log_u + (size_t)idx_rho * nT, *u2 = m.log_u + (size_t)(idx_rho + 1) * nT;
  if (u1[idx_u_1 + 1] != u1[idx_u_1])
    s.intp_u_1 = (log_u - u1[idx_u_1]) / (u1[idx_u_1 + 1] - u1[idx_u_1]);
  else
    s.intp_u_1 = 1.f;
  if (u2[idx_u_2 + 1] != u2[idx_u_2])
    s.intp_u_2 = (log_u - u2[idx_u_2]) / (u2[idx_u_2 + 1] - u2[idx_u_2]);
  else
    s.intp_u_2 = 1.f;
  s.idx_rho = idx_rho;
  s.idx_u_1 = idx_u_1;
  s.idx_u_2 = idx_u_2;
  return s;
}

// SESAME_pressure_from_internal_energy, sesame.h:480-594
SWIFT_ET_HD float et_sesame_pressure(float density, float u, const EosTable &m) {
  if (u <= 0.f) return 0.f;
  EtSesameIdx s = et_sesame_index(density, u, m);
  const int nT = m.num_T;
  float P_1 = m.P[(size_t)s.idx_rho * nT + s.idx_u_1];
  float P_2 = m.P[(size_t)s.idx_rho * nT + s.idx_u_1 + 1];
  float P_3 = m.P[(size_t)(s.idx_rho + 1) * nT + s.idx_u_2];
  float P_4 = m.P[(size_t)(s.idx_rho + 1) * nT + s.idx_u_2 + 1];
  // below the minimum u at this rho: the lowest table values
  if ((s.idx_rho > 0.f) && ((s.intp_u_1 < 0.f) || (s.intp_u_2 < 0.f) || (P_1 > P_2) || (P_3 > P_4))) {
    s.intp_u_1 = 0;
    s.intp_u_2 = 0;
  }
  int num_non_pos = 0;
  if (P_1 <= 0.f) num_non_pos++;
  if (P_2 <= 0.f) num_non_pos++;
  if (P_3 <= 0.f) num_non_pos++;
  if (P_4 <= 0.f) num_non_pos++;
  if (num_non_pos > 0) {
    if ((num_non_pos > 2) || (m.P_tiny == 0.f) || (s.intp_rho < 0.f) || (s.intp_u_1 < 0.f) || (s.intp_u_2 < 0.f))
      return 0.f;
    if (P_1 <= 0.f) P_1 = m.P_tiny;
    if (P_2 <= 0.f) P_2 = m.P_tiny;
    if (P_3 <= 0.f) P_3 = m.P_tiny;
    if (P_4 <= 0.f) P_4 = m.P_tiny;
  }
  P_1 = et_logf(P_1);
  P_2 = et_logf(P_2);
  P_3 = et_logf(P_3);
  P_4 = et_logf(P_4);
  const float P = (1.f - s.intp_rho) * ((1.f - s.intp_u_1) * P_1 + s.intp_u_1 * P_2) +
                  s.intp_rho * ((1.f - s.intp_u_2) * P_3 + s.intp_u_2 * P_4);
  return et_expf(P);
}

// SESAME_soundspeed_from_internal_energy, sesame.h:605-721
SWIFT_ET_HD float et_sesame_soundspeed(float density, float u, const EosTable &m) {
  if (u <= 0.f) return 0.f;
  EtSesameIdx s = et_sesame_index(density, u, m);
  const int nT = m.num_T;
  float c_1 = m.c[(size_t)s.idx_rho * nT + s.idx_u_1];
  float c_2 = m.c[(size_t)s.idx_rho * nT + s.idx_u_1 + 1];
  float c_3 = m.c[(size_t)(s.idx_rho + 1) * nT + s.idx_u_2];
  float c_4 = m.c[(size_t)(s.idx_rho + 1) * nT + s.idx_u_2 + 1];
  int num_non_pos = 0;
  if (c_1 <= 0.f) num_non_pos++;
  if (c_2 <= 0.f) num_non_pos++;
  if (c_3 <= 0.f) num_non_pos++;
  if (c_4 <= 0.f) num_non_pos++;
  if (num_non_pos > 2) {
    return m.c_tiny;
  } else if (num_non_pos > 0) {
    if ((s.intp_rho < 0.f) || (s.intp_u_1 < 0.f) || (s.intp_u_2 < 0.f)) return m.c_tiny;
    if (c_1 <= 0.f) c_1 = m.c_tiny;
    if (c_2 <= 0.f) c_2 = m.c_tiny;
    if (c_3 <= 0.f) c_3 = m.c_tiny;
    if (c_4 <= 0.f) c_4 = m.c_tiny;
  }
  c_1 = et_logf(c_1);
  c_2 = et_logf(c_2);
  c_3 = et_logf(c_3);
  c_4 = et_logf(c_4);
  // below the minimum u at this rho: the lowest table values (here AFTER the logs, as the reference has it)
  if ((s.idx_rho > 0.f) && ((s.intp_u_1 < 0.f) || (s.intp_u_2 < 0.f) || (c_1 > c_2) || (c_3 > c_4))) {
    s.intp_u_1 = 0;
    s.intp_u_2 = 0;
  }
  const float c = (1.f - s.intp_rho) * ((1.f - s.intp_u_1) * c_1 + s.intp_u_1 * c_2) +
                  s.intp_rho * ((1.f - s.intp_u_2) * c_3 + s.intp_u_2 * c_4);
  return et_expf(c);
}

// HM80_pressure_from_internal_energy, hm80.h:224-268: a regular grid in (log rho, log u), the table holds log P
SWIFT_ET_HD float et_hm80_pressure(float density, float u, const EosTable &m) {
  if (u <= 0.f) return 0.f;
  const float log_rho = et_logf(density);
  const float log_u = et_logf(u);
  int idx_rho = (int)floor((double)((log_rho - m.log_rho_min) * m.inv_log_rho_step));
  int idx_u = (int)floor((double)((log_u - m.log_u_min) * m.inv_log_u_step));
  if (idx_rho <= -1) idx_rho = 0;
  else if (idx_rho >= m.num_rho - 1) idx_rho = m.num_rho - 2;
  if (idx_u <= -1) idx_u = 0;
  else if (idx_u >= m.num_T - 1) idx_u = m.num_T - 2;
  const float intp_rho = (log_rho - m.log_rho_min - idx_rho * m.log_rho_step) * m.inv_log_rho_step;
  const float intp_u = (log_u - m.log_u_min - idx_u * m.log_u_step) * m.inv_log_u_step;
  const int nu = m.num_T;
  const float log_P_1 = m.P[(size_t)idx_rho * nu + idx_u];
  const float log_P_2 = m.P[(size_t)idx_rho * nu + idx_u + 1];
  const float log_P_3 = m.P[(size_t)(idx_rho + 1) * nu + idx_u];
  const float log_P_4 = m.P[(size_t)(idx_rho + 1) * nu + idx_u + 1];
  const float log_P = (1.f - intp_rho) * ((1.f - intp_u) * log_P_1 + intp_u * log_P_2) +
                      intp_rho * ((1.f - intp_u) * log_P_3 + intp_u * log_P_4);
  return et_expf(log_P);
}

// HM80_soundspeed_from_internal_energy, hm80.h:281-300 (hydro_gamma = 5/3, src/adiabatic_index.h)
SWIFT_ET_HD float et_hm80_soundspeed(float density, float u, const EosTable &m) {
  const float hydro_gamma = 1.66666666666666667f;
  float c;
  if (m.bulk_mod != 0) {
    c = sqrtf(m.bulk_mod / density);
  } else {
    const float P = et_hm80_pressure(density, u, m);
    c = sqrtf(hydro_gamma * P / density);
    if (c <= 0) c = sqrtf(hydro_gamma * m.P_min_for_c_min / density);
  }
  return c;
}

// gas_pressure_from_internal_energy / gas_soundspeed_from_internal_energy of the planetary dispatcher for a
// material that lives in a table (src/equation_of_state/planetary/equation_of_state.h:910-1146, 1305-1457)
SWIFT_ET_HD void et_eval(const EosTable &m, float density, float u, float &P, float &cs) {
  if (m.kind == kEosTableHM80) {
    P = et_hm80_pressure(density, u, m);
    cs = et_hm80_soundspeed(density, u, m);
  } else {
    P = et_sesame_pressure(density, u, m);
    cs = et_sesame_soundspeed(density, u, m);
  }
}

}  // namespace swiftcuda
