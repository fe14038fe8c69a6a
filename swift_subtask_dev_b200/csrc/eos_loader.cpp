// Host side of the table equations of state (include/swift_cuda.h, "Table equations of state"; SURVEY.md 8(f)
// N4): reads a table file in the reference's text format and turns it into what the reference holds after
//   load_table_SESAME + prepare_table_SESAME + convert_units_SESAME   src/equation_of_state/planetary/sesame.h:149-352
//   load_table_HM80 + prepare_table_HM80 + convert_units_HM80         planetary/hm80.h:76-178
// (same parsing -- fscanf("%f") --, same float operations in the same order, libm's logf as the reference
// calls it), plus a flat binary form of the result.  Pinned array for array against the reference's own loaders
// on the same files: tests/test_eos_tables.py.  No device code here; swift_cuda_set_eos_table is in swift_cuda.cu.
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/swift_cuda.h"

namespace swiftcuda {
int set_error(const std::string &m);  // swift_cuda.cu: what swift_cuda_last_error returns
}
using swiftcuda::set_error;

namespace {

// skip_line / skip_lines, sesame.h:109-131
int skip_lines(FILE *f, int n) {
  int c = 0;
  for (int i = 0; i < n; i++) {
    do {
      c = fgetc(f);
    } while (c != '\n' && c != EOF);
  }
  return c;
}

struct FileCloser {
  FILE *f;
  ~FileCloser() {
    if (f) fclose(f);
  }
};

float *alloc_f(size_t n) { return (float *)malloc((n ? n : 1) * sizeof(float)); }

}  // namespace

extern "C" {

void swift_cuda_eos_table_free(swift_cuda_eos_table *t) {
  if (!t) return;
  free(t->log_rho);
  free(t->log_u);
  free(t->P);
  free(t->c);
  memset(t, 0, sizeof(*t));
}

int swift_cuda_eos_table_load_sesame(const char *path, int mat_id, int version_date, const swift_cuda_eos_units *units,
                                     swift_cuda_eos_table *out) {
  if (!path || !units || !out) return set_error("eos_table_load_sesame: null argument");
  memset(out, 0, sizeof(*out));
  FILE *f = fopen(path, "r");
  if (!f) return set_error(std::string("Failed to open the SESAME EoS file '") + path + "'");
  FileCloser closer{f};
  const std::string bad = std::string("Failed to read the SESAME EoS table ") + path;
  // ---- load_table_SESAME, sesame.h:149-226
  skip_lines(f, 12);
  int date = 0, num_rho = 0, num_T = 0;
  if (fscanf(f, "%d", &date) != 1) return set_error(bad);
  if (date != version_date && version_date != 0)
    return set_error(std::string("EoS file ") + path + " version_date " + std::to_string(date) +
                     " does not match expected " + std::to_string(version_date) + " (YYYYMMDD)");
  if (fscanf(f, "%d %d", &num_rho, &num_T) != 2) return set_error(bad);
  // the first elements, rho = 0 and T = 0, are ignored
  num_rho--;
  num_T--;
  if (num_rho < 2 || num_T < 2) return set_error(bad + ": fewer than two densities or temperatures");
  out->kind = 1;
  out->mat_id = mat_id;
  out->num_rho = num_rho;
  out->num_T = num_T;
  const size_t n2 = (size_t)num_rho * num_T;
  out->log_rho = alloc_f(num_rho);
  out->log_u = alloc_f(n2);
  out->P = alloc_f(n2);
  out->c = alloc_f(n2);
  float *s = alloc_f(n2);  // sp. entropies: parsed as the reference does, not kept (no hot-path function reads them)
  struct Guard {
    swift_cuda_eos_table *t;
    float *s;
    bool ok;
    ~Guard() {
      free(s);
      if (!ok) swift_cuda_eos_table_free(t);
    }
  } guard{out, s, false};
  if (!out->log_rho || !out->log_u || !out->P || !out->c || !s) return set_error("eos_table_load_sesame: out of memory");
  float ignore;
  for (int i_rho = -1; i_rho < num_rho; i_rho++)
    if (fscanf(f, "%f", i_rho == -1 ? &ignore : &out->log_rho[i_rho]) != 1) return set_error(bad);
  for (int i_T = -1; i_T < num_T; i_T++)
    if (fscanf(f, "%f", &ignore) != 1) return set_error(bad);
  for (int i_T = -1; i_T < num_T; i_T++) {
    for (int i_rho = -1; i_rho < num_rho; i_rho++) {
      if (i_T == -1 || i_rho == -1) {
        if (fscanf(f, "%f %f %f %f", &ignore, &ignore, &ignore, &ignore) != 4) return set_error(bad);
      } else {
        const size_t k = (size_t)i_rho * num_T + i_T;
        if (fscanf(f, "%f %f %f %f", &out->log_u[k], &out->P[k], &out->c[k], &s[k]) != 4) return set_error(bad);
      }
    }
  }
  // ---- prepare_table_SESAME, sesame.h:229-302
  for (int i = 0; i < num_rho; i++) out->log_rho[i] = logf(out->log_rho[i]);
  float u_tiny = FLT_MAX, P_tiny = FLT_MAX, c_tiny = FLT_MAX;
  for (int i_rho = 0; i_rho < num_rho; i_rho++) {
    float *u = out->log_u + (size_t)i_rho * num_T;
    const float *P = out->P + (size_t)i_rho * num_T, *c = out->c + (size_t)i_rho * num_T;
    for (int i_T = num_T - 1; i_T > 0; i_T--) {
      // the slices of u are made monotonic: a lower-T value above this one is replaced by this one
      if (u[i_T] < u[i_T - 1]) u[i_T - 1] = u[i_T];
      // smallest positive values (the loop stops at i_T = 1, as the reference's does)
      if (u[i_T] < u_tiny && u[i_T] > 0) u_tiny = u[i_T];
      if (P[i_T] < P_tiny && P[i_T] > 0) P_tiny = P[i_T];
      if (c[i_T] < c_tiny && c[i_T] > 0) c_tiny = c[i_T];
    }
  }
  u_tiny *= 1e-3f;
  P_tiny *= 1e-3f;
  c_tiny *= 1e-3f;
  for (size_t k = 0; k < n2; k++) {
    if (out->log_u[k] <= 0) out->log_u[k] = u_tiny;
    out->log_u[k] = logf(out->log_u[k]);
  }
  // ---- convert_units_SESAME, sesame.h:305-352
  for (int i = 0; i < num_rho; i++) out->log_rho[i] += logf(units->density);
  for (size_t k = 0; k < n2; k++) {
    out->log_u[k] += logf(units->energy_per_unit_mass);
    out->P[k] *= units->pressure;
    out->c[k] *= units->speed;
  }
  P_tiny *= units->pressure;
  c_tiny *= units->speed;
  out->P_tiny = P_tiny;
  out->c_tiny = c_tiny;
  guard.ok = true;
  return 0;
}

int swift_cuda_eos_table_load_hm80(const char *path, int mat_id, const swift_cuda_eos_units *units,
                                   swift_cuda_eos_table *out) {
  if (!path || !units || !out) return set_error("eos_table_load_hm80: null argument");
  memset(out, 0, sizeof(*out));
  // set_HM80_HHe / _ice / _rock, hm80.h:53-73 (SI)
  float bulk_mod, P_min_for_c_min;
  switch (mat_id) {
    case 200: bulk_mod = 0.f; P_min_for_c_min = 1e3f; break;
    case 201: bulk_mod = 2.0e9f; P_min_for_c_min = 0.f; break;
    case 202: bulk_mod = 3.49e10f; P_min_for_c_min = 0.f; break;
    default: return set_error("eos_table_load_hm80: mat_id must be 200 (HHe), 201 (ice) or 202 (rock)");
  }
  const int expected_date = 20201003;
  FILE *f = fopen(path, "r");
  if (!f) return set_error(std::string("Failed to open the HM80 EoS file '") + path + "'");
  FileCloser closer{f};
  const std::string bad = std::string("Failed to read the HM80 EoS table ") + path;
  // ---- load_table_HM80, hm80.h:76-135
  char buffer[100];
  for (int i = 0; i < 5; i++)
    if (fgets(buffer, 100, f) == NULL) return set_error(bad + " (header)");
  int date = 0, num_rho = 0, num_u = 0;
  float log_rho_min, log_rho_max, log_u_min, log_u_max;
  if (fscanf(f, "%d", &date) != 1) return set_error(bad);
  if (date != expected_date)
    return set_error(std::string("EoS file ") + path + " date " + std::to_string(date) + " does not match expected " +
                     std::to_string(expected_date));
  if (fscanf(f, "%f %f %d %f %f %d", &log_rho_min, &log_rho_max, &num_rho, &log_u_min, &log_u_max, &num_u) != 6)
    return set_error(bad);
  if (num_rho < 2 || num_u < 2) return set_error(bad + ": fewer than two densities or energies");
  out->kind = 2;
  out->mat_id = mat_id;
  out->num_rho = num_rho;
  out->num_T = num_u;
  out->log_rho_step = (log_rho_max - log_rho_min) / (num_rho - 1);
  out->log_u_step = (log_u_max - log_u_min) / (num_u - 1);
  out->inv_log_rho_step = 1.f / out->log_rho_step;
  out->inv_log_u_step = 1.f / out->log_u_step;
  const size_t n2 = (size_t)num_rho * num_u;
  out->P = alloc_f(n2);
  if (!out->P) return set_error("eos_table_load_hm80: out of memory");
  for (size_t k = 0; k < n2; k++)
    if (fscanf(f, "%f", &out->P[k]) != 1) {
      swift_cuda_eos_table_free(out);
      return set_error(bad);
    }
  // ---- prepare_table_HM80, hm80.h:138-147
  for (size_t k = 0; k < n2; k++) out->P[k] = logf(out->P[k]);
  // ---- convert_units_HM80, hm80.h:150-178 (log_*_max are not used after the steps are formed)
  log_rho_min += logf(units->density);
  log_u_min += logf(units->energy_per_unit_mass);
  for (size_t k = 0; k < n2; k++) out->P[k] += logf(units->pressure);
  bulk_mod *= units->pressure;
  P_min_for_c_min *= units->pressure;
  out->log_rho_min = log_rho_min;
  out->log_u_min = log_u_min;
  out->bulk_mod = bulk_mod;
  out->P_min_for_c_min = P_min_for_c_min;
  return 0;
}

// ---- flat binary form ----------------------------------------------------------------------------------
// 64-byte header: magic "SWCUEOS1", kind, mat_id, num_rho, num_T, then the ten float parameters in the order of
// swift_cuda_eos_table; then log_rho [num_rho], log_u, P, c [num_rho * num_T each] for kind 1, P alone for kind 2.
namespace {
struct BinHeader {
  char magic[8];
  int32_t kind, mat_id, num_rho, num_T;
  float par[10];
};
static_assert(sizeof(BinHeader) == 64, "header layout");
const char kMagic[8] = {'S', 'W', 'C', 'U', 'E', 'O', 'S', '1'};
}  // namespace

int swift_cuda_eos_table_write_bin(const char *path, const swift_cuda_eos_table *t) {
  if (!path || !t) return set_error("eos_table_write_bin: null argument");
  if (t->kind != 1 && t->kind != 2) return set_error("eos_table_write_bin: not a table");
  FILE *f = fopen(path, "wb");
  if (!f) return set_error(std::string("eos_table_write_bin: cannot create '") + path + "'");
  FileCloser closer{f};
  BinHeader h;
  memcpy(h.magic, kMagic, 8);
  h.kind = t->kind;
  h.mat_id = t->mat_id;
  h.num_rho = t->num_rho;
  h.num_T = t->num_T;
  const float par[10] = {t->P_tiny,        t->c_tiny,     t->log_rho_min, t->log_rho_step,   t->inv_log_rho_step,
                         t->log_u_min,     t->log_u_step, t->inv_log_u_step, t->bulk_mod,    t->P_min_for_c_min};
  memcpy(h.par, par, sizeof(par));
  const size_t n2 = (size_t)t->num_rho * t->num_T;
  bool ok = fwrite(&h, sizeof(h), 1, f) == 1;
  if (t->kind == 1) {
    ok = ok && fwrite(t->log_rho, sizeof(float), t->num_rho, f) == (size_t)t->num_rho;
    ok = ok && fwrite(t->log_u, sizeof(float), n2, f) == n2;
    ok = ok && fwrite(t->P, sizeof(float), n2, f) == n2;
    ok = ok && fwrite(t->c, sizeof(float), n2, f) == n2;
  } else {
    ok = ok && fwrite(t->P, sizeof(float), n2, f) == n2;
  }
  if (!ok) return set_error(std::string("eos_table_write_bin: short write to '") + path + "'");
  return 0;
}

int swift_cuda_eos_table_read_bin(const char *path, swift_cuda_eos_table *out) {
  if (!path || !out) return set_error("eos_table_read_bin: null argument");
  memset(out, 0, sizeof(*out));
  FILE *f = fopen(path, "rb");
  if (!f) return set_error(std::string("eos_table_read_bin: cannot open '") + path + "'");
  FileCloser closer{f};
  BinHeader h;
  if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, kMagic, 8) != 0)
    return set_error(std::string("eos_table_read_bin: '") + path + "' is not a table file of this library");
  if ((h.kind != 1 && h.kind != 2) || h.num_rho < 2 || h.num_T < 2 || h.num_rho > (1 << 20) || h.num_T > (1 << 20))
    return set_error(std::string("eos_table_read_bin: bad header in '") + path + "'");
  out->kind = h.kind;
  out->mat_id = h.mat_id;
  out->num_rho = h.num_rho;
  out->num_T = h.num_T;
  out->P_tiny = h.par[0];
  out->c_tiny = h.par[1];
  out->log_rho_min = h.par[2];
  out->log_rho_step = h.par[3];
  out->inv_log_rho_step = h.par[4];
  out->log_u_min = h.par[5];
  out->log_u_step = h.par[6];
  out->inv_log_u_step = h.par[7];
  out->bulk_mod = h.par[8];
  out->P_min_for_c_min = h.par[9];
  const size_t n2 = (size_t)h.num_rho * h.num_T;
  bool ok = true;
  if (h.kind == 1) {
    out->log_rho = alloc_f(h.num_rho);
    out->log_u = alloc_f(n2);
    out->P = alloc_f(n2);
    out->c = alloc_f(n2);
    ok = out->log_rho && out->log_u && out->P && out->c;
    ok = ok && fread(out->log_rho, sizeof(float), h.num_rho, f) == (size_t)h.num_rho;
    ok = ok && fread(out->log_u, sizeof(float), n2, f) == n2;
    ok = ok && fread(out->P, sizeof(float), n2, f) == n2;
    ok = ok && fread(out->c, sizeof(float), n2, f) == n2;
  } else {
    out->P = alloc_f(n2);
    ok = out->P && fread(out->P, sizeof(float), n2, f) == n2;
  }
  if (ok && fgetc(f) != EOF) ok = false;  // trailing bytes: not what write_bin wrote
  if (!ok) {
    swift_cuda_eos_table_free(out);
    return set_error(std::string("eos_table_read_bin: '") + path + "' is truncated or has trailing bytes");
  }
  return 0;
}

}  // extern "C"
