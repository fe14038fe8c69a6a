// Flat binary particle files (include/swift_cuda.h, "Flat binary particle files"): the field
// set of the reference's HDF5 initial conditions and snapshots for gas (src/hydro/Planetary/hydro_io.h:55-84,
// 190-228) without HDF5.  Host only.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/swift_cuda.h"

namespace swiftcuda {
int set_error(const std::string &m);
}
using swiftcuda::set_error;

namespace {

struct Header {
  char magic[8];
  int64_t n;
  double dim[3];
  double time;
  int32_t fields;
  int32_t version;
  char pad[128 - 8 - 8 - 24 - 8 - 4 - 4];
};
static_assert(sizeof(Header) == 128, "header layout");
const char kMagic[8] = {'S', 'W', 'C', 'U', 'P', 'R', 'T', '1'};

struct Field {
  int bit;
  size_t elem;   // bytes per particle
  size_t offset; // of the pointer inside swift_cuda_particle_file
  const char *name;
};
#define FIELD(bit, elem, member, name) {bit, elem, offsetof(swift_cuda_particle_file, member), name}
const Field kFields[] = {
    FIELD(SWIFT_CUDA_FIELD_COORDINATES, 3 * sizeof(double), coordinates, "Coordinates"),
    FIELD(SWIFT_CUDA_FIELD_VELOCITIES, 3 * sizeof(float), velocities, "Velocities"),
    FIELD(SWIFT_CUDA_FIELD_MASSES, sizeof(float), masses, "Masses"),
    FIELD(SWIFT_CUDA_FIELD_SMOOTHING_LENGTHS, sizeof(float), smoothing_lengths, "SmoothingLengths"),
    FIELD(SWIFT_CUDA_FIELD_INTERNAL_ENERGIES, sizeof(float), internal_energies, "InternalEnergies"),
    FIELD(SWIFT_CUDA_FIELD_PARTICLE_IDS, sizeof(long long), particle_ids, "ParticleIDs"),
    FIELD(SWIFT_CUDA_FIELD_MATERIAL_IDS, sizeof(int), material_ids, "MaterialIDs"),
    FIELD(SWIFT_CUDA_FIELD_DENSITIES, sizeof(float), densities, "Densities"),
    FIELD(SWIFT_CUDA_FIELD_ACCELERATIONS, 3 * sizeof(float), accelerations, "Accelerations"),
    FIELD(SWIFT_CUDA_FIELD_PRESSURES, sizeof(float), pressures, "Pressures"),
};
const int kAllFields = 1023;

void *&member(swift_cuda_particle_file *p, const Field &f) { return *(void **)((char *)p + f.offset); }
const void *member(const swift_cuda_particle_file *p, const Field &f) {
  return *(void *const *)((const char *)p + f.offset);
}
size_t padded(size_t bytes) { return (bytes + 15) & ~(size_t)15; }

struct FileCloser {
  FILE *f;
  ~FileCloser() {
    if (f) fclose(f);
  }
};

}  // namespace

extern "C" {

void swift_cuda_particles_free(swift_cuda_particle_file *p) {
  if (!p) return;
  for (const Field &f : kFields) free(member(p, f));
  memset(p, 0, sizeof(*p));
}

int swift_cuda_particles_write_bin(const char *path, const swift_cuda_particle_file *p) {
  if (!path || !p) return set_error("particles_write_bin: null argument");
  if (p->n < 0) return set_error("particles_write_bin: negative particle count");
  if ((p->fields & SWIFT_CUDA_FIELDS_COMPULSORY) != SWIFT_CUDA_FIELDS_COMPULSORY)
    return set_error("particles_write_bin: Coordinates, Velocities, Masses, SmoothingLengths, InternalEnergies and "
                     "ParticleIDs are compulsory (hydro_io.h:61-72)");
  if (p->fields & ~kAllFields) return set_error("particles_write_bin: unknown field bit");
  for (const Field &f : kFields)
    if ((p->fields & f.bit) && p->n > 0 && member(p, f) == nullptr)
      return set_error(std::string("particles_write_bin: field ") + f.name + " is marked present but its array is NULL");
  FILE *fp = fopen(path, "wb");
  if (!fp) return set_error(std::string("particles_write_bin: cannot create '") + path + "'");
  FileCloser closer{fp};
  Header h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, kMagic, 8);
  h.n = p->n;
  for (int k = 0; k < 3; k++) h.dim[k] = p->dim[k];
  h.time = p->time;
  h.fields = p->fields;
  h.version = 1;
  bool ok = fwrite(&h, sizeof(h), 1, fp) == 1;
  const char zeros[16] = {0};
  for (const Field &f : kFields) {
    if (!(p->fields & f.bit)) continue;
    const size_t bytes = (size_t)p->n * f.elem;
    if (bytes) ok = ok && fwrite(member(p, f), 1, bytes, fp) == bytes;
    const size_t fill = padded(bytes) - bytes;
    if (fill) ok = ok && fwrite(zeros, 1, fill, fp) == fill;
  }
  if (!ok) return set_error(std::string("particles_write_bin: short write to '") + path + "'");
  return 0;
}

int swift_cuda_particles_read_bin(const char *path, swift_cuda_particle_file *out) {
  if (!path || !out) return set_error("particles_read_bin: null argument");
  memset(out, 0, sizeof(*out));
  FILE *fp = fopen(path, "rb");
  if (!fp) return set_error(std::string("particles_read_bin: cannot open '") + path + "'");
  FileCloser closer{fp};
  Header h;
  if (fread(&h, sizeof(h), 1, fp) != 1 || memcmp(h.magic, kMagic, 8) != 0)
    return set_error(std::string("particles_read_bin: '") + path + "' is not a particle file of this library");
  if (h.version != 1 || h.n < 0 || (h.fields & ~kAllFields) ||
      (h.fields & SWIFT_CUDA_FIELDS_COMPULSORY) != SWIFT_CUDA_FIELDS_COMPULSORY)
    return set_error(std::string("particles_read_bin: bad header in '") + path + "'");
  // the size the header promises, before anything is allocated
  size_t want = sizeof(Header);
  for (const Field &f : kFields)
    if (h.fields & f.bit) want += padded((size_t)h.n * f.elem);
  if (fseek(fp, 0, SEEK_END) != 0) return set_error("particles_read_bin: seek failed");
  const long have = ftell(fp);
  if (have < 0 || (size_t)have != want)
    return set_error(std::string("particles_read_bin: '") + path + "' has " + std::to_string(have) + " bytes, its header " +
                     "promises " + std::to_string(want) + " (truncated or trailing bytes)");
  if (fseek(fp, (long)sizeof(Header), SEEK_SET) != 0) return set_error("particles_read_bin: seek failed");
  out->n = h.n;
  for (int k = 0; k < 3; k++) out->dim[k] = h.dim[k];
  out->time = h.time;
  out->fields = h.fields;
  char skip[16];
  for (const Field &f : kFields) {
    if (!(h.fields & f.bit)) continue;
    const size_t bytes = (size_t)h.n * f.elem;
    void *a = malloc(bytes ? bytes : 1);
    member(out, f) = a;
    bool ok = a != nullptr && (bytes == 0 || fread(a, 1, bytes, fp) == bytes);
    const size_t fill = padded(bytes) - bytes;
    if (ok && fill) ok = fread(skip, 1, fill, fp) == fill;
    if (!ok) {
      swift_cuda_particles_free(out);
      return set_error(std::string("particles_read_bin: reading field ") + f.name + " of '" + path + "' failed");
    }
  }
  return 0;
}

}  // extern "C"
