// C ABI of the backend (include/swift_cuda.h): context, device buffers, the
// launch sequence of one density -> ghost -> force step.  No CPU fallback: every
// compute entry needs a CUDA device and fails loudly without one.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "leaf_kernels.cuh"
#include "list_kernels.cuh"
#include "force_fast.cuh"
#include "pack.cuh"
#include "integrate.cuh"
#include "sphenix.cuh"

using namespace swiftcuda;

static thread_local std::string g_err;
static int fail(const std::string &m) {
  g_err = m;
  return 1;
}
namespace swiftcuda {
int set_error(const std::string &m) { return fail(m); }  // for the host-only sources (eos_loader.cpp)
}
#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return fail(std::string(#call) + ": " + cudaGetErrorString(e_) + " (" + __FILE__ + ":" + \
                  std::to_string(__LINE__) + ")");                                            \
  } while (0)

struct swift_cuda_ctx {
  swift_cuda_params prm;
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t io = nullptr;      // copy stream of swift_cuda_step_io
  cudaEvent_t io_ev[4] = {};
  cudaEvent_t ev[6] = {};
  Tree tree;
  WorkLists wl;
  Dev dev;
  int64_t n = 0;
  int64_t n_io = -1;   // swift_cuda_set_transfer_count: particles moved by the host<->device copies (-1: all)
  bool built = false;
  std::vector<void *> allocs;
  // device-only helpers
  int32_t *d_seg_cell = nullptr, *d_seg_sid = nullptr;
  int64_t *d_seg_off = nullptr, *d_scratch_off = nullptr;
  unsigned long long *d_scratch = nullptr;
  int n_seg = 0;
  swift_cuda_counters counters = {};
  swift_cuda_timings timings = {};
  long long rec_capacity = 0;
  float4 *d_dens0_first = nullptr, *d_dens1_first = nullptr;  // snapshot of the first-pass sums
  float *d_h0 = nullptr;     // h as last uploaded (swift_cuda_restore_particles)
  int grid_density = 0, grid_redo = 0, grid_force = 0;  // persistent grids: one resident wave of CTAs
  // caller-order staging on the device (K1 pack / K7 unpack)
  double *s_x = nullptr;
  float *s_v = nullptr, *s_mass = nullptr, *s_h = nullptr, *s_u = nullptr, *s_carry[5] = {};
  int8_t *s_tb = nullptr;
  int32_t *s_mat = nullptr, *s_order = nullptr;
  int64_t *s_id = nullptr;
  float *s_out = nullptr;    // 20 floats per particle
  int8_t *s_out_bin = nullptr;
  // time integration (swift_cuda_integrator_init)
  bool all_cells_active = false;   // every cell is in the work lists
  bool integ_on = false;
  IntegDev integ = {};
  IntegPrm iprm = {};
  uint8_t *d_sort_need = nullptr;  // [ncells*13]
  int *d_unit_list = nullptr, *d_unit_count = nullptr;
  bool marks_fresh = false;        // k_mark_sorts has run for the current active set
  float integ_ms[4] = {};          // last drift, kick1, kick2, timestep (CUDA events on the stream)
  // which neighbour-loop kernels run: bit 0 leaf-resident first density pass, bit 1 fused ghost + redo,
  // bit 2 leaf-resident force (FAST field arithmetic), bit 3 ghost redo sweeps from recorded neighbour
  // lists (list_kernels.cuh).  0: the first-generation warp-per-unit kernels.
  // swift_cuda_step_parts: the caller's AoS records
  swift_cuda_part_layout lay = {};
  bool has_lay = false;
  // swift_cuda_step_parts hands the force loop's results back in `pieces` pieces (units of consecutive leaves
  // = consecutive records) when the caller's records are in cell order, as s->parts is after space_rebuild
  bool records_in_cell_order = false;
  int pieces = 12;
  std::vector<cudaEvent_t> piece_ev;
  std::vector<int64_t> seg_end;    // last slot + 1 of every sort segment's cell (segments are kept in this order)
  struct PiecePlan {               // built on the first step that runs in pieces
    int P = 0;
    std::vector<int> leaf;         // [P+1] leaf slots
    std::vector<int64_t> rec;      // [P+1] cell-order slots (= records)
    std::vector<int64_t> dep_hi;   // [P] every cell a piece's interactions read ends at or before this slot
  } plan;
  unsigned char *d_aos = nullptr;
  bool no_prune = false;            // build the work lists with every cell, whatever the time bins
  std::vector<int32_t> slot_of;     // host copy of Dev::part_slot (caller order)
  float dx_max_part = 0.f;          // largest |x_diff| since the build (swift_cuda_drift), for k_check_lists
  int path = 8;                     // default: ghost redo sweeps from recorded neighbour lists (8).  Bit 5 (32): the
                                    // density walks as distance filters with the decision made at consumption --
                                    // bit-exact, measured 8 % SLOWER than the windowed test (spills), left off
  // SPHENIX (sphenix.cuh): the scheme's hydro_props, staging of its per-particle state, the gradient loop's grid
  swift_cuda_sphenix_props sprops = {2.0f, 0.0f, 0.05f, 0.0f, 1.0f, 1.0f, 0.0f, 1e-3};
  float *s_sph_in[3] = {};
  float *s_sph_out = nullptr;
  int grid_gradient = 0;
  int grid_limiter = 0;        // the time-step limiter loop, k_density<0, false, 2>
  float limiter_ms[2] = {0.f, 0.f};  // device time of the last limiter loop / k_limiter
  int8_t *d_lim_tmp = nullptr; // n bytes: time bins in / wakeup out in caller order
  // table equations of state (swift_cuda_set_eos_table): device arrays that outlive rebuilds
  EosTable tabs[8] = {};
  int n_tab = 0;
  std::vector<void *> eos_allocs;
  long long n_gradient = 0;
  int sph_prep_mode = 0;   // k_sphenix_prep: swift_cuda_step_parts calls that run without the ghost / the extra ghost
  int fast_qd = kQD;   // rows of the hit FIFO k_force_fast uses (SWIFT_CUDA_FAST_QD)
  int force_slow = 0;  // test hook (SWIFT_CUDA_SLOW=1): every hit takes the exact sorted-axis-cut path
};

// leaf-resident kernel configurations (shared-memory tile capacity, queue depth)
constexpr int kDensCap = 960, kDensQD = 80;
typedef LeafTile<kDensCap, kDensQD, 1> DensTileT;
constexpr int kGhostCap = 1152, kGhostQD = 112;
typedef LeafTile<kGhostCap, kGhostQD, 2> GhostTileT;

template <typename T>
static int dmalloc(swift_cuda_ctx *c, T **p, size_t count) {
  void *q = nullptr;
  CK(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
  c->allocs.push_back(q);
  *p = reinterpret_cast<T *>(q);
  return 0;
}
template <typename T>
static int upload(swift_cuda_ctx *c, T **p, const T *src, size_t count) {
  if (dmalloc(c, p, count)) return 1;
  if (count) CK(cudaMemcpyAsync(*p, src, count * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  return 0;
}

static void free_device(swift_cuda_ctx *c) {
  for (void *p : c->allocs) cudaFree(p);
  c->allocs.clear();
  c->built = false;
  c->dx_max_part = 0.f;
  // the time-integration state lived in those allocations
  c->integ_on = false;
  c->marks_fresh = false;
  c->integ = IntegDev{};
  c->d_sort_need = nullptr;
  c->d_unit_list = nullptr;
  c->d_unit_count = nullptr;
  c->dev.wakeup = nullptr;
  c->d_lim_tmp = nullptr;
}

extern "C" void swift_cuda_destroy(swift_cuda_ctx *c);

extern "C" {

int swift_cuda_abi_version(void) { return SWIFT_CUDA_ABI_VERSION; }
size_t swift_cuda_sizeof(int which) {
  switch (which) {
    case 0: return sizeof(swift_cuda_params);
    case 1: return sizeof(swift_cuda_cell);
    case 2: return sizeof(swift_cuda_task);
    case 3: return sizeof(swift_cuda_parts_in);
    case 4: return sizeof(swift_cuda_parts_out);
    case 5: return sizeof(swift_cuda_counters);
    case 6: return sizeof(swift_cuda_timings);
    case 7: return sizeof(swift_cuda_til);
    case 8: return sizeof(swift_cuda_integrator);
    case 9: return sizeof(swift_cuda_state_out);
    case 10: return sizeof(swift_cuda_part_layout);
    case 11: return sizeof(swift_cuda_sphenix_props);
    case 12: return sizeof(swift_cuda_sphenix_in);
    case 13: return sizeof(swift_cuda_sphenix_out);
    default: return 0;
  }
}
const char *swift_cuda_last_error(void) { return g_err.c_str(); }

static int create_resources(swift_cuda_ctx *c) {
  CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  for (auto &e : c->ev) CK(cudaEventCreate(&e));
  CK(cudaStreamCreateWithFlags(&c->io, cudaStreamNonBlocking));
  for (auto &e : c->io_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  CK(cudaFuncSetAttribute(k_density_leaf<kDensCap, kDensQD>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          (int)sizeof(DensTileT)));
  CK(cudaFuncSetAttribute(k_ghost_leaf<kGhostCap, kGhostQD>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          (int)sizeof(GhostTileT)));
  return 0;
}

int swift_cuda_create(const swift_cuda_params *params, int device, swift_cuda_ctx **out) {
  if (!params || !out) return fail("swift_cuda_create: null argument");
  // a_factor_Balsara_eps, fac_mu and the a^-2 of the curl are not implemented: refuse rather than return
  // wrong Balsara switches (src/hydro/Minimal/hydro.h:700-720, hydro_iact.h:268-300)
  if (params->cosmo_a != 1.f || params->cosmo_H != 0.f)
    return fail("swift_cuda_create: cosmological factors are not supported (needs cosmo_a = 1, cosmo_H = 0)");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail("swift_cuda_create: no CUDA device (this backend has no CPU fallback)");
  if (device < 0 || device >= ndev) return fail("swift_cuda_create: bad device index");
  CK(cudaSetDevice(device));
  swift_cuda_ctx *c = new swift_cuda_ctx();
  c->prm = *params;
  if (const char *e = getenv("SWIFT_CUDA_FIELDS")) c->prm.fields_mode = atoi(e);
  if (c->prm.fields_mode != SWIFT_CUDA_FIELDS_EXACT && c->prm.fields_mode != SWIFT_CUDA_FIELDS_FAST) {
    delete c;
    return fail("swift_cuda_create: unknown fields_mode");
  }
  if (c->prm.scheme < SWIFT_CUDA_SCHEME_MINIMAL || c->prm.scheme > SWIFT_CUDA_SCHEME_SPHENIX) {
    delete c;
    return fail("swift_cuda_create: unknown scheme");
  }
  c->device = device;
  std::memset(&c->dev, 0, sizeof(Dev));
  if (const char *e = getenv("SWIFT_CUDA_PATH")) c->path = atoi(e);
  // the experimental leaf-resident / filter kernels carry their own copies of the two-loop ghost and bodies
  if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX) c->path &= 8;
  if (const char *e = getenv("SWIFT_CUDA_SLOW")) c->force_slow = atoi(e);
  if (const char *e = getenv("SWIFT_CUDA_FAST_QD")) c->fast_qd = std::min(std::max(atoi(e), 40), kQD);
  if (create_resources(c)) {
    swift_cuda_destroy(c);  // releases whatever was created; the error text is kept
    return 1;
  }
  *out = c;
  return 0;
}

void swift_cuda_destroy(swift_cuda_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  free_device(c);
  for (void *q : c->eos_allocs) cudaFree(q);
  for (auto &e : c->ev)
    if (e) cudaEventDestroy(e);
  if (c->stream) cudaStreamDestroy(c->stream);
  for (auto &e : c->io_ev)
    if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : c->piece_ev) cudaEventDestroy(e);
  if (c->io) cudaStreamDestroy(c->io);
  delete c;
}

}  // extern "C"

static int upload_consts(swift_cuda_ctx *c) {
  const swift_cuda_params &p = c->prm;
  Consts k;
  std::memset(&k, 0, sizeof(k));
  // src/kernel_hydro.h:52-53, 205-241 -- same expressions, evaluated by the host compiler
  const float kg = (float)(1.825742);
  k.gamma = kg;
  k.gamma2 = (float)(kg * kg);
  k.gamma_inv = (float)(1. / kg);
  k.kconst = (float)(16. * 0.31830988618379067154);  // 16 * M_1_PI
  k.ginv_dim = (float)(1. / (kg * kg * kg));
  k.ginv_dim_p1 = (float)(1. / (kg * kg * kg * kg));
  k.kroot = (float)(0.5f) * k.kconst * k.ginv_dim;
  k.eta_dim = p.eta_neighbours * p.eta_neighbours * p.eta_neighbours;
  k.h_tol = p.h_tolerance;
  k.h_max = p.h_max;
  k.h_min = p.h_min;
  k.alpha = p.viscosity_alpha;
  k.beta = p.scheme == SWIFT_CUDA_SCHEME_PLANETARY ? 4.f : 3.f;  // hydro_parameters.h:51 of each scheme
  k.a = p.cosmo_a;
  k.H = p.cosmo_H;
  k.max_active_bin = p.max_active_bin;
  k.scheme = p.scheme;
  k.max_iter = p.max_smoothing_iterations;
  k.n_til = p.n_til;
  for (int i = 0; i < 8; i++) k.til[i] = p.til[i];
  for (int i = 0; i < 4; i++) k.idg_gamma[i] = p.idg_gamma[i];
  for (int s = 0; s < 13; s++)
    for (int d = 0; d < 3; d++) k.axis[s][d] = kAxis[s][d];
  k.visc_alpha_max = c->sprops.viscosity_alpha_max;
  k.visc_alpha_min = c->sprops.viscosity_alpha_min;
  k.visc_length = c->sprops.viscosity_length;
  k.diff_beta = c->sprops.diffusion_beta;
  k.diff_alpha_max = c->sprops.diffusion_alpha_max;
  k.diff_alpha_min = c->sprops.diffusion_alpha_min;
  k.time_base = c->sprops.time_base;
  k.n_tab = c->n_tab;
  for (int i = 0; i < 8; i++) k.tab[i] = c->tabs[i];
  CK(cudaMemcpyToSymbolAsync(kc, &k, sizeof(k), 0, cudaMemcpyHostToDevice, c->stream));
  return 0;
}

// Copies the caller's arrays to the device as they are and permutes them into the cell-ordered
// SoA layout there (k_pack).  with_static: also mat_id / id / the order table (build only).
static int upload_and_pack(swift_cuda_ctx *c, const swift_cuda_parts_in *in, bool with_static) {
  c->marks_fresh = false;  // time bins may change
  const int64_t n = c->n;
  const int64_t m = (with_static || c->n_io < 0) ? n : c->n_io;  // particles on the bus
  Dev &d = c->dev;
  cudaStream_t st = c->stream;
  RawIn r;
  std::memset(&r, 0, sizeof(r));
  CK(cudaMemcpyAsync(c->s_x, in->x, m * 24, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(c->s_v, in->v, m * 12, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(c->s_mass, in->mass, m * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(c->s_h, in->h, m * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(c->s_u, in->u, m * 4, cudaMemcpyHostToDevice, st));
  r.x = c->s_x; r.v = c->s_v; r.mass = c->s_mass; r.h = c->s_h; r.u = c->s_u;
  if (in->time_bin) {
    CK(cudaMemcpyAsync(c->s_tb, in->time_bin, m, cudaMemcpyHostToDevice, st));
    r.time_bin = c->s_tb;
  }
  const float *carry[5] = {in->rho, in->pressure, in->f, in->soundspeed, in->balsara};
  const float **rc[5] = {&r.rho, &r.pressure, &r.f, &r.soundspeed, &r.balsara};
  for (int k = 0; k < 5; k++)
    if (carry[k]) {
      if (!c->s_carry[k] && dmalloc(c, &c->s_carry[k], (size_t)n)) return 1;
      CK(cudaMemcpyAsync(c->s_carry[k], carry[k], m * 4, cudaMemcpyHostToDevice, st));
      *rc[k] = c->s_carry[k];
    }
  if (with_static) {
    if (in->mat_id) {
      CK(cudaMemcpyAsync(c->s_mat, in->mat_id, n * 4, cudaMemcpyHostToDevice, st));
      r.mat_id = c->s_mat;
    }
    if (in->id) {
      CK(cudaMemcpyAsync(c->s_id, in->id, n * 8, cudaMemcpyHostToDevice, st));
      r.id = c->s_id;
    }
    if (!c->tree.order.empty())
      CK(cudaMemcpyAsync(c->s_order, c->tree.order.data(), n * 4, cudaMemcpyHostToDevice, st));
  }
  r.order = c->tree.order.empty() ? nullptr : c->s_order;
  PackDst o;
  o.pos = const_cast<PosD *>(d.pos);
  o.vm = d.vm; o.fin = d.fin; o.hb = d.hb; o.h = d.h; o.h0 = c->d_h0; o.u = const_cast<float *>(d.u);
  o.time_bin = const_cast<int8_t *>(d.time_bin);
  o.mat_id = const_cast<int32_t *>(d.mat_id);
  o.id = const_cast<int64_t *>(d.id);
  if (n > 0) k_pack<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, m, r, o, with_static ? 1 : 0, 3);
  CK(cudaGetLastError());
  return 0;
}

// Allocates the device arrays, packs the particles (device side) and uploads the tree, work
// lists and sort segment tables.
static int pack_and_upload(swift_cuda_ctx *c, const swift_cuda_parts_in *in) {
  const int64_t n = in->n;
  c->n = n;
  c->n_io = -1;
  Dev &d = c->dev;
  d.n = n;
  for (int k = 0; k < 3; k++) d.dim[k] = c->prm.dim[k];
  PosD *dpos;
  float *du; int8_t *dtb; int32_t *dmat; int64_t *did;
  if (dmalloc(c, &dpos, n) || dmalloc(c, &d.vm, n) || dmalloc(c, &d.h, n) ||
      dmalloc(c, &c->d_h0, n) || dmalloc(c, &d.fin, n) || dmalloc(c, &d.hb, n) || dmalloc(c, &du, n) ||
      dmalloc(c, &dtb, n) || dmalloc(c, &dmat, n) || dmalloc(c, &did, n))
    return 1;
  d.pos = dpos; d.u = du; d.time_bin = dtb; d.mat_id = dmat; d.id = did;
  if (dmalloc(c, &c->s_x, (size_t)n * 3) || dmalloc(c, &c->s_v, (size_t)n * 3) || dmalloc(c, &c->s_mass, n) ||
      dmalloc(c, &c->s_h, n) || dmalloc(c, &c->s_u, n) || dmalloc(c, &c->s_tb, n) || dmalloc(c, &c->s_mat, n) ||
      dmalloc(c, &c->s_id, n) || dmalloc(c, &c->s_order, n))
    return 1;
  for (int k = 0; k < 5; k++) c->s_carry[k] = nullptr;
  c->s_out = nullptr;
  c->s_out_bin = nullptr;
  if (upload_and_pack(c, in, true)) return 1;
  if (dmalloc(c, &d.dens0, n) || dmalloc(c, &d.dens1, n) ||
      dmalloc(c, &d.acc, n) || dmalloc(c, &d.h_dt, n) || dmalloc(c, &d.v_sig, n) || dmalloc(c, &d.min_bin, n) ||
      dmalloc(c, &d.left, n) || dmalloc(c, &d.right, n) || dmalloc(c, &d.gstate, n) ||
      dmalloc(c, &c->d_dens0_first, n) || dmalloc(c, &c->d_dens1_first, n))
    return 1;
  CK(cudaMemsetAsync(d.acc, 0, n * sizeof(float4), c->stream));
  CK(cudaMemsetAsync(d.h_dt, 0, n * sizeof(float), c->stream));
  CK(cudaMemsetAsync(d.v_sig, 0, n * sizeof(float), c->stream));
  // SPHENIX: the state the scheme carries, at the values hydro_convert_quantities gives a fresh particle
  std::vector<float4> sph_init;
  d.sph_state = d.sph_grad = d.gin = d.sx = nullptr;
  d.rec_grad = nullptr;
  for (int k = 0; k < 3; k++) c->s_sph_in[k] = nullptr;
  c->s_sph_out = nullptr;
  if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX) {
    sph_init.assign((size_t)n, make_float4(c->prm.viscosity_alpha, c->sprops.diffusion_alpha, 0.f, 0.f));
    if (upload(c, &d.sph_state, sph_init.data(), (size_t)n) || dmalloc(c, &d.sph_grad, n) || dmalloc(c, &d.gin, n) ||
        dmalloc(c, &d.sx, n))
      return 1;
    CK(cudaMemsetAsync(d.sph_grad, 0, n * sizeof(float4), c->stream));
    CK(cudaMemsetAsync(d.gin, 0, n * sizeof(float4), c->stream));
    CK(cudaMemsetAsync(d.sx, 0, n * sizeof(float4), c->stream));
  }

  // tree
  const size_t nc = c->tree.cells.size();
  std::vector<CellDev> cells(nc);
  for (size_t k = 0; k < nc; k++) {
    const swift_cuda_cell &sc = c->tree.cells[k];
    for (int a = 0; a < 3; a++) cells[k].loc[a] = sc.loc[a];
    cells[k].first = sc.first;
    cells[k].count = sc.count;
    cells[k].width = (float)std::max(sc.width[0], std::max(sc.width[1], sc.width[2]));
    cells[k].pad = 0;
  }
  d.ncells = (int)nc;
  CellDev *dcells; float *dhm_old;
  if (upload(c, &dcells, cells.data(), nc) || upload(c, &d.cell_hmax, c->tree.h_max.data(), nc) ||
      upload(c, &dhm_old, c->tree.h_max.data(), nc))
    return 1;
  d.cells = dcells; d.cell_hmax_old = dhm_old;
  {
    float *drl;
    if (upload(c, &drl, c->wl.recurse_limit.data(), nc) || dmalloc(c, &d.stale, 1)) return 1;
    d.recurse_limit = drl;
  }
  {
    std::vector<int32_t> par(nc);
    std::vector<uint8_t> spl(nc);
    for (size_t k = 0; k < nc; k++) {
      par[k] = c->tree.cells[k].parent;
      spl[k] = c->tree.cells[k].split ? 1 : 0;
    }
    int32_t *dpar; uint8_t *dspl;
    if (upload(c, &dpar, par.data(), nc) || upload(c, &dspl, spl.data(), nc) || dmalloc(c, &d.cell_box, nc * 6) ||
        dmalloc(c, &d.cell_box_u, nc * 6))
      return 1;
    d.cell_parent = dpar; d.cell_split = dspl;
  }


  // work lists
  WorkLists &wl = c->wl;
  d.n_leaves = (int)wl.leaves.size();
  d.n_units = (int)wl.unit_leaf.size();
  int32_t *dlv, *dib, *drb, *dul, *duo, *dlp; Item *dit, *drt; int64_t *dsl;
  if (upload(c, &dlv, wl.leaves.data(), wl.leaves.size()) ||
      upload(c, &dib, wl.item_begin.data(), wl.item_begin.size()) ||
      upload(c, &dit, wl.items.data(), wl.items.size()) ||
      upload(c, &drb, wl.redo_begin.data(), wl.redo_begin.size()) ||
      upload(c, &drt, wl.redo_items.data(), wl.redo_items.size()) ||
      upload(c, &dul, wl.unit_leaf.data(), wl.unit_leaf.size()) ||
      upload(c, &duo, wl.unit_begin.data(), wl.unit_begin.size()) ||
      upload(c, &dlp, wl.lane_part.data(), wl.lane_part.size()) ||
      upload(c, &dsl, wl.sort_lookup.data(), wl.sort_lookup.size()))
    return 1;
  d.leaves = dlv; d.item_begin = dib; d.items = dit; d.redo_begin = drb; d.redo_items = drt;
  d.unit_leaf = dul; d.unit_begin = duo; d.lane_part = dlp; d.sort_lookup = dsl;
  {
    std::vector<int4> groups;
    for (size_t sl = 0; sl < wl.leaves.size(); sl++)
      for (int u = wl.unit_begin[sl]; u < wl.unit_begin[sl + 1]; u += kLW)
        groups.push_back(make_int4((int)sl, u, std::min(kLW, wl.unit_begin[sl + 1] - u), 0));
    int4 *dg;
    if (upload(c, &dg, groups.data(), groups.size())) return 1;
    d.groups = dg;
    d.n_groups = (int)groups.size();
    std::vector<int32_t> part_lane((size_t)c->n, -1);
    for (size_t k = 0; k < wl.lane_part.size(); k++)
      if (wl.lane_part[k] >= 0) part_lane[wl.lane_part[k]] = (int32_t)k;
    int32_t *dpl;
    if (upload(c, &dpl, part_lane.data(), part_lane.size())) return 1;
    d.part_lane = dpl;
    CK(cudaStreamSynchronize(c->stream));  // groups, part_lane go out of scope
    d.nl = nullptr;
    d.nl_grow = 1.0005f;
    d.nl_slope = 0.f;
    if (const char *e = getenv("SWIFT_CUDA_GROW")) d.nl_grow = (float)atof(e);
    if (const char *e = getenv("SWIFT_CUDA_SLOPE")) d.nl_slope = (float)atof(e);
    if (c->path & 8) {
      if (dmalloc(c, &d.nl, (size_t)d.n_units * kLD * 32) || dmalloc(c, &d.nl_cnt, (size_t)d.n_units * 32) ||
          dmalloc(c, &d.nl_R, (size_t)d.n_units * 32) || dmalloc(c, &d.h_before, (size_t)c->n))
        return 1;
    }
  }
  {
    std::vector<int32_t> slot_of((size_t)c->n, -1);
    for (size_t sl = 0; sl < wl.leaves.size(); sl++) {
      const swift_cuda_cell &leaf = c->tree.cells[wl.leaves[sl]];
      std::fill(slot_of.begin() + leaf.first, slot_of.begin() + leaf.first + leaf.count, (int32_t)sl);
    }
    // caller-order copy for swift_cuda_update_particles' check of the active set
    c->slot_of.assign((size_t)c->n, -1);
    for (int64_t k = 0; k < c->n; k++) c->slot_of[c->tree.order.empty() ? k : c->tree.order[k]] = slot_of[k];
    int32_t *dps;
    if (upload(c, &dps, slot_of.data(), slot_of.size()) || dmalloc(c, &d.leaf_mark, wl.leaves.size()) ||
        dmalloc(c, &d.leaf_list, wl.leaves.size()) ||
        dmalloc(c, &d.leaf_count, (size_t)c->prm.max_smoothing_iterations + 2))
      return 1;
    d.part_slot = dps;
    CK(cudaStreamSynchronize(c->stream));  // slot_of goes out of scope
  }
  if (dmalloc(c, &d.sort, (size_t)wl.n_sort_entries)) return 1;
  c->n_seg = (int)wl.sort_cells.size();
  {
    // segments in the order their cells are complete when the records arrive front to back
    // (swift_cuda_step_parts sorts a piece's cells while the next piece is on the bus)
    std::vector<int32_t> perm(c->n_seg);
    for (int s = 0; s < c->n_seg; s++) perm[s] = s;
    auto end_of = [&](int s) {
      const swift_cuda_cell &cl = c->tree.cells[wl.sort_cells[s]];
      return (int64_t)cl.first + cl.count;
    };
    std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return end_of(a) < end_of(b); });
    std::vector<int32_t> sc(c->n_seg), ss(c->n_seg);
    std::vector<int64_t> so(c->n_seg);
    c->seg_end.resize(c->n_seg);
    for (int s = 0; s < c->n_seg; s++) {
      sc[s] = wl.sort_cells[perm[s]];
      ss[s] = wl.sort_sids[perm[s]];
      so[s] = wl.sort_offset[perm[s]];
      c->seg_end[s] = end_of(perm[s]);
    }
    wl.sort_cells.swap(sc);
    wl.sort_sids.swap(ss);
    wl.sort_offset.swap(so);
    c->plan.P = 0;
  }
  std::vector<int64_t> scratch_off(c->n_seg, 0);
  int64_t scratch_total = 0;
  for (int s = 0; s < c->n_seg; s++) {
    const int cnt = c->tree.cells[wl.sort_cells[s]].count;
    int npad = 1;
    while (npad < cnt) npad <<= 1;
    if (npad > kSortSmem) {
      scratch_off[s] = scratch_total;
      scratch_total += npad;
    }
  }
  if (upload(c, &c->d_seg_cell, wl.sort_cells.data(), wl.sort_cells.size()) ||
      upload(c, &c->d_seg_sid, wl.sort_sids.data(), wl.sort_sids.size()) ||
      upload(c, &c->d_seg_off, wl.sort_offset.data(), wl.sort_offset.size()) ||
      upload(c, &c->d_scratch_off, scratch_off.data(), scratch_off.size()) ||
      dmalloc(c, &c->d_scratch, (size_t)scratch_total))
    return 1;
  if (dmalloc(c, &d.hits, 4) || dmalloc(c, &d.redo_count, (size_t)c->prm.max_smoothing_iterations + 2) ||
      dmalloc(c, &d.walk_count, (size_t)c->prm.max_smoothing_iterations + 2) ||
      dmalloc(c, &d.rec_count, 4))
    return 1;
  CK(cudaMemsetAsync(d.rec_count, 0, 4 * sizeof(unsigned long long), c->stream));
  d.rec_capacity = 0;
  {
    // persistent neighbour-loop kernels: one resident wave of CTAs, one hit FIFO per warp
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, c->device));
    int o0 = 0, o1 = 0, o2 = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o0, k_density<0>, kBlock, 0));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, k_density<1>, kBlock, 0));
    int o3 = 0;
    if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX) {
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, k_force<SWIFT_CUDA_SCHEME_SPHENIX>, kBlock, 0));
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o3, k_density<0, false, 1>, kBlock, 0));
    } else if (c->prm.scheme == SWIFT_CUDA_SCHEME_PLANETARY)
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, k_force<SWIFT_CUDA_SCHEME_PLANETARY>, kBlock, 0));
    else
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, k_force<SWIFT_CUDA_SCHEME_MINIMAL>, kBlock, 0));
    const int sms = prop.multiProcessorCount;
    auto grid_for = [&](int occ, int work_units) {
      const int want = (work_units + kWarps - 1) / kWarps;
      return std::max(1, std::min(want, sms * std::max(occ, 1)));
    };
    c->grid_density = grid_for(o0, d.n_units);
    c->grid_redo = grid_for(o1, d.n_units);
    c->grid_force = grid_for(o2, d.n_units);
    c->grid_gradient = grid_for(o3, d.n_units);
    int o4 = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o4, k_density<0, false, 2>, kBlock, 0));
    c->grid_limiter = grid_for(o4, d.n_units);
    const size_t qwarps =
        (size_t)std::max(std::max(std::max(c->grid_density, c->grid_gradient), c->grid_limiter),
                         std::max(c->grid_redo, c->grid_force)) *
        kWarps;
    // unit counters: [0] first density pass, [1 + sweep] redo sweeps, [1 + max_iter] force, [2 + max_iter] gradient
    if (dmalloc(c, &d.gq, qwarps * kQD * 32) || dmalloc(c, &d.work_ctr, (size_t)c->prm.max_smoothing_iterations + 3))
      return 1;
  }
  for (int k = 0; k < 3; k++) d.rec_pairs[k] = nullptr;
  if (upload_consts(c)) return 1;
  CK(cudaStreamSynchronize(c->stream));
  c->built = true;
  return 0;
}

static void cells_active(const swift_cuda_ctx *c, const swift_cuda_parts_in *in, std::vector<uint8_t> &act) {
  const size_t nc = c->tree.cells.size();
  act.assign(nc, 1);
  if (!in->time_bin || c->no_prune) return;
  const std::vector<int32_t> &ord = c->tree.order;
  for (size_t k = 0; k < nc; k++) {
    const swift_cuda_cell &cell = c->tree.cells[k];
    uint8_t a = 0;
    for (int j = 0; j < cell.count && !a; j++) {
      const int64_t s = ord.empty() ? (cell.first + j) : ord[cell.first + j];
      a = in->time_bin[s] <= c->prm.max_active_bin;
    }
    act[k] = a;
  }
}

extern "C" {

int swift_cuda_build_from_particles(swift_cuda_ctx *c, const swift_cuda_parts_in *in) {
  if (!c || !in || !in->x || !in->v || !in->mass || !in->h || !in->u) return fail("build: null argument");
  CK(cudaSetDevice(c->device));
  free_device(c);
  c->tree = Tree();
  const int rc = build_tree(c->prm, in->n, in->x, in->h, c->tree);
  if (rc == 1) return fail("build: top-level grid has a zero dimension");
  if (rc == 2) return fail("build: periodic box needs at least 3 top-level cells per dimension (h_max too large)");
  std::vector<float> hs(in->n);
  gather_f32(in->h, c->tree.order.data(), in->n, hs.data());
  compute_cell_hmax(c->tree, hs.data());
  build_tasks(c->prm, c->tree);
  std::vector<uint8_t> act;
  cells_active(c, in, act);
  c->all_cells_active = true;  // empty cells have nothing to contribute either way
  for (size_t k = 0; k < act.size(); k++)
    if (!act[k] && c->tree.cells[k].count > 0) c->all_cells_active = false;
  c->integ_on = false;
  c->marks_fresh = false;
  build_worklists(c->prm, c->tree, act, c->wl);
  assign_lanes(c->tree, in->x, c->wl);
  return pack_and_upload(c, in);
}

int swift_cuda_build_from_cells(swift_cuda_ctx *c, const swift_cuda_parts_in *in, int ncells,
                                const swift_cuda_cell *cells, int ntasks, const swift_cuda_task *tasks) {
  if (!c || !in || !cells || (!tasks && ntasks > 0)) return fail("build_from_cells: null argument");
  CK(cudaSetDevice(c->device));
  free_device(c);
  c->tree = Tree();
  c->tree.cells.assign(cells, cells + ncells);
  c->tree.tasks.assign(tasks, tasks + ntasks);
  for (int k = 0; k < ncells; k++)
    if (cells[k].first < 0 || (int64_t)cells[k].first + cells[k].count > in->n)
      return fail("build_from_cells: cell particle range outside the particle arrays");
  compute_cell_hmax(c->tree, in->h);
  std::vector<uint8_t> act;
  cells_active(c, in, act);
  c->all_cells_active = true;  // empty cells have nothing to contribute either way
  for (size_t k = 0; k < act.size(); k++)
    if (!act[k] && c->tree.cells[k].count > 0) c->all_cells_active = false;
  c->integ_on = false;
  c->marks_fresh = false;
  build_worklists(c->prm, c->tree, act, c->wl);
  assign_lanes(c->tree, in->x, c->wl);
  return pack_and_upload(c, in);
}

int swift_cuda_update_particles(swift_cuda_ctx *c, const swift_cuda_parts_in *in) {
  if (!c || !c->built) return fail("update_particles: context not built");
  if (!in || in->n != c->n) return fail("update_particles: particle count changed; rebuild");
  if (!in->x || !in->v || !in->mass || !in->h || !in->u) return fail("update_particles: null argument");
  CK(cudaSetDevice(c->device));
  if (!c->all_cells_active && in->time_bin) {
    // the work lists were built without the cells that held no active particle at the build
    const int64_t m = c->n_io < 0 ? c->n : c->n_io;
    for (int64_t k = 0; k < m; k++)
      if (in->time_bin[k] <= c->prm.max_active_bin && c->slot_of[k] < 0)
        return fail("update_particles: particle " + std::to_string(k) +
                    " is active but its cell was inactive at the build and has no work list; rebuild");
  }
  if (upload_and_pack(c, in, false)) return 1;
  // cell h_max follows the particles (c->hydro.h_max after a drift)
  Dev &d = c->dev;
  if (d.ncells > 0)
    k_cell_hmax_init<<<(d.ncells * 32 + 255) / 256, 256, 0, c->stream>>>(d, d.cell_hmax,
                                                                         const_cast<float *>(d.cell_hmax_old));
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int swift_cuda_set_transfer_count(swift_cuda_ctx *c, int64_t count) {
  if (!c || !c->built) return fail("set_transfer_count: context not built");
  if (count > c->n) return fail("set_transfer_count: more than the particles of the context");
  c->n_io = count < 0 ? -1 : count;
  return 0;
}

int swift_cuda_restore_particles(swift_cuda_ctx *c) {
  if (!c || !c->built) return fail("restore_particles: context not built");
  CK(cudaSetDevice(c->device));
  CK(cudaMemcpyAsync(c->dev.h, c->d_h0, c->n * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
  CK(cudaMemcpyAsync(c->dev.cell_hmax, c->dev.cell_hmax_old, (size_t)c->dev.ncells * sizeof(float),
                     cudaMemcpyDeviceToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

// Sustained FP32 FMA rate of this GPU (the roofline denominator of the pair kernels; the
// driver-written MEASURED_PEAKS.json has no FP32 entry).  Returns TFLOP/s, < 0 on error.
namespace swiftcuda {
__global__ void k_fma_peak(float *out, int iters) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f,
        a6 = a0 + 6.f, a7 = a0 + 7.f;
  const float b = 0.999f, c = 1e-3f;
  for (int i = 0; i < iters; i++) {
    a0 = __fmaf_rn(a0, b, c); a1 = __fmaf_rn(a1, b, c); a2 = __fmaf_rn(a2, b, c); a3 = __fmaf_rn(a3, b, c);
    a4 = __fmaf_rn(a4, b, c); a5 = __fmaf_rn(a5, b, c); a6 = __fmaf_rn(a6, b, c); a7 = __fmaf_rn(a7, b, c);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
}  // namespace swiftcuda

double swift_cuda_measure_fp32_peak(swift_cuda_ctx *c) {
  if (!c) return -1.0;
  if (cudaSetDevice(c->device) != cudaSuccess) return -1.0;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, c->device));
  const int blocks = prop.multiProcessorCount * 16, threads = 256, iters = 1 << 16;
  float *buf = nullptr;
  if (cudaMalloc(&buf, (size_t)blocks * threads * sizeof(float)) != cudaSuccess) return -1.0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k_fma_peak<<<blocks, threads, 0, c->stream>>>(buf, 1 << 10);  // warm-up
  float best = 1e30f;
  for (int r = 0; r < 3; r++) {
    cudaEventRecord(e0, c->stream);
    k_fma_peak<<<blocks, threads, 0, c->stream>>>(buf, iters);
    cudaEventRecord(e1, c->stream);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    best = std::min(best, ms);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(buf);
  if (cudaGetLastError() != cudaSuccess) return -1.0;
  return 2.0 * 8.0 * (double)iters * blocks * threads / (best * 1e-3) / 1e12;
}

// ---- phases ----------------------------------------------------------------
// Time-integration mode: the active set changes per step on the device.  k_mark_sorts flags the sorted
// lists this step reads and lists the units that hold an active particle; both stay valid until the
// active set changes (set_time, timestep, a build).
static int launch_mark(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if (!c->integ_on) {
    d.sort_need = nullptr;
    d.unit_list = nullptr;
    return 0;
  }
  if (c->marks_fresh) return 0;
  d.sort_need = c->d_sort_need;
  d.unit_list = c->d_unit_list;
  d.unit_count = c->d_unit_count;
  CK(cudaMemsetAsync(d.sort_need, 0, (size_t)d.ncells * 13, c->stream));
  CK(cudaMemsetAsync(d.unit_count, 0, sizeof(int), c->stream));
  if (d.n_leaves > 0) {
    k_mark_sorts<<<(d.n_leaves * 32 + 255) / 256, 256, 0, c->stream>>>(d);
    c->counters.n_kernel_launches += 1;
  }
  CK(cudaGetLastError());
  c->marks_fresh = true;
  return 0;
}

static int launch_sort(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if (launch_mark(c)) return 1;
  if (c->n_seg > 0) {
    const int grid = d.sort_need ? (c->n_seg + 255) / 256 : c->n_seg;
    k_sort<<<grid, 256, 0, c->stream>>>(c->dev, c->n_seg, c->d_seg_cell, c->d_seg_sid, c->d_seg_off, c->d_scratch,
                                        c->d_scratch_off);
  }
  c->counters.n_kernel_launches += 1;
  CK(cudaGetLastError());
  return 0;
}

// bounding boxes of the cells' particles (positions may have moved since the tree was built)
static int launch_cell_box(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if (d.ncells == 0) return 0;
  k_cell_box_init<<<(d.ncells * 6 + 255) / 256, 256, 0, c->stream>>>(d);
  k_cell_box<<<(d.ncells * 32 + 255) / 256, 256, 0, c->stream>>>(d);
  k_cell_box_finish<<<(d.ncells * 6 + 255) / 256, 256, 0, c->stream>>>(d);
  c->counters.n_kernel_launches += 3;
  CK(cudaGetLastError());
  return 0;
}

// What every step needs before the first neighbour loop: counters, cell boxes, hydro_init_part.
static int launch_density_prep(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if (launch_mark(c)) return 1;
  const int nb = (int)((c->n + 255) / 256);
  CK(cudaMemsetAsync(d.hits, 0, 4 * sizeof(unsigned long long), c->stream));
  CK(cudaMemsetAsync(d.redo_count, 0, ((size_t)c->prm.max_smoothing_iterations + 2) * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.walk_count, 0, ((size_t)c->prm.max_smoothing_iterations + 2) * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.leaf_count, 0, ((size_t)c->prm.max_smoothing_iterations + 2) * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.leaf_mark, 0, (size_t)std::max(d.n_leaves, 1) * sizeof(int), c->stream));
  CK(cudaMemcpyAsync(d.cell_hmax, d.cell_hmax_old, (size_t)d.ncells * sizeof(float), cudaMemcpyDeviceToDevice,
                     c->stream));
  CK(cudaMemsetAsync(d.work_ctr, 0, ((size_t)c->prm.max_smoothing_iterations + 3) * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.stale, 0, sizeof(int), c->stream));
  if (d.ncells > 0) k_check_lists<<<(d.ncells + 255) / 256, 256, 0, c->stream>>>(d, c->dx_max_part);
  if (launch_cell_box(c)) return 1;
  if (d.nl) CK(cudaMemsetAsync(d.nl_cnt, 0xff, (size_t)d.n_units * 32 * sizeof(unsigned short), c->stream));
  k_init_step<<<nb, 256, 0, c->stream>>>(d);
  c->counters.n_kernel_launches += 1;
  if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX && c->n > 0) {
    k_sphenix_prep<<<nb, 256, 0, c->stream>>>(d, c->sph_prep_mode);
    c->counters.n_kernel_launches += 1;
  }
  CK(cudaGetLastError());
  return 0;
}

static int launch_density(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if (launch_density_prep(c)) return 1;
  if ((c->path & 1) && !c->integ_on) {
    if (d.n_groups > 0)
      k_density_leaf<kDensCap, kDensQD><<<d.n_groups, kLT, sizeof(DensTileT), c->stream>>>(d, c->force_slow);
  } else if (d.n_units > 0) {
    if ((c->path & 32) && !c->integ_on) k_density<0, true><<<c->grid_density, kBlock, 0, c->stream>>>(d, 0, c->force_slow);
    else k_density<0><<<c->grid_density, kBlock, 0, c->stream>>>(d, 0);
  }
  c->counters.n_kernel_launches += 1;
  CK(cudaGetLastError());
  // keep the raw first-pass sums for inspection (parity tests); not in time-integration mode, where a
  // step may touch a handful of particles and two O(n) copies would be a third of it
  if (!c->integ_on) {
    CK(cudaMemcpyAsync(c->d_dens0_first, d.dens0, c->n * sizeof(float4), cudaMemcpyDeviceToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_dens1_first, d.dens1, c->n * sizeof(float4), cudaMemcpyDeviceToDevice, c->stream));
  }
  return 0;
}

static int launch_ghost(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if ((c->path & 2) && !c->integ_on) {
    // the whole smoothing-length iteration of a leaf inside one CTA (leaf_kernels.cuh)
    if (d.n_groups > 0)
      k_ghost_leaf<kGhostCap, kGhostQD><<<d.n_groups, kLT, sizeof(GhostTileT), c->stream>>>(d, c->force_slow);
    k_cell_hmax<<<(d.ncells * 32 + 255) / 256, 256, 0, c->stream>>>(d);
    c->counters.n_kernel_launches += 2;
    CK(cudaGetLastError());
    return 0;
  }
  const int nb = (int)std::min<int64_t>((c->n + 255) / 256, 148 * 16);  // grid-stride kernel
  // Fixed launch sequence, no host round trip: every kernel looks at the device-side
  // redo counter of the sweep before and leaves at once when nothing is pending.
  // The iteration normally ends within ~5 sweeps; max_smoothing_iterations is 30.  After sweep 5, 9, 13, ...
  // the host reads ONE counter and stops enqueueing when nothing is pending (one ~15 us sync instead of
  // up to 48 empty launches, which dominate a step with few active particles).
  for (int sweep = 0; sweep < c->prm.max_smoothing_iterations; sweep++) {
    k_ghost<<<nb, 256, 0, c->stream>>>(d, sweep);
    c->counters.n_kernel_launches += 1;
    if (sweep >= 5 && (sweep - 5) % 4 == 0) {
      int pending = 0;
      CK(cudaMemcpyAsync(&pending, d.redo_count + sweep, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      if (pending == 0) break;
    }
    if (d.nl && sweep > 0 && d.n_units > 0) {
      k_density_list<<<(d.n_units + kWarps - 1) / kWarps, kBlock, 0, c->stream>>>(d, sweep, c->force_slow);
      c->counters.n_kernel_launches += 1;
    }
    if (d.n_leaves > 0) {
      if ((c->path & 32) && !c->integ_on) k_density<1, true><<<c->grid_redo, kBlock, 0, c->stream>>>(d, sweep, c->force_slow);
      else k_density<1><<<c->grid_redo, kBlock, 0, c->stream>>>(d, sweep);
    }
    c->counters.n_kernel_launches += 1;
  }
  k_cell_hmax<<<(d.ncells * 32 + 255) / 256, 256, 0, c->stream>>>(d);
  c->counters.n_kernel_launches += 1;
  CK(cudaGetLastError());
  return 0;
}

// SPHENIX: the gradient loop (the first-pass walk with the converged h and runner_iact_nonsym_gradient)
static int launch_gradient(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if (c->prm.scheme != SWIFT_CUDA_SCHEME_SPHENIX) return fail("the gradient loop exists in the SPHENIX scheme only");
  CK(cudaMemsetAsync(d.work_ctr + 2 + c->prm.max_smoothing_iterations, 0, sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.hits + 3, 0, sizeof(unsigned long long), c->stream));
  if (d.n_units > 0) k_density<0, false, 1><<<c->grid_gradient, kBlock, 0, c->stream>>>(d, 0);
  c->counters.n_kernel_launches += 1;
  CK(cudaGetLastError());
  return 0;
}
// SPHENIX: runner_do_extra_ghost
static int launch_extra_ghost(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  if (c->prm.scheme != SWIFT_CUDA_SCHEME_SPHENIX) return fail("the extra ghost exists in the SPHENIX scheme only");
  if (c->n > 0) k_extra_ghost<<<(unsigned)((c->n + 255) / 256), 256, 0, c->stream>>>(d);
  c->counters.n_kernel_launches += 1;
  CK(cudaGetLastError());
  return 0;
}

static int launch_force(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  // runner_do_end_hydro_force is the epilogue of the force kernel
  CK(cudaMemsetAsync(d.work_ctr + 1 + c->prm.max_smoothing_iterations, 0, sizeof(int), c->stream));
  if (d.n_units > 0) {
    const int nb = c->grid_force;
    const bool fast = c->prm.fields_mode == SWIFT_CUDA_FIELDS_FAST;
    if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX) {
      if (fast) k_force_fast<SWIFT_CUDA_SCHEME_SPHENIX><<<nb, kBlock, 0, c->stream>>>(d, c->force_slow, c->fast_qd);
      else k_force<SWIFT_CUDA_SCHEME_SPHENIX, false><<<nb, kBlock, 0, c->stream>>>(d);
    } else if (c->prm.scheme == SWIFT_CUDA_SCHEME_PLANETARY) {
      if (fast) k_force_fast<SWIFT_CUDA_SCHEME_PLANETARY><<<nb, kBlock, 0, c->stream>>>(d, c->force_slow, c->fast_qd);
      else k_force<SWIFT_CUDA_SCHEME_PLANETARY, false><<<nb, kBlock, 0, c->stream>>>(d);
    } else {
      if (fast) k_force_fast<SWIFT_CUDA_SCHEME_MINIMAL><<<nb, kBlock, 0, c->stream>>>(d, c->force_slow, c->fast_qd);
      else k_force<SWIFT_CUDA_SCHEME_MINIMAL, false><<<nb, kBlock, 0, c->stream>>>(d);
    }
  }
  c->counters.n_kernel_launches += 1;
  CK(cudaGetLastError());
  return 0;
}

static int finish_counters(swift_cuda_ctx *c) {
  Dev &d = c->dev;
  unsigned long long hits[4];
  std::vector<int> redo((size_t)c->prm.max_smoothing_iterations + 2);
  CK(cudaMemcpyAsync(hits, d.hits, sizeof(hits), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(redo.data(), d.redo_count, redo.size() * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  int stale = 0;
  CK(cudaMemcpyAsync(&stale, d.stale, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (stale > 0)
    return fail("work lists are stale: " + std::to_string(stale) + " cell(s) the sub-tasks recursed through now have "
                "gamma * h_max + dx_max_part >= dmin / 2 (cell_can_recurse_in_*_hydro_task fails: the reference would "
                "stop recursing there); rebuild the context (swift_cuda_build_*)");
#ifdef SWIFT_CUDA_GHOST_STATS
  {
    int st[12];
    CK(cudaMemcpy(st, d.work_ctr, sizeof(st), cudaMemcpyDeviceToHost));
    fprintf(stderr, "ghost stats: rebuilds %d cta-sweeps %d multichunk %d overflow %d groups %d | kcycles(thread0) stage %d filter %d drain %d | tile used avg %.0f q0 avg %.1f\n", st[4], st[5], st[2], st[3], d.n_groups, st[6] / 16, st[7] / 16, st[8] / 16, st[9] / (double)std::max(st[4], 1), st[10] / (double)std::max(st[4], 1));
  }
#endif
  c->counters.n_density = (long long)hits[0];
  c->counters.n_density_redo = (long long)hits[1];
  c->counters.n_force = (long long)hits[2];
  c->n_gradient = (long long)hits[3];
  int sweeps = 0;
  for (int s = 0; s < c->prm.max_smoothing_iterations; s++)
    if (redo[s] > 0) sweeps = s + 1;
  c->counters.ghost_sweeps = sweeps;
  const int last = c->prm.max_smoothing_iterations - 1;
  c->counters.ghost_unconverged = last >= 0 ? redo[last] : 0;
  c->counters.n_leaf_interactions = c->wl.n_interactions;
  c->counters.n_sort_entries = c->wl.n_sort_entries;
  return 0;
}

#define NEED_BUILT(c)                                         \
  if (!(c) || !(c)->built) return fail("context not built"); \
  CK(cudaSetDevice((c)->device));

int swift_cuda_do_sort(swift_cuda_ctx *c) {
  NEED_BUILT(c);
  if (launch_sort(c)) return 1;
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}
int swift_cuda_do_density(swift_cuda_ctx *c) {
  NEED_BUILT(c);
  if (launch_density(c)) return 1;
  return finish_counters(c);
}
int swift_cuda_do_ghost(swift_cuda_ctx *c) {
  NEED_BUILT(c);
  if (launch_ghost(c)) return 1;
  if (finish_counters(c)) return 1;
  if (c->counters.ghost_unconverged > 0)
    return fail("Smoothing length failed to converge on " + std::to_string(c->counters.ghost_unconverged) +
                " particles.");
  return 0;
}
int swift_cuda_do_force(swift_cuda_ctx *c) {
  NEED_BUILT(c);
  if (launch_force(c)) return 1;
  return finish_counters(c);
}
int swift_cuda_do_gradient(swift_cuda_ctx *c) {
  NEED_BUILT(c);
  if (launch_gradient(c)) return 1;
  return finish_counters(c);
}
int swift_cuda_do_extra_ghost(swift_cuda_ctx *c) {
  NEED_BUILT(c);
  if (launch_extra_ghost(c)) return 1;
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int swift_cuda_step(swift_cuda_ctx *c, int flags) {
  NEED_BUILT(c);
  c->counters.n_kernel_launches = 0;
  CK(cudaEventRecord(c->ev[0], c->stream));
  if (flags & SWIFT_CUDA_STEP_RESTORE) {
    CK(cudaMemcpyAsync(c->dev.h, c->d_h0, c->n * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
  }
  if ((flags & SWIFT_CUDA_STEP_SORT) && launch_sort(c)) return 1;
  CK(cudaEventRecord(c->ev[1], c->stream));
  if (launch_density(c)) return 1;
  CK(cudaEventRecord(c->ev[2], c->stream));
  if (launch_ghost(c)) return 1;
  // SPHENIX: gradient loop and extra ghost (timed with the ghost phase)
  if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX && (launch_gradient(c) || launch_extra_ghost(c))) return 1;
  CK(cudaEventRecord(c->ev[3], c->stream));
  if (launch_force(c)) return 1;
  CK(cudaEventRecord(c->ev[4], c->stream));
  if (finish_counters(c)) return 1;
  CK(cudaEventElapsedTime(&c->timings.sort_ms, c->ev[0], c->ev[1]));
  CK(cudaEventElapsedTime(&c->timings.density_ms, c->ev[1], c->ev[2]));
  CK(cudaEventElapsedTime(&c->timings.ghost_ms, c->ev[2], c->ev[3]));
  CK(cudaEventElapsedTime(&c->timings.force_ms, c->ev[3], c->ev[4]));
  CK(cudaEventElapsedTime(&c->timings.total_ms, c->ev[0], c->ev[4]));
  if (c->counters.ghost_unconverged > 0)
    return fail("Smoothing length failed to converge on " + std::to_string(c->counters.ghost_unconverged) +
                " particles.");
  return 0;
}

// ---- results ---------------------------------------------------------------
// K7 for a subset of the fields: scatter to caller order on the device (stream `ks`), then one
// copy per requested field on stream `cs` (after event `ready` when the streams differ).
// which: 1 = what the ghost leaves (h, rho, P, c, f, balsara, first-pass sums), 2 = what the force
// loop leaves (a, u_dt, h_dt, v_sig, min_ngb_time_bin).
static int enqueue_results(swift_cuda_ctx *c, const swift_cuda_parts_out *out, int which, cudaStream_t ks,
                           cudaStream_t cs, cudaEvent_t ready) {
  const int64_t n = c->n;
  Dev &d = c->dev;
  if (!c->s_out) {
    if (dmalloc(c, &c->s_out, (size_t)n * 20) || dmalloc(c, &c->s_out_bin, (size_t)n)) return 1;
  }
  struct Field { float *host; int width; float **dev; int phase; };
  RawOut r;
  std::memset(&r, 0, sizeof(r));
  Field fields[] = {{out->h, 1, &r.h, 1}, {out->rho, 1, &r.rho, 1}, {out->pressure, 1, &r.pressure, 1},
                    {out->soundspeed, 1, &r.soundspeed, 1}, {out->f, 1, &r.f, 1}, {out->balsara, 1, &r.balsara, 1},
                    {out->a, 3, &r.a, 2}, {out->u_dt, 1, &r.u_dt, 2}, {out->h_dt, 1, &r.h_dt, 2},
                    {out->v_sig, 1, &r.v_sig, 2},
                    {out->d_rho, 1, &r.d_rho, 1}, {out->d_rho_dh, 1, &r.d_rho_dh, 1}, {out->d_wcount, 1, &r.d_wcount, 1},
                    {out->d_wcount_dh, 1, &r.d_wcount_dh, 1}, {out->d_div_v, 1, &r.d_div_v, 1},
                    {out->d_rot_v, 3, &r.d_rot_v, 1}};
  int64_t off = 0;
  for (Field &f : fields) {
    if (f.host && (f.phase & which)) *f.dev = c->s_out + off;
    off += (int64_t)f.width * n;
  }
  const int64_t m = c->n_io < 0 ? n : c->n_io;  // particles on the bus
  const bool bin = out->min_ngb_time_bin && (which & 2);
  if (bin) r.min_bin = c->s_out_bin;
  r.order = c->tree.order.empty() ? nullptr : c->s_order;
  if (n > 0) k_unpack<<<(unsigned)((n + 255) / 256), 256, 0, ks>>>(d, r, c->d_dens0_first, c->d_dens1_first);
  CK(cudaGetLastError());
  if (cs != ks) {
    CK(cudaEventRecord(ready, ks));
    CK(cudaStreamWaitEvent(cs, ready, 0));
  }
  for (Field &f : fields)
    if (f.host && (f.phase & which))
      CK(cudaMemcpyAsync(f.host, *f.dev, (size_t)f.width * m * 4, cudaMemcpyDeviceToHost, cs));
  if (bin) CK(cudaMemcpyAsync(out->min_ngb_time_bin, c->s_out_bin, (size_t)m, cudaMemcpyDeviceToHost, cs));
  return 0;
}

int swift_cuda_get_results(swift_cuda_ctx *c, const swift_cuda_parts_out *out) {
  NEED_BUILT(c);
  if (!out) return fail("get_results: null argument");
  if (enqueue_results(c, out, 3, c->stream, c->stream, nullptr)) return 1;
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

// update_particles + step + get_results as ONE call with the transfers overlapped with the kernels:
// the positions go up first and the sorts start while v, m, h, u are still on the bus; what the
// ghost leaves (h, rho, P, c, f, balsara) goes down while the force loop runs.
int swift_cuda_step_io(swift_cuda_ctx *c, const swift_cuda_parts_in *in, int flags, const swift_cuda_parts_out *out) {
  NEED_BUILT(c);
  if (!in || !out) return fail("step_io: null argument");
  if (in->n != c->n) return fail("step_io: particle count changed; rebuild");
  if (!in->x || !in->v || !in->mass || !in->h || !in->u) return fail("step_io: null argument");
  const int64_t n = c->n;
  const int64_t m = c->n_io < 0 ? n : c->n_io;  // particles on the bus
  Dev &d = c->dev;
  c->counters.n_kernel_launches = 0;
  c->marks_fresh = false;  // time bins may change
  RawIn r;
  std::memset(&r, 0, sizeof(r));
  // io stream: positions, then the rest
  CK(cudaMemcpyAsync(c->s_x, in->x, m * 24, cudaMemcpyHostToDevice, c->io));
  CK(cudaEventRecord(c->io_ev[0], c->io));
  CK(cudaMemcpyAsync(c->s_v, in->v, m * 12, cudaMemcpyHostToDevice, c->io));
  CK(cudaMemcpyAsync(c->s_mass, in->mass, m * 4, cudaMemcpyHostToDevice, c->io));
  CK(cudaMemcpyAsync(c->s_h, in->h, m * 4, cudaMemcpyHostToDevice, c->io));
  CK(cudaMemcpyAsync(c->s_u, in->u, m * 4, cudaMemcpyHostToDevice, c->io));
  r.x = c->s_x; r.v = c->s_v; r.mass = c->s_mass; r.h = c->s_h; r.u = c->s_u;
  if (in->time_bin) {
    CK(cudaMemcpyAsync(c->s_tb, in->time_bin, m, cudaMemcpyHostToDevice, c->io));
    r.time_bin = c->s_tb;
  }
  const float *carry[5] = {in->rho, in->pressure, in->f, in->soundspeed, in->balsara};
  const float **rc[5] = {&r.rho, &r.pressure, &r.f, &r.soundspeed, &r.balsara};
  for (int k = 0; k < 5; k++)
    if (carry[k]) {
      if (!c->s_carry[k] && dmalloc(c, &c->s_carry[k], (size_t)n)) return 1;
      CK(cudaMemcpyAsync(c->s_carry[k], carry[k], m * 4, cudaMemcpyHostToDevice, c->io));
      *rc[k] = c->s_carry[k];
    }
  CK(cudaEventRecord(c->io_ev[1], c->io));
  r.order = c->tree.order.empty() ? nullptr : c->s_order;
  PackDst o;
  o.pos = const_cast<PosD *>(d.pos);
  o.vm = d.vm; o.fin = d.fin; o.hb = d.hb; o.h = d.h; o.h0 = c->d_h0; o.u = const_cast<float *>(d.u);
  o.time_bin = const_cast<int8_t *>(d.time_bin);
  o.mat_id = const_cast<int32_t *>(d.mat_id);
  o.id = const_cast<int64_t *>(d.id);
  const unsigned nb = (unsigned)((n + 255) / 256);
  // compute stream
  CK(cudaEventRecord(c->ev[0], c->stream));
  CK(cudaStreamWaitEvent(c->stream, c->io_ev[0], 0));
  if (n > 0) k_pack<<<nb, 256, 0, c->stream>>>(n, m, r, o, 0, 1);
  if ((flags & SWIFT_CUDA_STEP_SORT) && launch_sort(c)) return 1;
  CK(cudaStreamWaitEvent(c->stream, c->io_ev[1], 0));
  if (n > 0) k_pack<<<nb, 256, 0, c->stream>>>(n, m, r, o, 0, 2);
  if (d.ncells > 0)
    k_cell_hmax_init<<<(d.ncells * 32 + 255) / 256, 256, 0, c->stream>>>(d, d.cell_hmax,
                                                                         const_cast<float *>(d.cell_hmax_old));
  CK(cudaGetLastError());
  CK(cudaEventRecord(c->ev[1], c->stream));
  if (launch_density(c)) return 1;
  CK(cudaEventRecord(c->ev[2], c->stream));
  if (launch_ghost(c)) return 1;
  // SPHENIX: the gradient loop and the extra ghost leave what the ghost wrote (h, rho, P, c, f, balsara) alone:
  // those fields still go out under the loops that follow
  if (enqueue_results(c, out, 1, c->stream, c->io, c->io_ev[2])) return 1;
  if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX && (launch_gradient(c) || launch_extra_ghost(c))) return 1;
  CK(cudaEventRecord(c->ev[3], c->stream));
  if (launch_force(c)) return 1;
  CK(cudaEventRecord(c->ev[4], c->stream));
  if (enqueue_results(c, out, 2, c->stream, c->stream, nullptr)) return 1;
  if (finish_counters(c)) return 1;
  CK(cudaStreamSynchronize(c->io));
  CK(cudaEventElapsedTime(&c->timings.sort_ms, c->ev[0], c->ev[1]));
  CK(cudaEventElapsedTime(&c->timings.density_ms, c->ev[1], c->ev[2]));
  CK(cudaEventElapsedTime(&c->timings.ghost_ms, c->ev[2], c->ev[3]));
  CK(cudaEventElapsedTime(&c->timings.force_ms, c->ev[3], c->ev[4]));
  CK(cudaEventElapsedTime(&c->timings.total_ms, c->ev[0], c->ev[4]));
  if (c->counters.ghost_unconverged > 0)
    return fail("Smoothing length failed to converge on " + std::to_string(c->counters.ghost_unconverged) +
                " particles.");
  return 0;
}

}  // extern "C"

// ---- the caller's AoS records (struct part) ---------------------------------------------------
namespace {
struct HostSoA {
  std::vector<double> x;
  std::vector<float> v, mass, h, u, rho, P, f, cs, bal;
  std::vector<int8_t> tb;
  std::vector<int32_t> mat;
  std::vector<int64_t> id;
};
template <typename T>
static inline T rec_get(const unsigned char *r, int off) {
  T v;
  std::memcpy(&v, r + off, sizeof(T));
  return v;
}
// a SPHENIX context needs the scheme's fields in the record, the two-loop schemes must not name them
static int check_sphenix_layout(const swift_cuda_ctx *c, const swift_cuda_part_layout *L) {
  if (!c || !L) return fail("parts: null argument");
  const int sx[] = {L->alpha_visc, L->div_v_previous_step, L->div_v_dt, L->alpha_diff, L->laplace_u,
                    L->alpha_visc_max_ngb};
  if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX) {
    for (int o : sx)
      if (o < 0 || o + 4 > L->stride) return fail("parts: the SPHENIX scheme needs the offsets of p->viscosity / "
                                                   "p->diffusion / p->force.alpha_visc_max_ngb in the record layout");
  }
  return 0;
}

static void aos_to_soa(const swift_cuda_part_layout &L, const void *parts, int64_t n, HostSoA &o,
                       swift_cuda_parts_in &in) {
  o.x.resize(3 * n); o.v.resize(3 * n); o.mass.resize(n); o.h.resize(n); o.u.resize(n);
  o.rho.resize(n); o.P.resize(n); o.f.resize(n); o.cs.resize(n); o.bal.resize(n);
  o.tb.resize(n); o.mat.resize(n); o.id.resize(n);
  const unsigned char *base = static_cast<const unsigned char *>(parts);
  for (int64_t k = 0; k < n; k++) {
    const unsigned char *r = base + (size_t)k * L.stride;
    for (int a = 0; a < 3; a++) {
      o.x[3 * k + a] = rec_get<double>(r, L.x + 8 * a);
      o.v[3 * k + a] = rec_get<float>(r, L.v + 4 * a);
    }
    o.mass[k] = rec_get<float>(r, L.mass);
    o.h[k] = rec_get<float>(r, L.h);
    o.u[k] = rec_get<float>(r, L.u);
    o.rho[k] = rec_get<float>(r, L.rho);
    o.P[k] = rec_get<float>(r, L.pressure);
    o.f[k] = rec_get<float>(r, L.f);
    o.cs[k] = rec_get<float>(r, L.soundspeed);
    o.bal[k] = rec_get<float>(r, L.balsara);
    o.tb[k] = rec_get<int8_t>(r, L.time_bin);
    o.mat[k] = L.mat_id >= 0 ? rec_get<int32_t>(r, L.mat_id) : 0;
    o.id[k] = L.id >= 0 ? rec_get<int64_t>(r, L.id) : k;
  }
  std::memset(&in, 0, sizeof(in));
  in.n = n;
  in.x = o.x.data(); in.v = o.v.data(); in.mass = o.mass.data(); in.h = o.h.data(); in.u = o.u.data();
  in.mat_id = o.mat.data(); in.time_bin = o.tb.data(); in.id = o.id.data();
  in.rho = o.rho.data(); in.pressure = o.P.data(); in.f = o.f.data(); in.soundspeed = o.cs.data();
  in.balsara = o.bal.data();
}
static int check_layout(const swift_cuda_part_layout *L) {
  if (!L || L->stride <= 0) return fail("parts: bad record layout");
  const int need[] = {L->x, L->v, L->a_hydro, L->mass, L->h, L->u, L->u_dt, L->rho, L->rho_dh, L->wcount,
                      L->wcount_dh, L->div_v, L->rot_v, L->f, L->pressure, L->soundspeed, L->v_sig, L->h_dt,
                      L->balsara, L->time_bin};
  for (int o : need)
    if (o < 0 || o >= L->stride) return fail("parts: a required field offset lies outside the record");
  if (L->x % 8 || L->v % 4) return fail("parts: misaligned field offset");
  return 0;
}
}  // namespace

static int finish_parts_build(swift_cuda_ctx *c, const swift_cuda_part_layout *L) {
  c->lay = *L;
  c->has_lay = true;
  // records in cell order (always so through the cells entry: a cell's `first` indexes the caller's array)
  // and leaves in the order of their records?
  bool ident = true;
  for (size_t k = 0; k < c->tree.order.size() && ident; k++) ident = c->tree.order[k] == (int32_t)k;
  for (size_t sl = 1; sl < c->wl.leaves.size() && ident; sl++)
    ident = c->tree.cells[c->wl.leaves[sl]].first >= c->tree.cells[c->wl.leaves[sl - 1]].first;
  c->records_in_cell_order = ident;
  if (const char *e = getenv("SWIFT_CUDA_PIECES")) c->pieces = std::max(1, atoi(e));
  return dmalloc(c, &c->d_aos, (size_t)std::max<int64_t>(c->n, 1) * L->stride);
}

// Pieces of consecutive leaves (= consecutive units = consecutive records) of about equal unit counts,
// and for every piece the last record any of its first-pass / force interactions reads.
static void build_piece_plan(swift_cuda_ctx *c, int P) {
  const WorkLists &wl = c->wl;
  swift_cuda_ctx::PiecePlan &pl = c->plan;
  const int nl = (int)wl.leaves.size();
  const int64_t n = c->n;
  pl.P = P;
  pl.leaf.assign(P + 1, 0);
  pl.rec.assign(P + 1, 0);
  pl.dep_hi.assign(P, 0);
  for (int k = 1; k < P; k++) {
    const int32_t want = (int32_t)((int64_t)c->dev.n_units * k / P);
    int sl = (int)(std::lower_bound(wl.unit_begin.begin(), wl.unit_begin.begin() + nl, want) - wl.unit_begin.begin());
    sl = std::max(sl, pl.leaf[k - 1]);
    pl.leaf[k] = sl;
    pl.rec[k] = sl < nl ? (int64_t)c->tree.cells[wl.leaves[sl]].first : n;
  }
  pl.leaf[P] = nl;
  pl.rec[P] = n;
  auto end_of = [&](int32_t cell) {
    const swift_cuda_cell &cl = c->tree.cells[cell];
    return (int64_t)cl.first + cl.count;
  };
  for (int k = 0; k < P; k++) {
    int64_t hi = pl.rec[k + 1];
    for (int32_t it = wl.item_begin[pl.leaf[k]]; it < wl.item_begin[pl.leaf[k + 1]]; it++) {
      hi = std::max(hi, end_of(wl.items[it].own));
      hi = std::max(hi, end_of(wl.items[it].partner));
    }
    pl.dep_hi[k] = hi;
  }
}

// The front of swift_cuda_step_parts with the records in cell order: the records go up in pieces (copy
// stream); as a piece lands its records are unpacked, its particles armed, the cells it completes get
// their h_max, boxes and sorted lists, and the first density pass runs over every piece of units whose
// interactions read nothing beyond what has landed -- under the copy of the next piece.
static int parts_front_in_pieces(swift_cuda_ctx *c, const void *parts, const PackDst &o, const AosDst &a) {
  Dev &d = c->dev;
  const swift_cuda_part_layout L = c->lay;
  const swift_cuda_ctx::PiecePlan &pl = c->plan;
  const int P = pl.P;
  const int64_t n = c->n;
  const size_t nit = (size_t)c->prm.max_smoothing_iterations + 2;
  // events: [0, P) results of a force piece written back, [P, 2P) a piece of records has landed, [2P] start
  while ((int)c->piece_ev.size() < 2 * P + 1) {
    cudaEvent_t e;
    CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    c->piece_ev.push_back(e);
  }
  if (launch_mark(c)) return 1;
  // what launch_density_prep does before any particle is looked at
  CK(cudaMemsetAsync(d.hits, 0, 4 * sizeof(unsigned long long), c->stream));
  CK(cudaMemsetAsync(d.redo_count, 0, nit * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.walk_count, 0, nit * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.leaf_count, 0, nit * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.leaf_mark, 0, (size_t)std::max(d.n_leaves, 1) * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.work_ctr, 0, nit * sizeof(int), c->stream));
  CK(cudaMemsetAsync(d.stale, 0, sizeof(int), c->stream));
  if (d.nl) CK(cudaMemsetAsync(d.nl_cnt, 0xff, (size_t)d.n_units * 32 * sizeof(unsigned short), c->stream));
  const int nbc = (d.ncells * 32 + 255) / 256, nb6 = (d.ncells * 6 + 255) / 256;
  k_cell_box_init<<<nb6, 256, 0, c->stream>>>(d);
  c->counters.n_kernel_launches += 1;
  // the copy stream: every piece of records, front to back, each with its event
  CK(cudaEventRecord(c->piece_ev[2 * P], c->stream));
  CK(cudaStreamWaitEvent(c->io, c->piece_ev[2 * P], 0));
  for (int j = 0; j < P; j++) {
    const int64_t r0 = pl.rec[j], r1 = pl.rec[j + 1];
    if (r1 > r0)
      CK(cudaMemcpyAsync(c->d_aos + (size_t)r0 * L.stride, (const unsigned char *)parts + (size_t)r0 * L.stride,
                         (size_t)(r1 - r0) * L.stride, cudaMemcpyHostToDevice, c->io));
    CK(cudaEventRecord(c->piece_ev[P + j], c->io));
  }
  std::vector<char> launched(P, 0);
  size_t seg_lo = 0;
  for (int j = 0; j < P; j++) {
    const int64_t r0 = pl.rec[j], r1 = pl.rec[j + 1];
    CK(cudaStreamWaitEvent(c->stream, c->piece_ev[P + j], 0));
    if (r1 > r0) {
      const unsigned nb = (unsigned)((r1 - r0 + 255) / 256);
      k_pack_aos<<<nb, 256, 0, c->stream>>>(r1, c->d_aos, L, nullptr, o, a, 0, 0, 0, r0);
      k_init_step<<<nb, 256, 0, c->stream>>>(d, r0, r1);
      k_cell_box<<<nbc, 256, 0, c->stream>>>(d, r0, r1);
      c->counters.n_kernel_launches += 3;
    }
    // the cells this piece completes (a parent is complete with its last leaf)
    k_cell_hmax_init<<<nbc, 256, 0, c->stream>>>(d, d.cell_hmax, const_cast<float *>(d.cell_hmax_old),
                                                 j == 0 ? (int64_t)-1 : r0, j == P - 1 ? INT64_MAX : r1);
    k_cell_box_finish<<<nb6, 256, 0, c->stream>>>(d);
    c->counters.n_kernel_launches += 2;
    const size_t seg_hi = j == P - 1 ? (size_t)c->n_seg
                                     : (size_t)(std::upper_bound(c->seg_end.begin(), c->seg_end.end(), r1) - c->seg_end.begin());
    if (seg_hi > seg_lo) {
      k_sort<<<(unsigned)(seg_hi - seg_lo), 256, 0, c->stream>>>(d, (int)(seg_hi - seg_lo), c->d_seg_cell + seg_lo,
                                                                  c->d_seg_sid + seg_lo, c->d_seg_off + seg_lo, c->d_scratch,
                                                                  c->d_scratch_off + seg_lo);
      c->counters.n_kernel_launches += 1;
    }
    seg_lo = seg_hi;
    CK(cudaGetLastError());
    if (j == P - 1) CK(cudaEventRecord(c->ev[1], c->stream));
    // the first density pass over every piece of units that reads nothing beyond what has landed
    for (int k = 0; k < P; k++) {
      if (launched[k] || (pl.dep_hi[k] > r1 && j < P - 1)) continue;
      launched[k] = 1;
      d.unit_lo = c->wl.unit_begin[pl.leaf[k]];
      d.unit_hi = c->wl.unit_begin[pl.leaf[k + 1]];
      if (d.unit_hi > d.unit_lo) {
        CK(cudaMemsetAsync(d.work_ctr, 0, sizeof(int), c->stream));
        k_density<0><<<c->grid_density, kBlock, 0, c->stream>>>(d, 0);
        c->counters.n_kernel_launches += 1;
        CK(cudaGetLastError());
      }
    }
  }
  d.unit_lo = d.unit_hi = 0;
  if (d.ncells > 0) k_check_lists<<<(d.ncells + 255) / 256, 256, 0, c->stream>>>(d, c->dx_max_part);
  CK(cudaMemcpyAsync(c->d_dens0_first, d.dens0, n * sizeof(float4), cudaMemcpyDeviceToDevice, c->stream));
  CK(cudaMemcpyAsync(c->d_dens1_first, d.dens1, n * sizeof(float4), cudaMemcpyDeviceToDevice, c->stream));
  CK(cudaGetLastError());
  return 0;
}

extern "C" {

int swift_cuda_build_from_cells_parts(swift_cuda_ctx *c, const swift_cuda_part_layout *L, const void *parts, int64_t n,
                                      int ncells, const swift_cuda_cell *cells, int ntasks,
                                      const swift_cuda_task *tasks) {
  if (!c || !parts) return fail("build_from_cells_parts: null argument");
  if (check_sphenix_layout(c, L)) return 1;
  if (check_layout(L)) return 1;
  HostSoA soa;
  swift_cuda_parts_in in;
  aos_to_soa(*L, parts, n, soa, in);
  c->no_prune = true;  // which particles are active is decided per call (swift_cuda_set_max_active_bin)
  const int rc = swift_cuda_build_from_cells(c, &in, ncells, cells, ntasks, tasks);
  c->no_prune = false;
  if (rc) return rc;
  return finish_parts_build(c, L);
}

int swift_cuda_build_from_parts(swift_cuda_ctx *c, const swift_cuda_part_layout *L, const void *parts, int64_t n) {
  if (!c || !parts) return fail("build_from_parts: null argument");
  if (check_sphenix_layout(c, L)) return 1;
  if (check_layout(L)) return 1;
  HostSoA soa;
  swift_cuda_parts_in in;
  aos_to_soa(*L, parts, n, soa, in);
  c->no_prune = true;
  const int rc = swift_cuda_build_from_particles(c, &in);
  c->no_prune = false;
  if (rc) return rc;
  return finish_parts_build(c, L);
}

int swift_cuda_set_max_active_bin(swift_cuda_ctx *c, int max_active_bin) {
  NEED_BUILT(c);
  if (!c->all_cells_active)
    return fail("set_max_active_bin: the work lists were built for a fixed active set (cells without an active "
                "particle were left out); build with max_active_bin = 56 or through the struct part entries");
  c->prm.max_active_bin = max_active_bin;
  c->marks_fresh = false;
  if (upload_consts(c)) return 1;
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int swift_cuda_step_parts(swift_cuda_ctx *c, void *parts, int phases) {
  NEED_BUILT(c);
  if (!c->has_lay) return fail("step_parts: the context was not built from struct part records");
  if (!parts) return fail("step_parts: null argument");
  if (c->integ_on) return fail("step_parts: not in time-integration mode");
  const int64_t n = c->n;
  Dev &d = c->dev;
  const swift_cuda_part_layout L = c->lay;
  const bool do_dens = phases & SWIFT_CUDA_PHASE_DENSITY, do_ghost = phases & SWIFT_CUDA_PHASE_GHOST;
  const bool do_force = phases & SWIFT_CUDA_PHASE_FORCE, do_end = phases & SWIFT_CUDA_PHASE_END_FORCE;
  const bool sphenix = c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX;
  const bool do_grad = phases & SWIFT_CUDA_PHASE_GRADIENT, do_extra = phases & SWIFT_CUDA_PHASE_EXTRA_GHOST;
  if ((do_grad || do_extra) && !sphenix)
    return fail("step_parts: SWIFT_CUDA_PHASE_GRADIENT / _EXTRA_GHOST exist in the SPHENIX scheme only");
  const bool keep = (phases & SWIFT_CUDA_PHASE_KEEP) || (do_ghost && !do_dens);
  c->counters.n_kernel_launches = 0;
  const WorkLists &wl = c->wl;
  const int P = c->pieces;
  // records in cell order and enough of them: the step runs in P pieces of consecutive leaves at both ends
  const bool pieces_ok = !sphenix && P > 1 && c->records_in_cell_order && n >= (1 << 18) && d.n_units >= 64 * P &&
                         (int64_t)wl.unit_begin.size() == (int64_t)wl.leaves.size() + 1;
  if (pieces_ok && c->plan.P != P) build_piece_plan(c, P);
  CK(cudaEventRecord(c->ev[0], c->stream));
  PackDst o;
  o.pos = const_cast<PosD *>(d.pos);
  o.vm = d.vm; o.fin = d.fin; o.hb = d.hb; o.h = d.h; o.h0 = c->d_h0; o.u = const_cast<float *>(d.u);
  o.time_bin = const_cast<int8_t *>(d.time_bin);
  o.mat_id = const_cast<int32_t *>(d.mat_id);
  o.id = const_cast<int64_t *>(d.id);
  AosDst a;
  a.dens0 = d.dens0; a.dens1 = d.dens1; a.acc = d.acc; a.h_dt = d.h_dt; a.v_sig = d.v_sig; a.min_bin = d.min_bin;
  a.sph_state = d.sph_state; a.sph_grad = d.sph_grad;  // null unless SPHENIX
  // SPHENIX phases in separate calls: the neighbour records of ACTIVE particles come from the uploaded state when
  // the phase that would write them does not run here
  c->sph_prep_mode = sphenix ? ((do_ghost ? 0 : 1) | (do_extra ? 0 : 2)) : 0;
  const int32_t *order = (c->tree.order.empty() || c->records_in_cell_order) ? nullptr : c->s_order;
  const unsigned nb = (unsigned)((n + 255) / 256);
  const int load_force = (!do_dens && !do_ghost) ? 1 : 0;
  c->marks_fresh = false;
  // (without density and ghost in this call the running force sums of the records must survive k_init_step too)
  d.keep_sums = (keep || load_force) ? 1 : 0;
  d.no_end_force = do_end ? 0 : 1;
  const bool front_in_pieces = pieces_ok && (phases & SWIFT_CUDA_PHASE_SORT) && do_dens && !keep && !(c->path & 33);
  if (front_in_pieces) {
    if (parts_front_in_pieces(c, parts, o, a)) return 1;   // records ev[1] after the last sort
  } else {
    CK(cudaMemcpyAsync(c->d_aos, parts, (size_t)n * L.stride, cudaMemcpyHostToDevice, c->stream));
    if (n > 0) k_pack_aos<<<nb, 256, 0, c->stream>>>(n, c->d_aos, L, order, o, a, 0, keep ? 1 : 0, load_force);
    if (d.ncells > 0)
      k_cell_hmax_init<<<(d.ncells * 32 + 255) / 256, 256, 0, c->stream>>>(d, d.cell_hmax,
                                                                           const_cast<float *>(d.cell_hmax_old));
    c->counters.n_kernel_launches += 2;
    CK(cudaGetLastError());
    if ((phases & SWIFT_CUDA_PHASE_SORT) && launch_sort(c)) return 1;
    CK(cudaEventRecord(c->ev[1], c->stream));
    if (do_dens) {
      if (launch_density(c)) return 1;
    } else if (launch_density_prep(c)) {
      return 1;
    }
  }
  CK(cudaEventRecord(c->ev[2], c->stream));
  if (do_ghost && launch_ghost(c)) return 1;
  if (do_grad && launch_gradient(c)) return 1;
  if (do_extra && launch_extra_ghost(c)) return 1;
  c->sph_prep_mode = 0;
  CK(cudaEventRecord(c->ev[3], c->stream));
  const int what = (do_force || do_end) ? (do_ghost ? 16 : 8) : (do_ghost ? 4 : (do_dens ? 2 : 0));
  // SPHENIX: density alone leaves the raw sums (the common code below); any later phase writes its own fields
  const int ran = phases & (SWIFT_CUDA_PHASE_GHOST | SWIFT_CUDA_PHASE_GRADIENT | SWIFT_CUDA_PHASE_EXTRA_GHOST |
                            SWIFT_CUDA_PHASE_FORCE | SWIFT_CUDA_PHASE_END_FORCE);
  const bool in_pieces = do_force && pieces_ok;
  if (in_pieces) {
    // The force loop over the units of consecutive leaves, piece by piece: the records of a finished piece
    // are written back and copied to the host (copy stream) while the loop runs over the next piece.
    const swift_cuda_ctx::PiecePlan &pl = c->plan;
    while ((int)c->piece_ev.size() < 2 * P + 1) {
      cudaEvent_t e;
      CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      c->piece_ev.push_back(e);
    }
    for (int k = 0; k < P; k++) {
      const int64_t p_prev = pl.rec[k], p_next = pl.rec[k + 1];
      d.unit_lo = wl.unit_begin[pl.leaf[k]];
      d.unit_hi = wl.unit_begin[pl.leaf[k + 1]];
      if (d.unit_hi > d.unit_lo && launch_force(c)) return 1;
      if (k == P - 1) CK(cudaEventRecord(c->ev[4], c->stream));
      if (p_next > p_prev) {
        const unsigned nbp = (unsigned)((p_next - p_prev + 255) / 256);
        k_unpack_aos<<<nbp, 256, 0, c->stream>>>(d, c->d_aos, L, nullptr, what, c->prm.max_active_bin, p_prev, p_next);
        c->counters.n_kernel_launches += 1;
        CK(cudaGetLastError());
        CK(cudaEventRecord(c->piece_ev[k], c->stream));
        CK(cudaStreamWaitEvent(c->io, c->piece_ev[k], 0));
        CK(cudaMemcpyAsync((unsigned char *)parts + (size_t)p_prev * L.stride, c->d_aos + (size_t)p_prev * L.stride,
                           (size_t)(p_next - p_prev) * L.stride, cudaMemcpyDeviceToHost, c->io));
      }
    }
    d.unit_lo = d.unit_hi = 0;
    d.keep_sums = 0;
    d.no_end_force = 0;
    CK(cudaStreamSynchronize(c->io));
  } else {
    if (do_force) {
      if (launch_force(c)) return 1;
    } else if (do_end && n > 0) {
      k_end_force<<<nb, 256, 0, c->stream>>>(d);
      c->counters.n_kernel_launches += 1;
    }
    CK(cudaEventRecord(c->ev[4], c->stream));
    d.keep_sums = 0;
    d.no_end_force = 0;
    if ((sphenix ? (what | ran) : what) && n > 0) {
      if (sphenix && ran)
        k_unpack_aos_sphenix<<<nb, 256, 0, c->stream>>>(d, c->d_aos, L, order, ran, c->prm.max_active_bin, 0, n);
      else
        k_unpack_aos<<<nb, 256, 0, c->stream>>>(d, c->d_aos, L, order, what, c->prm.max_active_bin, 0, n);
      c->counters.n_kernel_launches += 1;
      CK(cudaGetLastError());
      CK(cudaMemcpyAsync(parts, c->d_aos, (size_t)n * L.stride, cudaMemcpyDeviceToHost, c->stream));
    }
  }
  if (finish_counters(c)) return 1;
  CK(cudaEventElapsedTime(&c->timings.sort_ms, c->ev[0], c->ev[1]));
  CK(cudaEventElapsedTime(&c->timings.density_ms, c->ev[1], c->ev[2]));
  CK(cudaEventElapsedTime(&c->timings.ghost_ms, c->ev[2], c->ev[3]));
  CK(cudaEventElapsedTime(&c->timings.force_ms, c->ev[3], c->ev[4]));
  CK(cudaEventElapsedTime(&c->timings.total_ms, c->ev[0], c->ev[4]));
  if (do_ghost && c->counters.ghost_unconverged > 0)
    return fail("Smoothing length failed to converge on " + std::to_string(c->counters.ghost_unconverged) +
                " particles.");
  return 0;
}

int swift_cuda_get_counters(swift_cuda_ctx *c, swift_cuda_counters *out) {
  if (!c || !out) return fail("get_counters: null argument");
  *out = c->counters;
  return 0;
}
int swift_cuda_get_timings(swift_cuda_ctx *c, swift_cuda_timings *out) {
  if (!c || !out) return fail("get_timings: null argument");
  *out = c->timings;
  return 0;
}

int swift_cuda_get_tree_size(swift_cuda_ctx *c, int *ncells, int *ntasks, int cdim[3]) {
  if (!c) return fail("get_tree_size: null context");
  if (ncells) *ncells = (int)c->tree.cells.size();
  if (ntasks) *ntasks = (int)c->tree.tasks.size();
  if (cdim)
    for (int k = 0; k < 3; k++) cdim[k] = c->tree.cdim[k];
  return 0;
}
int swift_cuda_get_tree(swift_cuda_ctx *c, swift_cuda_cell *cells, swift_cuda_task *tasks, int32_t *order) {
  if (!c) return fail("get_tree: null context");
  if (cells) std::memcpy(cells, c->tree.cells.data(), c->tree.cells.size() * sizeof(swift_cuda_cell));
  if (tasks) std::memcpy(tasks, c->tree.tasks.data(), c->tree.tasks.size() * sizeof(swift_cuda_task));
  if (order) std::memcpy(order, c->tree.order.data(), c->tree.order.size() * sizeof(int32_t));
  return 0;
}

// ---- host-only planning ------------------------------------------------------
int swift_cuda_host_plan(const swift_cuda_params *params, const swift_cuda_parts_in *in, swift_cuda_cell **cells,
                         int *ncells, swift_cuda_task **tasks, int *ntasks, int32_t *order, int cdim[3],
                         long long *n_interactions, long long *n_items) {
  if (!params || !in) return fail("host_plan: null argument");
  Tree tree;
  const int rc = build_tree(*params, in->n, in->x, in->h, tree);
  if (rc) return fail("host_plan: cannot build the top-level grid (code " + std::to_string(rc) + ")");
  std::vector<float> hs(in->n);
  gather_f32(in->h, tree.order.data(), in->n, hs.data());
  compute_cell_hmax(tree, hs.data());
  build_tasks(*params, tree);
  std::vector<uint8_t> act(tree.cells.size(), 1);
  WorkLists wl;
  build_worklists(*params, tree, act, wl);
  if (cells) {
    *cells = (swift_cuda_cell *)malloc(sizeof(swift_cuda_cell) * std::max<size_t>(tree.cells.size(), 1));
    std::memcpy(*cells, tree.cells.data(), sizeof(swift_cuda_cell) * tree.cells.size());
  }
  if (tasks) {
    *tasks = (swift_cuda_task *)malloc(sizeof(swift_cuda_task) * std::max<size_t>(tree.tasks.size(), 1));
    std::memcpy(*tasks, tree.tasks.data(), sizeof(swift_cuda_task) * tree.tasks.size());
  }
  if (ncells) *ncells = (int)tree.cells.size();
  if (ntasks) *ntasks = (int)tree.tasks.size();
  if (order) std::memcpy(order, tree.order.data(), sizeof(int32_t) * tree.order.size());
  if (cdim)
    for (int k = 0; k < 3; k++) cdim[k] = tree.cdim[k];
  if (n_interactions) *n_interactions = wl.n_interactions;
  if (n_items) *n_items = (long long)wl.items.size();
  return 0;
}
void swift_cuda_host_free(void *p) { free(p); }

// FNV-1a over everything the host builder hands to the device for these particles (tree, tasks, order, the two
// item lists, leaves, units, lanes, sort segments, recursion limits): what a test compares between the threaded
// builder and SWIFT_CUDA_HOST_THREADS=1.
int swift_cuda_host_plan_digest(const swift_cuda_params *params, const swift_cuda_parts_in *in,
                                unsigned long long *digest) {
  if (!params || !in || !digest) return fail("host_plan_digest: null argument");
  // SWIFT_CUDA_HOST_TIMING=1: the stages' wall-clock times on stderr (what scripts/host_plan_time.py reads)
  const bool timing = getenv("SWIFT_CUDA_HOST_TIMING") != nullptr;
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto t_last = now();
  auto lap = [&](const char *what) {
    if (!timing) return;
    const auto t = now();
    fprintf(stderr, "host_plan %-16s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t - t_last).count());
    t_last = t;
  };
  Tree tree;
  const int rc = build_tree(*params, in->n, in->x, in->h, tree);
  if (rc) return fail("host_plan_digest: cannot build the top-level grid (code " + std::to_string(rc) + ")");
  lap("build_tree");
  std::vector<float> hs(in->n);
  gather_f32(in->h, tree.order.data(), in->n, hs.data());
  compute_cell_hmax(tree, hs.data());
  lap("cell_hmax");
  build_tasks(*params, tree);
  lap("build_tasks");
  std::vector<uint8_t> act;
  {
    swift_cuda_ctx tmp;
    tmp.prm = *params;
    tmp.tree = tree;
    cells_active(&tmp, in, act);
  }
  lap("cells_active");
  WorkLists wl;
  build_worklists(*params, tree, act, wl);
  lap("build_worklists");
  assign_lanes(tree, in->x, wl);
  lap("assign_lanes");
  unsigned long long hsh = 1469598103934665603ULL;
  auto mix = [&](const void *p, size_t bytes) {
    const unsigned char *b = static_cast<const unsigned char *>(p);
    for (size_t k = 0; k < bytes; k++) hsh = (hsh ^ b[k]) * 1099511628211ULL;
  };
  auto mixv = [&](const auto &v) {
    const unsigned long long n = v.size();
    mix(&n, sizeof(n));
    if (n) mix(v.data(), n * sizeof(v[0]));
  };
  for (const swift_cuda_cell &c : tree.cells) {  // field by field: the struct has padding
    mix(c.loc, sizeof(c.loc)); mix(c.width, sizeof(c.width));
    const int ints[5] = {c.first, c.count, c.parent, c.depth, c.split};
    mix(ints, sizeof(ints)); mix(c.progeny, sizeof(c.progeny));
  }
  for (const swift_cuda_task &t : tree.tasks) {
    const int ints[3] = {t.type, t.ci, t.cj};
    mix(ints, sizeof(ints));
  }
  mixv(tree.order); mixv(tree.h_max);
  for (const std::vector<Item> *lst : {&wl.items, &wl.redo_items})
    for (const Item &it : *lst) {
      const int v[5] = {it.partner, it.own, it.kind, it.sid, it.shift};
      mix(v, sizeof(v));
    }
  mixv(wl.leaves); mixv(wl.item_begin); mixv(wl.redo_begin); mixv(wl.unit_leaf); mixv(wl.unit_begin);
  mixv(wl.lane_part); mixv(wl.sort_cells); mixv(wl.sort_sids); mixv(wl.sort_offset); mixv(wl.sort_lookup);
  mixv(wl.recurse_limit);
  mix(&wl.n_sort_entries, sizeof(wl.n_sort_entries));
  mix(&wl.n_interactions, sizeof(wl.n_interactions));
  *digest = hsh;
  return 0;
}

// ---- pair recording (test hook) ---------------------------------------------
int swift_cuda_record_pairs(swift_cuda_ctx *c, long long capacity) {
  NEED_BUILT(c);
  Dev &d = c->dev;
  if (capacity > 0 && d.rec_pairs[0] == nullptr) {
    for (int k = 0; k < 3; k++)
      if (dmalloc(c, &d.rec_pairs[k], (size_t)capacity * 2)) return 1;
    if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX && dmalloc(c, &d.rec_grad, (size_t)capacity * 2)) return 1;
    c->rec_capacity = capacity;
  }
  d.rec_capacity = capacity > 0 ? c->rec_capacity : 0;
  CK(cudaMemsetAsync(d.rec_count, 0, 4 * sizeof(unsigned long long), c->stream));
  return 0;
}
int swift_cuda_fetch_pairs(swift_cuda_ctx *c, int loop, long long *pairs, long long capacity, long long *count) {
  NEED_BUILT(c);
  if (loop < 0 || loop > 3 || (loop == 3 && c->prm.scheme != SWIFT_CUDA_SCHEME_SPHENIX))
    return fail("fetch_pairs: loop must be 0, 1 or 2 (3: the gradient loop of the SPHENIX scheme)");
  Dev &d = c->dev;
  unsigned long long cnt[4];
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaMemcpy(cnt, d.rec_count, sizeof(cnt), cudaMemcpyDeviceToHost));
  if (count) *count = (long long)cnt[loop];
  if ((long long)cnt[loop] > c->rec_capacity) return fail("fetch_pairs: device pair buffer overflowed");
  if (pairs) {
    const long long m = std::min<long long>(capacity, (long long)cnt[loop]);
    if (m > 0)
      CK(cudaMemcpy(pairs, loop == 3 ? d.rec_grad : d.rec_pairs[loop], (size_t)m * 2 * sizeof(long long),
                    cudaMemcpyDeviceToHost));
  }
  return 0;
}

// Marks (caller order) every particle that lies in a cell some interaction of an ACTIVE leaf
// reads: the set a multi-GPU rank has to keep fresh in its halo.
int swift_cuda_get_touched(swift_cuda_ctx *c, uint8_t *mask) {
  if (!c || !mask) return fail("get_touched: null argument");
  const int64_t n = (int64_t)(c->tree.order.empty() ? c->n : (int64_t)c->tree.order.size());
  std::vector<uint8_t> cell_mark(c->tree.cells.size(), 0);
  for (const Item &it : c->wl.items) cell_mark[it.partner] = cell_mark[it.own] = 1;
  for (const Item &it : c->wl.redo_items) cell_mark[it.partner] = cell_mark[it.own] = 1;
  std::memset(mask, 0, (size_t)n);
  for (size_t k = 0; k < cell_mark.size(); k++) {
    if (!cell_mark[k]) continue;
    const swift_cuda_cell &cell = c->tree.cells[k];
    for (int j = 0; j < cell.count; j++) {
      const int64_t slot = (int64_t)cell.first + j;
      mask[c->tree.order.empty() ? slot : c->tree.order[slot]] = 1;
    }
  }
  return 0;
}

// c->hydro.h_max of every cell from the device copy of h: after_ghost = 0 sets it (and h_max_old)
// from scratch (after an xv halo refresh), 1 raises it to cover the current h (after a rho refresh).
int swift_cuda_refresh_cell_hmax(swift_cuda_ctx *c, int after_ghost) {
  NEED_BUILT(c);
  Dev &d = c->dev;
  if (d.ncells > 0) {
    const int nb = (d.ncells * 32 + 255) / 256;
    if (after_ghost)
      k_cell_hmax<<<nb, 256, 0, c->stream>>>(d);
    else
      k_cell_hmax_init<<<nb, 256, 0, c->stream>>>(d, d.cell_hmax, const_cast<float *>(d.cell_hmax_old));
  }
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

// ---- halo buffers (multi-GPU) -------------------------------------------------
}  // extern "C"

namespace swiftcuda {
// phase 0 "xv": x, y, z (f64), vx, vy, vz, m, h (f32) = 44 B; phase 1 "rho": h, rho, P, f, c, balsara = 24 B
struct HaloXV {
  double x, y, z;
  float vx, vy, vz, m, h;
  float pad;
};
struct HaloRho {
  float h, rho, P, f, c, bal;
};
__global__ void k_halo_pack_xv(Dev d, const int32_t *idx, int64_t n, HaloXV *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t p = idx[k];
  const float4 vm = d.vm[p];
  const PosD q = d.pos[p];
  HaloXV r = {q.x, q.y, q.z, vm.x, vm.y, vm.z, vm.w, d.h[p], 0.f};
  buf[k] = r;
}
__global__ void k_halo_unpack_xv(Dev d, const int32_t *idx, int64_t n, const HaloXV *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t p = idx[k];
  const HaloXV r = buf[k];
  PosD q;
  q.x = r.x; q.y = r.y; q.z = r.z; q.pad = 0.0;
  const_cast<PosD *>(d.pos)[p] = q;
  d.vm[p] = make_float4(r.vx, r.vy, r.vz, r.m);
  d.h[p] = r.h;
}
__global__ void k_halo_pack_rho(Dev d, const int32_t *idx, int64_t n, HaloRho *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t p = idx[k];
  const float4 f = d.fin[p];
  HaloRho r = {d.h[p], f.x, f.y, f.z, f.w, d.hb[p].y};
  buf[k] = r;
}
__global__ void k_halo_unpack_rho(Dev d, const int32_t *idx, int64_t n, const HaloRho *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t p = idx[k];
  const HaloRho r = buf[k];
  d.h[p] = r.h;
  d.fin[p] = make_float4(r.rho, r.P, r.f, r.c);
  d.hb[p] = make_float4(r.h, r.bal, __int_as_float((int)d.time_bin[p] & kBinMask), r.P / (r.rho * r.rho));
}
// SPHENIX: the rho halo also carries what a gradient-loop neighbour is read for (u, the viscosity alpha of
// BEFORE the extra ghost); a third exchange after the extra ghost carries the evolved alphas the force loop
// reads (the reference's send_rho / send_gradient, src/engine_maketasks.c:141-324 under EXTRA_HYDRO_LOOP)
struct HaloRhoS {
  float h, rho, P, f, c, bal, u, alpha_visc;
};
struct HaloGrad {
  float alpha_visc, alpha_diff;
};
__global__ void k_halo_pack_rho_sphenix(Dev d, const int32_t *idx, int64_t n, HaloRhoS *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t p = idx[k];
  const float4 f = d.fin[p];
  HaloRhoS r = {d.h[p], f.x, f.y, f.z, f.w, d.hb[p].y, d.u[p], d.sph_state[p].x};
  buf[k] = r;
}
__global__ void k_halo_unpack_rho_sphenix(Dev d, const int32_t *idx, int64_t n, const HaloRhoS *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t p = idx[k];
  const HaloRhoS r = buf[k];
  d.h[p] = r.h;
  d.fin[p] = make_float4(r.rho, r.P, r.f, r.c);
  d.hb[p] = make_float4(r.h, r.bal, __int_as_float((int)d.time_bin[p] & kBinMask), r.P / (r.rho * r.rho));
  const_cast<float *>(d.u)[p] = r.u;
  d.gin[p] = make_float4(r.u, r.rho, r.c, r.alpha_visc);
}
__global__ void k_halo_pack_grad(Dev d, const int32_t *idx, int64_t n, HaloGrad *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const float4 st = d.sph_state[idx[k]];
  HaloGrad r = {st.x, st.y};
  buf[k] = r;
}
__global__ void k_halo_unpack_grad(Dev d, const int32_t *idx, int64_t n, const HaloGrad *buf) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t p = idx[k];
  const HaloGrad r = buf[k];
  float4 st = d.sph_state[p];
  st.x = r.alpha_visc;
  st.y = r.alpha_diff;
  d.sph_state[p] = st;
  d.sx[p] = make_float4(r.alpha_visc, r.alpha_diff, d.u[p], 0.f);
}
}  // namespace swiftcuda

// which halo phases a context's scheme has: 0 xv and 1 rho (two-loop schemes); 0 xv, 3 rho and 2 gradient (SPHENIX)
static int halo_phase_ok(const swift_cuda_ctx *c, int phase) {
  const bool sphenix = c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX;
  if (phase == 0 || (sphenix ? (phase == 2 || phase == 3) : phase == 1)) return 0;
  return fail(sphenix ? "halo: a SPHENIX context exchanges phases 0 (xv), 3 (rho, u, alpha) and 2 (alphas after the "
                        "extra ghost)"
                      : "halo: phases 0 (xv) and 1 (rho)");
}

extern "C" {

size_t swift_cuda_halo_bytes_per_particle(int phase) {
  switch (phase) {
    case 0: return sizeof(HaloXV);
    case 1: return sizeof(HaloRho);
    case 2: return sizeof(HaloGrad);
    case 3: return sizeof(HaloRhoS);
    default: return 0;
  }
}

int swift_cuda_halo_pack(swift_cuda_ctx *c, int phase, const int32_t *d_index, int64_t count, void *d_buffer,
                         void *stream) {
  NEED_BUILT(c);
  if (halo_phase_ok(c, phase)) return 1;
  if (count <= 0) return 0;
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  const int nb = (int)((count + 255) / 256);
  if (phase == 0)
    k_halo_pack_xv<<<nb, 256, 0, s>>>(c->dev, d_index, count, (HaloXV *)d_buffer);
  else if (phase == 1)
    k_halo_pack_rho<<<nb, 256, 0, s>>>(c->dev, d_index, count, (HaloRho *)d_buffer);
  else if (phase == 2)
    k_halo_pack_grad<<<nb, 256, 0, s>>>(c->dev, d_index, count, (HaloGrad *)d_buffer);
  else
    k_halo_pack_rho_sphenix<<<nb, 256, 0, s>>>(c->dev, d_index, count, (HaloRhoS *)d_buffer);
  CK(cudaGetLastError());
  return 0;
}
int swift_cuda_halo_unpack(swift_cuda_ctx *c, int phase, const int32_t *d_index, int64_t count,
                           const void *d_buffer, void *stream) {
  NEED_BUILT(c);
  if (halo_phase_ok(c, phase)) return 1;
  if (count <= 0) return 0;
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  const int nb = (int)((count + 255) / 256);
  if (phase == 0)
    k_halo_unpack_xv<<<nb, 256, 0, s>>>(c->dev, d_index, count, (const HaloXV *)d_buffer);
  else if (phase == 1)
    k_halo_unpack_rho<<<nb, 256, 0, s>>>(c->dev, d_index, count, (const HaloRho *)d_buffer);
  else if (phase == 2)
    k_halo_unpack_grad<<<nb, 256, 0, s>>>(c->dev, d_index, count, (const HaloGrad *)d_buffer);
  else
    k_halo_unpack_rho_sphenix<<<nb, 256, 0, s>>>(c->dev, d_index, count, (const HaloRhoS *)d_buffer);
  CK(cudaGetLastError());
  return 0;
}

// ---- time integration (integrate.cuh) ---------------------
#define NEED_INTEG(c)  \
  NEED_BUILT(c);       \
  if (!(c)->integ_on) return fail("swift_cuda_integrator_init has not been called on this context");

int swift_cuda_integrator_init(swift_cuda_ctx *c, const swift_cuda_integrator *p) {
  NEED_BUILT(c);
  if (!p) return fail("integrator_init: null argument");
  if (!c->all_cells_active)
    return fail("integrator_init: build the context with every particle active (max_active_bin = 56), so that "
                "every cell is in the work lists; the active set then changes per step on the device");
  if (!(p->time_base > 0.0)) return fail("integrator_init: time_base must be positive");
  IntegDev &g = c->integ;
  const int64_t n = c->n;
  if (g.vu_full == nullptr &&  // a second call on the same build re-initialises in place
      (dmalloc(c, &g.vu_full, (size_t)n) || dmalloc(c, &g.xdiff, (size_t)n) || dmalloc(c, &g.xdiff_sort, (size_t)n) ||
      dmalloc(c, &g.dx2_max, 2) || dmalloc(c, &g.ti_end_min, 1) || dmalloc(c, &g.updated, 1) ||
      dmalloc(c, &g.too_small, 1) || dmalloc(c, &c->d_sort_need, (size_t)c->dev.ncells * 13) ||
      dmalloc(c, &c->d_unit_list, (size_t)std::max(c->dev.n_units, 1)) || dmalloc(c, &c->d_unit_count, 1)))
    return 1;
  c->iprm.cfl = p->cfl_condition;
  c->iprm.log_max_h_change = p->log_max_h_change;
  c->iprm.min_u = (float)((double)p->minimal_internal_energy / 1.0);  // / cosmo->a_factor_internal_energy
  c->iprm.dt_min = p->dt_min;
  c->iprm.dt_max = p->dt_max;
  c->iprm.time_base = p->time_base;
  c->iprm.time_base_inv = 1.0 / p->time_base;
  c->iprm.ti_current = 0;
  if (c->prm.scheme == SWIFT_CUDA_SCHEME_SPHENIX) {
    // the scheme's drift / kick / time-step hooks are Minimal's arithmetic (src/hydro/SPHENIX/hydro.h:464-480,
    // 966-1146 with the floors NONE; v_sig lives in p->viscosity); the extra ghost's dt_alpha is
    // get_timestep(p->time_bin, e->time_base) of the same time line
    c->sprops.time_base = p->time_base;
    if (upload_consts(c)) return 1;
  }
  if (n > 0) k_integ_init<<<(int)((n + 255) / 256), 256, 0, c->stream>>>(c->dev, g);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  c->integ_on = true;
  c->marks_fresh = false;
  return 0;
}

int swift_cuda_set_time(swift_cuda_ctx *c, long long ti_current, int max_active_bin) {
  NEED_INTEG(c);
  c->iprm.ti_current = ti_current;
  c->prm.max_active_bin = max_active_bin;
  c->marks_fresh = false;
  if (upload_consts(c)) return 1;
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int swift_cuda_drift(swift_cuda_ctx *c, long long ti_old, float dx_max[2]) {
  NEED_INTEG(c);
  if (ti_old > c->iprm.ti_current) return fail("drift: attempt to drift to the past");
  Dev &d = c->dev;
  const double dt = (double)(c->iprm.ti_current - ti_old) * c->iprm.time_base;
  CK(cudaMemsetAsync(c->integ.dx2_max, 0, 2 * sizeof(unsigned), c->stream));
  CK(cudaEventRecord(c->ev[0], c->stream));
  if (c->n > 0) k_drift<<<(int)((c->n + 255) / 256), 256, 0, c->stream>>>(d, c->integ, c->iprm, dt);
  // c->hydro.h_max follows the particles (cell_drift.c:318, 367)
  if (d.ncells > 0)
    k_cell_hmax_init<<<(d.ncells * 32 + 255) / 256, 256, 0, c->stream>>>(d, d.cell_hmax,
                                                                         const_cast<float *>(d.cell_hmax_old));
  CK(cudaGetLastError());
  CK(cudaEventRecord(c->ev[5], c->stream));
  float dx2[2] = {0.f, 0.f};
  CK(cudaMemcpyAsync(dx2, c->integ.dx2_max, sizeof(dx2), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaEventElapsedTime(&c->integ_ms[0], c->ev[0], c->ev[5]));
  c->dx_max_part = sqrtf(dx2[0]);
  if (dx_max) {
    dx_max[0] = sqrtf(dx2[0]);  // cell_drift.c:359-360
    dx_max[1] = sqrtf(dx2[1]);
  }
  return 0;
}

int swift_cuda_kick(swift_cuda_ctx *c, int which) {
  NEED_INTEG(c);
  if (which != 1 && which != 2) return fail("kick: which must be 1 (kick1) or 2 (kick2)");
  CK(cudaEventRecord(c->ev[0], c->stream));
  if (c->n > 0) k_kick<<<(int)((c->n + 255) / 256), 256, 0, c->stream>>>(c->dev, c->integ, c->iprm, which);
  CK(cudaGetLastError());
  CK(cudaEventRecord(c->ev[5], c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaEventElapsedTime(&c->integ_ms[which], c->ev[0], c->ev[5]));
  return 0;
}

int swift_cuda_timestep(swift_cuda_ctx *c, long long *ti_end_min, long long *updated) {
  NEED_INTEG(c);
  const unsigned long long big = (unsigned long long)kMaxNrTimesteps;
  CK(cudaMemcpyAsync(c->integ.ti_end_min, &big, sizeof(big), cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(c->integ.updated, 0, sizeof(unsigned long long), c->stream));
  CK(cudaMemsetAsync(c->integ.too_small, 0, sizeof(int), c->stream));
  CK(cudaEventRecord(c->ev[0], c->stream));
  c->marks_fresh = false;  // time bins change
  if (c->n > 0) k_timestep<<<(int)((c->n + 255) / 256), 256, 0, c->stream>>>(c->dev, c->integ, c->iprm);
  CK(cudaGetLastError());
  CK(cudaEventRecord(c->ev[5], c->stream));
  unsigned long long t = 0, u = 0;
  int small = 0;
  CK(cudaMemcpyAsync(&t, c->integ.ti_end_min, sizeof(t), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(&u, c->integ.updated, sizeof(u), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(&small, c->integ.too_small, sizeof(small), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaEventElapsedTime(&c->integ_ms[3], c->ev[0], c->ev[5]));
  if (ti_end_min) *ti_end_min = (long long)t;
  if (updated) *updated = (long long)u;
  if (small > 0)
    return fail(std::to_string(small) + " particle(s) want a time-step below dt_min (" +
                std::to_string(c->iprm.dt_min) + ")");
  return 0;
}

// ---- table equations of state: include/swift_cuda.h, "Table equations of state" (loaders: eos_loader.cpp) ----
int swift_cuda_set_eos_table(swift_cuda_ctx *c, const swift_cuda_eos_table *t) {
  if (!c || !t) return fail("set_eos_table: null argument");
  if (c->prm.scheme != SWIFT_CUDA_SCHEME_PLANETARY) return fail("set_eos_table: table materials exist in the Planetary scheme only");
  if (t->kind != kEosTableSesame && t->kind != kEosTableHM80) return fail("set_eos_table: kind must be 1 (SESAME-style) or 2 (HM80)");
  if (t->mat_id < 200) return fail("set_eos_table: mat_id below 200 is an ideal gas or a Tillotson material");
  if (t->num_rho < 2 || t->num_T < 2) return fail("set_eos_table: fewer than two densities or energies");
  if (!t->P || (t->kind == kEosTableSesame && (!t->log_rho || !t->log_u || !t->c))) return fail("set_eos_table: null array");
  CK(cudaSetDevice(c->device));
  int slot = c->n_tab;
  for (int k = 0; k < c->n_tab; k++)
    if (c->tabs[k].mat_id == t->mat_id) slot = k;  // a second table for a material replaces the first
  if (slot >= 8) return fail("set_eos_table: at most 8 tables per context");
  const size_t n2 = (size_t)t->num_rho * t->num_T;
  auto up = [&](const float *src, size_t count, const float **dst) -> int {
    float *p = nullptr;
    if (cudaMalloc(&p, count * sizeof(float)) != cudaSuccess) return fail("set_eos_table: out of device memory");
    c->eos_allocs.push_back(p);
    if (cudaMemcpy(p, src, count * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess)
      return fail("set_eos_table: copy to the device failed");
    *dst = p;
    return 0;
  };
  EosTable e = {};
  e.kind = t->kind;
  e.mat_id = t->mat_id;
  e.num_rho = t->num_rho;
  e.num_T = t->num_T;
  if (up(t->P, n2, &e.P)) return 1;
  if (t->kind == kEosTableSesame) {
    if (up(t->log_rho, (size_t)t->num_rho, &e.log_rho) || up(t->log_u, n2, &e.log_u) || up(t->c, n2, &e.c)) return 1;
  }
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
  c->tabs[slot] = e;
  if (slot == c->n_tab) c->n_tab++;
  if (upload_consts(c)) return 1;
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

// ---- the time-step limiter (swift --limiter): include/swift_cuda.h, "Time-step limiter" ----------------
int swift_cuda_do_limiter_loop(swift_cuda_ctx *c, const signed char *time_bin, signed char *wakeup) {
  NEED_BUILT(c);
  Dev &d = c->dev;
  const int64_t n = c->n;
  if (d.wakeup == nullptr && (dmalloc(c, &d.wakeup, (size_t)std::max<int64_t>(n, 1)) ||
                              dmalloc(c, &c->d_lim_tmp, (size_t)std::max<int64_t>(n, 1))))
    return 1;
  const int nb = (int)((n + 255) / 256);
  const int32_t *order = c->tree.order.empty() ? nullptr : c->s_order;
  if (time_bin != nullptr) {
    // a caller that runs the time-step task itself hands the new bins over (caller order, all n)
    if (c->integ_on) return fail("do_limiter_loop: in the time-integration mode the bins are those of swift_cuda_timestep");
    if (c->n_io >= 0 && c->n_io < n) return fail("do_limiter_loop: not with a reduced transfer count (halo copies)");
    CK(cudaMemcpyAsync(c->d_lim_tmp, time_bin, (size_t)n, cudaMemcpyHostToDevice, c->stream));
    if (n > 0) k_set_bins<<<nb, 256, 0, c->stream>>>(n, order, c->d_lim_tmp, const_cast<int8_t *>(d.time_bin), d.hb);
    c->marks_fresh = false;
  }
  // the units that hold a starting particle: those that held an active one (a particle that was active gets a
  // bin that starts now, make_integer_timestep; the others keep theirs); after swift_cuda_timestep the marks
  // are taken again on the new bins
  if (launch_mark(c)) return 1;
  if (n > 0) k_wakeup_reset<<<nb, 256, 0, c->stream>>>(n, d.wakeup);
  CK(cudaMemsetAsync(d.work_ctr + 2 + c->prm.max_smoothing_iterations, 0, sizeof(int), c->stream));
  CK(cudaEventRecord(c->ev[0], c->stream));
  if (d.n_units > 0) k_density<0, false, 2><<<c->grid_limiter, kBlock, 0, c->stream>>>(d, 0);
  CK(cudaEventRecord(c->ev[5], c->stream));
  c->counters.n_kernel_launches += 1;
  CK(cudaGetLastError());
  if (wakeup != nullptr) {
    if (n > 0) k_scatter_wakeup<<<nb, 256, 0, c->stream>>>(n, order, d.wakeup, c->d_lim_tmp);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(wakeup, c->d_lim_tmp, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaEventElapsedTime(&c->limiter_ms[0], c->ev[0], c->ev[5]));
  return 0;
}

int swift_cuda_do_limiter(swift_cuda_ctx *c, long long *ti_end_min, long long *limited) {
  NEED_INTEG(c);
  Dev &d = c->dev;
  if (d.wakeup == nullptr) return fail("do_limiter: swift_cuda_do_limiter_loop has not run");
  const unsigned long long big = (unsigned long long)kMaxNrTimesteps;
  CK(cudaMemcpyAsync(c->integ.ti_end_min, &big, sizeof(big), cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(c->integ.updated, 0, sizeof(unsigned long long), c->stream));
  CK(cudaEventRecord(c->ev[0], c->stream));
  c->marks_fresh = false;  // time bins change
  if (c->n > 0) k_limiter<<<(int)((c->n + 255) / 256), 256, 0, c->stream>>>(d, c->integ, c->iprm);
  CK(cudaGetLastError());
  CK(cudaEventRecord(c->ev[5], c->stream));
  unsigned long long t = 0, u = 0;
  CK(cudaMemcpyAsync(&t, c->integ.ti_end_min, sizeof(t), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(&u, c->integ.updated, sizeof(u), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaEventElapsedTime(&c->limiter_ms[1], c->ev[0], c->ev[5]));
  if (ti_end_min) *ti_end_min = (long long)t;
  if (limited) *limited = (long long)u;
  return 0;
}

int swift_cuda_get_limiter_timings(swift_cuda_ctx *c, float ms[2]) {
  NEED_BUILT(c);
  if (!ms) return fail("get_limiter_timings: null argument");
  ms[0] = c->limiter_ms[0];
  ms[1] = c->limiter_ms[1];
  return 0;
}

int swift_cuda_get_integrator_timings(swift_cuda_ctx *c, float ms[4]) {
  NEED_INTEG(c);
  if (!ms) return fail("get_integrator_timings: null argument");
  for (int k = 0; k < 4; k++) ms[k] = c->integ_ms[k];
  return 0;
}

int swift_cuda_reset_offsets(swift_cuda_ctx *c, int what) {
  NEED_INTEG(c);
  CK(cudaMemsetAsync(c->integ.xdiff_sort, 0, (size_t)c->n * sizeof(float4), c->stream));
  if (what >= 2) CK(cudaMemsetAsync(c->integ.xdiff, 0, (size_t)c->n * sizeof(float4), c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int swift_cuda_get_state(swift_cuda_ctx *c, const swift_cuda_state_out *o) {
  NEED_INTEG(c);
  if (!o) return fail("get_state: null argument");
  const int64_t n = c->n;
  if (n == 0) return 0;
  Dev &d = c->dev;
  const IntegDev &g = c->integ;
  void *tmp = nullptr;  // inspection path: one scratch buffer, fields one after the other
  CK(cudaMalloc(&tmp, (size_t)n * 3 * sizeof(double)));
  const int32_t *order = c->tree.order.empty() ? nullptr : c->s_order;
  const int nb = (int)((n + 255) / 256);
  int rc = 0;
  auto f32 = [&](float *host, const void *src, int stride, int width) {
    if (!host || rc) return;
    k_scatter_f32<<<nb, 256, 0, c->stream>>>(n, order, (const float *)src, stride, width, (float *)tmp);
    if (cudaMemcpyAsync(host, tmp, (size_t)n * width * sizeof(float), cudaMemcpyDeviceToHost, c->stream) !=
            cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess)
      rc = 1;
  };
  if (o->x) {
    k_scatter_pos<<<nb, 256, 0, c->stream>>>(n, order, d.pos, (double *)tmp);
    if (cudaMemcpyAsync(o->x, tmp, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream) !=
            cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess)
      rc = 1;
  }
  f32(o->v, d.vm, 4, 3);
  f32(o->v_full, g.vu_full, 4, 3);
  f32(o->u, d.u, 1, 1);
  f32(o->u_full, (const float *)g.vu_full + 3, 4, 1);
  f32(o->h, d.h, 1, 1);
  f32(o->rho, d.fin, 4, 1);
  f32(o->pressure, (const float *)d.fin + 1, 4, 1);
  f32(o->soundspeed, (const float *)d.fin + 3, 4, 1);
  f32(o->v_sig, d.v_sig, 1, 1);
  f32(o->a, d.acc, 4, 3);
  f32(o->u_dt, (const float *)d.acc + 3, 4, 1);
  f32(o->h_dt, d.h_dt, 1, 1);
  f32(o->x_diff, g.xdiff, 4, 3);
  f32(o->x_diff_sort, g.xdiff_sort, 4, 3);
  if (!rc && (o->time_bin || o->min_ngb_time_bin)) {
    int8_t *tb = (int8_t *)tmp, *mb = tb + n;
    k_scatter_bins<<<nb, 256, 0, c->stream>>>(n, order, d.time_bin, d.min_bin, tb, mb);
    if (o->time_bin && cudaMemcpyAsync(o->time_bin, tb, (size_t)n, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess)
      rc = 1;
    if (o->min_ngb_time_bin &&
        cudaMemcpyAsync(o->min_ngb_time_bin, mb, (size_t)n, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess)
      rc = 1;
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) rc = 1;
  }
  const cudaError_t e = cudaGetLastError();
  cudaFree(tmp);
  if (rc || e != cudaSuccess) return fail(std::string("get_state: ") + cudaGetErrorString(e));
  return 0;
}

}  // extern "C"

// ---- SPHENIX: props, carried state, results (include/swift_cuda.h, "SPHENIX") -----------------------
extern "C" {

int swift_cuda_sphenix_set_props(swift_cuda_ctx *c, const swift_cuda_sphenix_props *p) {
  if (!c || !p) return fail("sphenix_set_props: null argument");
  if (c->prm.scheme != SWIFT_CUDA_SCHEME_SPHENIX) return fail("sphenix_set_props: the context's scheme is not SPHENIX");
  // (no range checks: the reference's own unit tests run with e->time_base = 0, i.e. dt_alpha = 0)
  CK(cudaSetDevice(c->device));
  c->sprops = *p;
  if (c->built) {  // the constants of a built context follow at once; the particles' state is left alone
    if (upload_consts(c)) return 1;
    CK(cudaStreamSynchronize(c->stream));
  }
  return 0;
}

int swift_cuda_sphenix_set_state(swift_cuda_ctx *c, const swift_cuda_sphenix_in *in) {
  NEED_BUILT(c);
  if (!in) return fail("sphenix_set_state: null argument");
  if (c->prm.scheme != SWIFT_CUDA_SCHEME_SPHENIX) return fail("sphenix_set_state: the context's scheme is not SPHENIX");
  const int64_t n = c->n;
  const float *src[3] = {in->alpha_visc, in->alpha_diff, in->div_v_previous_step};
  const float *dev[3] = {nullptr, nullptr, nullptr};
  for (int k = 0; k < 3; k++)
    if (src[k]) {
      if (!c->s_sph_in[k] && dmalloc(c, &c->s_sph_in[k], (size_t)n)) return 1;
      CK(cudaMemcpyAsync(c->s_sph_in[k], src[k], (size_t)n * sizeof(float), cudaMemcpyHostToDevice, c->stream));
      dev[k] = c->s_sph_in[k];
    }
  if (n > 0)
    k_sphenix_set_state<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(
        c->dev, c->tree.order.empty() ? nullptr : c->s_order, dev[0], dev[1], dev[2]);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int swift_cuda_sphenix_get_results(swift_cuda_ctx *c, const swift_cuda_sphenix_out *out) {
  NEED_BUILT(c);
  if (!out) return fail("sphenix_get_results: null argument");
  if (c->prm.scheme != SWIFT_CUDA_SCHEME_SPHENIX) return fail("sphenix_get_results: the context's scheme is not SPHENIX");
  const int64_t n = c->n;
  if (!c->s_sph_out && dmalloc(c, &c->s_sph_out, (size_t)n * 7)) return 1;
  if (n > 0)
    k_sphenix_unpack<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(
        c->dev, c->tree.order.empty() ? nullptr : c->s_order, c->s_sph_out);
  CK(cudaGetLastError());
  float *dst[7] = {out->div_v, out->laplace_u, out->alpha_visc_max_ngb, out->alpha_visc, out->alpha_diff,
                   out->div_v_previous_step, out->div_v_dt};
  for (int k = 0; k < 7; k++)
    if (dst[k])
      CK(cudaMemcpyAsync(dst[k], c->s_sph_out + (size_t)k * n, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost,
                         c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int swift_cuda_sphenix_get_counter(swift_cuda_ctx *c, long long *n_gradient) {
  if (!c || !n_gradient) return fail("sphenix_get_counter: null argument");
  *n_gradient = c->n_gradient;
  return 0;
}

}  // extern "C"


// ---- allocation-free time-step entry (include/swift_cuda.h: swift_cuda_timestep_static) --------------
// No counterpart in the reference.  The arithmetic of
// k_timestep (integrate.cuh; runner_do_timestep, src/runner_time_integration.c:637-1000, get_part_timestep
// and make_integer_timestep, src/timestep.h:45-82, 141-222) on the caller's arrays, with a two-level
// reduction whose scratch lives in a file-scope device array instead of a cudaMalloc'ed buffer.
#include <mutex>

namespace swiftcuda {
constexpr int kTsBlock = 256;
constexpr int kTsMaxBlocks = 65536;
static_assert((long long)kTsBlock * kTsMaxBlocks == SWIFT_CUDA_TIMESTEP_STATIC_MAX, "header and kernel disagree");
struct TsPartial {  // one per block
  unsigned long long ti_end_min;
  unsigned updated, too_small;
};
struct TsHeader {  // the first 256 bytes of the scratch: zeroed before every launch
  unsigned long long ti_end_min, updated;
  unsigned too_small, ticket;
};
constexpr size_t kTsHeaderBytes = 256;
constexpr size_t kTsScratchBytes = kTsHeaderBytes + sizeof(TsPartial) * (size_t)kTsMaxBlocks;  // 1 048 832
// In HOST code this name is a stand-in, not a device address: the entry below asks the runtime for the
// address (cudaGetSymbolAddress, once per device) and hands that to the kernel.
static __device__ __align__(256) unsigned char g_ts_scratch[kTsScratchBytes];

struct TsArgs {
  int64_t n;
  const float *h, *v_sig, *h_dt;
  const int8_t *min_bin;
  int8_t *time_bin;
  long long ti_current;
  int max_active_bin;
  float gamma, cfl, log_max_h_change;
  double dt_min, dt_max, time_base_inv;
};

__global__ void __launch_bounds__(kTsBlock) k_timestep_static(TsArgs a, unsigned char *scratch) {
  TsHeader *const hdr = reinterpret_cast<TsHeader *>(scratch);
  TsPartial *const part = reinterpret_cast<TsPartial *>(scratch + kTsHeaderBytes);
  const int64_t p = blockIdx.x * (int64_t)kTsBlock + threadIdx.x;
  unsigned long long t_end = (unsigned long long)kMaxNrTimesteps;
  unsigned upd = 0, small = 0;
  if (p < a.n) {
    const int old_bin = a.time_bin[p];
    if (old_bin > a.max_active_bin) {
      t_end = (unsigned long long)int_time_end(a.ti_current, old_bin);
    } else {
      const float h = a.h[p];
      const float front = __fmul_rn(__fmul_rn(2.f, a.gamma), a.cfl);
      const float dt_cfl = (float)__ddiv_rn(__dmul_rn((double)front, (double)h), (double)a.v_sig[p]);
      const float h_dt = a.h_dt[p];
      const float dt_h_change = (h_dt != 0.0f) ? fabsf(__fmul_rn(a.log_max_h_change, h) / h_dt) : 3.402823466e+38f;
      float new_dt = dt_cfl < dt_h_change ? dt_cfl : dt_h_change;
      new_dt = (double)new_dt < a.dt_max ? new_dt : (float)a.dt_max;
      if ((double)new_dt < a.dt_min) small = 1;
      long long new_dti = (long long)__dmul_rn((double)new_dt, a.time_base_inv);
      int new_bin = (signed char)time_bin_of(new_dti);
      const int ngb = (int)a.min_bin[p] + kMaxDeltaBin;
      new_bin = (signed char)(new_bin < ngb ? new_bin : ngb);
      new_dti = int_timestep(new_bin);
      const long long current_dti = int_timestep(old_bin);
      const long long ti_end = int_time_end(a.ti_current, old_bin);
      if (old_bin > 0) new_dti = new_dti < 2 * current_dti ? new_dti : 2 * current_dti;
      long long dti_timeline = kMaxNrTimesteps;
      while (new_dti < dti_timeline) dti_timeline /= 2LL;
      new_dti = dti_timeline;
      if (new_dti > current_dti) {
        if ((kMaxNrTimesteps - ti_end) % new_dti > 0) new_dti = current_dti;
      }
      a.time_bin[p] = (int8_t)time_bin_of(new_dti);
      upd = 1;
      t_end = (unsigned long long)(a.ti_current + new_dti);
    }
  }
  // level 1: the block's minimum and counts -> its slot of the scratch
  __shared__ unsigned long long s_t[kTsBlock / 32];
  __shared__ unsigned s_u[kTsBlock / 32], s_s[kTsBlock / 32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, t_end, o);
    t_end = other < t_end ? other : t_end;
  }
  upd = __popc(__ballot_sync(0xffffffffu, upd));
  small = __popc(__ballot_sync(0xffffffffu, small));
  if (lane == 0) { s_t[warp] = t_end; s_u[warp] = upd; s_s[warp] = small; }
  __syncthreads();
  if (threadIdx.x == 0) {
    TsPartial r = {s_t[0], s_u[0], s_s[0]};
    for (int w = 1; w < kTsBlock / 32; w++) {
      r.ti_end_min = s_t[w] < r.ti_end_min ? s_t[w] : r.ti_end_min;
      r.updated += s_u[w];
      r.too_small += s_s[w];
    }
    part[blockIdx.x] = r;
    __threadfence();  // the slot is visible before the ticket is taken
    s_last = atomicAdd(&hdr->ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  // level 2: the last block to finish folds the slots of all blocks
  __threadfence();
  unsigned long long tmin = (unsigned long long)kMaxNrTimesteps, nu = 0;
  unsigned ns = 0;
  for (unsigned b = threadIdx.x; b < gridDim.x; b += kTsBlock) {
    const volatile TsPartial *q = part + b;
    const unsigned long long t = q->ti_end_min;
    tmin = t < tmin ? t : tmin;
    nu += q->updated;
    ns += q->too_small;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, tmin, o);
    tmin = other < tmin ? other : tmin;
    nu += __shfl_xor_sync(0xffffffffu, nu, o);
    ns += __shfl_xor_sync(0xffffffffu, ns, o);
  }
  __shared__ unsigned long long s_n[kTsBlock / 32];
  if (lane == 0) { s_t[warp] = tmin; s_n[warp] = nu; s_s[warp] = ns; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kTsBlock / 32; w++) {
      tmin = s_t[w] < tmin ? s_t[w] : tmin;
      nu += s_n[w];
      ns += s_s[w];
    }
    hdr->ti_end_min = tmin;
    hdr->updated = nu;
    hdr->too_small = ns;
    hdr->ticket = 0;
  }
}
}  // namespace swiftcuda

// Device address of g_ts_scratch, resolved once per device with that device current.
static std::mutex g_ts_mu;
static unsigned char *g_ts_ptr[64] = {};
static int ts_scratch_of(int device, unsigned char **out) {
  std::lock_guard<std::mutex> lock(g_ts_mu);
  if (g_ts_ptr[device] == nullptr) {
    void *p = nullptr;
    CK(cudaGetSymbolAddress(&p, swiftcuda::g_ts_scratch));
    if (p == nullptr) return fail("timestep_static: cudaGetSymbolAddress gave a null address");
    g_ts_ptr[device] = static_cast<unsigned char *>(p);
  }
  *out = g_ts_ptr[device];
  return 0;
}
static int ts_check_ptr(const void *p, const char *name, int device) {
  if (p == nullptr) return fail(std::string("timestep_static: ") + name + " is null");
  cudaPointerAttributes at;
  CK(cudaPointerGetAttributes(&at, p));
  if ((at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged) || (at.type == cudaMemoryTypeDevice && at.device != device))
    return fail(std::string("timestep_static: ") + name + " is not a device pointer of GPU " + std::to_string(device));
  return 0;
}

extern "C" int swift_cuda_timestep_static(int device, const swift_cuda_timestep_static_args *a, long long *ti_end_min,
                                          long long *updated, void *stream) {
  using namespace swiftcuda;
  if (!a) return fail("timestep_static: null argument");
  if (a->count < 0) return fail("timestep_static: negative count");
  if (a->count > SWIFT_CUDA_TIMESTEP_STATIC_MAX)
    return fail("timestep_static: count " + std::to_string((long long)a->count) + " exceeds the " +
                std::to_string((long long)SWIFT_CUDA_TIMESTEP_STATIC_MAX) +
                " particles the library's static scratch array (" + std::to_string(kTsScratchBytes) +
                " bytes) covers; use swift_cuda_timestep on a context, or split the call");
  if (!(a->integ.time_base > 0.0)) return fail("timestep_static: time_base must be positive");
  if (device < 0 || device >= 64) return fail("timestep_static: device out of range");
  CK(cudaSetDevice(device));
  unsigned long long t = (unsigned long long)kMaxNrTimesteps, u = 0;
  if (a->count > 0) {
    if (ts_check_ptr(a->h, "h", device) || ts_check_ptr(a->v_sig, "v_sig", device) ||
        ts_check_ptr(a->h_dt, "h_dt", device) || ts_check_ptr(a->min_ngb_time_bin, "min_ngb_time_bin", device) ||
        ts_check_ptr(a->time_bin, "time_bin", device))
      return 1;
    unsigned char *scratch = nullptr;
    if (ts_scratch_of(device, &scratch)) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    TsArgs k;
    k.n = a->count;
    k.h = a->h; k.v_sig = a->v_sig; k.h_dt = a->h_dt;
    k.min_bin = a->min_ngb_time_bin;
    k.time_bin = a->time_bin;
    k.ti_current = a->ti_current;
    k.max_active_bin = a->max_active_bin;
    k.gamma = (float)(1.825742);  // kernel_gamma, as upload_consts
    k.cfl = a->integ.cfl_condition;
    k.log_max_h_change = a->integ.log_max_h_change;
    k.dt_min = a->integ.dt_min;
    k.dt_max = a->integ.dt_max;
    k.time_base_inv = 1.0 / a->integ.time_base;
    // the array keeps what the last call left: the header (results, ticket) must start at zero
    CK(cudaMemsetAsync(scratch, 0, kTsHeaderBytes, s));
    const int nb = (int)((a->count + kTsBlock - 1) / kTsBlock);
    k_timestep_static<<<nb, kTsBlock, 0, s>>>(k, scratch);
    CK(cudaGetLastError());
    TsHeader hdr;
    CK(cudaMemcpyAsync(&hdr, scratch, sizeof(hdr), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    if (hdr.ticket != 0) return fail("timestep_static: the reduction did not complete (overlapping calls share the scratch)");
    t = hdr.ti_end_min;
    u = hdr.updated;
    if (hdr.too_small > 0)
      return fail(std::to_string(hdr.too_small) + " particle(s) want a time-step below dt_min (" +
                  std::to_string(a->integ.dt_min) + ")");
  }
  if (ti_end_min) *ti_end_min = (long long)t;
  if (updated) *updated = (long long)u;
  return 0;
}
