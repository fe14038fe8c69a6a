// See host_tree.hpp for what this reproduces and the reference lines.
#include "host_tree.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <thread>

namespace swiftcuda {

// runner_shift, src/sort_part.h:42-56 (unit vectors of the 13 pair axes)
const double kAxis[13][3] = {
    {5.773502691896258e-01, 5.773502691896258e-01, 5.773502691896258e-01},
    {7.071067811865475e-01, 7.071067811865475e-01, 0.0},
    {5.773502691896258e-01, 5.773502691896258e-01, -5.773502691896258e-01},
    {7.071067811865475e-01, 0.0, 7.071067811865475e-01},
    {1.0, 0.0, 0.0},
    {7.071067811865475e-01, 0.0, -7.071067811865475e-01},
    {5.773502691896258e-01, -5.773502691896258e-01, 5.773502691896258e-01},
    {7.071067811865475e-01, -7.071067811865475e-01, 0.0},
    {5.773502691896258e-01, -5.773502691896258e-01, -5.773502691896258e-01},
    {0.0, 7.071067811865475e-01, 7.071067811865475e-01},
    {0.0, 1.0, 0.0},
    {0.0, 7.071067811865475e-01, -7.071067811865475e-01},
    {0.0, 0.0, 1.0},
};

// sid_scale, src/sort_part.h:93-95
static const float kSidScale[13] = {0.1897f, 0.4025f, 0.1897f, 0.4025f, 0.5788f, 0.4025f, 0.1897f,
                                    0.4025f, 0.1897f, 0.4025f, 0.5788f, 0.4025f, 0.5788f};

namespace {

// A direction code packs the sign of the cell offset per axis as base-3 digits
// (x most significant).  Codes 0..12 are the "flipped" half: the pair is
// stored with the cells swapped so that cj always lies in the positive
// half-space; the sort axis of code k and of 26-k is the same
// (sortlistID / runner_flip, src/sort_part.h:58-90).
inline bool code_is_flipped(int code) { return code < 13; }
inline int code_to_sid(int code) { return code < 13 ? code : (code == 13 ? 0 : 26 - code); }
inline bool sid_is_corner(int sid) { return sid == 0 || sid == 2 || sid == 6 || sid == 8; }

struct PairGeom {
  int ci, cj, sid;
  double shift[3];
};

// space_getsid on cell indices.
PairGeom orient_pair(const swift_cuda_params &prm, const std::vector<swift_cuda_cell> &cells, int a, int b) {
  PairGeom g;
  double dx[3];
  int code = 0;
  for (int k = 0; k < 3; k++) {
    dx[k] = cells[b].loc[k] - cells[a].loc[k];
    if (prm.periodic && dx[k] < -prm.dim[k] / 2)
      g.shift[k] = prm.dim[k];
    else if (prm.periodic && dx[k] > prm.dim[k] / 2)
      g.shift[k] = -prm.dim[k];
    else
      g.shift[k] = 0.0;
    dx[k] += g.shift[k];
    code = 3 * code + ((dx[k] < 0.0) ? 0 : ((dx[k] > 0.0) ? 2 : 1));
  }
  if (code_is_flipped(code)) {
    g.ci = b;
    g.cj = a;
    for (int k = 0; k < 3; k++) g.shift[k] = -g.shift[k];
  } else {
    g.ci = a;
    g.cj = b;
  }
  g.sid = code_to_sid(code);
  return g;
}

// Octant pairs (of ci, of cj) that touch across a pair of direction `sid`:
// role of cell_split_pairs (src/cell.c:62).  Octant k has bit 4|2|1 set when it
// is the upper half along x|y|z (src/space_split.c:223-225).
struct OctPairs {
  int n;
  int pi[16], pj[16];
};
OctPairs make_oct_pairs(int sid) {
  OctPairs o;
  o.n = 0;
  const int code = 26 - sid;
  const int dir[3] = {code / 9 - 1, (code / 3) % 3 - 1, code % 3 - 1};
  for (int a = 0; a < 8; a++)
    for (int b = 0; b < 8; b++) {
      bool touch = true;
      for (int d = 0; d < 3; d++) {
        const int ba = (a >> (2 - d)) & 1, bb = (b >> (2 - d)) & 1;
        const int off = 2 * dir[d] + bb - ba;
        touch &= (off >= -1 && off <= 1);
      }
      if (touch) {
        o.pi[o.n] = a;
        o.pj[o.n] = b;
        o.n++;
      }
    }
  return o;
}
const OctPairs &oct_pairs(int sid) {
  struct Table {
    OctPairs t[13];
    Table() {
      for (int s = 0; s < 13; s++) t[s] = make_oct_pairs(s);
    }
  };
  static const Table table;  // initialised once, thread safe (the list builder runs on several host threads)
  return table.t[sid];
}

inline double cell_dmin(const swift_cuda_cell &c) { return std::min(c.width[0], std::min(c.width[1], c.width[2])); }

}  // namespace

// ---------------------------------------------------------------------------
// Tree
// ---------------------------------------------------------------------------
// Host threads of the builders below (SWIFT_CUDA_HOST_THREADS=1: everything on the calling thread, the form the
// threaded one is checked against, tests/test_cabi.py::test_host_builder_threads_give_the_same_lists).
static int host_threads() {
  if (const char *e = std::getenv("SWIFT_CUDA_HOST_THREADS")) return std::max(1, std::atoi(e));
  return (int)std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
}
template <typename F>
static void parallel_chunks(size_t n, size_t min_per_thread, F &&fn) {  // fn(thread, begin, end) on contiguous chunks
  const size_t nt = std::max<size_t>(1, std::min<size_t>((size_t)host_threads(), n / std::max<size_t>(min_per_thread, 1)));
  if (nt <= 1) {
    fn((size_t)0, (size_t)0, n);
    return;
  }
  std::vector<std::thread> th;
  for (size_t t = 0; t < nt; t++) th.emplace_back([&, t]() { fn(t, n * t / nt, n * (t + 1) / nt); });
  for (auto &t : th) t.join();
}

namespace {
struct SplitCtx {
  const double *x;
  std::vector<swift_cuda_cell> *cells;
  int32_t *order;
  int32_t *scratch;
};

inline int octant_of(const double *p, const double pivot[3]) {
  return (p[0] >= pivot[0]) * 4 + (p[1] >= pivot[1]) * 2 + (p[2] >= pivot[2]);
}

swift_cuda_cell blank_cell() {
  swift_cuda_cell c;
  std::memset(&c, 0, sizeof(c));
  c.parent = -1;
  for (int k = 0; k < 8; k++) c.progeny[k] = -1;
  return c;
}

// Splits one top-level cell into `local` (indices relative to local[0] = the top cell).
void split_top(const double *x, int32_t *order, int32_t *scratch, std::vector<swift_cuda_cell> &local) {
  // explicit stack, depth-first, children appended as they are created
  std::vector<int> stack;
  stack.push_back(0);
  while (!stack.empty()) {
    const int ci = stack.back();
    stack.pop_back();
    const swift_cuda_cell c = local[ci];
    if (c.count <= kSplitSize) continue;
    const double pivot[3] = {c.loc[0] + c.width[0] / 2, c.loc[1] + c.width[1] / 2, c.loc[2] + c.width[2] / 2};
    int bucket[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < c.count; k++) bucket[octant_of(&x[3 * (size_t)order[c.first + k]], pivot)]++;
    int start[9];
    start[0] = 0;
    for (int k = 0; k < 8; k++) start[k + 1] = start[k] + bucket[k];
    int fill[8];
    for (int k = 0; k < 8; k++) fill[k] = start[k];
    for (int k = 0; k < c.count; k++) {
      const int32_t o = order[c.first + k];
      scratch[c.first + fill[octant_of(&x[3 * (size_t)o], pivot)]++] = o;
    }
    std::memcpy(&order[c.first], &scratch[c.first], sizeof(int32_t) * (size_t)c.count);
    local[ci].split = 1;
    for (int k = 0; k < 8; k++) {
      if (bucket[k] == 0) continue;  // empty progeny are recycled (space_split.c:246-249)
      swift_cuda_cell p = blank_cell();
      for (int d = 0; d < 3; d++) {
        p.loc[d] = c.loc[d];
        p.width[d] = c.width[d] / 2;
      }
      if (k & 4) p.loc[0] += p.width[0];
      if (k & 2) p.loc[1] += p.width[1];
      if (k & 1) p.loc[2] += p.width[2];
      p.first = c.first + start[k];
      p.count = bucket[k];
      p.parent = ci;
      p.depth = c.depth + 1;
      const int pi = (int)local.size();
      local.push_back(p);
      local[ci].progeny[k] = pi;
      stack.push_back(pi);
    }
  }
}
}  // namespace

int build_tree(const swift_cuda_params &prm, int64_t n, const double *x, const float *h, Tree &tree) {
  float h_max = 0.f;
  {
    std::vector<float> part((size_t)host_threads(), 0.f);
    parallel_chunks((size_t)n, 1 << 16, [&](size_t t, size_t lo, size_t hi) {
      float m = 0.f;
      for (size_t k = lo; k < hi; k++) m = std::max(m, h[k]);
      part[t] = m;
    });
    for (float m : part) h_max = std::max(h_max, m);
  }
  h_max = std::min(h_max, prm.h_max);
  const double dmax = std::max(prm.dim[0], std::max(prm.dim[1], prm.dim[2]));
  const double cell_min = 0.99 * dmax / kMaxTopCells;  // src/space.c:1160
  int cdim[3];
  double width[3], iwidth[3];
  for (int d = 0; d < 3; d++) {
    cdim[d] = (int)std::floor(prm.dim[d] / std::fmax(h_max * kKernelGamma * kSpaceStretch, cell_min));
    if (cdim[d] < 1) return 1;
    if (prm.periodic && cdim[d] < 3) return 2;
    width[d] = prm.dim[d] / cdim[d];
    iwidth[d] = 1.0 / width[d];
    tree.cdim[d] = cdim[d];
  }
  const int ntop = cdim[0] * cdim[1] * cdim[2];

  // counting sort of the particles into the top-level cells, stable (a cell's particles stay in input order):
  // every host thread histograms a contiguous range of the particles, then scatters it behind the ranges before
  std::vector<int32_t> top_of(n);
  std::vector<int64_t> start(ntop + 1, 0);
  {
    const size_t T = (size_t)host_threads();
    std::vector<std::vector<int64_t>> hist(T);
    parallel_chunks((size_t)n, 1 << 16, [&](size_t t, size_t lo, size_t hi) {
      std::vector<int64_t> &hc = hist[t];
      hc.assign(ntop, 0);
      for (size_t k = lo; k < hi; k++) {
        int idx[3];
        for (int d = 0; d < 3; d++) {
          idx[d] = (int)(x[3 * k + d] * iwidth[d]);
          idx[d] = std::min(std::max(idx[d], 0), cdim[d] - 1);
        }
        top_of[k] = (idx[0] * cdim[1] + idx[1]) * cdim[2] + idx[2];
        hc[top_of[k]]++;
      }
    });
    for (int c = 0; c < ntop; c++) {
      int64_t cnt = 0;
      for (size_t t = 0; t < T; t++)
        if (!hist[t].empty()) {
          const int64_t mine = hist[t][c];
          hist[t][c] = cnt;  // offset of this thread's particles inside the cell
          cnt += mine;
        }
      start[c + 1] = start[c] + cnt;
    }
    tree.order.resize(n);
    parallel_chunks((size_t)n, 1 << 16, [&](size_t t, size_t lo, size_t hi) {
      std::vector<int64_t> &off = hist[t];
      for (size_t k = lo; k < hi; k++) tree.order[start[top_of[k]] + off[top_of[k]]++] = (int32_t)k;
    });
  }
  top_of.clear();
  top_of.shrink_to_fit();

  // split every top-level cell independently (threads), then concatenate
  std::vector<std::vector<swift_cuda_cell>> locals(ntop);
  std::vector<int32_t> scratch(n);
  auto work = [&](int t, int nt) {
    for (int c = t; c < ntop; c += nt) {
      swift_cuda_cell top = blank_cell();
      const int i = c / (cdim[1] * cdim[2]), j = (c / cdim[2]) % cdim[1], k = c % cdim[2];
      top.loc[0] = i * width[0];
      top.loc[1] = j * width[1];
      top.loc[2] = k * width[2];
      for (int d = 0; d < 3; d++) top.width[d] = width[d];
      top.first = (int)start[c];
      top.count = (int)(start[c + 1] - start[c]);
      locals[c].push_back(top);
      split_top(x, tree.order.data(), scratch.data(), locals[c]);
    }
  };
  const int nt = host_threads();
  if (n < 200000 || nt == 1) {
    work(0, 1);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++) th.emplace_back(work, t, nt);
    for (auto &t : th) t.join();
  }
  // top-level cells first (indices 0..ntop-1, as s->cells_top), progeny after
  size_t total = 0;
  for (auto &l : locals) total += l.size();
  tree.cells.assign(total, blank_cell());
  std::vector<int> base(ntop);
  int next = ntop;
  for (int c = 0; c < ntop; c++) {
    base[c] = next - 1;  // local index j>=1 maps to base+j
    next += (int)locals[c].size() - 1;
  }
  for (int c = 0; c < ntop; c++) {
    auto remap = [&](int j) { return j < 0 ? -1 : (j == 0 ? c : base[c] + j); };
    for (size_t j = 0; j < locals[c].size(); j++) {
      swift_cuda_cell cell = locals[c][j];
      cell.parent = (j == 0) ? -1 : remap(cell.parent);
      for (int k = 0; k < 8; k++) cell.progeny[k] = remap(cell.progeny[k]);
      tree.cells[remap((int)j)] = cell;
    }
  }
  return 0;
}

void gather_f32(const float *src, const int32_t *order, int64_t n, float *dst) {
  parallel_chunks((size_t)n, 1 << 16, [&](size_t, size_t lo, size_t hi) {
    for (size_t k = lo; k < hi; k++) dst[k] = src[order[k]];
  });
}

void compute_cell_hmax(Tree &tree, const float *h) {
  const size_t nc = tree.cells.size();
  tree.h_max.assign(nc, 0.f);
  // leaves from particles, parents from children (children have larger indices than
  // their parent in both the library tree and any top-down numbered tree; fall back
  // to a direct scan otherwise).  The scans first, shared out over the host threads (a maximum: any
  // order gives the same bits), then the parents from the last cell to the first.
  auto from_children = [&](size_t c) {
    const swift_cuda_cell &cell = tree.cells[c];
    if (!cell.split) return false;
    for (int k = 0; k < 8; k++)
      if (cell.progeny[k] >= 0 && (size_t)cell.progeny[k] <= c) return false;
    return true;
  };
  parallel_chunks(nc, 2048, [&](size_t, size_t lo, size_t hi) {
    for (size_t c = lo; c < hi; c++) {
      if (from_children(c)) continue;
      const swift_cuda_cell &cell = tree.cells[c];
      float m = 0.f;
      for (int k = 0; k < cell.count; k++) m = std::max(m, h[cell.first + k]);
      tree.h_max[c] = m;
    }
  });
  for (size_t c = nc; c-- > 0;) {
    if (!from_children(c)) continue;
    const swift_cuda_cell &cell = tree.cells[c];
    float m = 0.f;
    for (int k = 0; k < 8; k++)
      if (cell.progeny[k] >= 0) m = std::max(m, tree.h_max[cell.progeny[k]]);
    tree.h_max[c] = m;
  }
}

// ---------------------------------------------------------------------------
// Tasks
// ---------------------------------------------------------------------------
namespace {
struct TaskBuilder {
  const swift_cuda_params &prm;
  Tree &tree;
  // cell_can_split_{self,pair}_hydro_task, src/cell.h:899-931
  bool can_split(int c) const {
    const swift_cuda_cell &cell = tree.cells[c];
    return cell.split && (kSpaceStretch * kKernelGamma * tree.h_max[c] < 0.5f * cell_dmin(cell));
  }
  void add(int type, int ci, int cj) { tree.tasks.push_back(swift_cuda_task{type, ci, cj}); }

  void split_self(int c) {
    if (tree.cells[c].count == 0) return;
    if (!can_split(c)) {
      add(0, c, -1);
      return;
    }
    if (tree.cells[c].count < kSubSizeSelf) {
      add(2, c, -1);
      return;
    }
    const int *pr = tree.cells[c].progeny;
    for (int k = 0; k < 8; k++)
      if (pr[k] >= 0) split_self(pr[k]);
    for (int j = 0; j < 8; j++)
      for (int k = j + 1; k < 8; k++)
        if (pr[j] >= 0 && pr[k] >= 0) split_pair(pr[j], pr[k]);
  }
  void split_pair(int a, int b) {
    if (tree.cells[a].count == 0 || tree.cells[b].count == 0) return;
    const PairGeom g = orient_pair(prm, tree.cells, a, b);
    if (!(can_split(g.ci) && can_split(g.cj))) {
      add(1, g.ci, g.cj);
      return;
    }
    const int ni = tree.cells[g.ci].count, nj = tree.cells[g.cj].count;
    const bool small = ni * kSidScale[g.sid] < kSubSizePair / nj;  // src/scheduler.c:1307-1308
    if (small && !sid_is_corner(g.sid)) {
      add(3, g.ci, g.cj);
      return;
    }
    const OctPairs &op = oct_pairs(g.sid);
    for (int k = 0; k < op.n; k++) {
      const int pa = tree.cells[g.ci].progeny[op.pi[k]], pb = tree.cells[g.cj].progeny[op.pj[k]];
      if (pa >= 0 && pb >= 0) split_pair(pa, pb);
    }
  }
};
}  // namespace

void build_tasks(const swift_cuda_params &prm, Tree &tree) {
  tree.tasks.clear();
  TaskBuilder tb{prm, tree};
  const int *cdim = tree.cdim;
  for (int i = 0; i < cdim[0]; i++)
    for (int j = 0; j < cdim[1]; j++)
      for (int k = 0; k < cdim[2]; k++) {
        const int cid = (i * cdim[1] + j) * cdim[2] + k;
        if (tree.cells[cid].count == 0) continue;
        tb.split_self(cid);
        for (int ii = -1; ii <= 1; ii++) {
          int iii = i + ii;
          if (!prm.periodic && (iii < 0 || iii >= cdim[0])) continue;
          iii = (iii + cdim[0]) % cdim[0];
          for (int jj = -1; jj <= 1; jj++) {
            int jjj = j + jj;
            if (!prm.periodic && (jjj < 0 || jjj >= cdim[1])) continue;
            jjj = (jjj + cdim[1]) % cdim[1];
            for (int kk = -1; kk <= 1; kk++) {
              int kkk = k + kk;
              if (!prm.periodic && (kkk < 0 || kkk >= cdim[2])) continue;
              kkk = (kkk + cdim[2]) % cdim[2];
              const int cjd = (iii * cdim[1] + jjj) * cdim[2] + kkk;
              if (cid >= cjd || tree.cells[cjd].count == 0) continue;
              tb.split_pair(cid, cjd);
            }
          }
        }
      }
}

// ---------------------------------------------------------------------------
// Work lists
// ---------------------------------------------------------------------------
namespace {

struct SeqItem {
  Item it;
  int64_t seq;  // position in the reference's (single-threaded) execution order
};

inline int8_t shift_code(const swift_cuda_params &prm, const double shift[3]) {
  int code = 0, mul = 1;
  for (int k = 0; k < 3; k++) {
    const int s = shift[k] > 0.0 ? 1 : (shift[k] < 0.0 ? -1 : 0);
    code += (s + 1) * mul;
    mul *= 3;
  }
  (void)prm;
  return (int8_t)code;
}

struct Flattener {
  const swift_cuda_params &prm;
  const Tree &tree;
  const std::vector<uint8_t> &active;
  std::vector<std::vector<SeqItem>> per_cell;
  int64_t n_interactions = 0;
  int64_t seq = 0;
  std::vector<uint8_t> *recursed = nullptr;  // cells a sub-task recursed THROUGH (the lists depend on their h_max)

  // cell_can_recurse_in_{pair,self}_hydro_task, src/cell.h:761-782 (dx_max_part_old = 0 after a rebuild)
  bool recurse_pair(int c) const {
    const swift_cuda_cell &cell = tree.cells[c];
    return cell.split && ((kKernelGamma * tree.h_max[c] + 0.f) < 0.5f * cell_dmin(cell));
  }
  bool recurse_self(int c) const {
    const swift_cuda_cell &cell = tree.cells[c];
    return cell.split && (kKernelGamma * tree.h_max[c] < 0.5f * cell_dmin(cell));
  }

  static Item blank() {
    Item it;
    std::memset(&it, 0, sizeof(it));
    return it;
  }

  void direct_self(int c) {
    if (tree.cells[c].count == 0 || !active[c]) return;
    Item it = blank();
    it.partner = c;
    it.own = c;
    it.kind = kSelf;
    it.sid = -1;
    it.shift = 13;
    per_cell[c].push_back(SeqItem{it, seq++});
    n_interactions++;
  }
  void direct_pair(int a, int b) {
    if (tree.cells[a].count == 0 || tree.cells[b].count == 0) return;
    if (!active[a] && !active[b]) return;
    const PairGeom g = orient_pair(prm, tree.cells, a, b);
    Item it = blank();
    it.sid = (int8_t)g.sid;
    it.shift = shift_code(prm, g.shift);
    if (active[g.ci]) {
      it.own = g.ci;
      it.partner = g.cj;
      it.kind = kPairLeft;
      per_cell[g.ci].push_back(SeqItem{it, seq++});
    }
    if (active[g.cj]) {
      it.own = g.cj;
      it.partner = g.ci;
      it.kind = kPairRight;
      per_cell[g.cj].push_back(SeqItem{it, seq++});
    }
    n_interactions++;
  }
  // DOSUB_PAIR1/2, src/runner_doiact_functions_hydro.h:2524-2586, 2630-2692
  void sub_pair(int a, int b) {
    if (!active[a] && !active[b]) return;
    if (tree.cells[a].count == 0 || tree.cells[b].count == 0) return;
    const PairGeom g = orient_pair(prm, tree.cells, a, b);
    if (recurse_pair(g.ci) && recurse_pair(g.cj)) {
      if (recursed) (*recursed)[g.ci] = (*recursed)[g.cj] = 1;
      const OctPairs &op = oct_pairs(g.sid);
      for (int k = 0; k < op.n; k++) {
        const int pa = tree.cells[g.ci].progeny[op.pi[k]], pb = tree.cells[g.cj].progeny[op.pj[k]];
        if (pa >= 0 && pb >= 0) sub_pair(pa, pb);
      }
    } else {
      direct_pair(g.ci, g.cj);
    }
  }
  // DOSUB_SELF1/2, :2587-2617, 2693-2719
  void sub_self(int c) {
    if (tree.cells[c].count == 0 || !active[c]) return;
    if (recurse_self(c)) {
      if (recursed) (*recursed)[c] = 1;
      const int *pr = tree.cells[c].progeny;
      for (int k = 0; k < 8; k++)
        if (pr[k] >= 0) {
          sub_self(pr[k]);
          for (int j = k + 1; j < 8; j++)
            if (pr[j] >= 0) sub_pair(pr[k], pr[j]);
        }
    } else {
      direct_self(c);
    }
  }
};

// The order in which runner_do_ghost (src/runner_ghost.c:1503-1546) re-runs the
// density loop for the unconverged particles of one leaf: the leaf and then
// each of its parents, every density task linked to that cell newest first,
// each expanded by DOSUB_SUBSET (src/runner_doiact_functions_hydro.h:2721-2805).
struct RedoBuilder {
  const swift_cuda_params &prm;
  const Tree &tree;
  const std::vector<uint8_t> &active;
  const std::vector<std::vector<int32_t>> &cell_tasks;  // tasks touching a cell, creation order
  const Flattener &fl;
  int leaf = -1;
  std::vector<Item> *out = nullptr;

  bool holds_leaf(int c) const {
    const swift_cuda_cell &a = tree.cells[c], &l = tree.cells[leaf];
    return l.first >= a.first && l.first < a.first + a.count;
  }
  void self_subset(int c) {
    Item it = Flattener::blank();
    it.partner = c;
    it.own = c;
    it.kind = kSubSelf;
    it.sid = -1;
    it.shift = 13;
    out->push_back(it);
  }
  // DOPAIR_SUBSET_BRANCH, :884-937 (wraps without looking at s->periodic)
  void pair_subset_branch(int ci, int cj) {
    if (tree.cells[cj].count == 0) return;
    double shift[3] = {0.0, 0.0, 0.0};
    int code = 0;
    for (int k = 0; k < 3; k++) {
      const double dx = tree.cells[cj].loc[k] - tree.cells[ci].loc[k];
      if (dx < -prm.dim[k] / 2)
        shift[k] = prm.dim[k];
      else if (dx > prm.dim[k] / 2)
        shift[k] = -prm.dim[k];
      const double w = dx + shift[k];
      code = 3 * code + ((w < 0) ? 0 : ((w > 0) ? 2 : 1));
    }
    Item it = Flattener::blank();
    it.partner = cj;
    it.own = ci;
    it.kind = code_is_flipped(code) ? kSubPairFlip : kSubPair;
    it.sid = (int8_t)code_to_sid(code);
    it.shift = shift_code(prm, shift);
    out->push_back(it);
  }
  void sub_subset(int ci, int cj) {
    if (!active[ci] && (cj < 0 || !active[cj])) return;
    if (tree.cells[ci].count == 0 || (cj >= 0 && tree.cells[cj].count == 0)) return;
    int sub = -1;
    if (tree.cells[ci].split)
      for (int k = 0; k < 8; k++) {
        const int p = tree.cells[ci].progeny[k];
        if (p >= 0 && holds_leaf(p)) {
          sub = p;
          break;
        }
      }
    if (cj < 0) {
      if (fl.recurse_self(ci)) {
        sub_subset(sub, -1);
        for (int j = 0; j < 8; j++) {
          const int p = tree.cells[ci].progeny[j];
          if (p != sub && p >= 0) sub_subset(sub, p);
        }
      } else {
        self_subset(ci);
      }
    } else {
      if (fl.recurse_pair(ci) && fl.recurse_pair(cj)) {
        const PairGeom g = orient_pair(prm, tree.cells, ci, cj);
        const OctPairs &op = oct_pairs(g.sid);
        for (int k = 0; k < op.n; k++) {
          const int pa = tree.cells[g.ci].progeny[op.pi[k]], pb = tree.cells[g.cj].progeny[op.pj[k]];
          if (pa == sub && sub >= 0 && pb >= 0) sub_subset(pa, pb);
          if (pa >= 0 && pb == sub && sub >= 0) sub_subset(pb, pa);
        }
      } else if (active[ci] || active[cj]) {
        pair_subset_branch(ci, cj);
      }
    }
  }
  void build(int leaf_cell, std::vector<Item> &dst) {
    leaf = leaf_cell;
    out = &dst;
    for (int fi = leaf; fi >= 0; fi = tree.cells[fi].parent) {
      const std::vector<int32_t> &tl = cell_tasks[fi];
      for (size_t l = tl.size(); l-- > 0;) {
        const swift_cuda_task &t = tree.tasks[tl[l]];
        const int other = t.cj >= 0 ? (t.ci == fi ? t.cj : t.ci) : -1;
        switch (t.type) {
          case 0: self_subset(fi); break;
          case 1: pair_subset_branch(fi, other); break;
          case 2: sub_subset(fi, -1); break;
          default: sub_subset(fi, other); break;
        }
      }
    }
  }
};
}  // namespace

void build_worklists(const swift_cuda_params &prm, const Tree &tree, const std::vector<uint8_t> &cell_active,
                     WorkLists &wl) {
  Flattener f{prm, tree, cell_active, {}, 0, 0};
  const size_t nc = tree.cells.size();
  f.per_cell.resize(nc);
  wl = WorkLists();
  wl.recurse_limit.assign(nc, INFINITY);
  std::vector<uint8_t> recursed(nc, 0);
  f.recursed = &recursed;
  std::vector<std::vector<int32_t>> cell_tasks(nc);
  for (size_t t = 0; t < tree.tasks.size(); t++) {
    const swift_cuda_task &tk = tree.tasks[t];
    cell_tasks[tk.ci].push_back((int32_t)t);
    if (tk.cj >= 0) cell_tasks[tk.cj].push_back((int32_t)t);
    switch (tk.type) {
      case 0: f.direct_self(tk.ci); break;
      case 1: f.direct_pair(tk.ci, tk.cj); break;
      case 2: f.sub_self(tk.ci); break;
      default: f.sub_pair(tk.ci, tk.cj); break;
    }
  }
  wl.n_interactions = f.n_interactions;
  // the run-time recursion went through these cells because gamma * h_max (+ dx_max_part) was below half
  // their width (cell_can_recurse_in_{pair,self}_hydro_task, src/cell.h:761-782): the flattened lists are
  // those of the reference only while that still holds (cell_unskip.c:808-815 refreshes h_max_old and
  // dx_max_part_old every step)
  for (size_t c = 0; c < nc; c++)
    if (recursed[c]) wl.recurse_limit[c] = 0.5f * (float)cell_dmin(tree.cells[c]);
  wl.sort_lookup.assign(nc * 13, -1);
  auto need_sort = [&](int cell, int sid) {
    int64_t &slot = wl.sort_lookup[(size_t)cell * 13 + sid];
    if (slot < 0) {
      slot = wl.n_sort_entries;
      wl.sort_cells.push_back(cell);
      wl.sort_sids.push_back(sid);
      wl.sort_offset.push_back(wl.n_sort_entries);
      wl.n_sort_entries += tree.cells[cell].count;
    }
  };
  wl.item_begin.push_back(0);
  wl.redo_begin.push_back(0);
  // leaves in the order of their particles (`first`): walking the units in order then walks the particle
  // arrays in order, and a range of units is a range of records (swift_cuda_step_parts hands the force
  // results back in such ranges)
  std::vector<int32_t> visit;
  for (size_t c = 0; c < nc; c++) {
    const swift_cuda_cell &cell = tree.cells[c];
    if (!(cell.split || cell.count == 0 || !cell_active[c])) visit.push_back((int32_t)c);
  }
  std::stable_sort(visit.begin(), visit.end(),
                   [&](int32_t a, int32_t b) { return tree.cells[a].first < tree.cells[b].first; });
  // The lists of a leaf depend on nothing but the flattened interactions (read-only from here on): contiguous
  // chunks of `visit` go to the host threads, each with its own output vectors, stitched together in order below.
  struct Chunk {
    std::vector<Item> items, redo;
    std::vector<int32_t> n_items, n_redo;  // per visited leaf of the chunk
  };
  const size_t nv = visit.size();
  std::vector<Chunk> chunks((size_t)host_threads());
  std::vector<std::pair<size_t, size_t>> range(chunks.size(), {0, 0});
  parallel_chunks(nv, 512, [&](size_t t, size_t lo, size_t hi) {
    Chunk &ck = chunks[t];
    range[t] = {lo, hi};
    RedoBuilder rb{prm, tree, cell_active, cell_tasks, f};
    std::vector<const std::vector<SeqItem> *> chain;
    std::vector<SeqItem> merged;
    ck.n_items.reserve(hi - lo);
    ck.n_redo.reserve(hi - lo);
    for (size_t v = lo; v < hi; v++) {
      const int32_t c = visit[v];
      // every interaction that updates this leaf's particles: its own and its ancestors', in execution order
      chain.clear();
      for (int a = (int)c; a >= 0; a = tree.cells[a].parent)
        if (!f.per_cell[a].empty()) chain.push_back(&f.per_cell[a]);
      const size_t before = ck.items.size();
      if (chain.size() == 1) {
        for (const SeqItem &si : *chain[0]) ck.items.push_back(si.it);
      } else if (chain.size() > 1) {
        merged.clear();
        for (auto *vv : chain) merged.insert(merged.end(), vv->begin(), vv->end());
        std::sort(merged.begin(), merged.end(), [](const SeqItem &a, const SeqItem &b) { return a.seq < b.seq; });
        for (const SeqItem &si : merged) ck.items.push_back(si.it);
      }
      const size_t rbefore = ck.redo.size();
      rb.build((int)c, ck.redo);
      ck.n_items.push_back((int32_t)(ck.items.size() - before));
      ck.n_redo.push_back((int32_t)(ck.redo.size() - rbefore));
    }
  });
  size_t tot_items = 0, tot_redo = 0;
  for (const Chunk &ck : chunks) {
    tot_items += ck.items.size();
    tot_redo += ck.redo.size();
  }
  wl.items.reserve(tot_items);
  wl.redo_items.reserve(tot_redo);
  for (size_t t = 0; t < chunks.size(); t++) {
    const Chunk &ck = chunks[t];
    size_t ip = 0, rp = 0;
    for (size_t v = range[t].first; v < range[t].second; v++) {
      const size_t q = v - range[t].first;
      const int ni = ck.n_items[q], nr = ck.n_redo[q];
      if (ni == 0 && nr == 0) continue;
      const int32_t c = visit[v];
      const swift_cuda_cell &cell = tree.cells[c];
      for (int k = 0; k < ni; k++) {
        const Item &it = ck.items[ip + k];
        if (it.kind != kSelf) {
          need_sort(it.own, it.sid);
          need_sort(it.partner, it.sid);
        }
        wl.items.push_back(it);
      }
      for (int k = 0; k < nr; k++) {
        const Item &it = ck.redo[rp + k];
        if (it.kind != kSubSelf) need_sort(it.partner, it.sid);
        wl.redo_items.push_back(it);
      }
      ip += ni;
      rp += nr;
      const int32_t slot = (int32_t)wl.leaves.size();
      wl.leaves.push_back((int32_t)c);
      wl.item_begin.push_back((int32_t)wl.items.size());
      wl.redo_begin.push_back((int32_t)wl.redo_items.size());
      wl.unit_begin.push_back((int32_t)wl.unit_leaf.size());
      for (int off = 0; off < cell.count; off += 32) wl.unit_leaf.push_back(slot);
    }
  }
  wl.unit_begin.push_back((int32_t)wl.unit_leaf.size());
}

namespace {
struct LanePoint {
  float p[3];
  int32_t slot;
};
// Deals n particles out to `nunits` warp units (32 lanes each) by recursive bisection along the
// longest axis of their bounding box: every unit is a compact box of ~n / nunits particles, so the
// candidates one unit has to look at are few (the unit's box grown by the search radius).
void split_units(LanePoint *a, int n, int nunits, int32_t *dst) {
  if (nunits == 1) {
    for (int k = 0; k < n; k++) dst[k] = a[k].slot;
    return;
  }
  const int ul = nunits / 2, ur = nunits - ul;
  int nl = (int)((int64_t)n * ul / nunits);
  nl = std::max(nl, n - 32 * ur);
  nl = std::min(nl, 32 * ul);
  float lo[3] = {a[0].p[0], a[0].p[1], a[0].p[2]}, hi[3] = {a[0].p[0], a[0].p[1], a[0].p[2]};
  for (int k = 1; k < n; k++)
    for (int d = 0; d < 3; d++) {
      lo[d] = std::min(lo[d], a[k].p[d]);
      hi[d] = std::max(hi[d], a[k].p[d]);
    }
  int ax = 0;
  for (int d = 1; d < 3; d++)
    if (hi[d] - lo[d] > hi[ax] - lo[ax]) ax = d;
  std::nth_element(a, a + nl, a + n, [ax](const LanePoint &u, const LanePoint &v) {
    return u.p[ax] < v.p[ax] || (u.p[ax] == v.p[ax] && u.slot < v.slot);
  });
  split_units(a, nl, ul, dst);
  split_units(a + nl, n - nl, ur, dst + (size_t)ul * 32);
}
}  // namespace

void assign_lanes(const Tree &tree, const double *x, WorkLists &wl) {
  const size_t nl = wl.leaves.size();
  wl.lane_part.assign(wl.unit_leaf.size() * 32, -1);
  const std::vector<int32_t> &ord = tree.order;
  auto work = [&](size_t t, size_t nt) {
    std::vector<LanePoint> pts;
    for (size_t s = t; s < nl; s += nt) {
      const swift_cuda_cell &cell = tree.cells[wl.leaves[s]];
      if (cell.count == 0) continue;
      pts.resize(cell.count);
      for (int k = 0; k < cell.count; k++) {
        const int64_t slot = (int64_t)cell.first + k;
        const int64_t src = ord.empty() ? slot : ord[slot];
        for (int d = 0; d < 3; d++) pts[k].p[d] = (float)(x[3 * src + d] - cell.loc[d]);
        pts[k].slot = (int32_t)slot;
      }
      const int nunits = wl.unit_begin[s + 1] - wl.unit_begin[s];
      split_units(pts.data(), cell.count, nunits, &wl.lane_part[(size_t)wl.unit_begin[s] * 32]);
    }
  };
  const size_t nt = (size_t)host_threads();
  if (nl < 2000 || nt == 1) {
    work(0, 1);
  } else {
    std::vector<std::thread> th;
    for (size_t t = 0; t < nt; t++) th.emplace_back(work, t, nt);
    for (auto &t : th) t.join();
  }
}

}  // namespace swiftcuda
