// Host side of the backend: cell tree, hydro task list and the flattened
// per-cell interaction lists the kernels consume.  Plain C++17, no CUDA.
//
// What it reproduces (decisions only, no physics):
//   space_regrid            src/space_regrid.c:120-127, 220, 297-301   top-level grid
//   space_split_recursive   src/space_split.c:199-262, cell_split.c:72-110  octree, 400 parts / leaf
//   engine_make_hydroloop_tasks_mapper  src/engine_maketasks.c:4060-4182   self + 26-neighbour pair tasks
//   scheduler_splittask_hydro           src/scheduler.c:1154-1378          push tasks down / make sub tasks
//   DOSUB_SELF1/DOSUB_PAIR1             src/runner_doiact_functions_hydro.h:2524-2617  run-time recursion
//   space_getsid                        src/space_getsid.h:46-84           sid, shift, which cell is "cj"
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/swift_cuda.h"

namespace swiftcuda {

constexpr int kSplitSize = 400;          // space_splitsize_default, src/space.h:49
constexpr int kMaxTopCells = 12;         // space_max_top_level_cells_default, src/space.h:64
constexpr int kSubSizeSelf = 32000;      // space_subsize_self_hydro_default
constexpr int kSubSizePair = 256000000;  // space_subsize_pair_hydro_default
constexpr float kSpaceStretch = 1.10f;   // space_stretch, src/space.h:65
constexpr float kKernelGamma = (float)(1.825742);  // src/kernel_hydro.h:52

// Kinds of direct cell interaction, seen from the particles that get updated.
enum Kind : int {
  kSelf = 0,       // DOSELF1 / DOSELF2 inside cell `own` (f64 position differences)
  kPairLeft = 1,   // DOPAIR1 / DOPAIR2, the updated particles lie in the pair's "ci"
  kPairRight = 2,  // ... in the pair's "cj"
  kSubSelf = 3,    // DOSELF_SUBSET inside cell `own` (f32 cell-relative positions)
  kSubPair = 4,    // DOPAIR_SUBSET against `partner`, walked upwards from its first sorted particle
  kSubPairFlip = 5 // DOPAIR_SUBSET, flipped: walked downwards from the last one
};

// One direct cell interaction in the list of a leaf cell.  The updated
// particles are those of the LEAF; `own` is the leaf or the ancestor of it at
// whose level the reference evaluates the interaction (its loc is the frame
// origin, its h_max enters the sorted-axis cuts).  16 bytes.
struct Item {
  int32_t partner;  // the other cell (== own for the self kinds)
  int32_t own;
  int8_t kind;      // Kind
  int8_t sid;       // sort axis 0..12 (pairs)
  int8_t shift;     // periodic shift of ci: (sx+1) + 3 (sy+1) + 9 (sz+1), s in {-1,0,1} box lengths
  int8_t pad0;
  int32_t pad1;
};

struct Tree {
  std::vector<swift_cuda_cell> cells;
  std::vector<swift_cuda_task> tasks;
  std::vector<int32_t> order;  // order[k] = caller index of the particle at cell-order slot k
  std::vector<float> h_max;    // per cell, from the particles' h at build time (c->hydro.h_max_old)
  int cdim[3] = {0, 0, 0};
};

struct WorkLists {
  std::vector<int32_t> leaves;       // leaf cells (split == 0) that hold active particles and have work
  std::vector<int32_t> item_begin;   // [leaves+1] offsets into items
  std::vector<Item> items;           // first density pass and force loop, in the reference's task order
  std::vector<int32_t> redo_begin;   // [leaves+1] offsets into redo_items
  std::vector<Item> redo_items;      // ghost redo passes, in the order runner_do_ghost walks the task links
  std::vector<int32_t> unit_leaf;    // warp units (32 lanes): index into leaves
  std::vector<int32_t> unit_begin;   // [leaves+1] first unit of each leaf
  std::vector<int32_t> lane_part;    // [units*32] particle owned by each lane (cell-order index) or -1; the
                                     // particles of a leaf are dealt out in Morton order so that the lanes
                                     // of a warp are spatial neighbours (same candidates, same hit rate)
  std::vector<int32_t> sort_cells;   // (cell, sid) segments to sort: cell
  std::vector<int32_t> sort_sids;    //                              sid
  std::vector<int64_t> sort_offset;  // start of each segment in the flat sort arrays
  std::vector<int64_t> sort_lookup;  // [ncells*13] segment start or -1
  std::vector<float> recurse_limit;  // [ncells] 0.5 * dmin of the cells a sub-task recursed through, else +inf
  int64_t n_sort_entries = 0;
  int64_t n_interactions = 0;        // direct self + pair interactions
};

extern const double kAxis[13][3];

// Builds grid + octree; fills tree.cells / order / cdim / h_max.  Returns 0, or
// an error code (1: empty grid, 2: periodic box with fewer than 3 cells).
int build_tree(const swift_cuda_params &prm, int64_t n, const double *x, const float *h, Tree &tree);

// dst[k] = src[order[k]] on the host threads (the particles' h in cell order).
void gather_f32(const float *src, const int32_t *order, int64_t n, float *dst);

// Computes tree.h_max from particle h (cell order).
void compute_cell_hmax(Tree &tree, const float *h_cell_order);

// Creates and splits the hydro tasks for tree.cells.
void build_tasks(const swift_cuda_params &prm, Tree &tree);

// Replays the run-time recursion of every task and groups the resulting direct
// interactions by updated cell.  `active[c]` != 0 when cell c holds an active particle.
void build_worklists(const swift_cuda_params &prm, const Tree &tree, const std::vector<uint8_t> &cell_active,
                     WorkLists &wl);

// Fills wl.lane_part from the particle positions (cell order, [n][3] gathered through `order`
// when it is not empty).
void assign_lanes(const Tree &tree, const double *x, WorkLists &wl);

}  // namespace swiftcuda
