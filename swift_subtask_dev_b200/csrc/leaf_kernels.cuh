// sm_100a neighbour-loop kernels, second generation: one CTA owns the particles of ONE leaf
// (a group of up to kLW 32-particle units of it) and stages every candidate those particles
// can reach ONCE, in shared memory.
//
//   stage   the warps of the CTA take the leaf's direct cell interactions in turn (a warp per
//           interaction, so kLW dependent load chains are in flight); the partner's particles
//           are read in sorted-axis order, converted to the pair's f32 frame
//           (src/runner_doiact_functions_hydro.h:1138-1158) and kept when they lie inside the
//           group's box (its particles, each grown by its search radius).  Survivors are
//           compacted IN ORDER into the tile; what the interaction body needs of a neighbour
//           (v, m; force: rho, P, f, c, h, balsara, bin, P/rho^2) is copied global -> shared
//           with 16-byte cp.async (LDGSTS), never passing through registers.  Next to the exact
//           f32 frame position every candidate gets a HALF-precision position in the leaf's
//           own frame (centred and scaled to [-1, 1]), four candidates to a 24-byte packet.
//   filter  every warp (unit) walks the staged pieces in the reference's task order; a lane
//           tests four candidates with two broadcast loads and packed half2 arithmetic against
//           a threshold that is provably wider than its search radius (the half rounding
//           errors are bounded) and appends the 2-byte tile slot of every survivor to its
//           private queue in shared memory: ~7 instructions per candidate, no per-interaction
//           set-up.
//   body    every lane consumes its queue in order: the candidate is read back from the tile
//           (LDS), dx is re-formed from the staged f32 frame positions exactly as the reference
//           forms it, the reference's `r2 < h^2 gamma^2` decides, the sums stay in registers.
//           The sorted-axis cuts of the reference (`sort_j[pjd].d < di` ...) are implied by
//           r2 < h^2 gamma^2 except within rounding of the boundary: they are evaluated
//           exactly (slow path) only for hits with r within `delta` of the search radius.
//
// Order of summation is the reference's (task order; inside a pair the sorted-axis order;
// inside a cell the index order), and the density arithmetic is unfused IEEE, so rho, the
// kernel sums and therefore h come out bit-identical to the single-threaded reference: the
// smoothing lengths decide the neighbour sets of every later loop, which must be exact.
// The ghost (k_ghost_leaf) keeps the queue of a wider radius across the Newton-Raphson sweeps:
// a redo sweep re-tests ~1.4 x 48 queued candidates per particle against the new h instead of
// walking ~1000 staged candidates again.
// The force kernel computes the FAST field arithmetic (FMA contraction, reciprocals by MUFU,
// loop-1 and loop-2 hits in the order they are found): same pair set as the reference, fields
// within the tolerance of BASELINE.json.  The first-generation k_force is the EXACT form.
#pragma once
#include <cuda_fp16.h>

#include "kernels.cuh"

namespace swiftcuda {

constexpr int kLW = 6;          // warps per CTA
constexpr int kLT = 32 * kLW;   // threads per CTA
constexpr int kWin = 192;       // candidates of one partner cell a warp stages at once (6 rounds of 32)
constexpr int kWinR = kWin / 32;
constexpr int kVit = 32;        // (interaction, window) pieces per chunk
constexpr int kSlotBits = 11;   // queue entry: piece << 11 | tile slot
constexpr float kHalfErr = 2.7e-3f;   // bound of the half-precision distance error in scaled units (see lane_half)
constexpr float kHalfRel = 1.002f;    // bound of the relative error of the half r2 arithmetic

struct LeafPiece {
  double fl[3];          // origin of the lanes' frame positions for this interaction
  double sh[3];          // periodic shift of the interaction
  float plo[3], phi[3];  // bounding box of the partner cell's particles in the leaf's frame (with margin)
  int begin, end;        // tile slots [begin, end), both multiples of 4, NaN padded
  int item;              // index into the item list
  int kind_ok;           // kind | ok << 8 | shift code << 16 (ok = 0: did not fit the tile, waits for the next chunk)
};

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ unsigned h2_bits(__half2 v) { return *reinterpret_cast<unsigned *>(&v); }
__device__ __forceinline__ __half2 bits_h2(unsigned v) { return *reinterpret_cast<__half2 *>(&v); }

// Four candidates of a packet against the lane's threshold(s): r2 (half2 x 2) < thr (half2 x 2);
// the queue address advances by one row (64 bytes) per survivor.  e0: queue entry of the first.
__device__ __forceinline__ void push4(unsigned &qaddr, unsigned r01, unsigned r23, unsigned t01, unsigned t23,
                                      unsigned e0) {
  asm volatile(
      "{\n\t"
      ".reg .pred p0, p1, p2, p3;\n\t"
      ".reg .b16 t;\n\t"
      ".reg .b32 e;\n\t"
      "setp.lt.f16x2 p0|p1, %1, %3;\n\t"
      "setp.lt.f16x2 p2|p3, %2, %4;\n\t"
      "cvt.u16.u32 t, %5;\n\t"
      "@p0 st.shared.u16 [%0], t;\n\t"
      "@p0 add.u32 %0, %0, 64;\n\t"
      "add.u32 e, %5, 1;\n\t"
      "cvt.u16.u32 t, e;\n\t"
      "@p1 st.shared.u16 [%0], t;\n\t"
      "@p1 add.u32 %0, %0, 64;\n\t"
      "add.u32 e, %5, 2;\n\t"
      "cvt.u16.u32 t, e;\n\t"
      "@p2 st.shared.u16 [%0], t;\n\t"
      "@p2 add.u32 %0, %0, 64;\n\t"
      "add.u32 e, %5, 3;\n\t"
      "cvt.u16.u32 t, e;\n\t"
      "@p3 st.shared.u16 [%0], t;\n\t"
      "@p3 add.u32 %0, %0, 64;\n\t"
      "}"
      : "+r"(qaddr)
      : "r"(r01), "r"(r23), "r"(t01), "r"(t23), "r"(e0)
      : "memory");
}

// Frames of one interaction: fo = origin of the staged candidates, fl = origin of the lanes'
// positions, sub = what DOPAIR_SUBSET takes off a lane's position first (the periodic shift),
// slack = extra f32 margin of the candidate box.
struct ItemFrames {
  double fo[3], fl[3], sub[3], sh[3];
  float slack;
  bool sorted, desc;
};
__device__ __forceinline__ void item_frames(const Dev &d, const Item &item, const CellDev &cA, const CellDev &cB,
                                            ItemFrames &f) {
  const int kind = item.kind;
  f.sorted = !(kind == kSelf || kind == kSubSelf);
  f.desc = (kind == kPairRight || kind == kSubPairFlip);
  f.slack = 0.f;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    f.sh[k] = f.sorted ? item_shift(d, item.shift, k) : 0.0;
    f.sub[k] = 0.0;
  }
  if (kind == kSelf || kind == kSubSelf) {
#pragma unroll
    for (int k = 0; k < 3; k++) f.fo[k] = f.fl[k] = cA.loc[k];
    if (kind == kSelf) f.slack = 1e-6f * cA.width;
  } else if (kind == kPairLeft) {
#pragma unroll
    for (int k = 0; k < 3; k++) { f.fo[k] = cB.loc[k]; f.fl[k] = __dadd_rn(cB.loc[k], f.sh[k]); }
  } else if (kind == kPairRight) {
#pragma unroll
    for (int k = 0; k < 3; k++) { f.fo[k] = __dadd_rn(cA.loc[k], f.sh[k]); f.fl[k] = cA.loc[k]; }
  } else {
#pragma unroll
    for (int k = 0; k < 3; k++) { f.fo[k] = f.fl[k] = cB.loc[k]; f.sub[k] = f.sh[k]; }
    f.slack = 2e-6f * fmaxf(cA.width, cB.width);
  }
}

// Table of the next (interaction, window) pieces of a leaf's item list, built by warp 0.
struct VitTable {
  int item[kVit], win[kVit], nwin[kVit];
  int nv;
};
__device__ __forceinline__ void build_vit(const Dev &d, const Item *__restrict__ items, int it0, int win0, int ie,
                                          VitTable &vt, int lane) {
  // lane k looks at item it0 + k; an item of n partner particles has ceil(n / kWin) windows
  const int it = it0 + lane;
  int nw = 0, nwin_all = 0;
  if (it < ie) {
    nwin_all = (d.cells[items[it].partner].count + kWin - 1) / kWin;
    nw = nwin_all - (lane == 0 ? win0 : 0);
    if (nw < 0) nw = 0;
  }
  int off = nw;
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(kFull, off, o);
    if (lane >= o) off += t;
  }
  const int first = off - nw;
  for (int k = 0; k < nw && first + k < kVit; k++) {
    vt.item[first + k] = it;
    vt.win[first + k] = (lane == 0 ? win0 : 0) + k;
    vt.nwin[first + k] = nwin_all;
  }
  const int total = __shfl_sync(kFull, off, 31);
  if (lane == 0) vt.nv = min(total, kVit);
}

// The group's geometry, the same for every thread of the CTA.
struct GroupGeom {
  float lo[3], hi[3];   // box of the group's particles, each grown by its search radius (leaf frame)
  float lo0[3], hi0[3]; // box of the bare positions
  float rmax;           // largest search radius of the group's particles
  float ctr[3];         // centre of the half-precision frame
  float scale;          // 1 / half extent
};

// ---------------------------------------------------------------------------
// Shared-memory layout.  PAY: 1 = first density pass (f32 frame position, v, m); 2 = ghost redo
// (f64 position, v, m); 3 = force (f32 frame position, v, m, fin, hb).
// ---------------------------------------------------------------------------
struct __align__(16) PosZK {
  double z;
  float key;
  int idx;
};
template <int CAP, int QD, int PAY>
struct LeafTile {
  float4 pos[PAY != 2 ? CAP : 1];      // frame position of the candidate, sort key in w
  double2 xy[PAY == 2 ? CAP : 1];      // ghost: the f64 position of the candidate (DOPAIR_SUBSET takes the
  PosZK zk[PAY == 2 ? CAP : 1];        //        difference in f64) and its sort key
  float4 vm[CAP];                      // v, m of the candidate (cp.async)
  float4 fin[PAY == 3 ? CAP : 1];      // rho, P, f, c
  float4 hb[PAY == 3 ? CAP : 1];       // h, balsara, time bin, P / rho^2
  int idx[PAY != 2 ? CAP : 1];         // particle index of the candidate (ghost: in zk)
  uint4 hA[CAP / 4];                   // packets of four: half x0 x1 | x2 x3 | y0 y1 | y2 y3 (leaf frame, scaled)
  uint2 hB[CAP / 4];                   //                   z0 z1 | z2 z3
  uint2 hH[PAY == 3 ? CAP / 4 : 1];    // force: the candidates' own thresholds t0 t1 | t2 t3
  unsigned short q[kLW][QD][32];       // per-lane queues of survivors (piece << 10 | slot)
  LeafPiece piece[kVit];
  VitTable vt;
  float wbox[kLW][13];
  int used;
  int flag[3];
};

// ---------------------------------------------------------------------------
// Staging of one piece by one warp.
// ---------------------------------------------------------------------------
template <int CAP, int QD, int PAY>
__device__ __forceinline__ void stage_piece(const Dev &d, const Item *__restrict__ items, int it, int win,
                                            const CellDev &cL, const GroupGeom &g, int lane,
                                            LeafTile<CAP, QD, PAY> &T, LeafPiece &out) {
  const float qnan = __int_as_float(0x7fc00000);
  const unsigned lt_mask = (1u << lane) - 1u;
  const Item item = items[it];
  const CellDev cA = d.cells[item.own];
  const CellDev cB = d.cells[item.partner];
  ItemFrames f;
  item_frames(d, item, cA, cB, f);
  // the group's box for this interaction, in the frame of the staged candidates
  float blo[3], bhi[3];
  float rcull = g.rmax;  // a candidate further than this from the box of the bare positions cannot hit
  if (PAY == 3) {
    // force: a hit needs r < max(H of the lane, H of the candidate)
    const float rb = __fmul_rn(d.cell_hmax[item.partner], kc.gamma) * 1.0001f;
    rcull = fmaxf(rcull, rb);
#pragma unroll
    for (int k = 0; k < 3; k++) { blo[k] = fminf(g.lo[k], g.lo0[k] - rb); bhi[k] = fmaxf(g.hi[k], g.hi0[k] + rb); }
  } else {
#pragma unroll
    for (int k = 0; k < 3; k++) { blo[k] = g.lo[k]; bhi[k] = g.hi[k]; }
  }
  StageBox sb, sb0;
  bool reach;
  double offl[3];  // leaf-frame position of a candidate = x_j + offl
  {
    double ol[3], og[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      ol[k] = cL.loc[k] - f.sub[k] - f.fl[k];
      og[k] = cB.loc[k] - f.fo[k];
      offl[k] = f.sub[k] + (f.fl[k] - f.fo[k]) - cL.loc[k];
    }
    stage_box(blo, bhi, ol, f.slack, sb);
    stage_box(g.lo0, g.hi0, ol, f.slack, sb0);
    reach = box_reaches_cell(sb, d.cell_box + (int64_t)item.partner * 6, og);
  }
  const float rc2 = rcull * rcull * 1.001f;
  int total = 0;
  float4 c[kWinR];
  int jj[kWinR];
  unsigned bal[kWinR];
  unsigned short hx[kWinR], hy[kWinR], hz[kWinR];
  double px[kWinR], py[kWinR], pz[kWinR];
#pragma unroll
  for (int r = 0; r < kWinR; r++) bal[r] = 0u;
  if (reach) {
    const int n = cB.count;
    const int k0 = win * kWin;
    const int64_t segB = f.sorted ? d.sort_lookup[(int64_t)item.partner * 13 + item.sid] : 0;
    SortEnt se[kWinR];
#pragma unroll
    for (int r = 0; r < kWinR; r++) {
      const int k = k0 + r * 32 + lane;
      se[r].d = 0.f;
      se[r].i = k;
      if (f.sorted && k < n) se[r] = d.sort[segB + (f.desc ? (n - 1 - k) : k)];
    }
#pragma unroll
    for (int r = 0; r < kWinR; r++) {
      const int k = k0 + r * 32 + lane;
      jj[r] = cB.first + se[r].i;
      px[r] = py[r] = pz[r] = 0.0;
      if (k < n) load_pos(d.pos, jj[r], px[r], py[r], pz[r]);
    }
#pragma unroll
    for (int r = 0; r < kWinR; r++) {
      const int k = k0 + r * 32 + lane;
      c[r] = make_float4(dsub_f(px[r], f.fo[0]), dsub_f(py[r], f.fo[1]), dsub_f(pz[r], f.fo[2]), se[r].d);
      // inside the grown box AND within rcull of the box of the bare positions (rounded corners)
      const float ex = fmaxf(fmaxf(sb0.lo[0] - c[r].x, c[r].x - sb0.hi[0]), 0.f);
      const float ey = fmaxf(fmaxf(sb0.lo[1] - c[r].y, c[r].y - sb0.hi[1]), 0.f);
      const float ez = fmaxf(fmaxf(sb0.lo[2] - c[r].z, c[r].z - sb0.hi[2]), 0.f);
      const bool keep = k < n && c[r].x >= sb.lo[0] && c[r].x <= sb.hi[0] && c[r].y >= sb.lo[1] &&
                        c[r].y <= sb.hi[1] && c[r].z >= sb.lo[2] && c[r].z <= sb.hi[2] &&
                        ex * ex + ey * ey + ez * ez <= rc2;
      bal[r] = __ballot_sync(kFull, keep);
      total += __popc(bal[r]);
      // half-precision position in the leaf's frame
      const float lx = (float)(px[r] + offl[0]), ly = (float)(py[r] + offl[1]), lz = (float)(pz[r] + offl[2]);
      hx[r] = __half_as_ushort(__float2half_rn((lx - g.ctr[0]) * g.scale));
      hy[r] = __half_as_ushort(__float2half_rn((ly - g.ctr[1]) * g.scale));
      hz[r] = __half_as_ushort(__float2half_rn((lz - g.ctr[2]) * g.scale));
    }
  }
  const int padded = (total + 3) & ~3;
  int base = 0;
  if (lane == 0 && padded > 0) base = atomicAdd(&T.used, padded);
  base = __shfl_sync(kFull, base, 0);
  const bool ok = base + padded <= CAP;
  if (ok && padded > 0) {
    unsigned short *const A16 = reinterpret_cast<unsigned short *>(T.hA);
    unsigned short *const B16 = reinterpret_cast<unsigned short *>(T.hB);
    unsigned short *const H16 = reinterpret_cast<unsigned short *>(T.hH);
    int off = base;
#pragma unroll
    for (int r = 0; r < kWinR; r++) {
      if (bal[r] & (1u << lane)) {
        const int s = off + __popc(bal[r] & lt_mask);
        if (PAY == 2) {
          T.xy[s] = make_double2(px[r], py[r]);
          PosZK zk;
          zk.z = pz[r]; zk.key = c[r].w; zk.idx = jj[r];
          T.zk[s] = zk;
        } else {
          T.pos[s] = c[r];
          T.idx[s] = jj[r];
        }
        cp_async16(&T.vm[s], &d.vm[jj[r]]);
        A16[(s >> 2) * 8 + (s & 3)] = hx[r];
        A16[(s >> 2) * 8 + 4 + (s & 3)] = hy[r];
        B16[(s >> 2) * 4 + (s & 3)] = hz[r];
        if (PAY == 3) {
          cp_async16(&T.fin[s], &d.fin[jj[r]]);
          cp_async16(&T.hb[s], &d.hb[jj[r]]);
          // the candidate's own search radius as a half threshold (rounded up)
          const float Rj = __fmul_rn(d.hb[jj[r]].x, kc.gamma) * 1.00001f * g.scale + kHalfErr;
          H16[(s >> 2) * 4 + (s & 3)] = __half_as_ushort(__float2half_ru(Rj * Rj * kHalfRel));
        }
      }
      off += __popc(bal[r]);
    }
    if (lane < padded - total) {
      const int s = base + total + lane;
      if (PAY != 2) {
        T.pos[s] = make_float4(qnan, qnan, qnan, qnan);
        T.idx[s] = -1;
      } else {
        PosZK zk;
        zk.z = 0.0; zk.key = qnan; zk.idx = -1;
        T.zk[s] = zk;
        T.xy[s] = make_double2(0.0, 0.0);
      }
      A16[(s >> 2) * 8 + (s & 3)] = 0x7e00;  // half NaN: never below a threshold
      A16[(s >> 2) * 8 + 4 + (s & 3)] = 0x7e00;
      B16[(s >> 2) * 4 + (s & 3)] = 0x7e00;
      if (PAY == 3) H16[(s >> 2) * 4 + (s & 3)] = 0;
    }
  }
  if (lane == 0) {
    out.fl[0] = f.fl[0]; out.fl[1] = f.fl[1]; out.fl[2] = f.fl[2];
    out.sh[0] = f.sh[0]; out.sh[1] = f.sh[1]; out.sh[2] = f.sh[2];
    // the partner's particle box in the leaf's frame: which units it can reach at all
    const float *bb = d.cell_box + (int64_t)item.partner * 6;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float o = (float)(cB.loc[k] + offl[k]);
      const float m = 4e-6f * (fabsf(o) + fabsf(bb[k]) + fabsf(bb[3 + k])) + f.slack;
      out.plo[k] = bb[k] + o - m;
      out.phi[k] = bb[3 + k] + o + m;
    }
    out.begin = ok ? base : 0;
    out.end = ok ? base + padded : 0;
    out.item = it;
    out.kind_ok = item.kind | (ok ? 256 : 0) | ((int)item.shift << 16);
  }
}

// Stages the next chunk of a leaf's items; returns the number of pieces in the tile [0, nvalid)
// and advances (it0, win0).  All threads of the CTA call it; barriers inside.
template <int CAP, int QD, int PAY>
__device__ __forceinline__ int stage_chunk(const Dev &d, const Item *__restrict__ items, int &it0, int &win0, int ie,
                                           int &nv_cap, const CellDev &cL, const GroupGeom &g,
                                           LeafTile<CAP, QD, PAY> &T) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) T.used = 0;
  if (w == 0) build_vit(d, items, it0, win0, ie, T.vt, lane);
  __syncthreads();
  const int nv = min(T.vt.nv, nv_cap);
  if (nv == 0) {  // 32 interactions without a partner particle: nothing to stage
    it0 = min(it0 + 32, ie);
    win0 = 0;
    __syncthreads();
    return 0;
  }
  for (int v = w; v < nv; v += kLW)
    stage_piece<CAP, QD, PAY>(d, items, T.vt.item[v], T.vt.win[v], cL, g, lane, T, T.piece[v]);
  cp_async_wait_all();
  __syncthreads();
  int nvalid = nv;
  for (int v0 = 0; v0 < nv; v0 += 32) {
    const bool bad = v0 + lane < nv && !(T.piece[v0 + lane].kind_ok & 256);
    const unsigned b = __ballot_sync(kFull, bad);
    if (b) {
      nvalid = v0 + __ffs(b) - 1;
      break;
    }
  }
  // the pieces are allocated in the order the warps get to them: should piece 0 itself find the
  // tile full, the chunk is staged again with that piece alone (a window always fits an empty tile)
  nv_cap = nvalid == 0 ? 1 : kVit;
  if (nvalid < nv) {
    it0 = T.vt.item[nvalid];
    win0 = T.vt.win[nvalid];
  } else {
    it0 = T.vt.item[nv - 1];
    win0 = T.vt.win[nv - 1] + 1;
    if (win0 >= T.vt.nwin[nv - 1]) { it0++; win0 = 0; }
  }
  return nvalid;
}

// The half-precision filter of one unit over the pieces [0, nvalid) of the tile.  thr: the lane's
// threshold (half2, both halves equal; negative: the lane takes no part).  `drain` is called when
// a queue may overflow.  ulo / uhi: the unit's box (leaf frame) for the per-piece reach test.
template <int CAP, int QD, int PAY, typename Drain>
__device__ __forceinline__ void filter_pieces(LeafTile<CAP, QD, PAY> &T, int nvalid, unsigned plx, unsigned ply,
                                              unsigned plz, unsigned &thr, const float ulo[3], const float uhi[3],
                                              unsigned short *qw, int &qcnt, Drain drain) {
  const __half2 px2 = bits_h2(plx), py2 = bits_h2(ply), pz2 = bits_h2(plz);
  const unsigned qbase = (unsigned)__cvta_generic_to_shared(qw);
  for (int v = 0; v < nvalid; v++) {
    const LeafPiece &pc = T.piece[v];
    const int pb = pc.begin, pe = pc.end;
    if (pb == pe) continue;
    const bool take = uhi[0] >= pc.plo[0] && ulo[0] <= pc.phi[0] && uhi[1] >= pc.plo[1] && ulo[1] <= pc.phi[1] &&
                      uhi[2] >= pc.plo[2] && ulo[2] <= pc.phi[2];
    if (!take) continue;
    const unsigned tag = (unsigned)v << kSlotBits;
    for (int b0 = pb; b0 < pe; b0 += 32) {
      const int n4 = min(32, pe - b0) >> 2;
      if (__any_sync(kFull, qcnt > QD - 32)) drain();
      unsigned qaddr = qbase + (unsigned)qcnt * 64u;
      const uint4 *__restrict__ pa = T.hA + (b0 >> 2);
      const uint2 *__restrict__ pbz = T.hB + (b0 >> 2);
      const uint2 *__restrict__ ph = T.hH + (PAY == 3 ? (b0 >> 2) : 0);
#pragma unroll 4
      for (int k = 0; k < n4; k++) {
        const uint4 A = pa[k];
        const uint2 B = pbz[k];
        const __half2 dx01 = __hsub2(px2, bits_h2(A.x)), dx23 = __hsub2(px2, bits_h2(A.y));
        const __half2 dy01 = __hsub2(py2, bits_h2(A.z)), dy23 = __hsub2(py2, bits_h2(A.w));
        const __half2 dz01 = __hsub2(pz2, bits_h2(B.x)), dz23 = __hsub2(pz2, bits_h2(B.y));
        const __half2 r01 = __hfma2(dz01, dz01, __hfma2(dy01, dy01, __hmul2(dx01, dx01)));
        const __half2 r23 = __hfma2(dz23, dz23, __hfma2(dy23, dy23, __hmul2(dx23, dx23)));
        unsigned t01 = thr, t23 = thr;
        if (PAY == 3) {
          const uint2 H = ph[k];
          t01 = h2_bits(__hmax2(bits_h2(thr), bits_h2(H.x)));
          t23 = h2_bits(__hmax2(bits_h2(thr), bits_h2(H.y)));
        }
        push4(qaddr, h2_bits(r01), h2_bits(r23), t01, t23, tag | (unsigned)(b0 + 4 * k));
      }
      qcnt = (int)((qaddr - qbase) >> 6);
    }
  }
}

// The lane's filter constants: its position in the half frame and its threshold for radius R.
// Error budget (scaled units, |coordinate| <= 1.05): both positions are rounded to half
// (<= 2^-11 relative each), the difference once more: |dx_half - dx| <= 1.5e-3 per axis, so
// |r_half - r| <= 2.6e-3 (kHalfErr); the three half operations of r2 add <= 3 * 2^-11 relative
// (kHalfRel); the threshold itself is rounded up.
__device__ __forceinline__ void lane_half(const GroupGeom &g, const float l[3], float R, bool on, unsigned &plx,
                                          unsigned &ply, unsigned &plz, unsigned &thr) {
  plx = h2_bits(__float2half2_rn((l[0] - g.ctr[0]) * g.scale));
  ply = h2_bits(__float2half2_rn((l[1] - g.ctr[1]) * g.scale));
  plz = h2_bits(__float2half2_rn((l[2] - g.ctr[2]) * g.scale));
  const float Rs = R * 1.00001f * g.scale + kHalfErr;
  const __half t = on ? __float2half_ru(Rs * Rs * kHalfRel) : __float2half_rn(-1.f);
  thr = h2_bits(__halves2half2(t, t));
}

// Box of the group from the per-warp boxes in shared memory, and the half frame that covers it.
// wbox[w]: lo xyz, hi xyz (grown), lo0 xyz, hi0 xyz (bare), largest radius.  extra: force, largest partner radius.
__device__ __forceinline__ bool group_geom(const float (*wbox)[13], float extra, GroupGeom &g) {
  g.rmax = 0.f;
  for (int ww = 0; ww < kLW; ww++) g.rmax = fmaxf(g.rmax, wbox[ww][12]);
#pragma unroll
  for (int k = 0; k < 3; k++) {
    g.lo[k] = g.lo0[k] = INFINITY;
    g.hi[k] = g.hi0[k] = -INFINITY;
    for (int ww = 0; ww < kLW; ww++) {
      g.lo[k] = fminf(g.lo[k], wbox[ww][k]);
      g.hi[k] = fmaxf(g.hi[k], wbox[ww][3 + k]);
      g.lo0[k] = fminf(g.lo0[k], wbox[ww][6 + k]);
      g.hi0[k] = fmaxf(g.hi0[k], wbox[ww][9 + k]);
    }
  }
  if (!(g.lo[0] <= g.hi[0])) return false;
  float ext = 0.f;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float lo = fminf(g.lo[k], g.lo0[k] - extra), hi = fmaxf(g.hi[k], g.hi0[k] + extra);
    g.ctr[k] = 0.5f * (lo + hi);
    ext = fmaxf(ext, 0.5f * (hi - lo));
  }
  g.scale = 1.f / (ext * 1.02f + 1e-30f);
  return true;
}

// ---------------------------------------------------------------------------
// K3': first density pass, DOSELF1 / DOPAIR1 arithmetic.  One CTA per group of units of a leaf.
// ---------------------------------------------------------------------------
template <int CAP, int QD>
__global__ void __launch_bounds__(kLT, 3) k_density_leaf(Dev d, int force_slow) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  typedef LeafTile<CAP, QD, 1> Tile;
  Tile &T = *reinterpret_cast<Tile *>(smem_raw);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int4 grp = d.groups[blockIdx.x];
  const int slot = grp.x, u0 = grp.y, nu = grp.z;
  const CellDev cL = d.cells[d.leaves[slot]];
  const int ib = d.item_begin[slot], ie = d.item_begin[slot + 1];
  const Item *__restrict__ items = d.items;
  const bool rec = d.rec_capacity > 0;

  int p = -1;
  if (w < nu) {
    const int q = d.lane_part[(int64_t)(u0 + w) * 32 + lane];
    if (q >= 0 && d.time_bin[q] <= kc.max_active_bin) p = q;
  }
  const bool mine = p >= 0;
  const int pp = mine ? p : cL.first;
  double xi, yi, zi;
  load_pos(d.pos, pp, xi, yi, zi);
  const float hi = d.h[pp];
  const float hig2 = sq_gamma2(hi);
  const float hi_inv = 1.f / hi;
  const float hig = __fmul_rn(hi, kc.gamma);
  const float4 vmi = d.vm[pp];
  const long long idi = rec ? d.id[pp] : 0;
  const float r2safe = safe_r2(hig, cut_delta(d), force_slow != 0);

  // box of the unit's particles, each grown by its search radius, in the leaf's frame
  const float l[3] = {dsub_f(xi, cL.loc[0]), dsub_f(yi, cL.loc[1]), dsub_f(zi, cL.loc[2])};
  float ulo[3], uhi[3];
  {
    const float ru = hig * 1.0001f;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      ulo[k] = warp_min(mine ? l[k] - ru : INFINITY);
      uhi[k] = warp_max(mine ? l[k] + ru : -INFINITY);
      const float b0 = warp_min(mine ? l[k] : INFINITY), b1 = warp_max(mine ? l[k] : -INFINITY);
      if (lane == 0) {
        T.wbox[w][k] = ulo[k];
        T.wbox[w][3 + k] = uhi[k];
        T.wbox[w][6 + k] = b0;
        T.wbox[w][9 + k] = b1;
      }
    }
    const float rm = warp_max(mine ? ru : 0.f);
    if (lane == 0) T.wbox[w][12] = rm;
  }
  __syncthreads();
  GroupGeom g;
  if (!group_geom(T.wbox, 0.f, g)) return;  // no active particle in the group
  const bool unit_on = uhi[0] >= ulo[0];
  unsigned plx, ply, plz, thr;
  lane_half(g, l, hig, mine, plx, ply, plz, thr);

  DensAcc acc;
  acc.zero();
  int qcnt = 0;
  unsigned long long nhits = 0;
  unsigned short *const qw = &T.q[w][0][lane];

  auto drain_all = [&]() {
    const int m = __reduce_max_sync(kFull, qcnt);
    for (int r = 0; r < m; r++) {
      if (r < qcnt) {
        const unsigned e = qw[r * 32];
        const int v = e >> kSlotBits, s = e & ((1 << kSlotBits) - 1);
        const float4 c = T.pos[s];
        const float4 vmj = T.vm[s];
        const LeafPiece &pc = T.piece[v];
        const int kind = pc.kind_ok & 255;
        float dx, dy, dz;
        bool ok = true;
        if (kind == kSelf) {
          // own cell (DOSELF1): the difference is taken in f64 (:2104-2105, 2176-2180)
          const int j = T.idx[s];
          double xj, yj, zj;
          load_pos(d.pos, j, xj, yj, zj);
          dx = dsub_f(xi, xj); dy = dsub_f(yi, yj); dz = dsub_f(zi, zj);
          ok = j != pp;
        } else {
          const float pfx = dsub_f(xi, pc.fl[0]), pfy = dsub_f(yi, pc.fl[1]), pfz = dsub_f(zi, pc.fl[2]);
          dx = __fsub_rn(pfx, c.x); dy = __fsub_rn(pfy, c.y); dz = __fsub_rn(pfz, c.z);
        }
        const float r2 = dist2(dx, dy, dz);
        if (ok && r2 < hig2) {
          if (!(r2 < r2safe)) ok = dens_cut_exact(cut_ctx(d), items[pc.item], xi, yi, zi, hig, c.w);
          if (ok) {
            iact_density(acc, r2, dx, dy, dz, hi_inv, vmi.x, vmi.y, vmi.z, vmj);
            nhits++;
            if (rec) record_pair(d, 0, idi, d.id[T.idx[s]]);
          }
        }
      }
    }
    qcnt = 0;
  };

  int it0 = ib, win0 = 0, nv_cap = kVit;
  while (it0 < ie) {
    const int nvalid = stage_chunk<CAP, QD, 1>(d, items, it0, win0, ie, nv_cap, cL, g, T);
    if (unit_on && nvalid > 0) {
      filter_pieces<CAP, QD, 1>(T, nvalid, plx, ply, plz, thr, ulo, uhi, qw, qcnt, drain_all);
      drain_all();
    }
    __syncthreads();
  }
  if (mine) {
    d.dens0[p] = make_float4(acc.rho, acc.rho_dh, acc.wcount, acc.wcount_dh);
    d.dens1[p] = make_float4(acc.div_v, acc.rx, acc.ry, acc.rz);
  }
  for (int o = 16; o > 0; o >>= 1) nhits += __shfl_down_sync(kFull, nhits, o);
  if (lane == 0 && nhits) atomicAdd(&d.hits[0], nhits);
}

// ---------------------------------------------------------------------------
// K4' + K5': the whole smoothing-length iteration of a group of units in ONE kernel
// (runner_do_ghost, src/runner_ghost.c:1148-1546, with the DOSELF_SUBSET / DOPAIR_SUBSET redo
// passes of :1503-1546).  A particle's density sums depend on its own h only, so the Newton-
// Raphson sweeps of a leaf need nothing from other CTAs.  After the first ghost update the
// candidates of every flagged particle within kGrow x its NEW search radius are staged and
// filtered once, in the order the redo passes visit them; every sweep then re-tests the ~1.3 x 48
// queued candidates of a lane exactly (f64 differences as the reference takes them) against the
// current h.  Should a particle outgrow its queue (h above kGrow x the radius the queue was built
// for), or a queue overflow, the group is staged and filtered again.
// ---------------------------------------------------------------------------
constexpr float kGrow = 1.06f;

template <int CAP, int QD>
__global__ void __launch_bounds__(kLT, 2) k_ghost_leaf(Dev d, int force_slow) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  typedef LeafTile<CAP, QD, 2> Tile;
  Tile &T = *reinterpret_cast<Tile *>(smem_raw);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int4 grp = d.groups[blockIdx.x];
  const int slot = grp.x, u0 = grp.y, nu = grp.z;
  const CellDev cL = d.cells[d.leaves[slot]];
  const int rb = d.redo_begin[slot], re = d.redo_begin[slot + 1];
  const Item *__restrict__ items = d.redo_items;
  const bool rec = d.rec_capacity > 0;
  const float delta = cut_delta(d);

  int p = -1;
  if (w < nu) {
    const int q = d.lane_part[(int64_t)(u0 + w) * 32 + lane];
    if (q >= 0 && d.gstate[q] == 1) p = q;
  }
  // sweep 0 of the ghost: hydro_end_density, the first Newton-Raphson step
  if (p >= 0) ghost_particle(d, 0, p);
  bool flagged = p >= 0 && d.gstate[p] == 2;
  const int pp = p >= 0 ? p : cL.first;
  double xi, yi, zi;
  load_pos(d.pos, pp, xi, yi, zi);
  const float4 vmi = d.vm[pp];
  const long long idi = rec ? d.id[pp] : 0;
  const float l[3] = {dsub_f(xi, cL.loc[0]), dsub_f(yi, cL.loc[1]), dsub_f(zi, cL.loc[2])};

  float hi = 0.f, hig = 0.f, hig2 = 0.f, hi_inv = 0.f, r2safe = -1.f;
  float Rq = 0.f;          // radius the lane's queue was built for
  bool resident = false;   // the tile and the queues of this CTA are valid for another sweep
  int nvalid_res = 0;      // pieces in the resident tile
  int qcnt = 0;
  DensAcc acc;
  unsigned long long nhits = 0;
  unsigned short *const qw = &T.q[w][0][lane];
  GroupGeom g;
  float ulo[3], uhi[3];

  // exact DOSELF_SUBSET / DOPAIR_SUBSET test of one staged candidate against the current h, and the body
  auto process = [&](int v, int s) {
    const double2 xy = T.xy[s];
    const PosZK zk = T.zk[s];
    const LeafPiece &pc = T.piece[v];
    const int kind = pc.kind_ok & 255;
    float dx, dy, dz;
    if (kind == kSubSelf) {
      // DOSELF_SUBSET (:974-1003): f32 positions relative to the cell
      const float pix = dsub_f(xi, pc.fl[0]), piy = dsub_f(yi, pc.fl[1]), piz = dsub_f(zi, pc.fl[2]);
      dx = __fsub_rn(pix, dsub_f(xy.x, pc.fl[0]));
      dy = __fsub_rn(piy, dsub_f(xy.y, pc.fl[1]));
      dz = __fsub_rn(piz, dsub_f(zk.z, pc.fl[2]));
    } else {
      // DOPAIR_SUBSET (:745-770): pix = x - shift in f64, dx = (float)(pix - pj.x)
      dx = dsub_f(__dsub_rn(xi, pc.sh[0]), xy.x);
      dy = dsub_f(__dsub_rn(yi, pc.sh[1]), xy.y);
      dz = dsub_f(__dsub_rn(zi, pc.sh[2]), zk.z);
    }
    const float r2 = dist2(dx, dy, dz);
    if (r2 < hig2) {
      const int j = zk.idx;
      bool ok = j != pp && j >= 0;
      if (ok && !(r2 < r2safe)) ok = dens_cut_exact(cut_ctx(d), items[pc.item], xi, yi, zi, hig, zk.key);
      if (ok) {
        iact_density(acc, r2, dx, dy, dz, hi_inv, vmi.x, vmi.y, vmi.z, T.vm[s]);
        nhits++;
        if (rec) record_pair(d, 1, idi, d.id[j]);
      }
    }
  };
  bool ovf = false;        // the lane's queue overflowed: it filters the tile again in every sweep
  unsigned thr = 0, plx = 0, ply = 0, plz = 0;
  const unsigned thr_off = h2_bits(__float2half2_rn(-1.f));
  // every queued candidate of the lanes with `on` set, in order
  auto run_queue = [&](bool on) {
    const int m = __reduce_max_sync(kFull, on ? qcnt : 0);
    for (int r = 0; r < m; r++) {
      if (on && r < qcnt) {
        const unsigned e = qw[r * 32];
        process(e >> kSlotBits, e & ((1 << kSlotBits) - 1));
      }
    }
  };
  auto drain = [&](int nvalid) {
    run_queue(flagged && !ovf);
    if (__any_sync(kFull, flagged && ovf)) {
      // lanes without a kept queue (far more candidates than a queue holds: h doubled): filter the
      // resident tile again with the current radius, consuming the queue whenever it fills up
      const bool on = flagged && ovf;
      unsigned thr2 = thr_off;
      if (on) {
        const float Rs = hig * 1.00001f * g.scale + kHalfErr;
        const __half t = __float2half_ru(Rs * Rs * kHalfRel);
        thr2 = h2_bits(__halves2half2(t, t));
        qcnt = 0;
      }
      auto mid = [&]() {
        const bool full = on && qcnt > QD - 32;
        run_queue(full);
        if (full) qcnt = 0;
      };
      filter_pieces<CAP, QD, 2>(T, nvalid, plx, ply, plz, thr2, ulo, uhi, qw, qcnt, mid);
      run_queue(on);
      if (on) qcnt = 0;
    }
  };
  // a lane about to overflow its queue stops queueing (the others keep theirs)
  auto on_full = [&]() {
    if (qcnt > QD - 32) {
      ovf = true;
      thr = thr_off;
      qcnt = 0;
    }
  };

  for (int sweep = 0; sweep < kc.max_iter; sweep++) {
    if (!__syncthreads_or(flagged ? 1 : 0)) break;
    if (flagged) {
      hi = d.h[p];
      hig2 = sq_gamma2(hi);
      hi_inv = 1.f / hi;
      hig = __fmul_rn(hi, kc.gamma);
      r2safe = safe_r2(hig, delta, force_slow != 0);
    }
    acc.zero();
#ifdef SWIFT_CUDA_GHOST_STATS
    if (threadIdx.x == 0) atomicAdd(&d.work_ctr[5], 1);
#endif
    const bool need = flagged && !(resident && hig <= Rq);
    if (__syncthreads_or(need ? 1 : 0)) {
      // stage and filter (again) for the particles still flagged, with room for h to grow
      Rq = flagged ? hig * kGrow : 0.f;
      {
        const float ru = Rq * 1.0001f;
#pragma unroll
        for (int k = 0; k < 3; k++) {
          ulo[k] = warp_min(flagged ? l[k] - ru : INFINITY);
          uhi[k] = warp_max(flagged ? l[k] + ru : -INFINITY);
          const float b0 = warp_min(flagged ? l[k] : INFINITY), b1 = warp_max(flagged ? l[k] : -INFINITY);
          if (lane == 0) {
            T.wbox[w][k] = ulo[k];
            T.wbox[w][3 + k] = uhi[k];
            T.wbox[w][6 + k] = b0;
            T.wbox[w][9 + k] = b1;
          }
        }
        const float rm = warp_max(flagged ? ru : 0.f);
        if (lane == 0) T.wbox[w][12] = rm;
      }
      __syncthreads();
      group_geom(T.wbox, 0.f, g);
      const bool unit_on = uhi[0] >= ulo[0];
      lane_half(g, l, Rq, flagged, plx, ply, plz, thr);
      const unsigned thr0 = thr;
      qcnt = 0;
      ovf = false;
      int chunks = 0;
      int it0 = rb, win0 = 0, nv_cap = kVit;
      while (it0 < re) {
#ifdef SWIFT_CUDA_GHOST_STATS
        long long t0 = clock64();
#endif
        const int nvalid = stage_chunk<CAP, QD, 2>(d, items, it0, win0, re, nv_cap, cL, g, T);
#ifdef SWIFT_CUDA_GHOST_STATS
        long long t1 = clock64();
        if (threadIdx.x == 0) { atomicAdd(&d.work_ctr[6], (int)((t1 - t0) >> 6)); atomicAdd(&d.work_ctr[9], T.used); }
#endif
        const bool last = !(it0 < re);
        if (nvalid > 0) chunks++;
        if (unit_on && nvalid > 0) {
          filter_pieces<CAP, QD, 2>(T, nvalid, plx, ply, plz, thr, ulo, uhi, qw, qcnt, on_full);
#ifdef SWIFT_CUDA_GHOST_STATS
          long long t2 = clock64();
          if (threadIdx.x == 0) { atomicAdd(&d.work_ctr[7], (int)((t2 - t1) >> 6)); atomicAdd(&d.work_ctr[10], qcnt); }
#endif
          drain(nvalid);
#ifdef SWIFT_CUDA_GHOST_STATS
          if (threadIdx.x == 0) atomicAdd(&d.work_ctr[8], (int)((clock64() - t2) >> 6));
#endif
          if (!(last && chunks == 1)) {  // the tile is about to be overwritten
            qcnt = 0;
            ovf = false;
            thr = thr0;
          }
        }
        nvalid_res = nvalid;
        if (!last) __syncthreads();
      }
      // the queues stay valid for the next sweep when the whole neighbourhood fitted one tile
      resident = chunks == 1;
      const bool overflowed = __any_sync(kFull, ovf);
#ifdef SWIFT_CUDA_GHOST_STATS
      if (threadIdx.x == 0) {
        atomicAdd(&d.work_ctr[4], 1);
        if (chunks > 1) atomicAdd(&d.work_ctr[2], 1);
        if (overflowed) atomicAdd(&d.work_ctr[3], 1);
      }
#endif
    } else {
#ifdef SWIFT_CUDA_GHOST_STATS
      long long t2 = clock64();
#endif
      drain(nvalid_res);
#ifdef SWIFT_CUDA_GHOST_STATS
      if (threadIdx.x == 0) atomicAdd(&d.work_ctr[8], (int)((clock64() - t2) >> 6));
#endif
    }
    if (flagged) {
      d.dens0[p] = make_float4(acc.rho, acc.rho_dh, acc.wcount, acc.wcount_dh);
      d.dens1[p] = make_float4(acc.div_v, acc.rx, acc.ry, acc.rz);
      if (sweep + 1 < kc.max_iter) {
        ghost_particle(d, sweep + 1, p);
        flagged = d.gstate[p] == 2;
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) nhits += __shfl_down_sync(kFull, nhits, o);
  if (lane == 0 && nhits) atomicAdd(&d.hits[1], nhits);
}

}  // namespace swiftcuda
