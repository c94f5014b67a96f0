// rmb_mobility.cu -- B200 (sm_100a) blob-mobility product Mf and its C ABI.
//
// Replaces the per-block PyCUDA kernels of the reference
// (/root/reference/mobility/mobility_pycuda.py:150-2081, host wrappers
// :2235-2746) with one templated N-body tile kernel:
//   prepare_kernel : O(N) -- height clamp, damping B, 1/a scaling, packing of
//                    position+vector into 16-byte aligned source records
//   mf_kernel      : O(N^2 n_boxes) -- one thread per target, source tiles
//                    streamed through shared memory by bulk async copies
//                    (cp.async.bulk + mbarrier, 2 stages); FP64 or FP32
//   finish_kernel  : O(N n_split) -- sum of source-split partials, self term,
//                    normalisation, damping B, in-plane mask
//   sym_kernel     : O(N^2 n_boxes / 2) -- the flagship (UF / UT / WF / WT and
//                    the fused UF+UT; fp64, wall or no
//                    wall, open or pseudo-periodic in x / y, all blobs targets):
//                    every unordered pair once, both blobs updated;
//                    sym_finish_kernel closes it
// See DESIGN.md for the layout and the roofline.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/rmb_mobility.h"
#include "mf_core.cuh"
#include "sym_schedule.cuh"
#include "tile_pipe.cuh"

using namespace rmb;

namespace {

// Tuning knobs (defaults = the configuration measured best on B200)
#ifndef RMB_TPB
#define RMB_TPB 128
#endif
#ifndef RMB_TILE
#define RMB_TILE 256
#endif
#ifndef RMB_UNROLL
#define RMB_UNROLL 4      // source-loop unroll of the open-boundary kernels
#endif
#ifndef RMB_UNROLL_PER
#define RMB_UNROLL_PER 1  // image-box loop of the pseudo-periodic kernels (measured best: 1)
#endif
#ifndef RMB_MINB
#define RMB_MINB 1
#endif
#ifndef RMB_SYM_UNROLL
#define RMB_SYM_UNROLL 1  // rotation-loop unroll of the pair-symmetric kernel
#endif
// Image-box loop of the pair-symmetric pseudo-periodic kernels (3 or 9 boxes).  Measured at config 3 (r02s,
// profiles/r02s_box_unroll_ab.jsonl): unroll 3 costs the fp64 kernels registers (UF 368 -> 335 G pairs/s, fused 228 -> 213)
// and gains the fp32 one 7 % (583 -> 624).
#ifndef RMB_SYM_BOX_UNROLL
#define RMB_SYM_BOX_UNROLL 1
#endif
#ifndef RMB_SYM_BOX_UNROLL_F32
#define RMB_SYM_BOX_UNROLL_F32 3
#endif
#ifndef RMB_SYM_NT
#define RMB_SYM_NT 1      // blobs per lane of the pair-symmetric kernel (1 or 2; TILE must be a multiple of 128 NT).
                          // Measured at N = 400k: 334.6 (1) vs 336.8 (2) G pairs/s, and 270 vs 252 at N = 20k
#endif
// Resident CTAs per SM the pair-symmetric kernels are compiled for (register cap 65536 / (128 * n)).  Measured
// (tools/tune.sh, N = 400k): UF 4 -> 5 CTAs +1.5 % (open), +2.5 % (9 boxes; uncapped 134 registers = 3 CTAs: -7 %);
// WF, WT +0.5..3 %; UT -1.7 %; the fused block is fastest uncapped (142 / 155 registers).
#ifndef RMB_SYM_MINB
#define RMB_SYM_MINB(BLOCK) (((BLOCK) == UF || (BLOCK) == WF || (BLOCK) == WT) ? 5 : 1)
#endif
// A launch of the pair-symmetric kernel (a rank's share of the triangular rows) is cut into equal-size items of
// `chunk` tiles x one row group (sym_schedule.cuh): chunk is chosen so that each of the RMB_SYM_RESIDENT CTA slots
// per SM gets >= RMB_SYM_ITEMS items, within [RMB_SYM_CHUNK_MIN, RMB_SYM_CHUNK_MAX] tiles.
#ifndef RMB_SYM_RESIDENT
#define RMB_SYM_RESIDENT 5
#endif
#ifndef RMB_SYM_ITEMS
#define RMB_SYM_ITEMS 128
#endif
#ifndef RMB_SYM_CHUNK_MIN
#define RMB_SYM_CHUNK_MIN 1   // r02n: 4 -> 1 gives N = 15,000 0.816 -> 0.731 ms (an item's prologue is ~2 us against 30-150 us
                              // per tile); nothing changes from N = 100k on (profiles/r02n_chunk_min_small_n.jsonl)
#endif
#ifndef RMB_SYM_CHUNK_MAX
#define RMB_SYM_CHUNK_MAX 16
#endif
#ifndef RMB_WAVES
#define RMB_WAVES 10
#endif
#ifndef RMB_MIN_SPLIT_SRC
#define RMB_MIN_SPLIT_SRC 64   // fewest sources one blockIdx.y of the row kernel handles (a multiple of 32; r01: 1024)
#endif
constexpr int TPB = RMB_TPB;    // targets (threads) per CTA
constexpr int TILE = RMB_TILE;  // sources per shared-memory tile
constexpr int STAGES = 2;
#define RMB_PRAGMA(x) _Pragma(#x)
#define RMB_UNROLL_N(n) RMB_PRAGMA(unroll n)

thread_local char g_err[512] = "";

int fail(int code, const char* msg)
{
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
int fail_cuda(cudaError_t e, const char* where)
{
  snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
  return (int)e;
}

// ---------------------------------------------------------------- prepare --
// mobility.py:52-84 fused on the device: r_eff (z <- a where z <= a), B_ii
// (z/a where z < a), B*vec; plus 1/a scaling (open-boundary kernels), the
// in-plane mask and the 1/a factor of the fused torque.  Records past n are
// zero so a full tile can always be copied.
template <typename T, int NV>
__global__ void prepare_kernel(long n, long n_pad, const double* __restrict__ r, const double* __restrict__ va,
                               const double* __restrict__ vb, double a, int wall, int scaled, int in_plane,
                               SrcRec<T, NV>* __restrict__ rec, double* __restrict__ damp)
{
  const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_pad) return;
  SrcRec<T, NV> o;
  constexpr int W = RecWords<T, NV>::value;
  T* w = reinterpret_cast<T*>(&o);
#pragma unroll
  for (int k = 0; k < W; k++) w[k] = T(0);
  if (j < n) {
    const double inva = 1.0 / a;
    const double s = scaled ? inva : 1.0;
    double z = r[3 * j + 2], B = 1.0;
    if (wall) {
      B = (z < a) ? z / a : 1.0;
      z = (z <= a) ? a : z;
    }
    damp[j] = B;
    o.x = T(r[3 * j] * s); o.y = T(r[3 * j + 1] * s); o.z = T(z * s);
#pragma unroll
    for (int c = 0; c < 3; c++) o.v[c] = T(B * va[3 * j + c]);
    if (NV == 2) {
#pragma unroll
      for (int c = 0; c < 3; c++) o.v[3 + c] = T(B * vb[3 * j + c] * inva);
    }
    if (in_plane) {
      o.v[2] = T(0);
      if (NV == 2) o.v[5] = T(0);
    }
  }
  rec[j] = o;
}

// ------------------------------------------------------------ main kernel --
template <typename T, int NV>
struct MfArgs {
  const SrcRec<T, NV>* src;  // n_tiles * TILE records (zero padded)
  long n_src, n_tgt, tgt_off;
  Periodic<T> P;
  T inva;
  int n_tiles;       // ceil(n_src / TILE); the records are zero-padded to n_tiles * TILE
  int src_per_split; // sources handled by one blockIdx.y (a multiple of 32; streamed in tiles of <= TILE records)
  int n_split;
  T* partial;        // [n_split][n_tgt][3] when n_split > 1
  double* out;       // [n_tgt][3], written here when n_split == 1
  const double* damp;
  double norm;
  int in_plane;
  BoxTable<T> box;   // image-box shifts / a (pseudo-periodic kernels)
  UfCtx<T> K;        // literal constants of pair_uf (host-filled: read from the constant bank, never materialised)
};

// Pairs of one target with the `cnt` sources of one shared-memory tile.
// DIAG: the tile may contain the target itself (central box only): that lane
// gets a finite dummy geometry and a zero vector; the closed-form self term
// is added once per target in the epilogue.
template <typename T, int BLOCK, bool WALL, bool PERIODIC, bool DIAG>
__device__ __forceinline__ void tile_pairs(const T* __restrict__ tile, int cnt, long j0, long gi, const Target<T>& t,
                                           const UfCtx<T>& C, const Periodic<T>& P, const BoxTable<T>& box, T inva,
                                           T& ux, T& uy, T& uz)
{
  constexpr int NV = (BLOCK == UF_UT) ? 2 : 1;
  constexpr int W = RecWords<T, NV>::value;
  constexpr int U_OPEN = RMB_UNROLL, U_PER = RMB_UNROLL_PER;
  if constexpr (!PERIODIC) {
#pragma unroll U_OPEN
    for (int j = 0; j < cnt; j++) {
      T w[W];
      load_rec<W>(tile + j * W, w);
      PairGeom<T> g;
      T r2, rxy2;
      pair_geometry<T, WALL, false>(t, w[0], w[1], w[2], P, T(0), T(0), T(0), inva, g, r2, rxy2);
      if (DIAG) {
        const bool self = (j0 + j == gi);
        if (self) { r2 = T(1); rxy2 = T(1); }
#pragma unroll
        for (int c = 3; c < 3 + 3 * NV; c++) w[c] = self ? T(0) : w[c];
      }
      const bool any_near = __any_sync(0xffffffffu, lt4(r2));
      pair_apply<T, BLOCK, WALL, true>(g, C, w + 3, any_near, r2, rxy2, ux, uy, uz);
    }
  } else {
    // image boxes innermost: record load, wrap and 1/a once per source, one shift add per box
#pragma unroll 1
    for (int j = 0; j < cnt; j++) {
      T w[W];
      load_rec<W>(tile + j * W, w);
      PairGeom<T> g0;
      periodic_base<T, WALL>(t, w[0], w[1], w[2], P, inva, g0);
      const bool self_j = DIAG && (j0 + j == gi);
#pragma unroll U_PER
      for (int b = 0; b < box.n; b++) {
        PairGeom<T> g;
        T r2, rxy2;
        periodic_box<T>(g0, P, box.s[b], g, r2, rxy2);
        T v[3 * NV];
#pragma unroll
        for (int c = 0; c < 3 * NV; c++) v[c] = w[3 + c];
        if (DIAG) {
          const bool self = self_j && (b == box.central);
          if (self) { r2 = T(1); rxy2 = T(1); }
#pragma unroll
          for (int c = 0; c < 3 * NV; c++) v[c] = self ? T(0) : v[c];
        }
        const bool any_near = __any_sync(0xffffffffu, lt4(r2));
        pair_apply<T, BLOCK, WALL, false>(g, C, v, any_near, r2, rxy2, ux, uy, uz);
      }
    }
  }
}

template <typename T, int BLOCK, bool WALL, bool PERIODIC, bool TMA>
__global__ void __launch_bounds__(TPB, RMB_MINB) mf_kernel(const __grid_constant__ MfArgs<T, (BLOCK == UF_UT) ? 2 : 1> A)
{
  constexpr int NV = (BLOCK == UF_UT) ? 2 : 1;
  constexpr int W = RecWords<T, NV>::value;
  __shared__ alignas(128) T tile[TMA ? STAGES : 1][TILE * W];
  __shared__ alignas(8) uint64_t full[STAGES];

  const int tid = threadIdx.x;
  const long li = (long)blockIdx.x * TPB + tid;
  const bool valid = li < A.n_tgt;
  const long gi = A.tgt_off + (valid ? li : A.n_tgt - 1);
  const long blk_g0 = A.tgt_off + (long)blockIdx.x * TPB, blk_g1 = blk_g0 + TPB;

  Target<T> t;
  {
    const SrcRec<T, NV>& me = A.src[gi];
    t.x = me.x; t.y = me.y; t.z = me.z;
    t.h = PERIODIC ? me.z * A.inva : me.z;
  }
  UfCtx<T> C = A.K;
  set_target_height(C, t.h);
  // this CTA's sources [s0, s1): tiles of TILE records, the last one possibly shorter (records are padded to n_tiles * TILE)
  const long s0 = (long)blockIdx.y * A.src_per_split;
  const long s1 = min(s0 + (long)A.src_per_split, (long)A.n_tiles * TILE);
  const int v0 = 0, v1 = (int)((s1 - s0 + TILE - 1) / TILE);
  const T* srcw = reinterpret_cast<const T*>(A.src) + (size_t)s0 * W;
  auto tile_records = [&](int v) { return (uint32_t)min((long)TILE, s1 - s0 - (long)v * TILE); };

  if (TMA) {
    if (tid == 0) {
#pragma unroll
      for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
      fence_barrier_init();
    }
    __syncthreads();
    if (tid == 0 && v0 < v1) {
      const uint32_t bytes = tile_records(v0) * W * sizeof(T);
      mbar_expect_tx(&full[0], bytes);
      bulk_g2s(tile[0], srcw + (size_t)v0 * TILE * W, bytes, &full[0]);
    }
  }

  T ux = 0, uy = 0, uz = 0;
  // fp32 variant: the pair arithmetic and the sum over ONE tile (<= 256 sources, or x 3^d images) are fp32, the
  // running sum over tiles is fp64 -- a plain fp32 sum over 10^6 sources costs 2e-5 at N = 1M
  // (profiles/r02j_fp32_error_vs_n.json), above the 1e-5 bar of north_star; 6 conversions/adds per tile are free
  [[maybe_unused]] double wx = 0, wy = 0, wz = 0;
  int stage = 0;
  uint32_t phases = 0;
  for (int v = v0; v < v1; v++) {
    const int tl = v;
    if (TMA) {
      // stage^1 was last read in iteration v-1, which ended with __syncthreads
      if (tid == 0 && v + 1 < v1) {
        const uint32_t bytes = tile_records(v + 1) * W * sizeof(T);
        mbar_expect_tx(&full[stage ^ 1], bytes);
        bulk_g2s(tile[stage ^ 1], srcw + (size_t)(v + 1) * TILE * W, bytes, &full[stage ^ 1]);
      }
      mbar_wait(&full[stage], (phases >> stage) & 1u);
      phases ^= (1u << stage);
    } else {
      const T* gsrc = srcw + (size_t)tl * TILE * W;
      const int words = (int)tile_records(v) * W;
      for (int k = tid; k < words; k += TPB) tile[0][k] = gsrc[k];
      __syncthreads();
    }
    const long j0 = s0 + (long)tl * TILE;
    const int cnt = (int)max(0L, min((long)tile_records(v), A.n_src - j0));
    const bool diag = (j0 < blk_g1) && (j0 + cnt > blk_g0);
    const T* tp = tile[TMA ? stage : 0];
    if (diag) tile_pairs<T, BLOCK, WALL, PERIODIC, true>(tp, cnt, j0, gi, t, C, A.P, A.box, A.inva, ux, uy, uz);
    else tile_pairs<T, BLOCK, WALL, PERIODIC, false>(tp, cnt, j0, gi, t, C, A.P, A.box, A.inva, ux, uy, uz);
    if constexpr (sizeof(T) == 4) {
      wx += (double)ux; wy += (double)uy; wz += (double)uz;
      ux = uy = uz = T(0);
    }
    __syncthreads();
    if (TMA) stage ^= 1;
  }

  if (!valid) return;
  if constexpr (sizeof(T) == 4) {
    if (A.n_split == 1) {
      // same fused epilogue as below, on the fp64 running sums
      const SrcRec<T, NV>& me = A.src[gi];
      double sv[3 * NV];
#pragma unroll
      for (int c = 0; c < 3 * NV; c++) sv[c] = (double)me.v[c];
      self_apply<double, BLOCK, WALL>((double)t.h, sv, wx, wy, wz);
      const double sc = A.norm * A.damp[gi];
      A.out[3 * li] = wx * sc;
      A.out[3 * li + 1] = wy * sc;
      A.out[3 * li + 2] = A.in_plane ? 0.0 : wz * sc;
    } else {
      T* p = A.partial + ((size_t)blockIdx.y * A.n_tgt + li) * 3;
      p[0] = T(wx); p[1] = T(wy); p[2] = T(wz);
    }
    return;
  }
  if (A.n_split == 1) {
    // epilogue fused: self term, normalisation, damping
    const SrcRec<T, NV>& me = A.src[gi];
    double sv[3 * NV];
#pragma unroll
    for (int c = 0; c < 3 * NV; c++) sv[c] = (double)me.v[c];
    double dx = (double)ux, dy = (double)uy, dz = (double)uz;
    self_apply<double, BLOCK, WALL>((double)t.h, sv, dx, dy, dz);
    const double sc = A.norm * A.damp[gi];
    A.out[3 * li] = dx * sc;
    A.out[3 * li + 1] = dy * sc;
    A.out[3 * li + 2] = A.in_plane ? 0.0 : dz * sc;
  } else {
    T* p = A.partial + ((size_t)blockIdx.y * A.n_tgt + li) * 3;
    p[0] = ux; p[1] = uy; p[2] = uz;
  }
}

// Sum of the source-split partials + self term + normalisation + damping.
template <typename T, int BLOCK, bool WALL, bool PERIODIC>
__global__ void finish_kernel(const MfArgs<T, (BLOCK == UF_UT) ? 2 : 1> A)
{
  constexpr int NV = (BLOCK == UF_UT) ? 2 : 1;
  const long li = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (li >= A.n_tgt) return;
  const long gi = A.tgt_off + li;
  double dx = 0, dy = 0, dz = 0;
  for (int s = 0; s < A.n_split; s++) {
    const T* p = A.partial + ((size_t)s * A.n_tgt + li) * 3;
    dx += (double)p[0]; dy += (double)p[1]; dz += (double)p[2];
  }
  const SrcRec<T, NV>& me = A.src[gi];
  double sv[3 * NV];
#pragma unroll
  for (int c = 0; c < 3 * NV; c++) sv[c] = (double)me.v[c];
  const double h = PERIODIC ? (double)me.z * (double)A.inva : (double)me.z;
  self_apply<double, BLOCK, WALL>(h, sv, dx, dy, dz);
  const double sc = A.norm * A.damp[gi];
  A.out[3 * li] = dx * sc;
  A.out[3 * li + 1] = dy * sc;
  A.out[3 * li + 2] = A.in_plane ? 0.0 : dz * sc;
}

// --------------------------------------------------- pair-symmetric kernel --
// All blobs both targets and sources (UF / UT / WF / WT / fused UF+UT, single
// wall or no wall, open or pseudo-periodic in x / y): every unordered pair is
// evaluated ONCE and updates both u_i and u_j (single-wall UF, pair_uf_wall_sym:
// 88 FP64 instructions for two interactions instead of 132).
//
// Blobs are cut into blocks of 32.  A warp owns one block I (lane l holds blob
// 32 I + l in registers) and sweeps the blocks J >= I of its source tiles.  In
// step s of a block, lane l meets j = 32 J + ((l + s) & 31): every lane reads a
// different record from the shared-memory tile (conflict-free LDS.128), adds
// M_ij f_j to its own u_i and M_ji f_i to a travelling accumulator that moves
// one lane per step (3 x SHFL.64), so after 32 steps lane l holds the sum for
// blob 32 J + l and flushes it with a global RED.  J == I is the diagonal block:
// steps 1..31, one direction only.  (Timing experiments: dropping the flush changes
// nothing, dropping the shuffles gains 3.5 %.)  A CTA = 4 warps = a row group of 128 blobs
// (blockIdx.x, scaled by the rank stride) x a range of source tiles
// (blockIdx.y); rows are triangular (tiles >= the row's own), long rows first.
// Sums are accumulated with atomics in acc[n_pad][3]; sym_finish_kernel adds the
// self term, the normalisation and the damping.  Padding records (zero force) are handled by the TAIL
// instantiations: a pair with a padding blob is given a finite dummy geometry.
template <typename T>
struct SymArgsT {
  const T* src;                   // n_tiles * TILE records of W words {x, y, z, v[3 NV], pad}; padding = zero vectors (see TAIL)
  double* acc;                    // [n_tiles * TILE][3], zeroed before the launch
  long n;                         // real blobs; records [n, n_tiles * TILE) are zero-force padding
  int n_tiles;
  int rows, chunk;                // row groups of this launch; tiles per work item (sym_item)
  int row_first, row_stride;      // row groups handled by this launch: row_first + k * row_stride, k < rows
  int pad_[3];                    // keeps K at an offset = 8 mod 16 in the parameter bank: with K 16-byte aligned nvcc 12.9
                                  // re-loads one more constant pair inside the pair loop (126 instead of 125 instructions;
                                  // measured 335.6 vs 338.3 G pairs/s, profiles/r02c_item_decode_ab.jsonl)
  UfCtx<T> K;
  // PER instantiations only (pseudo-periodic in x and/or y; records then hold unscaled positions)
  Periodic<T> P;
  BoxTable<T> box;
  T inva;
};
typedef SymArgsT<double> SymArgs;

// Registers of one lane: NT blobs, one from each of the warp's NT consecutive blocks.
// (T = float: fp32 pair arithmetic; positions and vectors are fp32, the sums u stay fp64)
template <typename T, int NT, int NV = 1>
struct SymLane {
  T x[NT], y[NT], h[NT], h2[NT], h4[NT], f[NT][3 * NV];
  double u[NT][3];
};

// vectors per blob and record words of a block (UF_UT: force and torque / a)
template <int BLOCK, typename T = double> struct SymRec {
  static constexpr int NV = (BLOCK == UF_UT) ? 2 : 1;
  static constexpr int W = RecWords<T, NV>::value;
};

// One block J against the lane's blobs k in [K0, K0 + NK): 32 rotation steps, both directions, then the
// travelling accumulators (home again) are flushed to acc.
// TAIL: the block pair touches the zero-force padding past the last real blob; a pair with a padding blob gets a
// finite dummy geometry, so that it contributes exactly 0 whatever the padding record holds.
// PER: pseudo-periodic in x and/or y (never z, so that |R|^2 = r^2 + 4 h_i h_j still holds).  The image boxes are
// looped innermost; box b seen from i is box -b seen from j (the wrap is odd), so one evaluation with (r, -r)
// serves both directions and the sum over b covers all boxes of both.
// One unordered pair of block BLOCK: u_i += M_ij v_j and (BOTH) u_j += M_ji v_i from shared factors (pair_math.cuh).
template <typename T, int BLOCK, bool WALL, bool BOTH>
__device__ __forceinline__ void sym_pair(T rx, T ry, T rz, T Rz, T hi, T hj, T hi2, T hi4, const UfCtx<T>& C,
                                         const T* vi, const T* vj, bool any_near, T r2, T* ui, T* uj)
{
  if (BLOCK == UF) {
    if (WALL) pair_uf_wall_sym<T, BOTH>(rx, ry, rz, Rz, hi, hj, hi2, hi4, C, vi, vj, any_near, r2, ui, uj);
    else pair_uf_free_sym<T, BOTH>(rx, ry, rz, C, vi, vj, any_near, r2, ui, uj);
  } else if (!WALL) {            // free space: UT and WF are the same function kf (v x r); fused = UF + that
    if (BLOCK == UF_UT) {
      pair_uf_free_sym<T, BOTH>(rx, ry, rz, C, vi, vj, any_near, r2, ui, uj);
      pair_coupling_free_sym<T, BOTH>(rx, ry, rz, C, vi + 3, vj + 3, any_near, r2, ui, uj);
    } else if (BLOCK == WT) {
      pair_wt_free_sym<T, BOTH>(rx, ry, rz, C, vi, vj, any_near, r2, ui, uj);
    } else {
      pair_coupling_free_sym<T, BOTH>(rx, ry, rz, C, vi, vj, any_near, r2, ui, uj);
    }
  } else if (BLOCK == UF_UT) {   // u = UF f + UT (t / a): the two share the geometry and both rsqrt (common subexpressions)
    pair_uf_wall_sym<T, BOTH>(rx, ry, rz, Rz, hi, hj, hi2, hi4, C, vi, vj, any_near, r2, ui, uj);
    pair_ut_wall_sym<T, BOTH>(rx, ry, rz, Rz, hi, hj, hi4, C, vi + 3, vj + 3, any_near, r2, ui, uj);
  } else if (BLOCK == WF) {
    pair_wf_wall_sym<T, BOTH>(rx, ry, rz, Rz, hi, hj, hi4, C, vi, vj, any_near, r2, ui, uj);
  } else if (BLOCK == UT) {
    pair_ut_wall_sym<T, BOTH>(rx, ry, rz, Rz, hi, hj, hi4, C, vi, vj, any_near, r2, ui, uj);
  } else {
    pair_wt_wall_sym<T, BOTH>(rx, ry, rz, Rz, hi4, hj, C, vi, vj, any_near, r2, ui, uj);
  }
}

template <typename T, bool PER>
__device__ __forceinline__ void sym_geometry(const SymArgsT<T>& A, T xi, T yi, T hi, const T* w, T& rx, T& ry, T& rz,
                                             T& hj)
{
  rx = xi - w[0]; ry = yi - w[1];
  if (PER) {
    if (A.P.per[0]) rx = wrap_shift(rx, A.P.L[0], A.P.invL[0], T(0));
    if (A.P.per[1]) ry = wrap_shift(ry, A.P.L[1], A.P.invL[1], T(0));
    rx *= A.inva; ry *= A.inva;
    hj = w[2] * A.inva;
  } else {
    hj = w[2];
  }
  rz = hi - hj;
}

// (r02: an instantiation that runs the non-central image boxes through far-field-only code -- legal when L >= 4a,
// since after the minimum-image wrap only the central box can hold an overlapping pair -- was measured at 348-360
// against 356-358 G pairs/s for this loop, profiles/r02a_per_far_boxes_ab.jsonl: the per-box vote + branch is not
// what bounds the kernel, and the variant was dropped.)
// T = float: the pair arithmetic, the travelling accumulator aj (32 terms) and this block's contribution to u_i
// (32 terms x boxes) are fp32; they are added to the fp64 sums (registers / global atomics) once per block.
template <typename T, int BLOCK, int NT, int K0, int NK, bool TAIL, bool WALL, bool PER>
__device__ __forceinline__ void sym_block(const T* __restrict__ blk, int lane,
                                          SymLane<T, NT, SymRec<BLOCK, T>::NV>& L, const SymArgsT<T>& A,
                                          double* __restrict__ acc_j, long i0, long j0)
{
  constexpr int W = SymRec<BLOCK, T>::W;
  constexpr bool F32 = sizeof(T) == 4;
  const UfCtx<T>& C = A.K;
  const long n = A.n;
  T aj[3] = {0, 0, 0};
  [[maybe_unused]] T ul[NT][3];                 // fp32 only: block-local part of u_i
  if constexpr (F32) {
#pragma unroll
    for (int k = 0; k < NT; k++) ul[k][0] = ul[k][1] = ul[k][2] = T(0);
  }
  int jl = lane;
  const int from = (lane + 1) & 31;
  constexpr int SU = RMB_SYM_UNROLL;
#pragma unroll SU
  for (int s = 0; s < 32; s++) {
    T w[W];
    load_rec<W>(blk + jl * W, w);
#pragma unroll
    for (int k = K0; k < K0 + NK; k++) {
      T rx, ry, rz, hj;
      T hi = L.h[k], hi2 = L.h2[k], hi4 = L.h4[k];
      sym_geometry<T, PER>(A, L.x[k], L.y[k], hi, w, rx, ry, rz, hj);
      if (TAIL && (j0 + jl >= n || i0 + 32 * k + lane >= n)) {   // rz = 3: no xy image shift can bring r to 0
        rx = T(PER ? 0 : 3); ry = T(0); rz = T(PER ? 3 : 0); hi = T(2); hj = T(2); hi2 = T(4); hi4 = T(8);
      }
      if (!PER) {
        const T r2 = RMB_FMA(rz, rz, RMB_FMA(ry, ry, rx * rx));
        const bool any_near = __any_sync(0xffffffffu, lt4(r2));
        if constexpr (F32) sym_pair<T, BLOCK, WALL, true>(rx, ry, rz, hi + hj, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, ul[k], aj);
        else sym_pair<T, BLOCK, WALL, true>(rx, ry, rz, hi + hj, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, L.u[k], aj);
      } else {
        const T Rz = hi + hj, rz2 = rz * rz;
        constexpr int BU = F32 ? RMB_SYM_BOX_UNROLL_F32 : RMB_SYM_BOX_UNROLL;
#pragma unroll BU
        for (int b = 0; b < A.box.n; b++) {
          const T bx = rx + A.box.s[b][0], by = ry + A.box.s[b][1];
          const T r2 = RMB_FMA(by, by, RMB_FMA(bx, bx, rz2));
          const bool any_near = __any_sync(0xffffffffu, lt4(r2));
          if constexpr (F32) sym_pair<T, BLOCK, WALL, true>(bx, by, rz, Rz, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, ul[k], aj);
          else sym_pair<T, BLOCK, WALL, true>(bx, by, rz, Rz, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, L.u[k], aj);
        }
      }
    }
    // the accumulator of blob j moves to the lane that meets j next
    aj[0] = __shfl_sync(0xffffffffu, aj[0], from);
    aj[1] = __shfl_sync(0xffffffffu, aj[1], from);
    aj[2] = __shfl_sync(0xffffffffu, aj[2], from);
    jl = (jl + 1) & 31;
  }
  if constexpr (F32) {
#pragma unroll
    for (int k = K0; k < K0 + NK; k++) { L.u[k][0] += (double)ul[k][0]; L.u[k][1] += (double)ul[k][1]; L.u[k][2] += (double)ul[k][2]; }
    atomicAdd(acc_j + 3 * lane, (double)aj[0]); atomicAdd(acc_j + 3 * lane + 1, (double)aj[1]); atomicAdd(acc_j + 3 * lane + 2, (double)aj[2]);
  } else {
    atomicAdd(acc_j + 3 * lane, aj[0]); atomicAdd(acc_j + 3 * lane + 1, aj[1]); atomicAdd(acc_j + 3 * lane + 2, aj[2]);
  }
}

// Diagonal block (J == the block of blob k): ordered pairs (l, (l + s) & 31), s = 1..31, one direction.
// PER: also s = 0, the blob's own images in the non-central boxes.
template <typename T, int BLOCK, int NT, bool TAIL, bool WALL, bool PER>
__device__ __forceinline__ void sym_diag(const T* __restrict__ blk, int lane, SymLane<T, NT, SymRec<BLOCK, T>::NV>& L, int k,
                                         const SymArgsT<T>& A, long i0)
{
  constexpr int W = SymRec<BLOCK, T>::W;
  constexpr bool F32 = sizeof(T) == 4;
  const UfCtx<T>& C = A.K;
  const long n = A.n;
  [[maybe_unused]] T ul[3] = {T(0), T(0), T(0)};   // fp32 only: this block's part of u_i
#pragma unroll 2
  for (int s = PER ? 0 : 1; s < 32; s++) {
    T w[W];
    const int jl = (lane + s) & 31;
    load_rec<W>(blk + jl * W, w);
    T rx, ry, rz, hj;
    T hi = L.h[k], hi2 = L.h2[k], hi4 = L.h4[k];
    sym_geometry<T, PER>(A, L.x[k], L.y[k], hi, w, rx, ry, rz, hj);
    if (TAIL && (i0 + jl >= n || i0 + lane >= n)) {
      rx = T(PER ? 0 : 3); ry = T(0); rz = T(PER ? 3 : 0); hi = T(2); hj = T(2); hi2 = T(4); hi4 = T(8);
    }
    if (!PER) {
      const T r2 = RMB_FMA(rz, rz, RMB_FMA(ry, ry, rx * rx));
      const bool any_near = __any_sync(0xffffffffu, lt4(r2));
      if constexpr (F32) sym_pair<T, BLOCK, WALL, false>(rx, ry, rz, hi + hj, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, ul, ul);
      else sym_pair<T, BLOCK, WALL, false>(rx, ry, rz, hi + hj, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, L.u[k], L.u[k]);
    } else {
      const T Rz = hi + hj, rz2 = rz * rz;
#pragma unroll 1
      for (int b = 0; b < A.box.n; b++) {
        if (s == 0 && b == A.box.central) continue;          // the self pair: closed form, added by sym_finish_kernel
        const T bx = rx + A.box.s[b][0], by = ry + A.box.s[b][1];
        const T r2 = RMB_FMA(by, by, RMB_FMA(bx, bx, rz2));
        const bool any_near = __any_sync(0xffffffffu, lt4(r2));
        if constexpr (F32) sym_pair<T, BLOCK, WALL, false>(bx, by, rz, Rz, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, ul, ul);
        else sym_pair<T, BLOCK, WALL, false>(bx, by, rz, Rz, hi, hj, hi2, hi4, C, L.f[k], w + 3, any_near, r2, L.u[k], L.u[k]);
      }
    }
  }
  if constexpr (F32) { L.u[k][0] += (double)ul[0]; L.u[k][1] += (double)ul[1]; L.u[k][2] += (double)ul[2]; }
}

template <int BLOCK, bool TMA, int NT, bool WALL, bool PER, typename T = double>
__global__ void __launch_bounds__(TPB, RMB_SYM_MINB(BLOCK)) sym_kernel(const __grid_constant__ SymArgsT<T> A)
{
  constexpr int W = SymRec<BLOCK, T>::W;
  constexpr uint32_t TILE_BYTES = TILE * W * sizeof(T);
  constexpr int ROW = TPB * NT;                            // blobs of one row group (CTA)
  __shared__ alignas(128) T tile[TMA ? STAGES : 1][TILE * W];
  __shared__ alignas(8) uint64_t full[STAGES];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // (broadcasting `warp` with __shfl_sync makes the block index provably uniform and removes the BRA.DIV / WARPSYNC
  // guards around the shuffles -- 125 -> 122 instructions per step -- but measured 3.6 % SLOWER: 338 -> 326 G pairs/s)
  // this CTA's work item: row group g of ROW blobs x the tiles [v0, v1) of its row [g * ROW / TILE, n_tiles)
  // Decoded by one thread and broadcast through shared memory: with the decode inlined in every thread nvcc 12.9
  // allocates 96 instead of 92 registers and schedules the pair loop 3 % slower (328 vs 338 G pairs/s at N = 400k,
  // profiles/r02c_item_decode_ab.jsonl) although the loop's instructions are the same.
  __shared__ SymItem s_item;
  if (tid == 0) s_item = sym_item((int)blockIdx.x, (int)blockIdx.y, A.rows, A.row_first, A.row_stride, ROW, TILE,
                                  A.n_tiles, A.chunk);
  __syncthreads();
  const SymItem item = s_item;
  if (!item.valid) return;
  const int g = (int)item.g;
  const long I0 = ((long)g * (TPB / 32) + warp) * NT;      // this warp's first block of 32

  constexpr int NV = SymRec<BLOCK, T>::NV;
  const T* srcw = A.src;
  SymLane<T, NT, NV> L;
#pragma unroll
  for (int k = 0; k < NT; k++) {
    const T* me = srcw + ((I0 + k) * 32 + lane) * W;
    L.x[k] = me[0]; L.y[k] = me[1]; L.h[k] = PER ? me[2] * A.inva : me[2];
    L.h2[k] = T(2.0) * L.h[k]; L.h4[k] = T(4.0) * L.h[k];
#pragma unroll
    for (int c = 0; c < 3 * NV; c++) L.f[k][c] = me[3 + c];
#pragma unroll
    for (int c = 0; c < 3; c++) L.u[k][c] = 0.0;
  }
  const int v0 = item.v0, v1 = item.v1;

  if (TMA) {
    if (tid == 0) {
#pragma unroll
      for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
      fence_barrier_init();
    }
    __syncthreads();
    if (tid == 0 && v0 < v1) {
      mbar_expect_tx(&full[0], TILE_BYTES);
      bulk_g2s(tile[0], srcw + (size_t)v0 * TILE * W, TILE_BYTES, &full[0]);
    }
  }

  int stage = 0;
  uint32_t phases = 0;
  for (int v = v0; v < v1; v++) {
    if (TMA) {
      if (tid == 0 && v + 1 < v1) {
        mbar_expect_tx(&full[stage ^ 1], TILE_BYTES);
        bulk_g2s(tile[stage ^ 1], srcw + (size_t)(v + 1) * TILE * W, TILE_BYTES, &full[stage ^ 1]);
      }
      mbar_wait(&full[stage], (phases >> stage) & 1u);
      phases ^= (1u << stage);
    } else {
      const T* gsrc = srcw + (size_t)v * TILE * W;
      for (int k = tid; k < TILE * W; k += TPB) tile[0][k] = gsrc[k];
      __syncthreads();
    }
    const T* tp = tile[TMA ? stage : 0];
    for (int jb = 0; jb < TILE / 32; jb++) {
      const long J = (long)v * (TILE / 32) + jb;           // warp-uniform case split on J against I0 .. I0 + NT - 1
      if (sym_block_case(I0, J) == 0) continue;            // J < I0: those pairs of blocks belong to another warp
      const T* blk = tp + jb * 32 * W;
      double* acc_j = A.acc + (size_t)J * 32 * 3;
      const long i0 = I0 * 32, j0 = J * 32;
      const bool tail = (j0 + 32 > A.n) || (i0 + 32 * NT > A.n);   // warp-uniform, true for the last blocks only
      if (sym_block_case(I0 + NT - 1, J) == 2) {           // the hot path: all NT blobs of the lane meet block J
        if (!tail) sym_block<T, BLOCK, NT, 0, NT, false, WALL, PER>(blk, lane, L, A, acc_j, i0, j0);
        else sym_block<T, BLOCK, NT, 0, NT, true, WALL, PER>(blk, lane, L, A, acc_j, i0, j0);
      } else if (sym_block_case(I0, J) == 1) {             // J == I0 (for NT == 2 the second blob skips: J < I0 + 1)
        if (!tail) sym_diag<T, BLOCK, NT, false, WALL, PER>(blk, lane, L, 0, A, i0);
        else sym_diag<T, BLOCK, NT, true, WALL, PER>(blk, lane, L, 0, A, i0);
      } else {                                             // NT == 2, J == I0 + 1
        sym_block<T, BLOCK, NT, 0, 1, true, WALL, PER>(blk, lane, L, A, acc_j, i0, j0);
        sym_diag<T, BLOCK, NT, true, WALL, PER>(blk, lane, L, NT - 1, A, i0 + 32 * (NT - 1));
      }
    }
    __syncthreads();
    if (TMA) stage ^= 1;
  }
#pragma unroll
  for (int k = 0; k < NT; k++) {
    double* a = A.acc + (size_t)((I0 + k) * 32 + lane) * 3;
    atomicAdd(a, L.u[k][0]); atomicAdd(a + 1, L.u[k][1]); atomicAdd(a + 2, L.u[k][2]);
  }
}

// out = B (acc + self term) / (8 pi eta a^k) for the targets [tgt_off, tgt_off + n_tgt).  Reads the caller's own
// arrays (positions, vectors), not the packed records of the partial step: the clamp, the damping B and B * v are
// O(1) per target (mobility.py:52-84), so the finish step depends on nothing but its arguments.
template <int BLOCK, bool WALL>
__global__ void sym_finish_kernel(const double* __restrict__ r, const double* __restrict__ va,
                                  const double* __restrict__ vb, const double* __restrict__ acc, long tgt_off,
                                  long n_tgt, double a, double norm, int in_plane, double* __restrict__ out)
{
  const long li = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (li >= n_tgt) return;
  const long gi = tgt_off + li;
  const double inva = 1.0 / a;
  double z = r[3 * gi + 2], B = 1.0;
  if (WALL) {
    B = (z < a) ? z / a : 1.0;
    z = (z <= a) ? a : z;
  }
  double sv[6] = {B * va[3 * gi], B * va[3 * gi + 1], in_plane ? 0.0 : B * va[3 * gi + 2], 0.0, 0.0, 0.0};
  if (SymRec<BLOCK>::NV == 2) {
    sv[3] = B * vb[3 * gi] * inva; sv[4] = B * vb[3 * gi + 1] * inva; sv[5] = in_plane ? 0.0 : B * vb[3 * gi + 2] * inva;
  }
  double dx = acc[3 * gi], dy = acc[3 * gi + 1], dz = acc[3 * gi + 2];
  self_apply<double, BLOCK, WALL>(z * inva, sv, dx, dy, dz);
  const double sc = norm * B;
  out[3 * li] = dx * sc;
  out[3 * li + 1] = dy * sc;
  out[3 * li + 2] = in_plane ? 0.0 : dz * sc;
}

// ---------------------------------------------------------- host dispatch --
int g_sm_count = 0;

int sm_count()
{
  if (g_sm_count > 0) return g_sm_count;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return fail(RMB_E_NODEVICE, "no CUDA device");
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
    return fail(RMB_E_NODEVICE, "cannot query the SM count");
  g_sm_count = n;
  return n;
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct Plan {
  int n_tiles, n_boxes, n_split, src_per_split, tgt_blocks;
  size_t rec_bytes, off_damp, off_partial, off_acc, total;
};

// One place decides the decomposition; rmb_workspace_bytes and the launch
// must agree.  Source splitting (blockIdx.y) is only used when the target
// blocks alone would give fewer than RMB_WAVES waves of resident CTAs.
int make_plan(int block, int precision, long n_tgt, long n_src, const double* L, Plan& p)
{
  const int sms = sm_count();
  if (sms <= 0) return sms;
  const int nv = (block == RMB_UF_UT) ? 2 : 1;
  const size_t word = precision == RMB_FP32 ? 4 : 8;
  const int W = precision == RMB_FP32 ? (nv == 2 ? 12 : 8) : (nv == 2 ? 10 : 6);
  p.n_tiles = (int)((n_src + TILE - 1) / TILE);
  p.n_boxes = 1;
  if (L) for (int d = 0; d < 3; d++) if (L[d] > 0) p.n_boxes *= 3;
  p.tgt_blocks = (int)((n_tgt + TPB - 1) / TPB);   // (image boxes are looped inside the pair loop, not as extra tiles)
  // Source splitting (blockIdx.y): used when the target CTAs alone would give fewer than RMB_WAVES waves of
  // resident CTAs (tail effect / idle SMs).  A split is a multiple of 32 sources and keeps >= RMB_MIN_SPLIT_SRC of
  // them: at N = 1,000 (8 target CTAs) that is 16 splits = 128 CTAs of 64 sources instead of 8 CTAs that each walk
  // all 1,000 sources at one warp per scheduler.
  const long n_pad = (long)p.n_tiles * TILE;
  const int resident = sms * (768 / TPB);
  long split = (RMB_WAVES * (long)resident + p.tgt_blocks - 1) / (p.tgt_blocks > 0 ? p.tgt_blocks : 1);
  if (split > n_pad / RMB_MIN_SPLIT_SRC) split = n_pad / RMB_MIN_SPLIT_SRC;
  if (split < 1) split = 1;
  long per = (n_pad + split - 1) / split;
  per = (per + 31) / 32 * 32;
  if (per < 32) per = 32;                         // empty source set
  p.src_per_split = (int)per;
  p.n_split = n_pad > 0 ? (int)((n_pad + per - 1) / per) : 1;
  p.rec_bytes = (size_t)p.n_tiles * TILE * W * word;
  p.off_damp = align_up(p.rec_bytes, 256);
  p.off_partial = align_up(p.off_damp + (size_t)n_src * sizeof(double), 256);
  p.total = p.off_partial + (p.n_split > 1 ? (size_t)p.n_split * n_tgt * 3 * word : 0);
  // accumulators of the pair-symmetric kernel (fp64, every blob a target; see use_symmetric)
  p.off_acc = align_up(p.total, 256);
  if (block >= RMB_UF && block <= RMB_UF_UT && n_tgt == n_src)      // fp64 sums for the fp32 pair arithmetic too
    p.total = p.off_acc + (size_t)p.n_tiles * TILE * 3 * sizeof(double);
  return 0;
}

// The pair-symmetric kernel exists for UF / UT / WF / WT / fused UF+UT, single wall or no wall, fp64, every blob a target, open
// or pseudo-periodic in x/y (a periodic z breaks |R|^2 = r^2 + 4 h_i h_j).  Forced by RMB_SYMMETRIC, chosen by default from
// RMB_SYMMETRIC_MIN_N blobs on (below that the row kernel is as fast), never with RMB_DETERMINISTIC.
bool use_symmetric(int block, int wall, int precision, int flags, long n_tgt, long n_src, const double* L)
{
  const bool blk = block >= RMB_UF && block <= RMB_UF_UT && (wall == RMB_SINGLE_WALL || wall == RMB_NO_WALL);
  const bool ok = blk && (precision == RMB_FP64 || precision == RMB_FP32) &&
                  !(L && L[2] > 0) && n_tgt == n_src && TILE % (TPB * RMB_SYM_NT) == 0;
  const bool want = (flags & RMB_SYMMETRIC) || (!(flags & RMB_DETERMINISTIC) && n_src >= RMB_SYMMETRIC_MIN_N);
  return ok && want;
}

struct Call {
  int block, wall, precision, flags;
  long n_tgt, n_src, tgt_off;
  const double *r, *va, *vb;
  double eta, a;
  double L[3];
  double* out;
  char* ws;
  cudaStream_t stream;
  Plan plan;
};

template <typename T, int BLOCK, bool WALL, bool PERIODIC>
int launch(const Call& c)
{
  constexpr int NV = (BLOCK == UF_UT) ? 2 : 1;
  const Plan& p = c.plan;
  SrcRec<T, NV>* rec = reinterpret_cast<SrcRec<T, NV>*>(c.ws);
  double* damp = reinterpret_cast<double*>(c.ws + p.off_damp);
  const long n_pad = (long)p.n_tiles * TILE;
  const int in_plane = (c.flags & RMB_IN_PLANE) ? 1 : 0;
  prepare_kernel<T, NV><<<(unsigned)((n_pad + 255) / 256), 256, 0, c.stream>>>(
      c.n_src, n_pad, c.r, c.va, c.vb, c.a, (WALL && BLOCK != UF_FS) ? 1 : 0, PERIODIC ? 0 : 1, in_plane, rec, damp);

  MfArgs<T, NV> A;
  A.src = rec;
  A.n_src = c.n_src; A.n_tgt = c.n_tgt; A.tgt_off = c.tgt_off;
  A.P.nbox = p.n_boxes;
  for (int d = 0; d < 3; d++) {
    A.P.per[d] = PERIODIC && c.L[d] > 0;
    A.P.L[d] = T(c.L[d]);
    A.P.invL[d] = A.P.per[d] ? T(1.0 / c.L[d]) : T(0);
  }
  A.inva = T(1.0 / c.a);
  A.box = make_box_table<T>(A.P, 1.0 / c.a);
  A.n_tiles = p.n_tiles; A.src_per_split = p.src_per_split; A.n_split = p.n_split;
  A.partial = reinterpret_cast<T*>(c.ws + p.off_partial);
  A.out = c.out;
  A.damp = damp;
  A.norm = block_norm(BLOCK, c.eta, c.a);
  A.in_plane = in_plane;
  A.K = make_uf_ctx<T>(T(0));

  const dim3 grid((unsigned)p.tgt_blocks, (unsigned)p.n_split);
  if (c.flags & RMB_NO_TMA) mf_kernel<T, BLOCK, WALL, PERIODIC, false><<<grid, TPB, 0, c.stream>>>(A);
  else mf_kernel<T, BLOCK, WALL, PERIODIC, true><<<grid, TPB, 0, c.stream>>>(A);
  if (p.n_split > 1)
    finish_kernel<T, BLOCK, WALL, PERIODIC><<<(unsigned)((c.n_tgt + 127) / 128), 128, 0, c.stream>>>(A);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(e, "rmb_mobility_product launch");
  return 0;
}

// Pair-symmetric path: every block, fp64, wall or no wall, not periodic in z, every blob a target.
// Step 1: records + raw pair sums of the row groups row_first, row_first + row_stride, ... into acc.
// fp32 pair arithmetic (precision = RMB_FP32): same schedule and the same fp64 accumulator, records and kernel in float.
int sym_partial_f32(const Call& c, double* acc, int row_first, int row_stride)
{
  typedef float T;
  const Plan& p = c.plan;
  double* damp = reinterpret_cast<double*>(c.ws + p.off_damp);
  const long n_pad = (long)p.n_tiles * TILE;
  const int in_plane = (c.flags & RMB_IN_PLANE) ? 1 : 0;
  const bool wall = c.wall == RMB_SINGLE_WALL;
  const bool per = c.L[0] > 0 || c.L[1] > 0;
  const unsigned prep_blocks = (unsigned)((n_pad + 255) / 256);
  if (c.block == RMB_UF_UT)
    prepare_kernel<T, 2><<<prep_blocks, 256, 0, c.stream>>>(c.n_src, n_pad, c.r, c.va, c.vb, c.a, wall ? 1 : 0, per ? 0 : 1,
                                                            in_plane, reinterpret_cast<SrcRec<T, 2>*>(c.ws), damp);
  else
    prepare_kernel<T, 1><<<prep_blocks, 256, 0, c.stream>>>(c.n_src, n_pad, c.r, c.va, c.vb, c.a, wall ? 1 : 0, per ? 0 : 1,
                                                            in_plane, reinterpret_cast<SrcRec<T, 1>*>(c.ws), damp);
  cudaError_t e = cudaMemsetAsync(acc, 0, (size_t)n_pad * 3 * sizeof(double), c.stream);
  if (e != cudaSuccess) return fail_cuda(e, "pair-symmetric product (fp32): memset");
  SymArgsT<T> A;
  A.src = reinterpret_cast<const T*>(c.ws); A.acc = acc; A.n = c.n_src; A.n_tiles = p.n_tiles;
  A.rows = 0; A.chunk = 1;
  A.row_first = row_first; A.row_stride = row_stride;
  A.K = make_uf_ctx<T>(T(0));
  A.P.nbox = p.n_boxes;
  for (int d = 0; d < 3; d++) {
    A.P.per[d] = d < 2 && c.L[d] > 0;
    A.P.L[d] = T(c.L[d]);
    A.P.invL[d] = A.P.per[d] ? T(1.0 / c.L[d]) : T(0);
  }
  A.inva = T(1.0 / c.a);
  A.box = make_box_table<T>(A.P, 1.0 / c.a);
  const long rows_all = n_pad / (TPB * RMB_SYM_NT);
  const long rows = (rows_all - row_first + row_stride - 1) / row_stride;
  if (rows > 0) {
    constexpr int NT = RMB_SYM_NT;
    A.rows = (int)rows;
    A.chunk = sym_chunk_for(rows, row_first, row_stride, TPB * NT, TILE, p.n_tiles, (long)sm_count() * RMB_SYM_RESIDENT,
                            RMB_SYM_ITEMS, RMB_SYM_CHUNK_MIN, RMB_SYM_CHUNK_MAX);
    const int gy = sym_grid_y(A.rows, row_first, row_stride, TPB * NT, TILE, p.n_tiles, A.chunk);
    if (gy > 65535) return fail(RMB_E_ARG, "pair-symmetric product: too many chunks per row pair");
    const dim3 grid((unsigned)((rows + 1) / 2), (unsigned)gy);
#define RMB_SYM_LAUNCH32(B)                                                                     \
    do {                                                                                        \
      if (wall && per) sym_kernel<B, true, NT, true, true, T><<<grid, TPB, 0, c.stream>>>(A);     \
      else if (wall) sym_kernel<B, true, NT, true, false, T><<<grid, TPB, 0, c.stream>>>(A);      \
      else if (per) sym_kernel<B, true, NT, false, true, T><<<grid, TPB, 0, c.stream>>>(A);       \
      else sym_kernel<B, true, NT, false, false, T><<<grid, TPB, 0, c.stream>>>(A);               \
    } while (0)
    if (c.block == RMB_UF_UT) {
      RMB_SYM_LAUNCH32(UF_UT);
    } else if (c.block == RMB_WT) {
      RMB_SYM_LAUNCH32(WT);
    } else if (c.block == RMB_UT || (c.block == RMB_WF && !wall)) {   // free space: WF is the same function as UT
      RMB_SYM_LAUNCH32(UT);
    } else if (c.block == RMB_WF) {
      if (per) sym_kernel<WF, true, NT, true, true, T><<<grid, TPB, 0, c.stream>>>(A);
      else sym_kernel<WF, true, NT, true, false, T><<<grid, TPB, 0, c.stream>>>(A);
    } else {
      RMB_SYM_LAUNCH32(UF);
    }
#undef RMB_SYM_LAUNCH32
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(e, "pair-symmetric product (fp32): launch");
  return 0;
}

int sym_partial(const Call& c, double* acc, int row_first, int row_stride)
{
  if (c.precision == RMB_FP32) return sym_partial_f32(c, acc, row_first, row_stride);
  const Plan& p = c.plan;
  double* damp = reinterpret_cast<double*>(c.ws + p.off_damp);
  const long n_pad = (long)p.n_tiles * TILE;
  const int in_plane = (c.flags & RMB_IN_PLANE) ? 1 : 0;
  const bool wall = c.wall == RMB_SINGLE_WALL;
  const bool per = c.L[0] > 0 || c.L[1] > 0;
  const unsigned prep_blocks = (unsigned)((n_pad + 255) / 256);
  if (c.block == RMB_UF_UT)
    prepare_kernel<double, 2><<<prep_blocks, 256, 0, c.stream>>>(c.n_src, n_pad, c.r, c.va, c.vb, c.a, wall ? 1 : 0, per ? 0 : 1,
                                                                 in_plane, reinterpret_cast<SrcRec<double, 2>*>(c.ws), damp);
  else
    prepare_kernel<double, 1><<<prep_blocks, 256, 0, c.stream>>>(c.n_src, n_pad, c.r, c.va, c.vb, c.a, wall ? 1 : 0, per ? 0 : 1,
                                                                 in_plane, reinterpret_cast<SrcRec<double, 1>*>(c.ws), damp);
  cudaError_t e = cudaMemsetAsync(acc, 0, (size_t)n_pad * 3 * sizeof(double), c.stream);
  if (e != cudaSuccess) return fail_cuda(e, "pair-symmetric product: memset");
  SymArgs A;
  A.src = reinterpret_cast<const double*>(c.ws); A.acc = acc; A.n = c.n_src; A.n_tiles = p.n_tiles;
  A.rows = 0; A.chunk = 1;
  A.row_first = row_first; A.row_stride = row_stride;
  A.K = make_uf_ctx<double>(0.0);
  A.P.nbox = p.n_boxes;
  for (int d = 0; d < 3; d++) {
    A.P.per[d] = d < 2 && c.L[d] > 0;
    A.P.L[d] = c.L[d];
    A.P.invL[d] = A.P.per[d] ? 1.0 / c.L[d] : 0.0;
  }
  A.inva = 1.0 / c.a;
  A.box = make_box_table<double>(A.P, 1.0 / c.a);
  const long rows_all = n_pad / (TPB * RMB_SYM_NT);
  const long rows = (rows_all - row_first + row_stride - 1) / row_stride;
  if (rows > 0) {
    // rows are triangular (1 .. n_tiles tiles): THIS launch (a rank's share of the rows) is cut into equal-size
    // items of `chunk` tiles, one CTA each, rows paired long + short so that the grid is rectangular
    constexpr int NT = RMB_SYM_NT;
    A.rows = (int)rows;
    A.chunk = sym_chunk_for(rows, row_first, row_stride, TPB * NT, TILE, p.n_tiles, (long)sm_count() * RMB_SYM_RESIDENT,
                            RMB_SYM_ITEMS, RMB_SYM_CHUNK_MIN, RMB_SYM_CHUNK_MAX);
    const int gy = sym_grid_y(A.rows, row_first, row_stride, TPB * NT, TILE, p.n_tiles, A.chunk);
    if (gy > 65535) return fail(RMB_E_ARG, "pair-symmetric product: too many chunks per row pair");
    const dim3 grid((unsigned)((rows + 1) / 2), (unsigned)gy);
#define RMB_SYM_LAUNCH(B)                                                                  \
    do {                                                                                   \
      if (wall && per) sym_kernel<B, true, NT, true, true><<<grid, TPB, 0, c.stream>>>(A);   \
      else if (wall) sym_kernel<B, true, NT, true, false><<<grid, TPB, 0, c.stream>>>(A);    \
      else if (per) sym_kernel<B, true, NT, false, true><<<grid, TPB, 0, c.stream>>>(A);     \
      else sym_kernel<B, true, NT, false, false><<<grid, TPB, 0, c.stream>>>(A);             \
    } while (0)
    if (c.block == RMB_UF_UT) {
      RMB_SYM_LAUNCH(UF_UT);
    } else if (c.block == RMB_WT) {
      RMB_SYM_LAUNCH(WT);
    } else if (c.block == RMB_UT || (c.block == RMB_WF && !wall)) {   // free space: WF is the same function as UT
      RMB_SYM_LAUNCH(UT);
    } else if (c.block == RMB_WF) {
      if (per) sym_kernel<WF, true, NT, true, true><<<grid, TPB, 0, c.stream>>>(A);
      else sym_kernel<WF, true, NT, true, false><<<grid, TPB, 0, c.stream>>>(A);
    } else if (per && wall) sym_kernel<UF, true, NT, true, true><<<grid, TPB, 0, c.stream>>>(A);
    else if (per) sym_kernel<UF, true, NT, false, true><<<grid, TPB, 0, c.stream>>>(A);
    else if (!wall) sym_kernel<UF, true, NT, false, false><<<grid, TPB, 0, c.stream>>>(A);
    else if (c.flags & RMB_NO_TMA) sym_kernel<UF, false, NT, true, false><<<grid, TPB, 0, c.stream>>>(A);
    else sym_kernel<UF, true, NT, true, false><<<grid, TPB, 0, c.stream>>>(A);
#undef RMB_SYM_LAUNCH
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(e, "pair-symmetric product: launch");
  return 0;
}

// Step 2: out = B (acc + self term) / (8 pi eta a^k) for a target range, from the caller's arrays and acc only.
int sym_finish(const Call& c, const double* acc)
{
  const unsigned nb = (unsigned)((c.n_tgt + 127) / 128);
  const int in_plane = (c.flags & RMB_IN_PLANE) ? 1 : 0;
  const double norm = block_norm(c.block, c.eta, c.a);
#define RMB_SYM_FINISH(B, W) sym_finish_kernel<B, W><<<nb, 128, 0, c.stream>>>( \
      c.r, c.va, c.vb, acc, c.tgt_off, c.n_tgt, c.a, norm, in_plane, c.out)
  const bool wall = c.wall == RMB_SINGLE_WALL;
  if (c.block == RMB_UF_UT) { if (wall) RMB_SYM_FINISH(UF_UT, true); else RMB_SYM_FINISH(UF_UT, false); }
  else if (c.block == RMB_UT) { if (wall) RMB_SYM_FINISH(UT, true); else RMB_SYM_FINISH(UT, false); }
  else if (c.block == RMB_WF) { if (wall) RMB_SYM_FINISH(WF, true); else RMB_SYM_FINISH(WF, false); }
  else if (c.block == RMB_WT) { if (wall) RMB_SYM_FINISH(WT, true); else RMB_SYM_FINISH(WT, false); }
  else if (wall) RMB_SYM_FINISH(UF, true);
  else RMB_SYM_FINISH(UF, false);
#undef RMB_SYM_FINISH
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(e, "pair-symmetric product: finish");
  return 0;
}

int launch_symmetric(const Call& c)
{
  double* acc = reinterpret_cast<double*>(c.ws + c.plan.off_acc);
  const int rc = sym_partial(c, acc, 0, 1);
  return rc != 0 ? rc : sym_finish(c, acc);
}

template <typename T, int BLOCK>
int launch_wp(const Call& c, bool per)
{
  if (c.wall == RMB_SINGLE_WALL) return per ? launch<T, BLOCK, true, true>(c) : launch<T, BLOCK, true, false>(c);
  return per ? launch<T, BLOCK, false, true>(c) : launch<T, BLOCK, false, false>(c);
}

template <typename T>
int launch_block(const Call& c, bool per)
{
  switch (c.block) {
    case RMB_UF: return launch_wp<T, UF>(c, per);
    case RMB_UT: return launch_wp<T, UT>(c, per);
    case RMB_WF: return launch_wp<T, WF>(c, per);
    case RMB_WT: return launch_wp<T, WT>(c, per);
    case RMB_UF_UT: return launch_wp<T, UF_UT>(c, per);
    case UF_FS: return per ? launch<T, UF_FS, true, true>(c) : launch<T, UF_FS, true, false>(c);
  }
  return fail(RMB_E_ARG, "unknown block");
}

// ------------------------------------------------------------ FMA probe ----
template <typename T>
__global__ void fma_peak_kernel(int iters, T* sink)
{
  T a0 = T(threadIdx.x) * T(1e-3), a1 = a0 + T(1), a2 = a0 + T(2), a3 = a0 + T(3);
  T a4 = a0 + T(4), a5 = a0 + T(5), a6 = a0 + T(6), a7 = a0 + T(7);
  const T m = T(0.999999), c = T(1e-7);
  for (int i = 0; i < iters; i++) {
    a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
    a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
  }
  const T s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == T(-1)) sink[0] = s;
}

}  // namespace

namespace rmb {
// rmb_last_error() is shared by every translation unit of the library
void set_last_error(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); }
}  // namespace rmb

// ------------------------------------------------------------------ C ABI --
extern "C" {

int rmb_version(void) { return RMB_VERSION; }
const char* rmb_last_error(void) { return g_err; }
int rmb_device_sm_count(void) { return sm_count(); }

size_t rmb_workspace_bytes(int block, int precision, long n_targets, long n_sources, const double* L)
{
  Plan p;
  if (block < 0 || block > RMB_UF_UT || n_targets < 0 || n_sources < 0) return 0;
  if (make_plan(block, precision, n_targets, n_sources, L, p) != 0) return 0;
  return p.total;
}

int rmb_launches_per_product(int block, int wall, int precision, int flags, long n_targets, long n_sources,
                             const double* L)
{
  Plan p;
  if (block < 0 || block > RMB_UF_UT || n_targets <= 0 || n_sources <= 0) return 0;
  if (make_plan(block, precision, n_targets, n_sources, L, p) != 0) return 0;
  if (use_symmetric(block, wall, precision, flags, n_targets, n_sources, L)) return 3;     // prepare_kernel + sym_kernel + sym_finish_kernel (+ one memset)
  return 2 + (p.n_split > 1 ? 1 : 0);   // prepare_kernel + mf_kernel (+ finish_kernel when sources are split)
}

int rmb_mobility_product(int block, int wall, int precision, int flags, long n_targets, long n_sources,
                         long target_offset, const double* r_sources, const double* vec_a, const double* vec_b,
                         double eta, double a, const double* L, double* out, void* workspace,
                         size_t workspace_bytes, void* stream)
{
  if (block < 0 || block > RMB_UF_UT) return fail(RMB_E_ARG, "block must be RMB_UF..RMB_UF_UT");
  if (wall != RMB_NO_WALL && wall != RMB_SINGLE_WALL && wall != RMB_FREE_SURFACE)
    return fail(RMB_E_ARG, "wall must be RMB_NO_WALL, RMB_SINGLE_WALL or RMB_FREE_SURFACE");
  if (wall == RMB_FREE_SURFACE && block != RMB_UF) return fail(RMB_E_ARG, "RMB_FREE_SURFACE exists for RMB_UF only");
  if (precision != RMB_FP64 && precision != RMB_FP32) return fail(RMB_E_ARG, "precision must be 0 or 1");
  if (n_targets < 0 || n_sources < 0 || target_offset < 0 || target_offset + n_targets > n_sources)
    return fail(RMB_E_ARG, "target range must lie inside [0, n_sources)");
  if (n_targets == 0 || n_sources == 0) return 0;
  if (!r_sources || !vec_a || !out || !workspace) return fail(RMB_E_ARG, "null device pointer");
  if (block == RMB_UF_UT && !vec_b) return fail(RMB_E_ARG, "RMB_UF_UT needs vec_b (torque)");
  if (!(a > 0) || !(eta > 0)) return fail(RMB_E_ARG, "a and eta must be positive");
  if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0) return fail(RMB_E_ARG, "workspace must be 16-byte aligned");
  Call c;
  c.block = (wall == RMB_FREE_SURFACE) ? (int)UF_FS : block; c.wall = wall; c.precision = precision; c.flags = flags;
  c.n_tgt = n_targets; c.n_src = n_sources; c.tgt_off = target_offset;
  c.r = r_sources; c.va = vec_a; c.vb = vec_b; c.eta = eta; c.a = a;
  for (int d = 0; d < 3; d++) c.L[d] = L ? L[d] : 0.0;
  c.out = out; c.ws = static_cast<char*>(workspace); c.stream = static_cast<cudaStream_t>(stream);
  const int rc = make_plan(block, precision, n_targets, n_sources, c.L, c.plan);
  if (rc != 0) return rc;
  if (workspace_bytes < c.plan.total) return fail(RMB_E_WORKSPACE, "workspace smaller than rmb_workspace_bytes()");
  const bool per = c.L[0] > 0 || c.L[1] > 0 || c.L[2] > 0;
  if (use_symmetric(block, wall, precision, flags, n_targets, n_sources, c.L)) return launch_symmetric(c);
  return precision == RMB_FP32 ? launch_block<float>(c, per) : launch_block<double>(c, per);
}

size_t rmb_symmetric_acc_len(long n)
{
  return n <= 0 ? 0 : (size_t)((n + TILE - 1) / TILE) * TILE * 3;
}

static int sym_call(Call& c, int block, int wall, int flags, long n, const double* r, const double* force,
                    const double* vec_b, double eta, double a,
                    const double* L, void* workspace, size_t workspace_bytes, void* stream)
{
  if (n <= 0) return fail(RMB_E_ARG, "n must be positive");
  if (block < RMB_UF || block > RMB_UF_UT) return fail(RMB_E_ARG, "block must be RMB_UF..RMB_UF_UT");
  if (wall != RMB_SINGLE_WALL && wall != RMB_NO_WALL)
    return fail(RMB_E_ARG, "the pair-symmetric product exists for RMB_SINGLE_WALL and RMB_NO_WALL");
  if (L && L[2] > 0) return fail(RMB_E_ARG, "the pair-symmetric product cannot be periodic in z");
  if (!workspace) return fail(RMB_E_ARG, "null device pointer");
  if (!(a > 0) || !(eta > 0)) return fail(RMB_E_ARG, "a and eta must be positive");
  if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0) return fail(RMB_E_ARG, "workspace must be 16-byte aligned");
  c.block = block; c.wall = wall; c.precision = RMB_FP64; c.flags = flags;
  c.n_tgt = n; c.n_src = n; c.tgt_off = 0;
  c.r = r; c.va = force; c.vb = vec_b; c.eta = eta; c.a = a;
  for (int d = 0; d < 3; d++) c.L[d] = L ? L[d] : 0.0;
  c.out = nullptr; c.ws = static_cast<char*>(workspace); c.stream = static_cast<cudaStream_t>(stream);
  const int rc = make_plan(block, RMB_FP64, n, n, c.L, c.plan);
  if (rc != 0) return rc;
  if (workspace_bytes < c.plan.total) return fail(RMB_E_WORKSPACE, "workspace smaller than rmb_workspace_bytes()");
  return 0;
}

int rmb_symmetric_partial(int block, int wall, int flags, long n, int rank, int world, const double* r,
                          const double* force, const double* vec_b, double eta, double a, const double* L, double* acc, void* workspace, size_t workspace_bytes,
                          void* stream)
{
  if (world < 1 || rank < 0 || rank >= world) return fail(RMB_E_ARG, "0 <= rank < world required");
  if (!r || !force || !acc) return fail(RMB_E_ARG, "null device pointer");
  if (block == RMB_UF_UT && !vec_b) return fail(RMB_E_ARG, "RMB_UF_UT needs vec_b (torque)");
  Call c;
  const int rc = sym_call(c, block, wall, flags, n, r, force, vec_b, eta, a, L, workspace, workspace_bytes, stream);
  return rc != 0 ? rc : sym_partial(c, acc, rank, world);
}

int rmb_symmetric_finish(int block, int wall, int flags, long n, long target_offset, long n_targets, const double* r,
                         const double* vec_a, const double* vec_b, const double* acc, double eta, double a, double* out,
                         void* stream)
{
  if (n <= 0) return fail(RMB_E_ARG, "n must be positive");
  if (block < RMB_UF || block > RMB_UF_UT) return fail(RMB_E_ARG, "block must be RMB_UF..RMB_UF_UT");
  if (wall != RMB_SINGLE_WALL && wall != RMB_NO_WALL)
    return fail(RMB_E_ARG, "the pair-symmetric product exists for RMB_SINGLE_WALL and RMB_NO_WALL");
  if (target_offset < 0 || n_targets < 0 || target_offset + n_targets > n)
    return fail(RMB_E_ARG, "target range must lie inside [0, n)");
  if (n_targets == 0) return 0;
  if (!r || !vec_a || !acc || !out) return fail(RMB_E_ARG, "null device pointer");
  if (block == RMB_UF_UT && !vec_b) return fail(RMB_E_ARG, "RMB_UF_UT needs vec_b (torque)");
  if (!(a > 0) || !(eta > 0)) return fail(RMB_E_ARG, "a and eta must be positive");
  Call c;
  c.block = block; c.wall = wall; c.precision = RMB_FP64; c.flags = flags;
  c.n_tgt = n_targets; c.n_src = n; c.tgt_off = target_offset;
  c.r = r; c.va = vec_a; c.vb = vec_b; c.eta = eta; c.a = a;
  for (int d = 0; d < 3; d++) c.L[d] = 0.0;
  c.out = out; c.ws = nullptr; c.stream = static_cast<cudaStream_t>(stream);
  return sym_finish(c, acc);
}

// ---- allocation-free variant (no counterpart in the reference) ----
// Scratch of rmb_mobility_product_static: ONE file-scope device array per GPU that loads the library, held until the
// process exits.  Host code must not use the NAME as a pointer (it is a host-side stand-in): the device address is
// fetched once per device with cudaGetSymbolAddress, with that device current, and cached.
#define RMB_STATIC_WORKSPACE_BYTES (64u << 20)
static __device__ __align__(256) unsigned char g_static_workspace[RMB_STATIC_WORKSPACE_BYTES];
static void* g_static_workspace_addr[64] = {nullptr};   // per device ordinal

static int static_workspace(void** ws)
{
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail_cuda(e, "rmb_mobility_product_static: cudaGetDevice");
  if (dev < 0 || dev >= 64) return fail(RMB_E_ARG, "rmb_mobility_product_static: device ordinal beyond 63");
  if (!g_static_workspace_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_workspace);
    if (e != cudaSuccess || !p) return fail_cuda(e, "rmb_mobility_product_static: cudaGetSymbolAddress");
    g_static_workspace_addr[dev] = p;
  }
  *ws = g_static_workspace_addr[dev];
  return 0;
}

size_t rmb_static_workspace_max(void) { return RMB_STATIC_WORKSPACE_BYTES; }

int rmb_mobility_product_static(int block, int wall, int precision, int flags, long n_targets, long n_sources,
                                long target_offset, const double* r_sources, const double* vec_a, const double* vec_b,
                                double eta, double a, const double* L, double* out, void* stream)
{
  if (block < 0 || block > RMB_UF_UT) return fail(RMB_E_ARG, "block must be RMB_UF..RMB_UF_UT");
  if (precision != RMB_FP64 && precision != RMB_FP32) return fail(RMB_E_ARG, "precision must be 0 or 1");
  if (n_targets < 0 || n_sources < 0) return fail(RMB_E_ARG, "negative size");
  if (n_targets == 0 || n_sources == 0) return 0;
  Plan p;
  const int rc = make_plan(block, precision, n_targets, n_sources, L, p);
  if (rc != 0) return rc;
  if (p.total > RMB_STATIC_WORKSPACE_BYTES) {
    snprintf(g_err, sizeof(g_err),
             "rmb_mobility_product_static: this product needs %zu bytes of scratch, the static array holds %u "
             "(about 450,000 blobs for an fp64 single-vector block): use rmb_mobility_product with a caller-owned workspace",
             p.total, (unsigned)RMB_STATIC_WORKSPACE_BYTES);
    return RMB_E_WORKSPACE;
  }
  void* ws = nullptr;
  const int rs = static_workspace(&ws);
  if (rs != 0) return rs;
  // everything that must start at zero is cleared by the product itself (the pair-symmetric accumulator by a
  // cudaMemsetAsync on `stream`; records, damping factors and partial sums are fully overwritten)
  return rmb_mobility_product(block, wall, precision, flags, n_targets, n_sources, target_offset, r_sources, vec_a,
                              vec_b, eta, a, L, out, ws, RMB_STATIC_WORKSPACE_BYTES, stream);
}

int rmb_fma_peak(int precision, int iters, double* flops, double* elapsed_ms)
{
  const int sms = sm_count();
  if (sms <= 0) return sms;
  if (iters <= 0 || !flops) return fail(RMB_E_ARG, "iters > 0 and flops != NULL required");
  void* sink = nullptr;
  cudaError_t e = cudaMalloc(&sink, 64);
  if (e != cudaSuccess) return fail_cuda(e, "cudaMalloc");
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = sms * 8, threads = 256;
  for (int rep = 0; rep < 2; rep++) {  // first pass warms up
    cudaEventRecord(e0);
    if (precision == RMB_FP32) fma_peak_kernel<float><<<blocks, threads>>>(iters, (float*)sink);
    else fma_peak_kernel<double><<<blocks, threads>>>(iters, (double*)sink);
    cudaEventRecord(e1);
    e = cudaEventSynchronize(e1);
    if (e != cudaSuccess) break;
  }
  float ms = 0;
  if (e == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(sink);
  if (e != cudaSuccess) return fail_cuda(e, "rmb_fma_peak");
  *flops = 2.0 * 8.0 * (double)iters * blocks * threads / (ms * 1e-3);
  if (elapsed_ms) *elapsed_ms = ms;
  return 0;
}

}  // extern "C"
