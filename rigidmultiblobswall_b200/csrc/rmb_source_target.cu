// rmb_source_target.cu -- UF product between separate source and target sets
// with per-blob radii (B200, sm_100a) and its C ABI entry
// rmb_mobility_source_target().
//
// Replaces mobility_numba.mobility_trans_times_force_source_target_numba
// (/root/reference/mobility/mobility_numba.py:1480-1658), its free-surface twin
// (:1941-2092), the PyCUDA kernel velocity_from_force_source_target
// (mobility_pycuda.py:1394-1560) and the clamp / damping wrappers of
// mobility/mobility.py:494-530, :551-607, :1407.
//   st_prepare_kernel : O(Ns) -- per-radius height clamp, damping B, packing
//   st_kernel         : O(Nt Ns n_boxes) -- one thread per target, source tiles
//                       streamed through shared memory by bulk async copies
//   st_finish_kernel  : O(Nt n_split) -- sum of source-split partials
// Same tile pipeline as rmb_mobility.cu (tile_pipe.cuh); per-pair algebra in
// st_math.cuh.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/rmb_mobility.h"
#include "mf_core.cuh"
#include "st_math.cuh"
#include "tile_pipe.cuh"

namespace rmb { void set_last_error(const char* msg); }  // rmb_mobility.cu
using namespace rmb;

namespace {

constexpr int TPB = 128;   // targets (threads) per CTA
constexpr int TILE = 256;  // sources per shared-memory tile
constexpr int STAGES = 2;
constexpr int W = 8;       // words per StRec

int fail(int code, const char* msg)
{
  set_last_error(msg);
  return code;
}

template <typename T>
__global__ void st_prepare_kernel(long n, long n_pad, const double* __restrict__ r, const double* __restrict__ f,
                                  const double* __restrict__ radius, int wall, StRec<T>* __restrict__ rec)
{
  const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_pad) return;
  StRec<T> o;
  if (j < n) {
    st_pack<T>(wall != 0, r[3 * j], r[3 * j + 1], r[3 * j + 2], radius[j], f[3 * j], f[3 * j + 1], f[3 * j + 2], o);
  } else {
    // padding: zero force at a finite distance from everything (radius 1, far below the targets)
    o.x = o.y = T(0); o.z = T(-1.0e30); o.fx = o.fy = o.fz = T(0); o.b = T(1); o.ib = T(1);
  }
  rec[j] = o;
}

template <typename T>
struct StArgs {
  const StRec<T>* src;   // n_tiles * TILE records
  const double* r_trg;   // [n_trg][3]
  const double* rad_trg; // [n_trg]
  long n_src, n_trg;
  Periodic<T> P;
  int n_tiles, n_vt, vt_per_split, n_split;
  T* partial;            // [n_split][n_trg][3] when n_split > 1
  double* out;           // [n_trg][3]
  double norm;           // 1/(8 pi eta)
  StConsts<T> K;         // literals of st_math.cuh (constant bank)
};

template <typename T, int WALLMODE, bool PERIODIC, bool TMA>
__global__ void __launch_bounds__(TPB) st_kernel(const __grid_constant__ StArgs<T> A)
{
  constexpr uint32_t TILE_BYTES = TILE * W * sizeof(T);
  __shared__ alignas(128) T tile[TMA ? STAGES : 1][TILE * W];
  __shared__ alignas(8) uint64_t full[STAGES];

  const int tid = threadIdx.x;
  const long li = (long)blockIdx.x * TPB + tid;
  const bool valid = li < A.n_trg;
  const long ti = valid ? li : A.n_trg - 1;

  StTarget<T> t;
  const double Bt = st_target<T>(WALLMODE == ST_WALL, A.r_trg[3 * ti], A.r_trg[3 * ti + 1], A.r_trg[3 * ti + 2],
                                 A.rad_trg[ti], t);
  const int v0 = blockIdx.y * A.vt_per_split;
  const int v1 = min(v0 + A.vt_per_split, A.n_vt);
  const T* srcw = reinterpret_cast<const T*>(A.src);

  if (TMA) {
    if (tid == 0) {
#pragma unroll
      for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
      fence_barrier_init();
    }
    __syncthreads();
    if (tid == 0 && v0 < v1) {
      mbar_expect_tx(&full[0], TILE_BYTES);
      bulk_g2s(tile[0], srcw + (size_t)(v0 % A.n_tiles) * TILE * W, TILE_BYTES, &full[0]);
    }
  }

  T ux = 0, uy = 0, uz = 0;
  [[maybe_unused]] double wx = 0, wy = 0, wz = 0;   // fp32: fp64 running sum over tiles (see mf_kernel)
  int stage = 0;
  uint32_t phases = 0;
  for (int v = v0; v < v1; v++) {
    const int box = v / A.n_tiles;
    const int tl = v - box * A.n_tiles;
    if (TMA) {
      if (tid == 0 && v + 1 < v1) {
        mbar_expect_tx(&full[stage ^ 1], TILE_BYTES);
        bulk_g2s(tile[stage ^ 1], srcw + (size_t)((v + 1) % A.n_tiles) * TILE * W, TILE_BYTES, &full[stage ^ 1]);
      }
      mbar_wait(&full[stage], (phases >> stage) & 1u);
      phases ^= (1u << stage);
    } else {
      const T* gsrc = srcw + (size_t)tl * TILE * W;
      for (int k = tid; k < TILE * W; k += TPB) tile[0][k] = gsrc[k];
      __syncthreads();
    }
    T sx = 0, sy = 0, sz = 0;
    bool central = true;
    if (PERIODIC) box_shift(A.P, box, sx, sy, sz, central);
    const int cnt = (int)min((long)TILE, A.n_src - (long)tl * TILE);
    const T* tp = tile[TMA ? stage : 0];
#pragma unroll 2
    for (int j = 0; j < cnt; j++) {
      T w[W];
      load_rec<W>(tp + j * W, w);
      T rx = t.x - w[0], ry = t.y - w[1], rz = t.z - w[2];
      if (PERIODIC) {
        if (A.P.per[0]) rx = wrap_shift(rx, A.P.L[0], A.P.invL[0], sx);
        if (A.P.per[1]) ry = wrap_shift(ry, A.P.L[1], A.P.invL[1], sy);
        if (A.P.per[2]) rz = wrap_shift(rz, A.P.L[2], A.P.invL[2], sz);
      }
      pair_st<T, WALLMODE>(A.K, t, rx, ry, rz, t.z + w[2], w[2], w[6], w[7], w[3], w[4], w[5], ux, uy, uz);
    }
    if constexpr (sizeof(T) == 4) {
      wx += (double)ux; wy += (double)uy; wz += (double)uz;
      ux = uy = uz = T(0);
    }
    __syncthreads();
    if (TMA) stage ^= 1;
  }

  if (!valid) return;
  if constexpr (sizeof(T) == 8) { wx = ux; wy = uy; wz = uz; }
  if (A.n_split == 1) {
    const double sc = A.norm * Bt;
    A.out[3 * li] = wx * sc;
    A.out[3 * li + 1] = wy * sc;
    A.out[3 * li + 2] = wz * sc;
  } else {
    T* p = A.partial + ((size_t)blockIdx.y * A.n_trg + li) * 3;
    p[0] = T(wx); p[1] = T(wy); p[2] = T(wz);
  }
}

template <typename T, int WALLMODE>
__global__ void st_finish_kernel(const StArgs<T> A)
{
  const long li = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (li >= A.n_trg) return;
  double dx = 0, dy = 0, dz = 0;
  for (int s = 0; s < A.n_split; s++) {
    const T* p = A.partial + ((size_t)s * A.n_trg + li) * 3;
    dx += (double)p[0]; dy += (double)p[1]; dz += (double)p[2];
  }
  double B = 1.0;
  if (WALLMODE == ST_WALL) {
    const double z = A.r_trg[3 * li + 2], a = A.rad_trg[li];
    B = (z < a) ? z / a : 1.0;
  }
  const double sc = A.norm * B;
  A.out[3 * li] = dx * sc; A.out[3 * li + 1] = dy * sc; A.out[3 * li + 2] = dz * sc;
}

// ---------------------------------------------------------- host dispatch --
struct StPlan {
  int n_tiles, n_boxes, n_vt, n_split, vt_per_split, tgt_blocks;
  size_t rec_bytes, off_partial, total;
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Same decomposition rule as rmb_mobility.cu: sources are split over blockIdx.y
// only when the target CTAs alone would give fewer than 10 waves of resident CTAs.
int make_plan(int precision, long n_trg, long n_src, const double* L, StPlan& p)
{
  const int sms = rmb_device_sm_count();
  if (sms <= 0) return sms;
  const size_t word = precision == RMB_FP32 ? 4 : 8;
  p.n_tiles = (int)((n_src + TILE - 1) / TILE);
  p.n_boxes = 1;
  if (L) for (int d = 0; d < 3; d++) if (L[d] > 0) p.n_boxes *= 3;
  p.n_vt = p.n_tiles * p.n_boxes;
  p.tgt_blocks = (int)((n_trg + TPB - 1) / TPB);
  const int resident = sms * (768 / TPB);
  int split = (10 * resident + p.tgt_blocks - 1) / (p.tgt_blocks > 0 ? p.tgt_blocks : 1);
  if (split > p.n_vt / 4) split = p.n_vt / 4;
  if (split < 1) split = 1;
  p.vt_per_split = (p.n_vt + split - 1) / split;
  if (p.vt_per_split < 1) p.vt_per_split = 1;   // empty source set
  p.n_split = (p.n_vt + p.vt_per_split - 1) / p.vt_per_split;
  p.rec_bytes = (size_t)p.n_tiles * TILE * W * word;
  p.off_partial = align_up(p.rec_bytes, 256);
  p.total = p.off_partial + (p.n_split > 1 ? (size_t)p.n_split * n_trg * 3 * word : 0);
  return 0;
}

struct StCall {
  int flags;
  long n_trg, n_src;
  const double *r_trg, *r_src, *force, *rad_trg, *rad_src;
  double eta;
  double L[3];
  double* out;
  char* ws;
  cudaStream_t stream;
  StPlan plan;
};

template <typename T, int WALLMODE, bool PERIODIC>
int launch(const StCall& c)
{
  const StPlan& p = c.plan;
  StRec<T>* rec = reinterpret_cast<StRec<T>*>(c.ws);
  const long n_pad = (long)p.n_tiles * TILE;
  st_prepare_kernel<T><<<(unsigned)((n_pad + 255) / 256), 256, 0, c.stream>>>(c.n_src, n_pad, c.r_src, c.force,
                                                                            c.rad_src, WALLMODE == ST_WALL, rec);
  StArgs<T> A;
  A.src = rec; A.r_trg = c.r_trg; A.rad_trg = c.rad_trg;
  A.n_src = c.n_src; A.n_trg = c.n_trg;
  A.P.nbox = p.n_boxes;
  for (int d = 0; d < 3; d++) {
    A.P.per[d] = PERIODIC && c.L[d] > 0;
    A.P.L[d] = T(c.L[d]);
    A.P.invL[d] = A.P.per[d] ? T(1.0 / c.L[d]) : T(0);
  }
  A.n_tiles = p.n_tiles; A.n_vt = p.n_vt; A.vt_per_split = p.vt_per_split; A.n_split = p.n_split;
  A.partial = reinterpret_cast<T*>(c.ws + p.off_partial);
  A.out = c.out;
  A.norm = 1.0 / (8.0 * 3.14159265358979323846 * c.eta);
  A.K = make_st_consts<T>();
  const dim3 grid((unsigned)p.tgt_blocks, (unsigned)p.n_split);
  if (c.flags & RMB_NO_TMA) st_kernel<T, WALLMODE, PERIODIC, false><<<grid, TPB, 0, c.stream>>>(A);
  else st_kernel<T, WALLMODE, PERIODIC, true><<<grid, TPB, 0, c.stream>>>(A);
  if (p.n_split > 1) st_finish_kernel<T, WALLMODE><<<(unsigned)((c.n_trg + 127) / 128), 128, 0, c.stream>>>(A);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    char msg[256];
    snprintf(msg, sizeof(msg), "rmb_mobility_source_target launch: %s", cudaGetErrorString(e));
    return fail((int)e, msg);
  }
  return 0;
}

template <typename T>
int launch_wall(const StCall& c, int wall, bool per)
{
  switch (wall) {
    case RMB_NO_WALL: return per ? launch<T, ST_NO_WALL, true>(c) : launch<T, ST_NO_WALL, false>(c);
    case RMB_SINGLE_WALL: return per ? launch<T, ST_WALL, true>(c) : launch<T, ST_WALL, false>(c);
    default: return per ? launch<T, ST_FREE_SURFACE, true>(c) : launch<T, ST_FREE_SURFACE, false>(c);
  }
}

}  // namespace

extern "C" {

size_t rmb_source_target_workspace_bytes(int precision, long n_targets, long n_sources, const double* L)
{
  StPlan p;
  if (n_targets < 0 || n_sources < 0) return 0;
  if (make_plan(precision, n_targets, n_sources, L, p) != 0) return 0;
  return p.total;
}

int rmb_mobility_source_target(int wall, int precision, int flags, long n_targets, long n_sources,
                               const double* r_targets, const double* r_sources, const double* force,
                               const double* radius_targets, const double* radius_sources, double eta,
                               const double* L, double* out, void* workspace, size_t workspace_bytes, void* stream)
{
  if (wall != RMB_NO_WALL && wall != RMB_SINGLE_WALL && wall != RMB_FREE_SURFACE)
    return fail(RMB_E_ARG, "wall must be RMB_NO_WALL, RMB_SINGLE_WALL or RMB_FREE_SURFACE");
  if (precision != RMB_FP64 && precision != RMB_FP32) return fail(RMB_E_ARG, "precision must be 0 or 1");
  if (n_targets < 0 || n_sources < 0) return fail(RMB_E_ARG, "negative size");
  if (n_targets == 0) return 0;
  if (!out) return fail(RMB_E_ARG, "null device pointer");
  if (n_sources == 0) {  // u = 0 (the reference returns zeros for an empty source set)
    const cudaError_t e = cudaMemsetAsync(out, 0, sizeof(double) * 3 * n_targets, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : (int)e;
  }
  if (!r_targets || !r_sources || !force || !radius_targets || !radius_sources || !workspace)
    return fail(RMB_E_ARG, "null device pointer");
  if (!(eta > 0)) return fail(RMB_E_ARG, "eta must be positive");
  if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0) return fail(RMB_E_ARG, "workspace must be 16-byte aligned");
  StCall c;
  c.flags = flags; c.n_trg = n_targets; c.n_src = n_sources;
  c.r_trg = r_targets; c.r_src = r_sources; c.force = force; c.rad_trg = radius_targets; c.rad_src = radius_sources;
  c.eta = eta;
  for (int d = 0; d < 3; d++) c.L[d] = L ? L[d] : 0.0;
  c.out = out; c.ws = static_cast<char*>(workspace); c.stream = static_cast<cudaStream_t>(stream);
  const int rc = make_plan(precision, n_targets, n_sources, c.L, c.plan);
  if (rc != 0) return rc;
  if (workspace_bytes < c.plan.total)
    return fail(RMB_E_WORKSPACE, "workspace smaller than rmb_source_target_workspace_bytes()");
  const bool per = c.L[0] > 0 || c.L[1] > 0 || c.L[2] > 0;
  return precision == RMB_FP32 ? launch_wall<float>(c, wall, per) : launch_wall<double>(c, wall, per);
}

}  // extern "C"
