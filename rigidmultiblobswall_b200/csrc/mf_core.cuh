// mf_core.cuh -- record layout, image wrap and the per-pair dispatch shared by
// the CUDA kernels (rmb_mobility.cu) and the host-side algebra check
// (tests/host_pair_harness.cpp, g++ -DRMB_HOST_ONLY).
#pragma once
#include "pair_math.cuh"

namespace rmb {

enum Block { UF = 0, UT = 1, WF = 2, WT = 3, UF_UT = 4, UF_FS = 5 /* UF above a free (stress-free) surface */ };

// One packed source blob in HBM / shared memory: position followed by NV
// 3-vectors (NV = 1: force or torque; NV = 2: force and torque/a for the
// fused UF+UT product), padded to a multiple of 16 bytes so that a record is
// read with 16-byte shared-memory broadcasts only:
//   fp64 NV=1: 6 words = 48 B (3 x LDS.128)   fp64 NV=2: 10 words = 80 B
//   fp32 NV=1: 8 words = 32 B (2 x LDS.128)   fp32 NV=2: 12 words = 48 B
// Positions are in units of a for the open-boundary kernels and unscaled for
// the pseudo-periodic ones (the wrap is done in the caller's length unit,
// exactly where the reference does it).
template <typename T, int NV> struct RecWords;
template <> struct RecWords<double, 1> { static constexpr int value = 6; };
template <> struct RecWords<double, 2> { static constexpr int value = 10; };
template <> struct RecWords<float, 1> { static constexpr int value = 8; };
template <> struct RecWords<float, 2> { static constexpr int value = 12; };

template <typename T, int NV>
struct alignas(16) SrcRec {
  T x, y, z;
  T v[RecWords<T, NV>::value - 3];
};

// Pseudo-periodic parameters (host-filled).  L[d] <= 0: not periodic in d.
template <typename T>
struct Periodic {
  T L[3];
  T invL[3];
  int per[3];
  int nbox;  // 3^(number of periodic directions)
};

// mobility_numba.py:184-192:  r -= int(r/L + 0.5*sign(r)) * L ; r += box*L.
// int() truncates toward zero; copysign(0.5, 0) = 0.5 truncates to 0 like
// sign(0) = 0 does.
template <typename T>
RMB_HD T wrap_shift(T r, T L, T invL, T shift)
{
#if defined(RMB_HOST_ONLY)
  const T n = std::trunc(RMB_FMA(r, invL, std::copysign(T(0.5), r)));
#else
  const T n = trunc(RMB_FMA(r, invL, copysign(T(0.5), r)));
#endif
  return RMB_FMA(-n, L, r) + shift;
}

// Decode image box index b in [0, nbox) into the shift vector (box * L).
// Ordering: x slowest, z fastest, each periodic direction running -1, 0, +1
// (same nesting as the reference loops).  central = all components zero.
template <typename T>
RMB_HD void box_shift(const Periodic<T>& P, int b, T& sx, T& sy, T& sz, bool& central)
{
  int k[3];
  for (int d = 2; d >= 0; --d) {
    if (P.per[d]) { k[d] = (b % 3) - 1; b /= 3; } else k[d] = 0;
  }
  sx = T(k[0]) * P.L[0]; sy = T(k[1]) * P.L[1]; sz = T(k[2]) * P.L[2];
  central = (k[0] == 0) && (k[1] == 0) && (k[2] == 0);
}

// Target-side loop invariants.
template <typename T>
struct Target {
  T x, y, z;  // same unit as the source records
  T h;        // z / a
};

// Geometry of one pair.  PERIODIC: positions unscaled, wrap, then 1/a.
template <typename T, bool WALL, bool PERIODIC>
RMB_HD void pair_geometry(const Target<T>& t, T xj, T yj, T zj, const Periodic<T>& P, T sx, T sy, T sz, T inva,
                          PairGeom<T>& g, T& r2, T& rxy2)
{
  g.rx = t.x - xj; g.ry = t.y - yj; g.rz = t.z - zj;
  if (PERIODIC) {
    if (P.per[0]) g.rx = wrap_shift(g.rx, P.L[0], P.invL[0], sx);
    if (P.per[1]) g.ry = wrap_shift(g.ry, P.L[1], P.invL[1], sy);
    if (P.per[2]) g.rz = wrap_shift(g.rz, P.L[2], P.invL[2], sz);
    g.rx *= inva; g.ry *= inva; g.rz *= inva;
    if (WALL) {
      g.Rz = (t.z + zj) * inva;       // no wrap, no shift (mobility_numba.py:242)
      g.h = zj * inva;
    }
  } else {
    if (WALL) {
      g.Rz = t.z + zj;
      g.h = zj;
    }
  }
  rxy2 = RMB_FMA(g.ry, g.ry, g.rx * g.rx);
  r2 = RMB_FMA(g.rz, g.rz, rxy2);
}

// Pseudo-periodic pair in two parts, so that the wrap, the 1/a scaling and the
// record load are paid once per (target, source) and not once per image box.
// Part 1, per source: minimum-image wrap in the caller's length unit (where the
// reference does it, mobility_numba.py:184-192), then 1/a.  The wall term uses
// the unwrapped z_i + z_j (:242).
template <typename T, bool WALL>
RMB_HD void periodic_base(const Target<T>& t, T xj, T yj, T zj, const Periodic<T>& P, T inva, PairGeom<T>& g0)
{
  g0.rx = t.x - xj; g0.ry = t.y - yj; g0.rz = t.z - zj;
  if (P.per[0]) g0.rx = wrap_shift(g0.rx, P.L[0], P.invL[0], T(0));
  if (P.per[1]) g0.ry = wrap_shift(g0.ry, P.L[1], P.invL[1], T(0));
  if (P.per[2]) g0.rz = wrap_shift(g0.rz, P.L[2], P.invL[2], T(0));
  g0.rx *= inva; g0.ry *= inva; g0.rz *= inva;
  g0.Rz = T(0); g0.h = T(0);
  if (WALL) {
    g0.Rz = (t.z + zj) * inva;
    g0.h = zj * inva;
  }
}

// All image boxes of the periodic directions: shift vectors box * L / a in the
// nesting order of the reference loops (x slowest, z fastest, -1, 0, +1 each).
template <typename T>
struct BoxTable {
  T s[27][3];
  int n, central;
};

template <typename T>
inline BoxTable<T> make_box_table(const Periodic<T>& P, double inva)
{
  BoxTable<T> B;
  B.n = P.nbox; B.central = 0;
  for (int b = 0; b < 27; b++) {
    int k[3] = {0, 0, 0}, c = b;
    if (b < P.nbox)
      for (int d = 2; d >= 0; --d)
        if (P.per[d]) { k[d] = (c % 3) - 1; c /= 3; }
    for (int d = 0; d < 3; d++) B.s[b][d] = T(k[d] * (double)P.L[d] * inva);
    if (b < P.nbox && k[0] == 0 && k[1] == 0 && k[2] == 0) B.central = b;
  }
  return B;
}

// Part 2, per image box: add the shift (already in units of a).
template <typename T>
RMB_HD void periodic_box(const PairGeom<T>& g0, const Periodic<T>& P, const T* shift, PairGeom<T>& g, T& r2, T& rxy2)
{
  (void)P;   // the shift of a non-periodic direction is 0: branch-free
  g.Rz = g0.Rz; g.h = g0.h;
  g.rx = g0.rx + shift[0];
  g.ry = g0.ry + shift[1];
  g.rz = g0.rz + shift[2];
  rxy2 = RMB_FMA(g.ry, g.ry, g.rx * g.rx);
  r2 = RMB_FMA(g.rz, g.rz, rxy2);
}

// One pair, one block.  v = the record's vectors (already zeroed by the
// caller for the self pair, whose geometry it also made finite).  g.h is the
// SOURCE height; C.ht is the TARGET height, which the transposed wall term of
// UT uses instead (mobility_numba.py:651).
template <typename T, int BLOCK, bool WALL, bool ZOPEN = false>
RMB_HD void pair_apply(const PairGeom<T>& g, const UfCtx<T>& C, const T* v, bool any_near, T r2, T rxy2, T& ux, T& uy,
                       T& uz)
{
  if (BLOCK == UF || BLOCK == UF_UT) pair_uf<T, WALL, ZOPEN>(g, C, v[0], v[1], v[2], any_near, r2, rxy2, ux, uy, uz);
  if (BLOCK == WF) pair_wf<T, WALL, ZOPEN>(g, C, v[0], v[1], v[2], any_near, r2, rxy2, ux, uy, uz);
  if (BLOCK == WT) pair_wt<T, WALL, ZOPEN>(g, C, v[0], v[1], v[2], any_near, r2, rxy2, ux, uy, uz);
  if (BLOCK == UF_FS) {
    // free surface (mobility_numba.py:1770-1937): RPY(r) f + RPY(R) f', R = (rx, ry, z_i+z_j), f' = (fx, fy, -fz)
    // (the image block enters with the sign pattern [+ + -; + + -; + + -] on its columns, :1918-1926)
    pair_uf<T, false>(g, C, v[0], v[1], v[2], any_near, r2, rxy2, ux, uy, uz);
    PairGeom<T> gi = g;
    gi.rz = g.Rz;
    pair_uf<T, false>(gi, C, v[0], v[1], -v[2], true, RMB_FMA(g.Rz, g.Rz, rxy2), rxy2, ux, uy, uz);
  }
  if (BLOCK == UT || BLOCK == UF_UT) {
    const T* t = (BLOCK == UF_UT) ? v + 3 : v;   // fused: second vector is torque / a
    pair_ut<T, WALL, ZOPEN>(g, C, t[0], t[1], t[2], any_near, r2, rxy2, ux, uy, uz);
  }
}

// Closed-form self term (i == j in the central box), h = z_i/a, v = own vectors.
template <typename T, int BLOCK, bool WALL>
RMB_HD void self_apply(T h, const T* v, T& ux, T& uy, T& uz)
{
  if (BLOCK == UF || BLOCK == UF_UT) self_uf<T, WALL>(h, v[0], v[1], v[2], ux, uy, uz);
  if (BLOCK == UT && WALL) self_ut<T>(h, v[0], v[1], ux, uy);
  if (BLOCK == WF && WALL) self_wf<T>(h, v[0], v[1], ux, uy);
  if (BLOCK == WT) self_wt<T, WALL>(h, v[0], v[1], v[2], ux, uy, uz);
  if (BLOCK == UF_UT && WALL) self_ut<T>(h, v[3], v[4], ux, uy);
  if (BLOCK == UF_FS) {
    // i == j: 4/3 f plus the blob's own image at distance 2h, which is an ordinary (far or near) pair
    self_uf<T, false>(h, v[0], v[1], v[2], ux, uy, uz);
    PairGeom<T> gi;
    gi.rx = T(0); gi.ry = T(0); gi.rz = h + h; gi.Rz = gi.rz; gi.h = h;
    pair_uf<T, false>(gi, make_uf_ctx<T>(h), v[0], v[1], -v[2], true, gi.rz * gi.rz, T(0), ux, uy, uz);
  }
}

// Normalisation 1/(8 pi eta a^k), k = 1 (UF, fused), 2 (UT, WF), 3 (WT).
inline double block_norm(int block, double eta, double a)
{
  const double pi = 3.14159265358979323846;
  double n = 1.0 / (8.0 * pi * eta * a);
  if (block == UT || block == WF) n = 1.0 / (8.0 * pi * eta * a * a);
  if (block == WT) n = 1.0 / (8.0 * pi * eta * a * a * a);
  return n;
}

}  // namespace rmb
