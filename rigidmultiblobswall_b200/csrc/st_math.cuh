// st_math.cuh -- per-pair arithmetic of the source-target UF kernels with
// per-blob radii (target radius a, source radius b), lengths in the caller's
// unit, accumulators in units of 1/(8 pi eta).
//
// What is computed follows /root/reference/mobility/mobility_numba.py
//   :1552-1566  RPY for unequal radii (Zuk et al. 2014): far, partial overlap,
//               complete overlap (which also covers r = 0: no self branch)
//   :1581-1647  no-slip wall correction, 5 groups of 9 entries up to r^-9
//   :2040-2081  free-surface image: RPY(R) with the z column sign-flipped
// rewritten as "coefficients times vectors" like pair_math.cuh: no 3x3
// matrix, no division, no sqrt.  Plain C++ templates: compiles under nvcc and,
// for tests/test_pair_math_host.py, under g++ -DRMB_HOST_ONLY.
#pragma once
#include "pair_math.cuh"

namespace rmb {

// Target-side loop invariants.
template <typename T>
struct StTarget {
  T x, y, z;      // z already clamped to >= a in the wall form
  T a, a2, ia;    // radius, radius^2, 1/radius
  T a_2, z_2;     // 2 radius, 2 z (loop invariants of the pair arithmetic)
};

// "is r2 <= s2 possible": conservative test on the high words only (off the
// FP64 pipe); used for the warp-uniform gate, the exact compare follows.
#if defined(RMB_HOST_ONLY)
RMB_HD bool maybe_le(double r2, double s2) { return r2 <= s2; }
RMB_HD bool maybe_le(float r2, float s2) { return r2 <= s2; }
RMB_HD double min_t(double a, double b) { return a < b ? a : b; }
RMB_HD float min_t(float a, float b) { return a < b ? a : b; }
RMB_HD bool warp_any(bool p) { return p; }
#else
RMB_HD bool warp_any(bool p) { return __any_sync(0xffffffffu, p); }  // all 32 lanes run the pair loop
RMB_HD bool maybe_le(double r2, double s2) { return __double2hiint(r2) <= __double2hiint(s2); }
RMB_HD bool maybe_le(float r2, float s2) { return r2 <= s2; }
RMB_HD double min_t(double a, double b) { return fmin(a, b); }
RMB_HD float min_t(float a, float b) { return fminf(a, b); }
#endif

// Literals with a non-zero low word (and the second literal of two-literal FMAs): host-filled, read from
// the kernel's constant bank so that the pair loop holds no constant-materialising moves.
template <typename T>
struct StConsts {
  T c13, c23, c43, c103, cm23, cm703, c203, cm132, k2, k3, k6, k10, khalf;
};
template <typename T>
RMB_HOST_DEVICE StConsts<T> make_st_consts()
{
  StConsts<T> K;
  K.c13 = T(1.0 / 3.0); K.c23 = T(2.0 / 3.0); K.c43 = T(4.0 / 3.0); K.c103 = T(10.0 / 3.0); K.cm23 = T(-2.0 / 3.0);
  K.cm703 = T(-70.0 / 3.0); K.c203 = T(20.0 / 3.0); K.cm132 = T(-1.0 / 32.0);
  K.k2 = T(2.0); K.k3 = T(3.0); K.k6 = T(6.0); K.k10 = T(10.0); K.khalf = T(0.5);
  return K;
}

// u += C1 f + C2 (r.f) r with (mobility_numba.py:1552-1566)
//   r > a+b      : C1 = (1 + S/(3r^2))/r,  C2 = (1 - S/r^2)/r^3,  S = a^2+b^2
//   |b-a| < r    : C1 = [ (a+b)/2 - (D+3r^2)^2/(32 r^3) ] 4/(3ab),  D = (b-a)^2
//                  C2 = (D-r^2)^2/(8 r^5 ab)
//   r <= |b-a|   : C1 = 4/(3 max(a,b)), C2 = 0
// sab2 = (a+b)^2; any_near = warp-uniform "some lane may have r2 <= sab2".
template <typename T>
RMB_HD void rpy_radii(const StConsts<T>& K, T rx, T ry, T rz, T r2, T a, T ia, T b, T ib, T S, T sab2, bool any_near,
                      T fx, T fy, T fz, T& ux, T& uy, T& uz)
{
  const T invr = rsqrt_full(r2);
  const T invr2 = invr * invr;
  const T invr3 = invr2 * invr;
  T C1 = RMB_FMA(K.c13 * S, invr3, invr);
  T C2 = RMB_FMA(-(S * invr2), invr3, invr3);
  if (any_near) {
    const T d = b - a;
    const T D = d * d;
    const T iab = ia * ib;
    const T g = RMB_FMA(K.k3, r2, D);
    const T C1n = RMB_FMA(K.cm132 * (g * g), invr3, K.khalf * (a + b)) * (K.c43 * iab);
    const T e = D - r2;
    const T C2n = (T(0.125) * iab) * (e * e) * (invr3 * invr2);
    const bool far = r2 > sab2;
    const bool inside = !(r2 > D);
    C1 = far ? C1 : (inside ? K.c43 * min_t(ia, ib) : C1n);
    C2 = far ? C2 : (inside ? T(0) : C2n);
  }
  const T rf = RMB_FMA(rz, fz, RMB_FMA(ry, fy, rx * fx));
  const T Br = C2 * rf;
  ux = RMB_FMA(C1, fx, ux); uy = RMB_FMA(C1, fy, uy); uz = RMB_FMA(C1, fz, uz);
  ux = RMB_FMA(Br, rx, ux); uy = RMB_FMA(Br, ry, uy); uz = RMB_FMA(Br, rz, uz);
}

// No-slip wall correction (mobility_numba.py:1581-1647).  With R = (rx, ry, Z),
// Z = x3 + y3 (x3 target height, y3 source height), q = 1/|R|, t = q^2,
// f' = (-fx, -fy, fz) (the reference subtracts the x and y columns and adds the
// z column), the 45 entry updates collapse to
//   du = al f' + (be (R.f') + ga fz) R + (de (R.f') + ep fz) z^
//   al = q + q^3 [ S/3 + 2 x3 y3 - 2 Z t W + (2P/3) t (5k - 1) ]
//   be = q^3 [ 1 - t (S + 6 x3 y3) + 10 Z t^2 W + (10P/3) t^2 (1 - 7k) ]
//   ga = q^3 [ 2 x3 ((a^2-b^2) t - 1) + (20P/3) Z t^2 ]
//   de = q^3 [ 2 y3 ((b^2-a^2) t - 1) + (20P/3) Z t^2 ]
//   ep = -2q - q^3 [ 2S/3 + (4P/3) t ]
// S = a^2+b^2, P = a^2 b^2, W = a^2 y3 + b^2 x3, k = Z^2 t.
template <typename T>
RMB_HD void wall_radii(const StConsts<T>& K, T rx, T ry, T Z, T rxy2, T x3, T x3_2, T y3, T a2, T b2, T S, T fx, T fy,
                       T fz, T& ux, T& uy, T& uz)
{
  const T q = rsqrt_full(RMB_FMA(Z, Z, rxy2));
  const T t = q * q;
  const T q3 = q * t;
  const T xy = x3 * y3;
  const T P = a2 * b2;
  const T W = RMB_FMA(a2, y3, b2 * x3);
  const T Zt = Z * t;
  const T k = Z * Zt;
  const T dab = a2 - b2;
  const T Pt = P * t;
  const T ZW = Zt * W;
  T al = RMB_FMA(K.c103 * Pt, k, K.cm23 * Pt);          // (2P/3) t (5k - 1)
  al = RMB_FMA(T(-2.0), ZW, al);
  al += RMB_FMA(K.k2, xy, K.c13 * S);
  al = RMB_FMA(q3, al, q);
  T be = RMB_FMA(K.cm703 * Pt, k, K.c103 * Pt);         // (10P/3) t (1 - 7k)
  be = RMB_FMA(K.k10, ZW, be);
  be -= RMB_FMA(K.k6, xy, S);
  be = RMB_FMA(be, t, T(1.0));
  const T c20 = K.c203 * Pt * Zt;                              // (20P/3) Z t^2
  const T ga = RMB_FMA(x3_2, RMB_FMA(dab, t, T(-1.0)), c20);
  const T de = RMB_FMA(y3 + y3, RMB_FMA(-dab, t, T(-1.0)), c20);
  const T ep = -RMB_FMA(q3, RMB_FMA(K.c43, Pt, K.c23 * S), q + q);
  const T Rf = RMB_FMA(Z, fz, -RMB_FMA(ry, fy, rx * fx));
  const T cR = RMB_FMA(be, Rf, ga * fz) * q3;
  const T cz = RMB_FMA(de * q3, Rf, ep * fz);
  ux = RMB_FMA(-al, fx, ux); uy = RMB_FMA(-al, fy, uy); uz = RMB_FMA(al, fz, uz);
  ux = RMB_FMA(cR, rx, ux); uy = RMB_FMA(cR, ry, uy); uz = RMB_FMA(cR, Z, uz);
  uz += cz;
}

enum StWall { ST_NO_WALL = 0, ST_WALL = 1, ST_FREE_SURFACE = 2 };

// One packed source blob: 8 words (fp64 64 B = 4 x LDS.128, fp32 32 B = 2 x LDS.128).
template <typename T>
struct alignas(16) StRec {
  T x, y, z, fx, fy, fz, b, ib;   // position, (damped) force, radius, 1/radius
};

// mobility.py:87-119 fused: z <- max(z, radius), B = z/radius where z < radius
// (wall form only); returns B.  A padding record (radius 1, zero force, far
// away is not needed: zero force contributes exactly 0 for any finite geometry).
template <typename T>
RMB_HD double st_pack(bool wall, double x, double y, double z, double radius, double fx, double fy, double fz,
                      StRec<T>& o)
{
  double B = 1.0;
  if (wall) {
    B = (z < radius) ? z / radius : 1.0;
    z = (z > radius) ? z : radius;
  }
  o.x = T(x); o.y = T(y); o.z = T(z);
  o.fx = T(B * fx); o.fy = T(B * fy); o.fz = T(B * fz);
  o.b = T(radius); o.ib = T(1.0 / radius);
  return B;
}

template <typename T>
RMB_HD double st_target(bool wall, double x, double y, double z, double radius, StTarget<T>& t)
{
  double B = 1.0;
  if (wall) {
    B = (z < radius) ? z / radius : 1.0;
    z = (z > radius) ? z : radius;
  }
  t.x = T(x); t.y = T(y); t.z = T(z);
  t.a = T(radius); t.a2 = T(radius * radius); t.ia = T(1.0 / radius);
  t.a_2 = T(2.0 * radius); t.z_2 = T(2.0 * z);
  return B;
}

// One pair.  (rx, ry, rz) after the image wrap, Z = z_i + z_j (no wrap, :1583),
// y3 = source height, b / ib = source radius and its reciprocal.
template <typename T, int WALLMODE>
RMB_HD void pair_st(const StConsts<T>& K, const StTarget<T>& t, T rx, T ry, T rz, T Z, T y3, T b, T ib, T fx, T fy, T fz, T& ux, T& uy,
                    T& uz)
{
  const T rxy2 = RMB_FMA(ry, ry, rx * rx);
  const T r2 = RMB_FMA(rz, rz, rxy2);
  const T b2 = b * b;
  const T S = t.a2 + b2;
  const T sab2 = RMB_FMA(t.a_2, b, S);          // (a + b)^2 = a^2 + b^2 + 2ab
  rpy_radii<T>(K, rx, ry, rz, r2, t.a, t.ia, b, ib, S, sab2, warp_any(maybe_le(r2, sab2)), fx, fy, fz, ux, uy, uz);
  if (WALLMODE == ST_WALL) wall_radii<T>(K, rx, ry, Z, rxy2, t.z, t.z_2, y3, t.a2, b2, S, fx, fy, fz, ux, uy, uz);
  if (WALLMODE == ST_FREE_SURFACE) {
    const T R2 = RMB_FMA(Z, Z, rxy2);
    rpy_radii<T>(K, rx, ry, Z, R2, t.a, t.ia, b, ib, S, sab2, warp_any(maybe_le(R2, sab2)), fx, fy, -fz, ux, uy, uz);
  }
}

}  // namespace rmb
