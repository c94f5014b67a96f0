// pair_math.cuh -- per-pair arithmetic of the blob mobility blocks.
//
// What each function computes is fixed by the reference
// (/root/reference/mobility/mobility_numba.py, lines cited per function); HOW
// it is computed is not: every block is rewritten as "coefficients times
// vectors" so that no 3x3 matrix is formed, there is no division and no sqrt
// (one reciprocal-square-root per distance, MUFU.RSQ64H + one cubic
// correction), and h_hat = h/Rz never appears on its own (h_hat*ez = h/|R|).
//
// All lengths are in units of the blob radius a.  Accumulators are NOT
// normalised: the 1/(8 pi eta a^k) factor is applied once per target.
// The self pair (i == j, central box) is never passed to these functions;
// its closed-form term is added once per target by self_term_*().
//
// The file is plain C++ templates: it compiles under nvcc (device code) and,
// for tests/test_pair_math_host.py only, under g++ (RMB_HOST_ONLY) so that the
// restructured algebra can be checked against the oracle without a GPU.
#pragma once

#if defined(RMB_HOST_ONLY)
#include <cmath>
#define RMB_HD inline
#define RMB_HOST_DEVICE inline
#define RMB_FMA(a, b, c) std::fma((a), (b), (c))
#else
#define RMB_HD __device__ __forceinline__
#define RMB_HOST_DEVICE __host__ __device__ __forceinline__
#define RMB_FMA(a, b, c) fma((a), (b), (c))
#endif

namespace rmb {

// ---------------------------------------------------------------- rsqrt ----
#if defined(RMB_HOST_ONLY)
// Host emulation of the device sequence: a seed with ~2^-20 relative error
// (what rsqrt.approx.ftz.f64 guarantees) followed by the same correction.
RMB_HD double rsqrt_seed(double x) { return (double)(float)(1.0 / std::sqrt(x)) * (1.0 + 9.0e-7); }
RMB_HD float rsqrt_fast(float x) { return 1.0f / std::sqrt(x); }
#else
RMB_HD double rsqrt_seed(double x)
{
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));  // MUFU.RSQ64H
  return y;
}
RMB_HD float rsqrt_fast(float x) { return rsqrtf(x); }  // MUFU.RSQ
#endif

// 1/sqrt(x) to < 1 ulp-ish: y0(1 + e/2 + 3e^2/8), e = 1 - x y0^2 (cubic
// convergence: seed error 2^-20 -> 5/16 e^3 ~ 2^-58).  5 FP64 instructions.
RMB_HD double rsqrt_full(double x)
{
  const double y0 = rsqrt_seed(x);
  const double t = x * y0;
  const double e = RMB_FMA(-t, y0, 1.0);
  const double p = RMB_FMA(0.375, e, 0.5);
  return RMB_FMA(y0, p * e, y0);
}
RMB_HD float rsqrt_full(float x) { return rsqrt_fast(x); }
// Same with the two literals of the correction handed in (from the kernel's constant bank: nvcc otherwise
// re-materialises 0.375 with two moves in every loop iteration).
RMB_HD double rsqrt_full(double x, double c38, double chalf)
{
  const double y0 = rsqrt_seed(x);
  const double t = x * y0;
  const double e = RMB_FMA(-t, y0, 1.0);
  const double p = RMB_FMA(c38, e, chalf);
  return RMB_FMA(y0, p * e, y0);
}
RMB_HD float rsqrt_full(float x, float, float) { return rsqrt_fast(x); }

// "is r2 < 4" without touching the FP64 pipe: compare the high word.
#if defined(RMB_HOST_ONLY)
RMB_HD bool lt4(double r2) { return r2 < 4.0; }
RMB_HD bool lt4(float r2) { return r2 < 4.0f; }
#else
RMB_HD bool lt4(double r2) { return __double2hiint(r2) < 0x40100000; }
RMB_HD bool lt4(float r2) { return r2 < 4.0f; }
#endif

// Geometry of one pair, lengths / a.  (rx,ry,rz) = x_i - x_j after the image
// wrap; Rz = z_i + z_j and h (height of source, or of target for UT) are
// only used by the wall corrections.
template <typename T>
struct PairGeom {
  T rx, ry, rz, Rz, h;
};

// ------------------------------------------------------------------- UF ----
// Free-space RPY translation-translation (mobility_numba.py:199-239):
//   far  (r > 2):  [ (1 + 2/(3r^2)) I + (1 - 2/r^2) rr/r^2 ] / r
//   near (r <= 2): 4/3 [ (1 - 9r/32) I + 3/(32 r) rr ]
// Wall correction (:242-276), Swan & Brady fact1..5 with e = R/|R|,
// R = (rx, ry, Rz):
//   u += fact1 f + e (fact2 (e.f) + fact3 f_z) + z^ (fact4 (e.f) + fact5 f_z)
// Rewritten with q = 1/|R|, t = q^2, alpha = Rz^2, beta = h_i h_j, k = alpha t
// (= ez^2), H = h_j; h_hat(1-h_hat)ez^2 = beta t and h_hat ez = H q, so
//   fact1 = q p1, fact2 = q p2, fact3 = q^2 p3, fact4 = q^2 p4, fact5 = q^3 p5
//   p1 = -1 - (2 beta + 2/3) t + (2 alpha + 2/3) t^2 - (10/3) k t^2
//   p2 = -1 + (6 beta + 2) t - (10 alpha + 10/3) t^2 + (70/3) k t^2
//   p3 = 2H - Rz t [ (12 beta + 4) - (20 alpha + 40/3) t + (140/3) k t ]
//   p4 = 2H - (20/3) Rz t^2
//   p5 = -4H^2 - (4 alpha + 8/3) t + 20 k t
//   u += q p1 f + q^3 [ (p2 (R.f) + p3 f_z) R + z^ (p4 (R.f) + p5 f_z) ]
// Evaluation order (every FMA has at most ONE literal operand, and the
// literals with a non-zero low word come from UfCtx, i.e. from the kernel's
// constant bank, so the loop holds no constant-materialising moves):
//   Aq = alpha + 1/3, X = Aq - (5/3) k, Y = Aq - (7/3) k, Wv = Aq + 1/3 - 5k
//   b2 = (2 h_i) H + 2/3                                  (2 h_i per target; 6 beta + 2 = 3 b2)
//   p1 = -1 + t (2t X - b2)          v = 3 b2 - 10 t Y = 3 v3,   p2 = -1 + t v
//   p3 = 2 [H - Rz t (v - (10/3) t)]                     p4 = 2 [H - Rz t (10/3) t]
//   p5 = -4 [H^2 + t Wv]
// and the factors 2 and 4 ride on 2 f_z: 45 FP64 instructions for the wall part.
// ZOPEN (no wrap in z, i.e. rz = z_i - z_j): |R|^2 = r^2 + 4 h_i h_j instead of rx^2 + ry^2 + Rz^2.
// any_near: warp-uniform flag "some lane has r2 < 4" -- the near
// coefficients are only evaluated when it is set.
template <typename T>
struct UfCtx {
  T c13, c23, cm53, cm73, c103, c43;   // 1/3, 2/3, -5/3, -7/3, 10/3, 4/3
  T k1, k2, k12, km20, km10, km3, khalf, c38;   // 1, 2, 12, -20, -10, -3, 1/2, 3/8 (second literals of two-literal FMAs)
  T ht, ht2, ht4, ht6;                 // target height h_i and its multiples (wall form)
};

template <typename T>
RMB_HOST_DEVICE UfCtx<T> make_uf_ctx(T ht)
{
  UfCtx<T> C;
  C.c13 = T(1.0 / 3.0); C.c23 = T(2.0 / 3.0); C.cm53 = T(-5.0 / 3.0); C.cm73 = T(-7.0 / 3.0);
  C.c103 = T(10.0 / 3.0); C.c43 = T(4.0 / 3.0);
  C.k1 = T(1.0); C.k2 = T(2.0); C.k12 = T(12.0); C.km20 = T(-20.0); C.km10 = T(-10.0); C.km3 = T(-3.0);
  C.khalf = T(0.5); C.c38 = T(0.375);
  C.ht = ht; C.ht2 = T(2.0) * ht; C.ht4 = T(4.0) * ht; C.ht6 = T(6.0) * ht;
  return C;
}
template <typename T>
RMB_HD void set_target_height(UfCtx<T>& C, T ht)
{
  C.ht = ht; C.ht2 = T(2.0) * ht; C.ht4 = T(4.0) * ht; C.ht6 = T(6.0) * ht;
}

template <typename T, bool WALL, bool ZOPEN = false>
RMB_HD void pair_uf(const PairGeom<T>& p, const UfCtx<T>& C, T fx, T fy, T fz, bool any_near, T r2, T rxy2, T& ux,
                    T& uy, T& uz)
{
  const T s = RMB_FMA(p.ry, fy, p.rx * fx);
  const T invr = rsqrt_full(r2, C.c38, C.khalf);
  const T invr2 = invr * invr;
  const T rf = RMB_FMA(p.rz, fz, s);
  const T invr3 = invr2 * invr;
  T A = RMB_FMA(C.c23, invr3, invr);                 // (1 + 2/(3 r^2)) / r
  const T rfi = rf * invr3;
  T B = RMB_FMA(T(-2.0), invr2 * rfi, rfi);          // (1 - 2/r^2) (r.f) / r^3
  if (any_near) {
    const T r = r2 * invr;
    const T An = RMB_FMA(T(-0.375), r, C.c43);
    const T Bn = (T(0.125) * invr) * rf;
    const bool near = lt4(r2);
    A = near ? An : A;
    B = near ? Bn : B;
  }
  if (!WALL) {
    ux = RMB_FMA(A, fx, ux); uy = RMB_FMA(A, fy, uy); uz = RMB_FMA(A, fz, uz);
    ux = RMB_FMA(B, p.rx, ux); uy = RMB_FMA(B, p.ry, uy); uz = RMB_FMA(B, p.rz, uz);
  } else {
    // |R|^2 = rx^2 + ry^2 + Rz^2.  Without a z wrap rz = z_i - z_j, so |R|^2 = r^2 + 4 h_i h_j (one FMA, no Rz^2).
    const T q = rsqrt_full(ZOPEN ? RMB_FMA(C.ht4, p.h, r2) : RMB_FMA(p.Rz, p.Rz, rxy2), C.c38, C.khalf);
    const T t = q * q;
    const T Aq = RMB_FMA(p.Rz, p.Rz, C.c13);             // alpha + 1/3
    const T Zt = p.Rz * t;
    const T t2 = t + t;
    const T k = p.Rz * Zt;                               // alpha t
    const T q3 = q * t;
    const T b2 = RMB_FMA(C.ht2, p.h, C.c23);             // 2 beta + 2/3
    const T X = RMB_FMA(C.cm53, k, Aq);
    const T Y = RMB_FMA(C.cm73, k, Aq);
    const T p1 = RMB_FMA(RMB_FMA(t2, X, -b2), t, T(-1.0));
    const T t103 = C.c103 * t;
    const T v3 = RMB_FMA(-t103, Y, b2);                  // v / 3, v = (6 beta + 2) - 10 t Y;  p2 = -1 + t v
    const T p2 = RMB_FMA(v3, T(3.0) * t, T(-1.0));
    const T p3h = RMB_FMA(-Zt, RMB_FMA(T(3.0), v3, -t103), p.h);   // p3 / 2 = H - Rz t (v - (10/3) t)
    const T p4h = RMB_FMA(-Zt, t103, p.h);               // p4 / 2
    const T Wv = RMB_FMA(T(-5.0), k, Aq + C.c13);
    const T p5q = RMB_FMA(t, Wv, p.h * p.h);             // -p5 / 4
    const T fz2 = fz + fz;
    const T Rf = RMB_FMA(p.Rz, fz, s);
    const T G = RMB_FMA(p2, Rf, p3h * fz2);              // p2 (R.f) + p3 f_z
    const T SZh = RMB_FMA(p4h, Rf, -(p5q * fz2));        // (p4 (R.f) + p5 f_z) / 2
    const T Af = RMB_FMA(q, p1, A);
    const T Bx = RMB_FMA(q3, G, B);
    ux = RMB_FMA(Af, fx, ux); uy = RMB_FMA(Af, fy, uy); uz = RMB_FMA(Af, fz, uz);
    ux = RMB_FMA(Bx, p.rx, ux); uy = RMB_FMA(Bx, p.ry, uy);
    uz = RMB_FMA(B, p.rz, uz); uz = RMB_FMA(q3, RMB_FMA(G, p.Rz, SZh + SZh), uz);
  }
}

// Both directions of one pair at once (single wall, no image wrap): u_i += M_ij f_j AND u_j += M_ji f_i.
// M_ji is M_ij with r -> -r and the source height h_j -> h_i, so the two evaluations share the geometry, both
// reciprocal square roots, the free-space coefficients and q, t, k, A_q, X, Y, b2, v, p1, p2, W_v: 44 common FP64
// instructions + 22 per direction = 88 for two interactions (2 x 66 when evaluated separately).
//   direction d (force f, source height H, sign s = +1 for i<-j, -1 for j<-i), with S = r_x f_x + r_y f_y:
//     B_f = b' (S + r_z f_z)              (free space, even in r)
//     R.f = s S + R_z f_z,   G = p2 (R.f) + 2 p3h f_z,   SZh = p4h (R.f) - 2 p5q f_z
//     u_xy += A_f f_xy + (B_f + s q^3 G) r_xy,   u_z += A_f f_z + B_f r_z + q^3 (G R_z + 2 SZh)
// hi2 = 2 h_i, hi4 = 4 h_i (per target, loop invariant); C: the literal constants.
template <typename T, int SIGN>
RMB_HD void uf_sym_direction(T rx, T ry, T rz, T Rz, T H, const T* f, T bp, T Af, T q3, T p2, T Zt, T vm, T ZtT, T tW,
                             T* u)
{
  const T S = RMB_FMA(ry, f[1], rx * f[0]);
  const T Bf = bp * RMB_FMA(rz, f[2], S);
  const T Rf = RMB_FMA(Rz, f[2], SIGN > 0 ? S : -S);
  const T p3h = RMB_FMA(-Zt, vm, H);
  const T p4h = H - ZtT;
  const T p5q = RMB_FMA(H, H, tW);
  const T fz2 = f[2] + f[2];
  const T G = RMB_FMA(p2, Rf, p3h * fz2);
  const T SZh = RMB_FMA(p4h, Rf, -(p5q * fz2));
  const T Bx = RMB_FMA(SIGN > 0 ? q3 : -q3, G, Bf);
  u[0] = RMB_FMA(Af, f[0], u[0]); u[1] = RMB_FMA(Af, f[1], u[1]); u[2] = RMB_FMA(Af, f[2], u[2]);
  u[0] = RMB_FMA(Bx, rx, u[0]); u[1] = RMB_FMA(Bx, ry, u[1]);
  u[2] = RMB_FMA(Bf, rz, u[2]); u[2] = RMB_FMA(q3, RMB_FMA(G, Rz, SZh + SZh), u[2]);
}

// (rx, ry, rz) = x_i - x_j, Rz = h_i + h_j (lengths / a); fi, fj: the two forces; BOTH: also update u_j.
template <typename T, bool BOTH>
RMB_HD void pair_uf_wall_sym(T rx, T ry, T rz, T Rz, T hi, T hj, T hi2, T hi4, const UfCtx<T>& C, const T* fi,
                             const T* fj, bool any_near, T r2, T* ui, T* uj)
{
  const T invr = rsqrt_full(r2, C.c38, C.khalf);
  const T invr2 = invr * invr;
  const T invr3 = invr2 * invr;
  T A = RMB_FMA(C.c23, invr3, invr);
  T bp = RMB_FMA(T(-2.0), invr2 * invr3, invr3);       // (1 - 2/r^2) / r^3
  if (any_near) {
    const T r = r2 * invr;
    const T An = RMB_FMA(T(-0.375), r, C.c43);
    const T bn = T(0.125) * invr;
    const bool near = lt4(r2);
    A = near ? An : A;
    bp = near ? bn : bp;
  }
  const T q = rsqrt_full(RMB_FMA(hi4, hj, r2), C.c38, C.khalf);
  const T t = q * q;
  const T Aq = RMB_FMA(Rz, Rz, C.c13);
  const T Zt = Rz * t;
  const T t2 = t + t;
  const T k = Rz * Zt;
  const T q3 = q * t;
  const T b2 = RMB_FMA(hi2, hj, C.c23);
  const T X = RMB_FMA(C.cm53, k, Aq);
  const T Y = RMB_FMA(C.cm73, k, Aq);
  const T p1 = RMB_FMA(RMB_FMA(t2, X, -b2), t, T(-1.0));
  const T t103 = C.c103 * t;
  const T v3 = RMB_FMA(-t103, Y, b2);
  const T p2 = RMB_FMA(v3, T(3.0) * t, T(-1.0));
  const T vm = RMB_FMA(T(3.0), v3, -t103);
  const T ZtT = Zt * t103;
  const T tW = t * RMB_FMA(T(-5.0), k, Aq + C.c13);
  const T Af = RMB_FMA(q, p1, A);
  uf_sym_direction<T, 1>(rx, ry, rz, Rz, hj, fj, bp, Af, q3, p2, Zt, vm, ZtT, tW, ui);
  if (BOTH) uf_sym_direction<T, -1>(rx, ry, rz, Rz, hi, fi, bp, Af, q3, p2, Zt, vm, ZtT, tW, uj);
}

// Free-space twin of pair_uf_wall_sym (no_wall UF): RPY is even in r, so the two directions share everything
// but the force: 16 common FP64 instructions + 10 per direction.
template <typename T, bool BOTH>
RMB_HD void pair_uf_free_sym(T rx, T ry, T rz, const UfCtx<T>& C, const T* fi, const T* fj, bool any_near, T r2, T* ui,
                             T* uj)
{
  const T invr = rsqrt_full(r2, C.c38, C.khalf);
  const T invr2 = invr * invr;
  const T invr3 = invr2 * invr;
  T A = RMB_FMA(C.c23, invr3, invr);
  T bp = RMB_FMA(T(-2.0), invr2 * invr3, invr3);
  if (any_near) {
    const T r = r2 * invr;
    const T An = RMB_FMA(T(-0.375), r, C.c43);
    const T bn = T(0.125) * invr;
    const bool near = lt4(r2);
    A = near ? An : A;
    bp = near ? bn : bp;
  }
  {
    const T Bf = bp * RMB_FMA(rz, fj[2], RMB_FMA(ry, fj[1], rx * fj[0]));
    ui[0] = RMB_FMA(A, fj[0], ui[0]); ui[1] = RMB_FMA(A, fj[1], ui[1]); ui[2] = RMB_FMA(A, fj[2], ui[2]);
    ui[0] = RMB_FMA(Bf, rx, ui[0]); ui[1] = RMB_FMA(Bf, ry, ui[1]); ui[2] = RMB_FMA(Bf, rz, ui[2]);
  }
  if (BOTH) {
    const T Bf = bp * RMB_FMA(rz, fi[2], RMB_FMA(ry, fi[1], rx * fi[0]));
    uj[0] = RMB_FMA(A, fi[0], uj[0]); uj[1] = RMB_FMA(A, fi[1], uj[1]); uj[2] = RMB_FMA(A, fi[2], uj[2]);
    uj[0] = RMB_FMA(Bf, rx, uj[0]); uj[1] = RMB_FMA(Bf, ry, uj[1]); uj[2] = RMB_FMA(Bf, rz, uj[2]);
  }
}

// Self term of UF (:203-208, :246-252); h = z_i / a (clamped), f = own force.
template <typename T, bool WALL>
RMB_HD void self_uf(T h, T fx, T fy, T fz, T& ux, T& uy, T& uz)
{
  T mxx = T(4.0 / 3.0), mzz = T(4.0 / 3.0);
  if (WALL) {
    const T iz = T(1.0) / h;
    const T iz3 = iz * iz * iz;
    const T iz5 = iz3 * iz * iz;
    mxx += -(T(9.0) * iz - T(2.0) * iz3 + iz5) / T(12.0);
    mzz += -(T(9.0) * iz - T(4.0) * iz3 + iz5) / T(6.0);
  }
  ux = RMB_FMA(mxx, fx, ux); uy = RMB_FMA(mxx, fy, uy); uz = RMB_FMA(mzz, fz, uz);
}

// -------------------------------------------------------------- UT / WF ----
// Free-space translation-rotation coupling (mobility_numba.py:611-648,
// :1001-1035): antisymmetric, M v = kf (v x r) with kf = 1/r^3 (r >= 2) or
// (1 - 3r/8)/2 (r < 2).
// Wall corrections: WF (:1037-1066) uses e = (rx,ry,Rz)/|R| and the SOURCE
// height; UT (:648-679) is the transposed call: e = (-rx,-ry,Rz)/|R| and the
// TARGET height.  fact1 = q^2, fact2 = 12 h_hat ez^2 q^2 + 2(1-10 ez^2) q^4,
// fact3 = -2 ez (3 h_hat q^2 - 5 q^4), fact4 = -2 ez (h_hat q^2 - q^4).
// In unnormalised form (q = 1/|R|, t = q^2, Zt = Rz t, k = Rz^2 t, h the
// height the reference uses) every entry is q^3 times a polynomial:
//   fact3 q^2        = q^3 t (10 Zt - 6h)
//   fact1 ez + fact4 = q^3 c14,  c14 = Rz - 2h + 2 Zt
//   fact1 + fact2    = q^2 c12,  c12 = 1 + 2t + 12 h Zt - 20 k t
//   fact3 ez + fact1 + fact2 = q^2 c12', c12' = 1 + 2t + 6 h Zt - 10 k t
// and free + wall collapse to (f = force, tq = torque, cz = (r x v)_z):
//   WF: w_x = f_y P1 - r_y P2, w_y = -f_x P1 + r_x P2, w_z = (kf - q^3)(r_y f_x - r_x f_y)
//       P1 = kf r_z - q^3 c14,  P2 = kf f_z - q^3 [t (10 Zt - 6h)(R.f) + c12 f_z]
//   UT: u_x = t_y P1 - r_y t_z (kf - q^3) - q^3 E r_x,  u_y = -t_x P1 + r_x t_z (kf - q^3) - q^3 E r_y,
//       u_z = cz (q^3 c12' - kf),  P1 = kf r_z + q^3 c14,  E = t (10 Zt - 6h) cz,  cz = r_x t_y - r_y t_x
// 50 (WF) / 47 (UT) FP64 instructions per far pair, was 70 / 68 in e-vector form.
template <typename T>
RMB_HD T coupling_k(T r2, bool any_near, const UfCtx<T>& C)
{
  const T invr = rsqrt_full(r2, C.c38, C.khalf);
  T k = invr * invr * invr;
  if (any_near) {
    const T kn = RMB_FMA(T(-0.1875), r2 * invr, T(0.5));
    k = lt4(r2) ? kn : k;
  }
  return k;
}

// rotation from force: w_i += WF_ij f_j   (units 1/(8 pi eta a^2)); p.h = SOURCE height
template <typename T, bool WALL, bool ZOPEN = false>
RMB_HD void pair_wf(const PairGeom<T>& p, const UfCtx<T>& C, T fx, T fy, T fz, bool any_near, T r2, T rxy2, T& ux,
                    T& uy, T& uz)
{
  const T kf = coupling_k(r2, any_near, C);
  if (!WALL) {
    ux = RMB_FMA(kf, RMB_FMA(p.rz, fy, -(p.ry * fz)), ux);
    uy = RMB_FMA(kf, RMB_FMA(p.rx, fz, -(p.rz * fx)), uy);
    uz = RMB_FMA(kf, RMB_FMA(p.ry, fx, -(p.rx * fy)), uz);
  } else {
    const T q = rsqrt_full(ZOPEN ? RMB_FMA(C.ht4, p.h, r2) : RMB_FMA(p.Rz, p.Rz, rxy2), C.c38, C.khalf);
    const T t = q * q;
    const T Zt = p.Rz * t;
    const T q3 = q * t;
    const T k = p.Rz * Zt;
    const T Rf = RMB_FMA(p.Rz, fz, RMB_FMA(p.ry, fy, p.rx * fx));
    const T e10 = RMB_FMA(T(10.0), Zt, T(-6.0) * p.h);
    const T E = (e10 * t) * Rf;
    const T c14 = RMB_FMA(T(2.0), Zt, ZOPEN ? p.rz : RMB_FMA(T(-2.0), p.h, p.Rz));   // Rz - 2 h_j = z_i - z_j
    const T c12 = RMB_FMA(RMB_FMA(C.km20, k, C.k2), t, RMB_FMA(C.k12, p.h * Zt, C.k1));
    const T D = RMB_FMA(c12, fz, E);
    const T P1 = RMB_FMA(-q3, c14, kf * p.rz);
    const T P2 = RMB_FMA(-q3, D, kf * fz);
    const T cz = RMB_FMA(p.ry, fx, -(p.rx * fy));
    ux = RMB_FMA(fy, P1, ux); ux = RMB_FMA(-p.ry, P2, ux);
    uy = RMB_FMA(-fx, P1, uy); uy = RMB_FMA(p.rx, P2, uy);
    uz = RMB_FMA(kf - q3, cz, uz);
  }
}

// translation from torque: u_i += UT_ij t_j   (units 1/(8 pi eta a^2)); the wall term uses the
// TARGET height C.ht (transposed call, :651); p.h (source height) only enters |R|^2.
template <typename T, bool WALL, bool ZOPEN = false>
RMB_HD void pair_ut(const PairGeom<T>& p, const UfCtx<T>& C, T tx, T ty, T tz, bool any_near, T r2, T rxy2, T& ux,
                    T& uy, T& uz)
{
  const T kf = coupling_k(r2, any_near, C);
  if (!WALL) {
    ux = RMB_FMA(kf, RMB_FMA(p.rz, ty, -(p.ry * tz)), ux);
    uy = RMB_FMA(kf, RMB_FMA(p.rx, tz, -(p.rz * tx)), uy);
    uz = RMB_FMA(kf, RMB_FMA(p.ry, tx, -(p.rx * ty)), uz);
  } else {
    const T q = rsqrt_full(ZOPEN ? RMB_FMA(C.ht4, p.h, r2) : RMB_FMA(p.Rz, p.Rz, rxy2), C.c38, C.khalf);
    const T t = q * q;
    const T Zt = p.Rz * t;
    const T q3 = q * t;
    const T k = p.Rz * Zt;
    const T cz = RMB_FMA(p.rx, ty, -(p.ry * tx));
    const T E = (RMB_FMA(T(10.0), Zt, -C.ht6) * t) * cz;
    const T c14 = RMB_FMA(T(2.0), Zt, ZOPEN ? -p.rz : p.Rz - C.ht2);                  // Rz - 2 h_i = z_j - z_i
    const T c12 = RMB_FMA(RMB_FMA(C.km10, k, C.k2), t, RMB_FMA(C.ht6, Zt, C.k1));
    const T P1 = RMB_FMA(q3, c14, kf * p.rz);
    const T dt = (kf - q3) * tz;
    const T qE = q3 * E;
    ux = RMB_FMA(ty, P1, ux); ux = RMB_FMA(-p.ry, dt, ux); ux = RMB_FMA(-qE, p.rx, ux);
    uy = RMB_FMA(-tx, P1, uy); uy = RMB_FMA(p.rx, dt, uy); uy = RMB_FMA(-qE, p.ry, uy);
    uz = RMB_FMA(cz, RMB_FMA(q3, c12, -kf), uz);
  }
}

// Self terms of the couplings (:653-657, :1041-1045): +-h^-4/8 on xy / yx.
template <typename T>
RMB_HD void self_wf(T h, T fx, T fy, T& ux, T& uy)
{
  const T iz = T(1.0) / h;
  const T iz2 = iz * iz;
  const T c = T(0.125) * iz2 * iz2;
  ux -= c * fy; uy += c * fx;
}
template <typename T>
RMB_HD void self_ut(T h, T tx, T ty, T& ux, T& uy)
{
  const T iz = T(1.0) / h;
  const T iz2 = iz * iz;
  const T c = T(0.125) * iz2 * iz2;
  ux += c * ty; uy -= c * tx;
}

// ------------------------------------------------------------------- WT ----
// Free space (:1252-1290): far (r >= 2) (-1/2 I + 3/2 rr/r^2)/r^3; near
// (1 - 27r/32 + 5r^3/64) I + (9/(32r) - 3r/64) rr.  Wall (:1302-1321):
// fact1 = (1-6ez^2)q^3/2, fact2 = -3q^3/2, fact3 = 3 q^3 ez, fact4 = 3 q^3.
template <typename T, bool WALL, bool ZOPEN = false>
RMB_HD void pair_wt(const PairGeom<T>& p, const UfCtx<T>& C, T tx, T ty, T tz, bool any_near, T r2, T rxy2, T& ux,
                    T& uy, T& uz)
{
  const T invr = rsqrt_full(r2, C.c38, C.khalf);
  const T invr2 = invr * invr;
  const T invr3 = invr2 * invr;
  const T s = RMB_FMA(p.ry, ty, p.rx * tx);
  const T rt = RMB_FMA(p.rz, tz, s);
  T A = T(-0.5) * invr3;
  T B = (T(1.5) * invr2) * invr3 * rt;
  if (any_near) {
    const T r = r2 * invr;
    const T An = RMB_FMA(RMB_FMA(T(0.078125), r2, T(-0.84375)), r, T(1.0));
    const T Bn = RMB_FMA(T(-0.046875), r, T(0.28125) * invr) * rt;
    const bool near = lt4(r2);
    A = near ? An : A;
    B = near ? Bn : B;
  }
  if (!WALL) {
    ux = RMB_FMA(A, tx, ux); uy = RMB_FMA(A, ty, uy); uz = RMB_FMA(A, tz, uz);
    ux = RMB_FMA(B, p.rx, ux); uy = RMB_FMA(B, p.ry, uy); uz = RMB_FMA(B, p.rz, uz);
  } else {
    // unnormalised wall term (cz = r_x t_y - r_y t_x, Rt = R.t, q5 = q^5):
    //   u_x += q^3 (1/2 - 3k) t_x - 3/2 q5 Rt r_x - 3 q5 cz r_y      u_y likewise with (+3 q5 cz r_x)
    //   u_z += q^3 (1/2 - 3k) t_z + 3/2 q5 Rt Rz
    const T q = rsqrt_full(ZOPEN ? RMB_FMA(C.ht4, p.h, r2) : RMB_FMA(p.Rz, p.Rz, rxy2), C.c38, C.khalf);
    const T t = q * q;
    const T q3 = q * t;
    const T k = (p.Rz * p.Rz) * t;
    const T q5 = q3 * t;
    const T Rt = RMB_FMA(p.Rz, tz, s);
    const T cz = RMB_FMA(p.rx, ty, -(p.ry * tx));
    const T Aa = RMB_FMA(RMB_FMA(C.km3, k, C.khalf), q3, A);
    const T m1 = (T(-1.5) * q5) * Rt;
    const T m2 = (T(3.0) * q5) * cz;
    const T Bm = B + m1;
    ux = RMB_FMA(Aa, tx, ux); ux = RMB_FMA(Bm, p.rx, ux); ux = RMB_FMA(-m2, p.ry, ux);
    uy = RMB_FMA(Aa, ty, uy); uy = RMB_FMA(Bm, p.ry, uy); uy = RMB_FMA(m2, p.rx, uy);
    uz = RMB_FMA(Aa, tz, uz); uz = RMB_FMA(B, p.rz, uz); uz = RMB_FMA(-m1, p.Rz, uz);
  }
}

template <typename T, bool WALL>
RMB_HD void self_wt(T h, T tx, T ty, T tz, T& ux, T& uy, T& uz)
{
  T mxx = T(1.0), mzz = T(1.0);
  if (WALL) {
    const T iz = T(1.0) / h;
    const T iz3 = iz * iz * iz;
    mxx -= T(0.3125) * iz3;
    mzz -= T(0.125) * iz3;
  }
  ux = RMB_FMA(mxx, tx, ux); uy = RMB_FMA(mxx, ty, uy); uz = RMB_FMA(mzz, tz, uz);
}

// ---------------------------------------------------------------------------------------------------------------
// Pair-symmetric forms of the rotation blocks (single wall, no z wrap).  The (j <- i) interaction is the (i <- j)
// one with r -> -r and the height that enters the wall term exchanged (WF: the source height, UT: the target
// height, WT: none), so the reciprocal square roots, q^3, t, k, R_z t and the h-free parts of the polynomials are
// shared.  The direction helpers take the SIGNED separation (target - source) and kfrz = k_f r_z with that sign.
template <typename T>
struct RotShared {
  T kf, q3, t, Zt, Zt10, ZtH, c12b, kq;   // ZtH = (12 | 6) R_z t for WF | UT; kq = kf - q3
};

template <typename T, bool IS_WF>
RMB_HD RotShared<T> rot_shared(T Rz, T hi4, T hj, const UfCtx<T>& C, bool any_near, T r2)
{
  RotShared<T> S;
  S.kf = coupling_k(r2, any_near, C);
  const T q = rsqrt_full(RMB_FMA(hi4, hj, r2), C.c38, C.khalf);
  S.t = q * q;
  S.Zt = Rz * S.t;
  S.q3 = q * S.t;
  const T k = Rz * S.Zt;
  S.Zt10 = T(10.0) * S.Zt;
  S.ZtH = (IS_WF ? C.k12 : T(6.0)) * S.Zt;
  S.c12b = RMB_FMA(RMB_FMA(IS_WF ? C.km20 : C.km10, k, C.k2), S.t, C.k1);
  S.kq = S.kf - S.q3;
  return S;
}

// w += M_WF(r, source height h) f      (pair_wf, wall)
template <typename T>
RMB_HD void wf_direction(T rx, T ry, T rz, T Rz, T h, T kfrz, const T* f, const RotShared<T>& S, T* w)
{
  const T Rf = RMB_FMA(Rz, f[2], RMB_FMA(ry, f[1], rx * f[0]));
  const T E = (RMB_FMA(T(-6.0), h, S.Zt10) * S.t) * Rf;
  const T c14 = RMB_FMA(T(2.0), S.Zt, rz);
  const T c12 = RMB_FMA(h, S.ZtH, S.c12b);
  const T D = RMB_FMA(c12, f[2], E);
  const T P1 = RMB_FMA(-S.q3, c14, kfrz);
  const T P2 = RMB_FMA(-S.q3, D, S.kf * f[2]);
  const T cz = RMB_FMA(ry, f[0], -(rx * f[1]));
  w[0] = RMB_FMA(f[1], P1, w[0]); w[0] = RMB_FMA(-ry, P2, w[0]);
  w[1] = RMB_FMA(-f[0], P1, w[1]); w[1] = RMB_FMA(rx, P2, w[1]);
  w[2] = RMB_FMA(S.kq, cz, w[2]);
}

// u += M_UT(r, target height h) tau    (pair_ut, wall)
template <typename T>
RMB_HD void ut_direction(T rx, T ry, T rz, T h, T kfrz, const T* tq, const RotShared<T>& S, T* u)
{
  const T cz = RMB_FMA(rx, tq[1], -(ry * tq[0]));
  const T E = (RMB_FMA(T(-6.0), h, S.Zt10) * S.t) * cz;
  const T c14 = RMB_FMA(T(2.0), S.Zt, -rz);
  const T c12 = RMB_FMA(h, S.ZtH, S.c12b);
  const T P1 = RMB_FMA(S.q3, c14, kfrz);
  const T dt = S.kq * tq[2];
  const T qE = S.q3 * E;
  u[0] = RMB_FMA(tq[1], P1, u[0]); u[0] = RMB_FMA(-ry, dt, u[0]); u[0] = RMB_FMA(-qE, rx, u[0]);
  u[1] = RMB_FMA(-tq[0], P1, u[1]); u[1] = RMB_FMA(rx, dt, u[1]); u[1] = RMB_FMA(-qE, ry, u[1]);
  u[2] = RMB_FMA(cz, RMB_FMA(S.q3, c12, -S.kf), u[2]);
}

template <typename T, bool BOTH>
RMB_HD void pair_wf_wall_sym(T rx, T ry, T rz, T Rz, T hi, T hj, T hi4, const UfCtx<T>& C, const T* fi, const T* fj,
                             bool any_near, T r2, T* wi, T* wj)
{
  const RotShared<T> S = rot_shared<T, true>(Rz, hi4, hj, C, any_near, r2);
  const T kfrz = S.kf * rz;
  wf_direction<T>(rx, ry, rz, Rz, hj, kfrz, fj, S, wi);
  if (BOTH) wf_direction<T>(-rx, -ry, -rz, Rz, hi, -kfrz, fi, S, wj);
}

template <typename T, bool BOTH>
RMB_HD void pair_ut_wall_sym(T rx, T ry, T rz, T Rz, T hi, T hj, T hi4, const UfCtx<T>& C, const T* ti, const T* tj,
                             bool any_near, T r2, T* ui, T* uj)
{
  const RotShared<T> S = rot_shared<T, false>(Rz, hi4, hj, C, any_near, r2);
  const T kfrz = S.kf * rz;
  ut_direction<T>(rx, ry, rz, hi, kfrz, tj, S, ui);
  if (BOTH) ut_direction<T>(-rx, -ry, -rz, hj, -kfrz, ti, S, uj);
}

// w += M_WT(r) tau for one direction; everything that does not contain the torque is shared (the wall term of WT
// has no height in it besides R_z)
template <typename T>
RMB_HD void wt_direction(T rx, T ry, T rz, T Rz, T Aa, T Bc, T q5a, T q5b, const T* tq, T* w)
{
  const T s = RMB_FMA(ry, tq[1], rx * tq[0]);
  const T B = Bc * RMB_FMA(rz, tq[2], s);
  const T m1 = q5a * RMB_FMA(Rz, tq[2], s);
  const T m2 = q5b * RMB_FMA(rx, tq[1], -(ry * tq[0]));
  const T Bm = B + m1;
  w[0] = RMB_FMA(Aa, tq[0], w[0]); w[0] = RMB_FMA(Bm, rx, w[0]); w[0] = RMB_FMA(-m2, ry, w[0]);
  w[1] = RMB_FMA(Aa, tq[1], w[1]); w[1] = RMB_FMA(Bm, ry, w[1]); w[1] = RMB_FMA(m2, rx, w[1]);
  w[2] = RMB_FMA(Aa, tq[2], w[2]); w[2] = RMB_FMA(B, rz, w[2]); w[2] = RMB_FMA(-m1, Rz, w[2]);
}

template <typename T, bool BOTH>
RMB_HD void pair_wt_wall_sym(T rx, T ry, T rz, T Rz, T hi4, T hj, const UfCtx<T>& C, const T* ti, const T* tj,
                             bool any_near, T r2, T* wi, T* wj)
{
  const T invr = rsqrt_full(r2, C.c38, C.khalf);
  const T invr2 = invr * invr;
  const T invr3 = invr2 * invr;
  T A = T(-0.5) * invr3;
  T Bc = (T(1.5) * invr2) * invr3;
  if (any_near) {
    const T r = r2 * invr;
    const T An = RMB_FMA(RMB_FMA(T(0.078125), r2, T(-0.84375)), r, T(1.0));
    const T Bn = RMB_FMA(T(-0.046875), r, T(0.28125) * invr);
    const bool near = lt4(r2);
    A = near ? An : A;
    Bc = near ? Bn : Bc;
  }
  const T q = rsqrt_full(RMB_FMA(hi4, hj, r2), C.c38, C.khalf);
  const T t = q * q;
  const T q3 = q * t;
  const T k = (Rz * Rz) * t;
  const T q5 = q3 * t;
  const T Aa = RMB_FMA(RMB_FMA(C.km3, k, C.khalf), q3, A);
  const T q5a = T(-1.5) * q5, q5b = T(3.0) * q5;
  wt_direction<T>(rx, ry, rz, Rz, Aa, Bc, q5a, q5b, tj, wi);
  if (BOTH) wt_direction<T>(-rx, -ry, -rz, Rz, Aa, Bc, q5a, q5b, ti, wj);
}

// Free-space twins (no_wall UT / WF / WT).  The coupling kf (v x r) is the same function for UT and WF; with
// w = kf r formed once, the two directions are u_i += v_j x w and u_j -= v_i x w: 3 + 6 + 6 FP64 instructions
// after the shared kf (row kernel: 9 per interaction plus its own kf).
template <typename T, bool BOTH>
RMB_HD void pair_coupling_free_sym(T rx, T ry, T rz, const UfCtx<T>& C, const T* vi, const T* vj, bool any_near, T r2,
                                   T* ui, T* uj)
{
  const T kf = coupling_k(r2, any_near, C);
  const T wx = kf * rx, wy = kf * ry, wz = kf * rz;
  ui[0] = RMB_FMA(vj[1], wz, ui[0]); ui[0] = RMB_FMA(-vj[2], wy, ui[0]);
  ui[1] = RMB_FMA(vj[2], wx, ui[1]); ui[1] = RMB_FMA(-vj[0], wz, ui[1]);
  ui[2] = RMB_FMA(vj[0], wy, ui[2]); ui[2] = RMB_FMA(-vj[1], wx, ui[2]);
  if (BOTH) {
    uj[0] = RMB_FMA(vi[2], wy, uj[0]); uj[0] = RMB_FMA(-vi[1], wz, uj[0]);
    uj[1] = RMB_FMA(vi[0], wz, uj[1]); uj[1] = RMB_FMA(-vi[2], wx, uj[1]);
    uj[2] = RMB_FMA(vi[1], wx, uj[2]); uj[2] = RMB_FMA(-vi[0], wy, uj[2]);
  }
}

// Free-space WT is even in r: w += A tau + Bc (r . tau) r in both directions with the same A, Bc.
template <typename T, bool BOTH>
RMB_HD void pair_wt_free_sym(T rx, T ry, T rz, const UfCtx<T>& C, const T* ti, const T* tj, bool any_near, T r2, T* wi,
                             T* wj)
{
  const T invr = rsqrt_full(r2, C.c38, C.khalf);
  const T invr2 = invr * invr;
  const T invr3 = invr2 * invr;
  T A = T(-0.5) * invr3;
  T Bc = (T(1.5) * invr2) * invr3;
  if (any_near) {
    const T r = r2 * invr;
    const T An = RMB_FMA(RMB_FMA(T(0.078125), r2, T(-0.84375)), r, T(1.0));
    const T Bn = RMB_FMA(T(-0.046875), r, T(0.28125) * invr);
    const bool near = lt4(r2);
    A = near ? An : A;
    Bc = near ? Bn : Bc;
  }
  {
    const T B = Bc * RMB_FMA(rz, tj[2], RMB_FMA(ry, tj[1], rx * tj[0]));
    wi[0] = RMB_FMA(A, tj[0], wi[0]); wi[1] = RMB_FMA(A, tj[1], wi[1]); wi[2] = RMB_FMA(A, tj[2], wi[2]);
    wi[0] = RMB_FMA(B, rx, wi[0]); wi[1] = RMB_FMA(B, ry, wi[1]); wi[2] = RMB_FMA(B, rz, wi[2]);
  }
  if (BOTH) {
    const T B = Bc * RMB_FMA(rz, ti[2], RMB_FMA(ry, ti[1], rx * ti[0]));
    wj[0] = RMB_FMA(A, ti[0], wj[0]); wj[1] = RMB_FMA(A, ti[1], wj[1]); wj[2] = RMB_FMA(A, ti[2], wj[2]);
    wj[0] = RMB_FMA(B, rx, wj[0]); wj[1] = RMB_FMA(B, ry, wj[1]); wj[2] = RMB_FMA(B, rz, wj[2]);
  }
}

}  // namespace rmb
