// tests/host_pair_harness.cpp -- TEST ONLY.
// Compiles the product's per-pair algebra (csrc/pair_math.cuh, mf_core.cuh)
// with g++ (-DRMB_HOST_ONLY) and drives it with a plain double loop, so the
// restructured formulas can be checked against the oracle on a box with no
// GPU.  It is not a CPU fallback: the package never loads it.
#include <cstdlib>
#include <vector>

#include "../rigidmultiblobswall_b200/csrc/mf_core.cuh"
#include "../rigidmultiblobswall_b200/csrc/st_math.cuh"

using namespace rmb;

template <typename T, int BLOCK, bool WALL, bool PERIODIC>
static void run(long n, const double* r, const double* va, const double* vb, double eta, double a,
                const double* L, int in_plane, double* out)
{
  constexpr int NV = (BLOCK == UF_UT) ? 2 : 1;
  std::vector<SrcRec<T, NV>> rec(n);
  std::vector<double> damp(n);
  const double inva = 1.0 / a;
  const double s = PERIODIC ? 1.0 : inva;
  for (long j = 0; j < n; j++) {
    double z = r[3 * j + 2], B = 1.0;
    if (WALL && BLOCK != UF_FS) { B = (z < a) ? z / a : 1.0; z = (z <= a) ? a : z; }
    damp[j] = B;
    rec[j].x = T(r[3 * j] * s); rec[j].y = T(r[3 * j + 1] * s); rec[j].z = T(z * s);
    for (int c = 0; c < 3; c++) rec[j].v[c] = T(B * va[3 * j + c]);
    if (NV == 2) for (int c = 0; c < 3; c++) rec[j].v[3 + c] = T(B * vb[3 * j + c] * inva);
    if (in_plane) { rec[j].v[2] = 0; if (NV == 2) rec[j].v[5] = 0; }
  }
  Periodic<T> P;
  P.nbox = 1;
  for (int d = 0; d < 3; d++) {
    P.per[d] = PERIODIC && L[d] > 0;
    P.L[d] = T(L[d]); P.invL[d] = P.per[d] ? T(1.0 / L[d]) : T(0);
    if (P.per[d]) P.nbox *= 3;
  }
  const double norm = block_norm(BLOCK, eta, a);
  for (long i = 0; i < n; i++) {
    Target<T> t;
    t.x = rec[i].x; t.y = rec[i].y; t.z = rec[i].z;
    t.h = PERIODIC ? T(rec[i].z * T(inva)) : rec[i].z;
    T ux = 0, uy = 0, uz = 0;
    const UfCtx<T> C = make_uf_ctx<T>(t.h);
    if (!PERIODIC) {
      for (long j = 0; j < n; j++) {
        if (j == i) continue;
        PairGeom<T> g; T r2, rxy2;
        pair_geometry<T, WALL, false>(t, rec[j].x, rec[j].y, rec[j].z, P, T(0), T(0), T(0), T(inva), g, r2, rxy2);
        pair_apply<T, BLOCK, WALL, true>(g, C, rec[j].v, lt4(r2), r2, rxy2, ux, uy, uz);
      }
    } else {
      // same two-part geometry as the kernel: wrap + 1/a per source, shift per image box
      const BoxTable<T> box = make_box_table<T>(P, inva);
      for (long j = 0; j < n; j++) {
        PairGeom<T> g0;
        periodic_base<T, WALL>(t, rec[j].x, rec[j].y, rec[j].z, P, T(inva), g0);
        for (int b = 0; b < box.n; b++) {
          if (b == box.central && j == i) continue;
          PairGeom<T> g; T r2, rxy2;
          periodic_box<T>(g0, P, box.s[b], g, r2, rxy2);
          pair_apply<T, BLOCK, WALL, false>(g, C, rec[j].v, lt4(r2), r2, rxy2, ux, uy, uz);
        }
      }
    }
    self_apply<T, BLOCK, WALL>(t.h, rec[i].v, ux, uy, uz);
    const double sc = norm * damp[i];
    out[3 * i] = double(ux) * sc; out[3 * i + 1] = double(uy) * sc;
    out[3 * i + 2] = in_plane ? 0.0 : double(uz) * sc;
  }
}

template <typename T, int BLOCK>
static void run_wp(int wall, bool per, long n, const double* r, const double* va, const double* vb, double eta,
                   double a, const double* L, int in_plane, double* out)
{
  if (wall && per) run<T, BLOCK, true, true>(n, r, va, vb, eta, a, L, in_plane, out);
  else if (wall) run<T, BLOCK, true, false>(n, r, va, vb, eta, a, L, in_plane, out);
  else if (per) run<T, BLOCK, false, true>(n, r, va, vb, eta, a, L, in_plane, out);
  else run<T, BLOCK, false, false>(n, r, va, vb, eta, a, L, in_plane, out);
}

template <typename T>
static void run_b(int block, int wall, bool per, long n, const double* r, const double* va, const double* vb,
                  double eta, double a, const double* L, int in_plane, double* out)
{
  switch (block) {
    case UF: run_wp<T, UF>(wall, per, n, r, va, vb, eta, a, L, in_plane, out); break;
    case UT: run_wp<T, UT>(wall, per, n, r, va, vb, eta, a, L, in_plane, out); break;
    case WF: run_wp<T, WF>(wall, per, n, r, va, vb, eta, a, L, in_plane, out); break;
    case WT: run_wp<T, WT>(wall, per, n, r, va, vb, eta, a, L, in_plane, out); break;
    case UF_UT: run_wp<T, UF_UT>(wall, per, n, r, va, vb, eta, a, L, in_plane, out); break;
    case UF_FS: run_wp<T, UF_FS>(1, per, n, r, va, vb, eta, a, L, in_plane, out); break;
  }
}

extern "C" void host_mf(int block, int wall, int fp32, long n, const double* r, const double* va, const double* vb,
                        double eta, double a, const double* L, int in_plane, double* out)
{
  const bool per = L[0] > 0 || L[1] > 0 || L[2] > 0;
  if (fp32) run_b<float>(block, wall, per, n, r, va, vb, eta, a, L, in_plane, out);
  else run_b<double>(block, wall, per, n, r, va, vb, eta, a, L, in_plane, out);
}

// ---- source-target kernels with per-blob radii (csrc/st_math.cuh) ----
template <typename T, int WALLMODE>
static void run_st(long nt, long ns, const double* trg, const double* src, const double* f, const double* rt,
                   const double* rs, double eta, const double* L, double* out)
{
  std::vector<StRec<T>> rec(ns);
  for (long j = 0; j < ns; j++)
    st_pack<T>(WALLMODE == ST_WALL, src[3 * j], src[3 * j + 1], src[3 * j + 2], rs[j], f[3 * j], f[3 * j + 1],
               f[3 * j + 2], rec[j]);
  Periodic<T> P;
  P.nbox = 1;
  for (int d = 0; d < 3; d++) {
    P.per[d] = L[d] > 0;
    P.L[d] = T(L[d]); P.invL[d] = P.per[d] ? T(1.0 / L[d]) : T(0);
    if (P.per[d]) P.nbox *= 3;
  }
  const double norm = 1.0 / (8.0 * 3.14159265358979323846 * eta);
  for (long i = 0; i < nt; i++) {
    StTarget<T> t;
    const double B = st_target<T>(WALLMODE == ST_WALL, trg[3 * i], trg[3 * i + 1], trg[3 * i + 2], rt[i], t);
    T ux = 0, uy = 0, uz = 0;
    for (int b = 0; b < P.nbox; b++) {
      T sx, sy, sz; bool central;
      box_shift(P, b, sx, sy, sz, central);
      for (long j = 0; j < ns; j++) {
        T rx = t.x - rec[j].x, ry = t.y - rec[j].y, rz = t.z - rec[j].z;
        if (P.per[0]) rx = wrap_shift(rx, P.L[0], P.invL[0], sx);
        if (P.per[1]) ry = wrap_shift(ry, P.L[1], P.invL[1], sy);
        if (P.per[2]) rz = wrap_shift(rz, P.L[2], P.invL[2], sz);
        pair_st<T, WALLMODE>(make_st_consts<T>(), t, rx, ry, rz, t.z + rec[j].z, rec[j].z, rec[j].b, rec[j].ib, rec[j].fx, rec[j].fy,
                             rec[j].fz, ux, uy, uz);
      }
    }
    out[3 * i] = double(ux) * norm * B; out[3 * i + 1] = double(uy) * norm * B; out[3 * i + 2] = double(uz) * norm * B;
  }
}

extern "C" void host_st(int wallmode, int fp32, long nt, long ns, const double* trg, const double* src,
                        const double* f, const double* rt, const double* rs, double eta, const double* L, double* out)
{
#define RMB_ST(T)                                                                        \
  switch (wallmode) {                                                                    \
    case ST_NO_WALL: run_st<T, ST_NO_WALL>(nt, ns, trg, src, f, rt, rs, eta, L, out); break; \
    case ST_WALL: run_st<T, ST_WALL>(nt, ns, trg, src, f, rt, rs, eta, L, out); break;   \
    default: run_st<T, ST_FREE_SURFACE>(nt, ns, trg, src, f, rt, rs, eta, L, out); break; \
  }
  if (fp32) { RMB_ST(float) } else { RMB_ST(double) }
#undef RMB_ST
}

// ---- pair-symmetric evaluation (pair_uf_wall_sym): every unordered pair once, both directions ----
extern "C" void host_uf_sym(int wall, long n, const double* r, const double* f, double eta, double a, double* out)
{
  typedef double T;
  std::vector<SrcRec<T, 1>> rec(n);
  std::vector<double> damp(n), u(3 * n, 0.0);
  const double inva = 1.0 / a;
  for (long j = 0; j < n; j++) {
    double z = r[3 * j + 2];
    const double B = (wall && z < a) ? z / a : 1.0;
    z = (wall && z <= a) ? a : z;
    damp[j] = B;
    rec[j].x = r[3 * j] * inva; rec[j].y = r[3 * j + 1] * inva; rec[j].z = z * inva;
    for (int c = 0; c < 3; c++) rec[j].v[c] = B * f[3 * j + c];
  }
  for (long i = 0; i < n; i++) {
    const UfCtx<T> C = make_uf_ctx<T>(rec[i].z);
    if (!wall) {
      for (long j = i + 1; j < n; j++) {
        const T rx = rec[i].x - rec[j].x, ry = rec[i].y - rec[j].y, rz = rec[i].z - rec[j].z;
        const T r2 = rx * rx + ry * ry + rz * rz;
        pair_uf_free_sym<T, true>(rx, ry, rz, C, rec[i].v, rec[j].v, lt4(r2), r2, &u[3 * i], &u[3 * j]);
      }
      continue;
    }
    for (long j = i + 1; j < n; j++) {
      const T rx = rec[i].x - rec[j].x, ry = rec[i].y - rec[j].y, rz = rec[i].z - rec[j].z;
      const T r2 = rx * rx + ry * ry + rz * rz;
      pair_uf_wall_sym<T, true>(rx, ry, rz, rec[i].z + rec[j].z, rec[i].z, rec[j].z, C.ht2, C.ht4, C, rec[i].v, rec[j].v,
                                lt4(r2), r2,
                                &u[3 * i], &u[3 * j]);
    }
  }
  const double norm = block_norm(UF, eta, a);
  for (long i = 0; i < n; i++) {
    T ux = u[3 * i], uy = u[3 * i + 1], uz = u[3 * i + 2];
    if (wall) self_apply<T, UF, true>(rec[i].z, rec[i].v, ux, uy, uz);
    else self_apply<T, UF, false>(rec[i].z, rec[i].v, ux, uy, uz);
    out[3 * i] = ux * norm * damp[i]; out[3 * i + 1] = uy * norm * damp[i]; out[3 * i + 2] = uz * norm * damp[i];
  }
}

// ---- the same with pseudo-periodic images in x / y: per unordered pair the wrap once, then every box with (r, -r);
// ---- box b seen from i is box -b seen from j, so the loop over b covers all boxes of both directions ----
extern "C" void host_sym_periodic(int block, int wall, long n, const double* r, const double* f, double eta, double a,
                                     const double* L, double* out)
{
  typedef double T;
  std::vector<SrcRec<T, 1>> rec(n);     // unscaled positions, as the periodic kernels keep them
  std::vector<double> damp(n), u(3 * n, 0.0);
  const double inva = 1.0 / a;
  for (long j = 0; j < n; j++) {
    double z = r[3 * j + 2];
    const double B = (wall && z < a) ? z / a : 1.0;
    z = (wall && z <= a) ? a : z;
    damp[j] = B;
    rec[j].x = r[3 * j]; rec[j].y = r[3 * j + 1]; rec[j].z = z;
    for (int c = 0; c < 3; c++) rec[j].v[c] = B * f[3 * j + c];
  }
  Periodic<T> P;
  P.nbox = 1;
  for (int d = 0; d < 3; d++) {
    P.per[d] = d < 2 && L[d] > 0;
    P.L[d] = L[d]; P.invL[d] = P.per[d] ? 1.0 / L[d] : 0.0;
    if (P.per[d]) P.nbox *= 3;
  }
  const BoxTable<T> box = make_box_table<T>(P, inva);
  for (long i = 0; i < n; i++) {
    const T hi = rec[i].z * inva;
    const UfCtx<T> C = make_uf_ctx<T>(hi);
    for (long j = i; j < n; j++) {
      T rx = rec[i].x - rec[j].x, ry = rec[i].y - rec[j].y;
      if (P.per[0]) rx = wrap_shift(rx, P.L[0], P.invL[0], T(0));
      if (P.per[1]) ry = wrap_shift(ry, P.L[1], P.invL[1], T(0));
      rx *= inva; ry *= inva;
      const T hj = rec[j].z * inva, rz = hi - hj;
      for (int b = 0; b < box.n; b++) {
        if (i == j && b == box.central) continue;
        const T bx = rx + box.s[b][0], by = ry + box.s[b][1];
        const T r2 = bx * bx + by * by + rz * rz;
        const bool nr = lt4(r2);
        T* ui = &u[3 * i];
        T* uj = &u[3 * j];
        const T *vi = rec[i].v, *vj = rec[j].v;
        if (i == j) {     // own image: one direction (the opposite one is box -b of the same loop)
          if (block == UF && wall) pair_uf_wall_sym<T, false>(bx, by, rz, hi + hj, hi, hj, C.ht2, C.ht4, C, vi, vj, nr, r2, ui, ui);
          else if (block == UF) pair_uf_free_sym<T, false>(bx, by, rz, C, vi, vj, nr, r2, ui, ui);
          else if (!wall && block == WT) pair_wt_free_sym<T, false>(bx, by, rz, C, vi, vj, nr, r2, ui, ui);
          else if (!wall) pair_coupling_free_sym<T, false>(bx, by, rz, C, vi, vj, nr, r2, ui, ui);
          else if (block == WF) pair_wf_wall_sym<T, false>(bx, by, rz, hi + hj, hi, hj, C.ht4, C, vi, vj, nr, r2, ui, ui);
          else if (block == UT) pair_ut_wall_sym<T, false>(bx, by, rz, hi + hj, hi, hj, C.ht4, C, vi, vj, nr, r2, ui, ui);
          else pair_wt_wall_sym<T, false>(bx, by, rz, hi + hj, C.ht4, hj, C, vi, vj, nr, r2, ui, ui);
        } else {
          if (block == UF && wall) pair_uf_wall_sym<T, true>(bx, by, rz, hi + hj, hi, hj, C.ht2, C.ht4, C, vi, vj, nr, r2, ui, uj);
          else if (block == UF) pair_uf_free_sym<T, true>(bx, by, rz, C, vi, vj, nr, r2, ui, uj);
          else if (!wall && block == WT) pair_wt_free_sym<T, true>(bx, by, rz, C, vi, vj, nr, r2, ui, uj);
          else if (!wall) pair_coupling_free_sym<T, true>(bx, by, rz, C, vi, vj, nr, r2, ui, uj);
          else if (block == WF) pair_wf_wall_sym<T, true>(bx, by, rz, hi + hj, hi, hj, C.ht4, C, vi, vj, nr, r2, ui, uj);
          else if (block == UT) pair_ut_wall_sym<T, true>(bx, by, rz, hi + hj, hi, hj, C.ht4, C, vi, vj, nr, r2, ui, uj);
          else pair_wt_wall_sym<T, true>(bx, by, rz, hi + hj, C.ht4, hj, C, vi, vj, nr, r2, ui, uj);
        }
      }
    }
  }
  const double norm = block_norm(block, eta, a);
  for (long i = 0; i < n; i++) {
    T ux = u[3 * i], uy = u[3 * i + 1], uz = u[3 * i + 2];
    const T h = rec[i].z * inva;
    switch (block * 2 + (wall ? 1 : 0)) {
      case UF * 2 + 1: self_apply<T, UF, true>(h, rec[i].v, ux, uy, uz); break;
      case UF * 2: self_apply<T, UF, false>(h, rec[i].v, ux, uy, uz); break;
      case UT * 2 + 1: self_apply<T, UT, true>(h, rec[i].v, ux, uy, uz); break;
      case WF * 2 + 1: self_apply<T, WF, true>(h, rec[i].v, ux, uy, uz); break;
      case WT * 2 + 1: self_apply<T, WT, true>(h, rec[i].v, ux, uy, uz); break;
      case WT * 2: self_apply<T, WT, false>(h, rec[i].v, ux, uy, uz); break;
      default: break;     // no_wall UT / WF: no self term
    }
    out[3 * i] = ux * norm * damp[i]; out[3 * i + 1] = uy * norm * damp[i]; out[3 * i + 2] = uz * norm * damp[i];
  }
}
