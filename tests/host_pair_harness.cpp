// tests/host_pair_harness.cpp -- TEST ONLY.
// Compiles the product's per-pair algebra (csrc/pair_math.cuh, mf_core.cuh)
// with g++ (-DRMB_HOST_ONLY) and drives it with a plain double loop, so the
// restructured formulas can be checked against the oracle on a box with no
// GPU.  It is not a CPU fallback: the package never loads it.
#include <cstdlib>
#include <vector>

#include "../rigidmultiblobswall_b200/csrc/mf_core.cuh"

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
    for (int b = 0; b < P.nbox; b++) {
      T sx, sy, sz; bool central;
      box_shift(P, b, sx, sy, sz, central);
      for (long j = 0; j < n; j++) {
        if (central && j == i) continue;
        PairGeom<T> g; T r2, rxy2;
        pair_geometry<T, WALL, PERIODIC>(t, rec[j].x, rec[j].y, rec[j].z, P, sx, sy, sz, T(inva), g, r2, rxy2);
        pair_apply<T, BLOCK, WALL>(g, t.h, rec[j].v, lt4(r2), r2, rxy2, ux, uy, uz);
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
