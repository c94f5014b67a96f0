/*
 * oracle/mobility_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C (gcc + OpenMP) restatement of the reference's CPU pair kernels in
 * mobility/mobility_numba.py.  It exists only to check the CUDA path
 * (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference
 * legs).  Nothing under rigidmultiblobswall_b200/ may import, link or call it.
 *
 * Parity status: PINNED.  tests/golden/*.npz hold outputs of the reference's
 * own numba functions (generated in the build container by
 * tests/golden/make_golden.py, which imports /root/reference/mobility in
 * place); tests/test_oracle.py checks this file against every one of them.
 *
 * The arithmetic below follows the reference expression by expression
 * (no algebraic restructuring) so that it is an independent check of the
 * restructured CUDA kernels.  Reference lines (mobility/mobility_numba.py):
 *   image wrap + box shift .......................... :55-66, :184-192
 *   free-space RPY, UF (self / far / near) .......... :68-107, :199-239
 *   Swan-Brady wall correction, UF .................. :242-276
 *   free UT/WF coupling (epsilon form) .............. :505-536, :611-648, :1001-1035
 *   wall correction UT (transposed call, h = z_i/a) . :648-679
 *   wall correction WF (h = z_j/a) .................. :1037-1066
 *   free WT (self / far / near) ..................... :1142-1178, :1252-1290
 *   wall correction WT .............................. :1292-1321
 *   normalisations 1/(8 pi eta a^k) ................. :27, :138, :562, :953, :1204
 * The python-side wrappers (height clamp z<-max(z,a), damping B) are
 * mobility/mobility.py:52-84 and :1132-1163; restated in rmb_oracle_wall_wrap.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

enum { BLK_UF = 0, BLK_UT = 1, BLK_WF = 2, BLK_WT = 3 };

typedef struct { double xx, xy, xz, yx, yy, yz, zx, zy, zz; } mat3;

/* mobility_numba.py:184-192 -- int() truncates toward zero */
static inline double wrap_shift(double r, double L, int box)
{
  if (L > 0) {
    r = r - (double)((long)(r / L + 0.5 * ((r > 0) - (r < 0)))) * L;
    r = r + box * L;
  }
  return r;
}

/* free-space UF block, lengths already in units of a (:199-239) */
static inline void free_uf(double rx, double ry, double rz, int self, mat3 *M)
{
  if (self) {
    M->xx = M->yy = M->zz = 4.0 / 3.0;
    M->xy = M->xz = M->yz = 0;
  } else {
    double r2 = rx * rx + ry * ry + rz * rz;
    double r = sqrt(r2);
    double invr = 1.0 / r;
    double invr2 = invr * invr;
    if (r > 2) {
      double c1 = 1.0 + 2.0 / (3.0 * r2);
      double c2 = (1.0 - 2.0 * invr2) * invr2;
      M->xx = (c1 + c2 * rx * rx) * invr;
      M->xy = (c2 * rx * ry) * invr;
      M->xz = (c2 * rx * rz) * invr;
      M->yy = (c1 + c2 * ry * ry) * invr;
      M->yz = (c2 * ry * rz) * invr;
      M->zz = (c1 + c2 * rz * rz) * invr;
    } else {
      double c1 = (4.0 / 3.0) * (1.0 - 0.28125 * r);
      double c2 = (4.0 / 3.0) * 0.09375 * invr;
      M->xx = c1 + c2 * rx * rx;
      M->xy = c2 * rx * ry;
      M->xz = c2 * rx * rz;
      M->yy = c1 + c2 * ry * ry;
      M->yz = c2 * ry * rz;
      M->zz = c1 + c2 * rz * rz;
    }
  }
  M->yx = M->xy; M->zx = M->xz; M->zy = M->yz;
}

/* wall correction UF (:242-276); rz here is (z_i+z_j)/a, hj = z_j/a */
static inline void wall_uf(double rx, double ry, double rz, double hj, int self, mat3 *M)
{
  if (self) {
    double invZi = 1.0 / hj;
    double invZi3 = invZi * invZi * invZi;
    double invZi5 = invZi3 * invZi * invZi;
    M->xx += -(9.0 * invZi - 2.0 * invZi3 + invZi5) / 12.0;
    M->yy += -(9.0 * invZi - 2.0 * invZi3 + invZi5) / 12.0;
    M->zz += -(9.0 * invZi - 4.0 * invZi3 + invZi5) / 6.0;
  } else {
    double h_hat = hj / rz;
    double invR = 1.0 / sqrt(rx * rx + ry * ry + rz * rz);
    double ex = rx * invR, ey = ry * invR, ez = rz * invR;
    double invR3 = invR * invR * invR;
    double invR5 = invR3 * invR * invR;
    double fact1 = -(3.0 * (1.0 + 2.0 * h_hat * (1.0 - h_hat) * ez * ez) * invR + 2.0 * (1.0 - 3.0 * ez * ez) * invR3 - 2.0 * (1.0 - 5.0 * ez * ez) * invR5) / 3.0;
    double fact2 = -(3.0 * (1.0 - 6.0 * h_hat * (1.0 - h_hat) * ez * ez) * invR - 6.0 * (1.0 - 5.0 * ez * ez) * invR3 + 10.0 * (1.0 - 7.0 * ez * ez) * invR5) / 3.0;
    double fact3 = ez * (3.0 * h_hat * (1.0 - 6.0 * (1.0 - h_hat) * ez * ez) * invR - 6.0 * (1.0 - 5.0 * ez * ez) * invR3 + 10.0 * (2.0 - 7.0 * ez * ez) * invR5) * 2.0 / 3.0;
    double fact4 = ez * (3.0 * h_hat * invR - 10.0 * invR5) * 2.0 / 3.0;
    double fact5 = -(3.0 * h_hat * h_hat * ez * ez * invR + 3.0 * ez * ez * invR3 + (2.0 - 15.0 * ez * ez) * invR5) * 4.0 / 3.0;
    M->xx += fact1 + fact2 * ex * ex;
    M->xy += fact2 * ex * ey;
    M->xz += fact2 * ex * ez + fact3 * ex;
    M->yx += fact2 * ey * ex;
    M->yy += fact1 + fact2 * ey * ey;
    M->yz += fact2 * ey * ez + fact3 * ey;
    M->zx += fact2 * ez * ex + fact4 * ex;
    M->zy += fact2 * ez * ey + fact4 * ey;
    M->zz += fact1 + fact2 * ez * ez + fact3 * ez + fact4 * ez + fact5;
  }
}

/* free-space translation-rotation coupling, shared by UT and WF (:611-648) */
static inline void free_coupling(double rx, double ry, double rz, int self, mat3 *M)
{
  M->xx = M->yy = M->zz = 0;
  if (self) {
    M->xy = M->xz = M->yz = 0;
  } else {
    double r2 = rx * rx + ry * ry + rz * rz;
    double r = sqrt(r2);
    double r3 = r2 * r;
    double invr3 = 1.0 / r3;
    if (r >= 2) {
      M->xy = rz * invr3;
      M->xz = -ry * invr3;
      M->yz = rx * invr3;
    } else {
      double c1 = 0.5 * (1.0 - 0.375 * r);
      M->xy = c1 * rz;
      M->xz = -c1 * ry;
      M->yz = c1 * rx;
    }
  }
  M->yx = -M->xy; M->zx = -M->xz; M->zy = -M->yz;
}

/* wall correction UT (:648-679): caller passes rx,ry already negated,
   rz = (z_i+z_j)/a and h = z_i/a (height of the TARGET) */
static inline void wall_ut(double rx, double ry, double rz, double h, int self, mat3 *M)
{
  if (self) {
    double invZi = 1.0 / h;
    double invZi4 = invZi * invZi * invZi * invZi;
    M->xy -= -invZi4 * 0.125;
    M->yx -= invZi4 * 0.125;
  } else {
    double h_hat = h / rz;
    double invR = 1.0 / sqrt(rx * rx + ry * ry + rz * rz);
    double invR2 = invR * invR;
    double invR4 = invR2 * invR2;
    double ex = rx * invR, ey = ry * invR, ez = rz * invR;
    double fact1 = invR2;
    double fact2 = (6.0 * h_hat * ez * ez * invR2 + (1.0 - 10.0 * ez * ez) * invR4) * 2.0;
    double fact3 = -ez * (3.0 * h_hat * invR2 - 5.0 * invR4) * 2.0;
    double fact4 = -ez * (h_hat * invR2 - invR4) * 2.0;
    M->xx -= -fact3 * ex * ey;
    M->xy -= -fact1 * ez + fact3 * ex * ex - fact4;
    M->xz -= fact1 * ey;
    M->yx -= fact1 * ez - fact3 * ey * ey + fact4;
    M->yy -= fact3 * ex * ey;
    M->yz -= -fact1 * ex;
    M->zx -= -fact1 * ey - fact2 * ey - fact3 * ey * ez;
    M->zy -= fact1 * ex + fact2 * ex + fact3 * ex * ez;
  }
}

/* wall correction WF (:1037-1066): rz = (z_i+z_j)/a, h = z_j/a */
static inline void wall_wf(double rx, double ry, double rz, double h, int self, mat3 *M)
{
  if (self) {
    double invZi = 1.0 / h;
    double invZi4 = invZi * invZi * invZi * invZi;
    M->xy += -invZi4 * 0.125;
    M->yx += invZi4 * 0.125;
  } else {
    double h_hat = h / rz;
    double invR = 1.0 / sqrt(rx * rx + ry * ry + rz * rz);
    double invR2 = invR * invR;
    double invR4 = invR2 * invR2;
    double ex = rx * invR, ey = ry * invR, ez = rz * invR;
    double fact1 = invR2;
    double fact2 = (6.0 * h_hat * ez * ez * invR2 + (1.0 - 10.0 * ez * ez) * invR4) * 2.0;
    double fact3 = -ez * (3.0 * h_hat * invR2 - 5.0 * invR4) * 2.0;
    double fact4 = -ez * (h_hat * invR2 - invR4) * 2.0;
    M->xx -= -fact3 * ex * ey;
    M->xy -= fact1 * ez - fact3 * ey * ey + fact4;
    M->xz -= -fact1 * ey - fact2 * ey - fact3 * ey * ez;
    M->yx -= -fact1 * ez + fact3 * ex * ex - fact4;
    M->yy -= fact3 * ex * ey;
    M->yz -= fact1 * ex + fact2 * ex + fact3 * ex * ez;
    M->zx -= fact1 * ey;
    M->zy -= -fact1 * ex;
  }
}

/* free-space WT block (:1252-1290) */
static inline void free_wt(double rx, double ry, double rz, int self, mat3 *M)
{
  if (self) {
    M->xx = M->yy = M->zz = 1.0;
    M->xy = M->xz = M->yz = 0;
  } else {
    double r2 = rx * rx + ry * ry + rz * rz;
    double r = sqrt(r2);
    double r3 = r2 * r;
    double invr = 1.0 / r;
    double invr2 = 1.0 / r2;
    double invr3 = 1.0 / r3;
    if (r >= 2) {
      double c1 = -0.5;
      double c2 = 1.5 * invr2;
      M->xx = (c1 + c2 * rx * rx) * invr3;
      M->xy = (c2 * rx * ry) * invr3;
      M->xz = (c2 * rx * rz) * invr3;
      M->yy = (c1 + c2 * ry * ry) * invr3;
      M->yz = (c2 * ry * rz) * invr3;
      M->zz = (c1 + c2 * rz * rz) * invr3;
    } else {
      double c1 = (1.0 - 0.84375 * r + 0.078125 * r3);
      double c2 = 0.28125 * invr - 0.046875 * r;
      M->xx = c1 + c2 * rx * rx;
      M->xy = c2 * rx * ry;
      M->xz = c2 * rx * rz;
      M->yy = c1 + c2 * ry * ry;
      M->yz = c2 * ry * rz;
      M->zz = c1 + c2 * rz * rz;
    }
  }
  M->yx = M->xy; M->zx = M->xz; M->zy = M->yz;
}

/* wall correction WT (:1292-1321): rz = (z_i+z_j)/a, h = z_j/a */
static inline void wall_wt(double rx, double ry, double rz, double h, int self, mat3 *M)
{
  if (self) {
    double invZi = 1.0 / h;
    double invZi3 = invZi * invZi * invZi;
    M->xx += -invZi3 * 0.3125;
    M->yy += -invZi3 * 0.3125;
    M->zz += -invZi3 * 0.125;
  } else {
    double invR = 1.0 / sqrt(rx * rx + ry * ry + rz * rz);
    double invR3 = invR * invR * invR;
    double ex = rx * invR, ey = ry * invR, ez = rz * invR;
    double fact1 = ((1.0 - 6.0 * ez * ez) * invR3) * 0.5;
    double fact2 = -(9.0 * invR3) / 6.0;
    double fact3 = (3.0 * invR3 * ez);
    double fact4 = (3.0 * invR3);
    M->xx += fact1 + fact2 * ex * ex + fact4 * ey * ey;
    M->xy += (fact2 - fact4) * ex * ey;
    M->xz += fact2 * ex * ez;
    M->yx += (fact2 - fact4) * ex * ey;
    M->yy += fact1 + fact2 * ey * ey + fact4 * ex * ex;
    M->yz += fact2 * ey * ez;
    M->zx += fact2 * ez * ex + fact3 * ex;
    M->zy += fact2 * ez * ey + fact3 * ey;
    M->zz += fact1 + fact2 * ez * ez + fact3 * ez;
  }
}

/*
 * One mobility block applied to a vector, for a subset of targets.
 *   block : BLK_UF / BLK_UT / BLK_WF / BLK_WT
 *   wall  : 0 = unbounded (no_wall_* functions), 1 = single wall, 2 = free surface (BLK_UF only)
 *   r     : n x 3 positions (already height-clamped when wall = 1)
 *   v     : n x 3 force or torque
 *   L     : periodic lengths, component <= 0 means not periodic
 *   tgt_begin, tgt_end : targets i in [tgt_begin, tgt_end) are computed,
 *            out holds 3*(tgt_end - tgt_begin) values (lets a bounded sample
 *            of a large problem be timed / checked)
 */
void rmb_oracle_block(int block, int wall, long n, const double *r, const double *v,
                      double eta, double a, const double *L,
                      long tgt_begin, long tgt_end, double *out)
{
  const double inva = 1.0 / a;
  double norm = 1.0 / (8.0 * M_PI * eta * a);
  if (block == BLK_UT || block == BLK_WF) norm = 1.0 / (8.0 * M_PI * eta * a * a);
  if (block == BLK_WT) norm = 1.0 / (8.0 * M_PI * eta * a * a * a);
  const int px = L[0] > 0, py = L[1] > 0, pz = L[2] > 0;

#pragma omp parallel for schedule(static)
  for (long i = tgt_begin; i < tgt_end; i++) {
    double ux = 0, uy = 0, uz = 0;
    const double xi = r[3 * i], yi = r[3 * i + 1], zi = r[3 * i + 2];
    for (int bx = -px; bx <= px; bx++)
      for (int by = -py; by <= py; by++)
        for (int bz = -pz; bz <= pz; bz++)
          for (long j = 0; j < n; j++) {
            double rx = wrap_shift(xi - r[3 * j], L[0], bx);
            double ry = wrap_shift(yi - r[3 * j + 1], L[1], by);
            double rz = wrap_shift(zi - r[3 * j + 2], L[2], bz);
            const int self = (i == j) && bx == 0 && by == 0 && bz == 0;
            rx *= inva; ry *= inva; rz *= inva;
            mat3 M;
            if (block == BLK_UF) free_uf(rx, ry, rz, self, &M);
            else if (block == BLK_WT) free_wt(rx, ry, rz, self, &M);
            else free_coupling(rx, ry, rz, self, &M);
            if (wall == 2) {
              /* free surface (mobility_numba.py:1770-1937): add the RPY block of the image pair, evaluated at
                 (rx, ry, (z_i+z_j)/a) WITHOUT a self branch, with the z column sign-flipped (:1918-1926) */
              mat3 I;
              free_uf(rx, ry, (zi + r[3 * j + 2]) * inva, 0, &I);
              M.xx += I.xx; M.xy += I.xy; M.xz -= I.xz;
              M.yx += I.xy; M.yy += I.yy; M.yz -= I.yz;
              M.zx += I.xz; M.zy += I.yz; M.zz -= I.zz;
            }
            if (wall == 1) {
              const double Rz = (zi + r[3 * j + 2]) * inva;
              if (block == BLK_UF) wall_uf(rx, ry, Rz, r[3 * j + 2] * inva, self, &M);
              else if (block == BLK_UT) wall_ut(-rx, -ry, Rz, zi * inva, self, &M);
              else if (block == BLK_WF) wall_wf(rx, ry, Rz, r[3 * j + 2] * inva, self, &M);
              else wall_wt(rx, ry, Rz, r[3 * j + 2] * inva, self, &M);
            }
            const double fx = v[3 * j], fy = v[3 * j + 1], fz = v[3 * j + 2];
            ux += (M.xx * fx + M.xy * fy + M.xz * fz) * norm;
            uy += (M.yx * fx + M.yy * fy + M.yz * fz) * norm;
            uz += (M.zx * fx + M.zy * fy + M.zz * fz) * norm;
          }
    out[3 * (i - tgt_begin)] = ux;
    out[3 * (i - tgt_begin) + 1] = uy;
    out[3 * (i - tgt_begin) + 2] = uz;
  }
}

/*
 * The python-level wrapper of the wall functions (mobility/mobility.py:52-84,
 * :1132-1163): r_eff = r with z <- a where z <= a; B_ii = z_i/a where z_i < a;
 * result = B * M(r_eff) * (B * v).  in_plane != 0 restates the in_plane_*
 * kernels (mobility_numba.py:291-435, :690-828): the z column and z row of
 * every pair block are zero, i.e. v_z is ignored and out_z = 0.
 */
void rmb_oracle_wall_wrap(int block, int in_plane, long n, const double *r, const double *v,
                          double eta, double a, const double *L,
                          long tgt_begin, long tgt_end, double *out)
{
  double *re = (double *)malloc(sizeof(double) * 3 * n);
  double *ve = (double *)malloc(sizeof(double) * 3 * n);
  double *B = (double *)malloc(sizeof(double) * n);
  for (long k = 0; k < n; k++) {
    const double z = r[3 * k + 2];
    B[k] = (z < a) ? z / a : 1.0;
    re[3 * k] = r[3 * k]; re[3 * k + 1] = r[3 * k + 1];
    re[3 * k + 2] = (z <= a) ? a : z;
    ve[3 * k] = B[k] * v[3 * k]; ve[3 * k + 1] = B[k] * v[3 * k + 1];
    ve[3 * k + 2] = in_plane ? 0.0 : B[k] * v[3 * k + 2];
  }
  rmb_oracle_block(block, 1, n, re, ve, eta, a, L, tgt_begin, tgt_end, out);
  for (long i = tgt_begin; i < tgt_end; i++) {
    double *o = out + 3 * (i - tgt_begin);
    o[0] *= B[i]; o[1] *= B[i]; o[2] *= B[i];
    if (in_plane) o[2] = 0.0;
  }
  free(re); free(ve); free(B);
}

int rmb_oracle_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/*
 * Timing variant of the flagship block only (single-wall UF, no periodic
 * images; mobility_numba.py:124-287): the same expressions as free_uf() +
 * wall_uf() above, written branch-free over a structure-of-arrays copy so
 * that gcc vectorises the source loop the way numba's fastmath LLVM build
 * does (the reference's own numba build measured ~400 Mpairs/s on 8 threads
 * in the build container; the scalar restatement above ~140).  Used by
 * bench.py for the CPU baseline; checked against the golden vectors in
 * tests/test_oracle.py like everything else in this file.
 */
void rmb_oracle_uf_wall_fast(long n, const double *r, const double *v, double eta, double a,
                             long tgt_begin, long tgt_end, double *out)
{
  const double inva = 1.0 / a;
  const double norm = 1.0 / (8.0 * M_PI * eta * a);
  double *X = (double *)malloc(sizeof(double) * 6 * n);
  double *Y = X + n, *Z = X + 2 * n, *FX = X + 3 * n, *FY = X + 4 * n, *FZ = X + 5 * n;
  for (long k = 0; k < n; k++) {
    const double z = r[3 * k + 2];
    const double B = (z < a) ? z / a : 1.0;
    X[k] = r[3 * k] * inva; Y[k] = r[3 * k + 1] * inva; Z[k] = ((z <= a) ? a : z) * inva;
    FX[k] = B * v[3 * k]; FY[k] = B * v[3 * k + 1]; FZ[k] = B * v[3 * k + 2];
  }
#pragma omp parallel for schedule(static)
  for (long i = tgt_begin; i < tgt_end; i++) {
    const double xi = X[i], yi = Y[i], zi = Z[i];
    double ux = 0, uy = 0, uz = 0;
#pragma omp simd reduction(+ : ux, uy, uz)
    for (long j = 0; j < n; j++) {
      const double self = (j == i) ? 1.0 : 0.0;
      const double rx = xi - X[j], ry = yi - Y[j], rz0 = zi - Z[j];
      const double r2 = rx * rx + ry * ry + rz0 * rz0 + self;   /* keep the self lane finite */
      const double rr = sqrt(r2);
      const double invr = 1.0 / rr;
      const double invr2 = invr * invr;
      const int far = rr > 2;
      const double c1 = far ? (1.0 + 2.0 / (3.0 * r2)) * invr : (4.0 / 3.0) * (1.0 - 0.28125 * rr);
      const double c2 = far ? (1.0 - 2.0 * invr2) * invr2 * invr : (4.0 / 3.0) * 0.09375 * invr;
      double Mxx = c1 + c2 * rx * rx, Mxy = c2 * rx * ry, Mxz = c2 * rx * rz0;
      double Myy = c1 + c2 * ry * ry, Myz = c2 * ry * rz0, Mzz = c1 + c2 * rz0 * rz0;
      double Myx = Mxy, Mzx = Mxz, Mzy = Myz;
      const double rz = zi + Z[j];
      const double hj = Z[j];
      const double h_hat = hj / rz;
      const double invR = 1.0 / sqrt(rx * rx + ry * ry + rz * rz);
      const double ex = rx * invR, ey = ry * invR, ez = rz * invR;
      const double invR3 = invR * invR * invR;
      const double invR5 = invR3 * invR * invR;
      const double fact1 = -(3.0 * (1.0 + 2.0 * h_hat * (1.0 - h_hat) * ez * ez) * invR + 2.0 * (1.0 - 3.0 * ez * ez) * invR3 - 2.0 * (1.0 - 5.0 * ez * ez) * invR5) / 3.0;
      const double fact2 = -(3.0 * (1.0 - 6.0 * h_hat * (1.0 - h_hat) * ez * ez) * invR - 6.0 * (1.0 - 5.0 * ez * ez) * invR3 + 10.0 * (1.0 - 7.0 * ez * ez) * invR5) / 3.0;
      const double fact3 = ez * (3.0 * h_hat * (1.0 - 6.0 * (1.0 - h_hat) * ez * ez) * invR - 6.0 * (1.0 - 5.0 * ez * ez) * invR3 + 10.0 * (2.0 - 7.0 * ez * ez) * invR5) * 2.0 / 3.0;
      const double fact4 = ez * (3.0 * h_hat * invR - 10.0 * invR5) * 2.0 / 3.0;
      const double fact5 = -(3.0 * h_hat * h_hat * ez * ez * invR + 3.0 * ez * ez * invR3 + (2.0 - 15.0 * ez * ez) * invR5) * 4.0 / 3.0;
      Mxx += fact1 + fact2 * ex * ex;
      Mxy += fact2 * ex * ey;
      Mxz += fact2 * ex * ez + fact3 * ex;
      Myx += fact2 * ey * ex;
      Myy += fact1 + fact2 * ey * ey;
      Myz += fact2 * ey * ez + fact3 * ey;
      Mzx += fact2 * ez * ex + fact4 * ex;
      Mzy += fact2 * ez * ey + fact4 * ey;
      Mzz += fact1 + fact2 * ez * ez + fact3 * ez + fact4 * ez + fact5;
      const double m = 1.0 - self;   /* the self pair is added in closed form below */
      const double fx = FX[j] * m, fy = FY[j] * m, fz = FZ[j] * m;
      ux += Mxx * fx + Mxy * fy + Mxz * fz;
      uy += Myx * fx + Myy * fy + Myz * fz;
      uz += Mzx * fx + Mzy * fy + Mzz * fz;
    }
    {
      const double invZi = 1.0 / zi;
      const double invZi3 = invZi * invZi * invZi;
      const double invZi5 = invZi3 * invZi * invZi;
      const double sxx = 4.0 / 3.0 - (9.0 * invZi - 2.0 * invZi3 + invZi5) / 12.0;
      const double szz = 4.0 / 3.0 - (9.0 * invZi - 4.0 * invZi3 + invZi5) / 6.0;
      ux += sxx * FX[i]; uy += sxx * FY[i]; uz += szz * FZ[i];
    }
    const double z = r[3 * i + 2];
    const double B = ((z < a) ? z / a : 1.0) * norm;
    out[3 * (i - tgt_begin)] = ux * B;
    out[3 * (i - tgt_begin) + 1] = uy * B;
    out[3 * (i - tgt_begin) + 2] = uz * B;
  }
  free(X);
}
