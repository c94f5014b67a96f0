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

/* torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU legs of bench.py set the thread count themselves */
void rmb_oracle_set_threads(int n)
{
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
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

/* ------------------------------------------------------------------------
 * Source-target kernels with per-blob radii.
 *   mobility_trans_times_force_source_target_numba ........ mobility_numba.py:1480-1658
 *   free_surface_mobility_trans_times_force_source_target_numba ... :1941-2092
 * Restated expression by expression; there is no self branch (the r = 0 pair
 * falls in the "complete overlap" case, :1563-1566).
 * ---------------------------------------------------------------------- */

/* Zuk et al. (2014) RPY coefficients for unequal radii (:1552-1566) */
static inline void rpy_radii_coeff(double r2, double a, double b, double *C1, double *C2)
{
  const double fourOverThree = 4.0 / 3.0;
  const double r = sqrt(r2);
  if (r > (a + b)) {
    const double a2 = a * a, b2 = b * b;
    *C1 = (1 + (b2 + a2) / (3 * r2)) / r;
    *C2 = ((1 - (b2 + a2) / r2) / r2) / r;
  } else if (r > fabs(b - a)) {
    const double r3 = r2 * r;
    const double bma2 = (b - a) * (b - a);
    *C1 = ((16 * (b + a) * r3 - (bma2 + 3 * r2) * (bma2 + 3 * r2)) / (32 * r3)) * fourOverThree / (b * a);
    *C2 = ((3 * (bma2 - r2) * (bma2 - r2) / (32 * r3)) / r2) * fourOverThree / (b * a);
  } else {
    const double largest_radius = a > b ? a : b;
    *C1 = fourOverThree / largest_radius;
    *C2 = 0;
  }
}

/*
 * wall : 0 = unbounded, 1 = no-slip wall correction (:1581-1647), 2 = free surface (:2040-2081).
 * No clamp / damping here: that is the python wrapper (mobility.py:551-585), restated in
 * rmb_oracle_source_target_wall_wrap below.
 */
void rmb_oracle_source_target(int wall, long n_trg, long n_src, const double *target, const double *source,
                              const double *force, const double *radius_target, const double *radius_source,
                              double eta, const double *L, double *out)
{
  const double norm_fact_f = 1.0 / (8.0 * M_PI * eta);
  const int px = L[0] > 0, py = L[1] > 0, pz = L[2] > 0;
#pragma omp parallel for schedule(static)
  for (long i = 0; i < n_trg; i++) {
    const double rxi = target[3 * i], ryi = target[3 * i + 1], rzi = target[3 * i + 2];
    const double a = radius_target[i];
    double ux = 0, uy = 0, uz = 0;
    for (int bx = -px; bx <= px; bx++)
      for (int by = -py; by <= py; by++)
        for (int bz = -pz; bz <= pz; bz++)
          for (long j = 0; j < n_src; j++) {
            const double b = radius_source[j];
            double rx = wrap_shift(rxi - source[3 * j], L[0], bx);
            double ry = wrap_shift(ryi - source[3 * j + 1], L[1], by);
            double rz = wrap_shift(rzi - source[3 * j + 2], L[2], bz);
            double r2 = rx * rx + ry * ry + rz * rz;
            double C1, C2;
            rpy_radii_coeff(r2, a, b, &C1, &C2);
            double Mxx = C1 + C2 * rx * rx, Mxy = C2 * rx * ry, Mxz = C2 * rx * rz;
            double Myy = C1 + C2 * ry * ry, Myz = C2 * ry * rz, Mzz = C1 + C2 * rz * rz;
            double Myx = Mxy, Mzx = Mxz, Mzy = Myz;
            if (wall == 2) {
              rz = rzi + source[3 * j + 2];
              r2 = rx * rx + ry * ry + rz * rz;
              rpy_radii_coeff(r2, a, b, &C1, &C2);
              const double Ixx = C1 + C2 * rx * rx, Ixy = C2 * rx * ry, Ixz = C2 * rx * rz;
              const double Iyy = C1 + C2 * ry * ry, Iyz = C2 * ry * rz, Izz = C1 + C2 * rz * rz;
              Mxx = Mxx + Ixx; Mxy = Mxy + Ixy; Mxz = Mxz - Ixz;
              Myx = Myx + Ixy; Myy = Myy + Iyy; Myz = Myz - Iyz;
              Mzx = Mzx + Ixz; Mzy = Mzy + Iyz; Mzz = Mzz - Izz;
            }
            if (wall == 1) {
              const double y3 = source[3 * j + 2];
              const double x3 = rzi;
              rz = rzi + source[3 * j + 2];
              r2 = rx * rx + ry * ry + rz * rz;
              const double r = sqrt(r2);
              const double a2 = a * a, b2 = b * b;
              const double r3 = r2 * r, r5 = r3 * r2, r7 = r5 * r2, r9 = r7 * r2;

              Mxx -= ((1 + (b2 + a2) / (3.0 * r2)) + (1 - (b2 + a2) / r2) * rx * rx / r2) / r;
              Mxy -= ((1 - (b2 + a2) / r2) * rx * ry / r2) / r;
              Mxz += ((1 - (b2 + a2) / r2) * rx * rz / r2) / r;
              Myx -= ((1 - (b2 + a2) / r2) * ry * rx / r2) / r;
              Myy -= ((1 + (b2 + a2) / (3.0 * r2)) + (1 - (b2 + a2) / r2) * ry * ry / r2) / r;
              Myz += ((1 - (b2 + a2) / r2) * ry * rz / r2) / r;
              Mzx -= ((1 - (b2 + a2) / r2) * rz * rx / r2) / r;
              Mzy -= ((1 - (b2 + a2) / r2) * rz * ry / r2) / r;
              Mzz += ((1 + (b2 + a2) / (3.0 * r2)) + (1 - (b2 + a2) / r2) * rz * rz / r2) / r;

              Mxx -= 2 * (x3 * y3 * (1.0 / r3 - 3 * rx * rx / r5));
              Mxy -= 2 * (x3 * y3 * (-3 * rx * ry / r5));
              Mxz += 2 * (-rx * x3 / r3 + x3 * y3 * (-3 * rx * rz / r5));
              Myx -= 2 * (x3 * y3 * (-3 * ry * rx / r5));
              Myy -= 2 * (x3 * y3 * (1.0 / r3 - 3 * ry * ry / r5));
              Myz += 2 * (-ry * x3 / r3 + x3 * y3 * (-3 * ry * rz / r5));
              Mzx -= 2 * (-y3 * rx / r3 + x3 * y3 * (-3 * rz * rx / r5));
              Mzy -= 2 * (-y3 * ry / r3 + x3 * y3 * (-3 * rz * ry / r5));
              Mzz += 2 * (-1.0 / r - rz * x3 / r3 - y3 * rz / r3 + x3 * y3 * (1.0 / r3 - 3 * rz * rz / r5));

              Mxx -= (2 * a2 / 3.0) * (-y3 * (3 * rz / r5 - 15 * rz * rx * rx / r7));
              Mxy -= (2 * a2 / 3.0) * (-y3 * (-15 * rz * rx * ry / r7));
              Mxz += (2 * a2 / 3.0) * (3 * rx * rz / r5 - y3 * (3 * rx / r5 - 15 * rz * rx * rz / r7));
              Myx -= (2 * a2 / 3.0) * (-y3 * (-15 * rz * ry * rx / r7));
              Myy -= (2 * a2 / 3.0) * (-y3 * (3 * rz / r5 - 15 * rz * ry * ry / r7));
              Myz += (2 * a2 / 3.0) * (3 * ry * rz / r5 - y3 * (3 * ry / r5 - 15 * rz * ry * rz / r7));
              Mzx -= (2 * a2 / 3.0) * (-y3 * (3 * rx / r5 - 15 * rz * rz * rx / r7));
              Mzy -= (2 * a2 / 3.0) * (-y3 * (3 * ry / r5 - 15 * rz * rz * ry / r7));
              Mzz += (2 * a2 / 3.0) * (-1.0 / r3 + 3 * rz * rz / r5 - y3 * (3 * rz / r5 + 3 * rz / r5 + 3 * rz / r5 - 15 * rz * rz * rz / r7));

              Mxx -= (2 * b2 / 3.0) * (-x3 * (3 * rz / r5 - 15 * rz * rx * rx / r7));
              Mxy -= (2 * b2 / 3.0) * (-x3 * (-15 * rz * rx * ry / r7));
              Mxz += (2 * b2 / 3.0) * (-x3 * (3 * rx / r5 - 15 * rz * rx * rz / r7));
              Myx -= (2 * b2 / 3.0) * (-x3 * (-15 * rz * ry * rx / r7));
              Myy -= (2 * b2 / 3.0) * (-x3 * (3 * rz / r5 - 15 * rz * ry * ry / r7));
              Myz += (2 * b2 / 3.0) * (-x3 * (3 * ry / r5 - 15 * rz * ry * rz / r7));
              Mzx -= (2 * b2 / 3.0) * (3 * rz * rx / r5 - x3 * (3 * rx / r5 - 15 * rz * rz * rx / r7));
              Mzy -= (2 * b2 / 3.0) * (3 * rz * ry / r5 - x3 * (3 * ry / r5 - 15 * rz * rz * ry / r7));
              Mzz += (2 * b2 / 3.0) * (-1.0 / r3 + 3 * rz * rz / r5 - x3 * (3 * rz / r5 + 3 * rz / r5 + 3 * rz / r5 - 15 * rz * rz * rz / r7));

              Mxx -= (2 * b2 * a2 / 3.0) * (-1.0 / r5 + 5 * rz * rz / r7 + 5 * rx * rx / r7 - 35 * rz * rz * rx * rx / r9);
              Mxy -= (2 * b2 * a2 / 3.0) * (5 * rx * ry / r7 - 35 * rz * rz * rx * ry / r9);
              Mxz += (2 * b2 * a2 / 3.0) * (5 * rx * rz / r7 + 5 * rx * rz / r7 + 5 * rx * rz / r7 - 35 * rz * rz * rx * rz / r9);
              Myx -= (2 * b2 * a2 / 3.0) * (5 * ry * rx / r7 - 35 * rz * rz * ry * rx / r9);
              Myy -= (2 * b2 * a2 / 3.0) * (-1.0 / r5 + 5 * rz * rz / r7 + 5 * ry * ry / r7 - 35 * rz * rz * ry * ry / r9);
              Myz += (2 * b2 * a2 / 3.0) * (5 * ry * rz / r7 + 5 * ry * rz / r7 + 5 * ry * rz / r7 - 35 * rz * rz * rz * ry / r9);
              Mzx -= (2 * b2 * a2 / 3.0) * (5 * rz * rx / r7 + 5 * rz * rx / r7 + 5 * rz * rx / r7 - 35 * rz * rz * rz * rx / r9);
              Mzy -= (2 * b2 * a2 / 3.0) * (5 * rz * ry / r7 + 5 * rz * ry / r7 + 5 * rz * ry / r7 - 35 * rz * rz * rz * ry / r9);
              Mzz += (2 * b2 * a2 / 3.0) * (-1.0 / r5 + 5 * rz * rz / r7 - 1.0 / r5 + 5 * rz * rz / r7 - 1.0 / r5 + 5 * rz * rz / r7 + 5 * rz * rz / r7 + 5 * rz * rz / r7 + 5 * rz * rz / r7 - 35 * rz * rz * rz * rz / r9);
            }
            const double fx = force[3 * j], fy = force[3 * j + 1], fz = force[3 * j + 2];
            ux += (Mxx * fx + Mxy * fy + Mxz * fz) * norm_fact_f;
            uy += (Myx * fx + Myy * fy + Myz * fz) * norm_fact_f;
            uz += (Mzx * fx + Mzy * fy + Mzz * fz) * norm_fact_f;
          }
    out[3 * i] = ux; out[3 * i + 1] = uy; out[3 * i + 2] = uz;
  }
}

/*
 * mobility.py:551-585 (single_wall_mobility_trans_times_force_source_target_numba):
 * heights clamped to the blob's own radius (shift_heights_different_radius, :87-99),
 * B_ii = z_i / radius_i where z_i < radius_i (damping_matrix_B_different_radius, :102-119),
 * result = B_target * M(y_eff, x_eff) * (B_source * force).
 */
void rmb_oracle_source_target_wall_wrap(long n_trg, long n_src, const double *target, const double *source,
                                        const double *force, const double *radius_target,
                                        const double *radius_source, double eta, const double *L, double *out)
{
  double *x = (double *)malloc(sizeof(double) * 3 * (n_trg > 0 ? n_trg : 1));
  double *y = (double *)malloc(sizeof(double) * 3 * (n_src > 0 ? n_src : 1));
  double *f = (double *)malloc(sizeof(double) * 3 * (n_src > 0 ? n_src : 1));
  for (long k = 0; k < n_trg; k++) {
    x[3 * k] = target[3 * k]; x[3 * k + 1] = target[3 * k + 1];
    x[3 * k + 2] = target[3 * k + 2] > radius_target[k] ? target[3 * k + 2] : radius_target[k];
  }
  for (long k = 0; k < n_src; k++) {
    const double z = source[3 * k + 2];
    const double B = z < radius_source[k] ? z / radius_source[k] : 1.0;
    y[3 * k] = source[3 * k]; y[3 * k + 1] = source[3 * k + 1];
    y[3 * k + 2] = z > radius_source[k] ? z : radius_source[k];
    f[3 * k] = B * force[3 * k]; f[3 * k + 1] = B * force[3 * k + 1]; f[3 * k + 2] = B * force[3 * k + 2];
  }
  rmb_oracle_source_target(1, n_trg, n_src, x, y, f, radius_target, radius_source, eta, L, out);
  for (long k = 0; k < n_trg; k++) {
    const double z = target[3 * k + 2];
    const double B = z < radius_target[k] ? z / radius_target[k] : 1.0;
    out[3 * k] *= B; out[3 * k + 1] *= B; out[3 * k + 2] *= B;
  }
  free(x); free(y); free(f);
}
