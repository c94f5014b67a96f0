// rmb_rigid.cu -- rigid-body operators around the blob mobility product
// (SURVEY.md 8f rank 1-2): blob positions from (location, quaternion,
// reference configuration), K and K^T products, and the dense per-body blob
// mobility blocks the block-diagonal preconditioner factorises.
//
// Reference (host Python loops over bodies, /root/reference):
//   body/body.py:64-80            get_r_vectors
//   body/body.py:83-115           calc_rot_matrix / calc_K_matrix  (K = [I, -(r-q)x])
//   multi_bodies/multi_bodies.py:328-347   K_matrix_vector_prod
//   multi_bodies/multi_bodies.py:350-371   K_matrix_T_vector_prod
//   multi_bodies/multi_bodies.py:766-785   per-body dense M_b (body.calc_mobility_blobs ->
//                                          mobility.single_wall_fluid_mobility, mobility.py:1018)
// Here: one thread per blob (K), one warp per body (K^T, segmented reduction),
// one thread per blob pair of a body (dense blocks).  All O(N) or O(sum Nb^2).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/rmb_mobility.h"
#include "../../include/rmb_rigid.h"
#include "mf_core.cuh"

using namespace rmb;

namespace {

thread_local char g_rerr[256] = "";
int rfail(int code, const char* msg)
{
  snprintf(g_rerr, sizeof(g_rerr), "%s", msg);
  return code;
}
int rcheck(const char* where)
{
  const cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) return 0;
  snprintf(g_rerr, sizeof(g_rerr), "%s: %s", where, cudaGetErrorString(e));
  return (int)e;
}

// quaternion (s, p) -> rotation matrix, quaternion_integrator/quaternion.py:41-50
__device__ __forceinline__ void rot_from_quat(const double* q, double R[9])
{
  const double s = q[0], p0 = q[1], p1 = q[2], p2 = q[3];
  const double d = s * s - 0.5;
  R[0] = 2.0 * (p0 * p0 + d);   R[1] = 2.0 * (p0 * p1 - s * p2); R[2] = 2.0 * (p0 * p2 + s * p1);
  R[3] = 2.0 * (p1 * p0 + s * p2); R[4] = 2.0 * (p1 * p1 + d);   R[5] = 2.0 * (p1 * p2 - s * p0);
  R[6] = 2.0 * (p2 * p0 - s * p1); R[7] = 2.0 * (p2 * p1 + s * p0); R[8] = 2.0 * (p2 * p2 + d);
}

__global__ void blobs_from_bodies_kernel(long n_blobs, const int* __restrict__ blob_body,
                                         const double* __restrict__ ref, const double* __restrict__ loc,
                                         const double* __restrict__ quat, double* __restrict__ r)
{
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_blobs) return;
  const int b = blob_body[i];
  double R[9];
  rot_from_quat(quat + 4 * (long)b, R);
  const double x = ref[3 * i], y = ref[3 * i + 1], z = ref[3 * i + 2];
  r[3 * i] = R[0] * x + R[1] * y + R[2] * z + loc[3 * (long)b];
  r[3 * i + 1] = R[3] * x + R[4] * y + R[5] * z + loc[3 * (long)b + 1];
  r[3 * i + 2] = R[6] * x + R[7] * y + R[8] * z + loc[3 * (long)b + 2];
}

// out_i = beta * out_i + alpha * (U_b + omega_b x (r_i - q_b))
__global__ void K_apply_kernel(long n_blobs, const int* __restrict__ blob_body, const double* __restrict__ r,
                               const double* __restrict__ loc, const double* __restrict__ U, double alpha,
                               double beta, double* __restrict__ out)
{
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_blobs) return;
  const long b = blob_body[i];
  const double dx = r[3 * i] - loc[3 * b], dy = r[3 * i + 1] - loc[3 * b + 1], dz = r[3 * i + 2] - loc[3 * b + 2];
  const double* u = U + 6 * b;
  const double vx = u[0] + (u[4] * dz - u[5] * dy);
  const double vy = u[1] + (u[5] * dx - u[3] * dz);
  const double vz = u[2] + (u[3] * dy - u[4] * dx);
  if (beta == 0.0) {
    out[3 * i] = alpha * vx; out[3 * i + 1] = alpha * vy; out[3 * i + 2] = alpha * vz;
  } else {
    out[3 * i] = beta * out[3 * i] + alpha * vx;
    out[3 * i + 1] = beta * out[3 * i + 1] + alpha * vy;
    out[3 * i + 2] = beta * out[3 * i + 2] + alpha * vz;
  }
}

// One warp per body: FT_b = alpha * sum_i [ lambda_i ; (r_i - q_b) x lambda_i ]
__global__ void KT_apply_kernel(long n_bodies, const long* __restrict__ body_offset, const double* __restrict__ r,
                                const double* __restrict__ loc, const double* __restrict__ lam, double alpha,
                                double* __restrict__ out)
{
  const long b = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= n_bodies) return;
  const long i0 = body_offset[b], i1 = body_offset[b + 1];
  const double qx = loc[3 * b], qy = loc[3 * b + 1], qz = loc[3 * b + 2];
  double s[6] = {0, 0, 0, 0, 0, 0};
  for (long i = i0 + lane; i < i1; i += 32) {
    const double lx = lam[3 * i], ly = lam[3 * i + 1], lz = lam[3 * i + 2];
    const double dx = r[3 * i] - qx, dy = r[3 * i + 1] - qy, dz = r[3 * i + 2] - qz;
    s[0] += lx; s[1] += ly; s[2] += lz;
    s[3] += dy * lz - dz * ly;
    s[4] += dz * lx - dx * lz;
    s[5] += dx * ly - dy * lx;
  }
#pragma unroll
  for (int c = 0; c < 6; c++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s[c] += __shfl_xor_sync(0xffffffffu, s[c], o);
  }
  if (lane < 6) {
    double v = s[0];
#pragma unroll
    for (int c = 1; c < 6; c++) v = (lane == c) ? s[c] : v;
    out[6 * b + lane] = alpha * v;
  }
}

// Dense blob mobility of every body (UF block only, what the preconditioner
// needs): M[b] is (3 nb_max) x (3 nb_max) row-major, identity-padded past the
// body's own 3 Nb rows so a batched Cholesky of the padded blocks is valid.
// Entry (i,j) 3x3 = B_i B_j Mtilde(r_eff_i, r_eff_j) / (8 pi eta a), built
// from the same pair algebra as the matrix-free product by applying it to
// the unit vectors.
template <bool WALL>
__global__ void body_blocks_kernel(long n_bodies, const long* __restrict__ body_offset, int nb_max,
                                   const double* __restrict__ r, double norm, double a, double* __restrict__ M)
{
  const long b = blockIdx.y;
  const long i0 = body_offset[b];
  const int nb = (int)(body_offset[b + 1] - i0);
  const int ld = 3 * nb_max;
  double* Mb = M + (size_t)b * ld * ld;
  const double inva = 1.0 / a;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < nb_max * nb_max; p += gridDim.x * blockDim.x) {
    const int i = p / nb_max, j = p - i * nb_max;
    double blk[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (i < nb && j < nb) {
      double zi = r[3 * (i0 + i) + 2], zj = r[3 * (i0 + j) + 2], Bi = 1.0, Bj = 1.0;
      if (WALL) {
        Bi = (zi < a) ? zi / a : 1.0; Bj = (zj < a) ? zj / a : 1.0;
        zi = (zi <= a) ? a : zi; zj = (zj <= a) ? a : zj;
      }
      Target<double> t;
      t.x = r[3 * (i0 + i)] * inva; t.y = r[3 * (i0 + i) + 1] * inva; t.z = zi * inva; t.h = t.z;
      Periodic<double> P = {};
      for (int c = 0; c < 3; c++) {
        double e[3] = {0, 0, 0};
        e[c] = 1.0;
        double ux = 0, uy = 0, uz = 0;
        if (i == j) {
          self_apply<double, UF, WALL>(t.h, e, ux, uy, uz);
        } else {
          PairGeom<double> g;
          double r2, rxy2;
          pair_geometry<double, WALL, false>(t, r[3 * (i0 + j)] * inva, r[3 * (i0 + j) + 1] * inva, zj * inva, P, 0.0,
                                             0.0, 0.0, inva, g, r2, rxy2);
          pair_apply<double, UF, WALL>(g, make_uf_ctx<double>(t.h), e, true, r2, rxy2, ux, uy, uz);
        }
        const double sc = norm * Bi * Bj;
        blk[c] = ux * sc; blk[3 + c] = uy * sc; blk[6 + c] = uz * sc;
      }
    } else if (i == j) {
      blk[0] = blk[4] = blk[8] = 1.0;
    }
    for (int rr = 0; rr < 3; rr++)
      for (int c = 0; c < 3; c++) Mb[(size_t)(3 * i + rr) * ld + 3 * j + c] = blk[3 * rr + c];
  }
}

}  // namespace

extern "C" {

const char* rmb_rigid_last_error(void) { return g_rerr; }

int rmb_blobs_from_bodies(long n_blobs, const int* blob_body, const double* ref_config, const double* location,
                          const double* quaternion, double* r_out, void* stream)
{
  if (n_blobs < 0) return rfail(RMB_E_ARG, "n_blobs < 0");
  if (n_blobs == 0) return 0;
  if (!blob_body || !ref_config || !location || !quaternion || !r_out) return rfail(RMB_E_ARG, "null device pointer");
  blobs_from_bodies_kernel<<<(unsigned)((n_blobs + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      n_blobs, blob_body, ref_config, location, quaternion, r_out);
  return rcheck("rmb_blobs_from_bodies");
}

int rmb_K_apply(long n_blobs, const int* blob_body, const double* r, const double* location, const double* U,
                double alpha, double beta, double* out, void* stream)
{
  if (n_blobs < 0) return rfail(RMB_E_ARG, "n_blobs < 0");
  if (n_blobs == 0) return 0;
  if (!blob_body || !r || !location || !U || !out) return rfail(RMB_E_ARG, "null device pointer");
  K_apply_kernel<<<(unsigned)((n_blobs + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n_blobs, blob_body, r, location,
                                                                                       U, alpha, beta, out);
  return rcheck("rmb_K_apply");
}

int rmb_KT_apply(long n_bodies, const long* body_offset, const double* r, const double* location,
                 const double* lambda, double alpha, double* out, void* stream)
{
  if (n_bodies < 0) return rfail(RMB_E_ARG, "n_bodies < 0");
  if (n_bodies == 0) return 0;
  if (!body_offset || !r || !location || !lambda || !out) return rfail(RMB_E_ARG, "null device pointer");
  const long threads = n_bodies * 32;
  KT_apply_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n_bodies, body_offset, r,
                                                                                        location, lambda, alpha, out);
  return rcheck("rmb_KT_apply");
}

int rmb_body_mobility_blocks(long n_bodies, const long* body_offset, int nb_max, const double* r, double eta,
                             double a, int wall, double* M, void* stream)
{
  if (n_bodies < 0 || nb_max <= 0) return rfail(RMB_E_ARG, "n_bodies >= 0 and nb_max > 0 required");
  if (n_bodies == 0) return 0;
  if (!body_offset || !r || !M) return rfail(RMB_E_ARG, "null device pointer");
  if (n_bodies > 65535) return rfail(RMB_E_ARG, "at most 65535 bodies per call (grid.y); call in chunks");
  const int pairs = nb_max * nb_max;
  const dim3 grid((unsigned)((pairs + 127) / 128 > 64 ? 64 : (pairs + 127) / 128), (unsigned)n_bodies);
  if (wall) body_blocks_kernel<true><<<grid, 128, 0, (cudaStream_t)stream>>>(n_bodies, body_offset, nb_max, r, block_norm(UF, eta, a), a, M);
  else body_blocks_kernel<false><<<grid, 128, 0, (cudaStream_t)stream>>>(n_bodies, body_offset, nb_max, r, block_norm(UF, eta, a), a, M);
  return rcheck("rmb_body_mobility_blocks");
}

}  // extern "C"
